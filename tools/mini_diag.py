import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import pyamg_b200 as amg
g = int(sys.argv[1]) if len(sys.argv) > 1 else 12
A = amg.gallery.poisson((g, g, g))
As = A.to_scipy()
np.random.seed(1)
x = np.random.rand(A.shape[0])
y = A.matvec(torch.from_numpy(x).cuda()).cpu().numpy()
ref = As @ x
bad = np.where(y != ref)[0]
print('n', A.shape[0], 'bad rows', bad.size, bad[:20], bad[-5:] if bad.size else '')
for i in bad[:6]:
    print(i, y[i], ref[i], 'rowlen', As.indptr[i+1]-As.indptr[i])
# which partial sums would give y?
if bad.size:
    i = bad[0]
    cols = As.indices[As.indptr[i]:As.indptr[i+1]]; vals = As.data[As.indptr[i]:As.indptr[i+1]]
    print('cols', cols, 'vals', vals, 'prods', vals * x[cols])
yy = torch.full((A.shape[0],), float('nan'), dtype=torch.float64, device='cuda')
A.matvec(torch.from_numpy(x).cuda(), out=yy)
yy = yy.cpu().numpy()
nanrows = np.where(np.isnan(yy))[0]
wrong = np.where((yy != ref) & ~np.isnan(yy))[0]
def ranges(v):
    out=[]; 
    if v.size==0: return out
    s=p=v[0]
    for q in v[1:]:
        if q!=p+1: out.append((int(s),int(p))); s=q
        p=q
    out.append((int(s),int(p))); return out
print('unwritten', ranges(nanrows)); print('wrong', ranges(wrong))
