"""tiny GPU sanity run for compute-sanitizer / debugging: SpMV, residual and a 2-cycle solve."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import pyamg_b200 as amg

g = int(sys.argv[1]) if len(sys.argv) > 1 else 20
A = amg.gallery.poisson((g, g, g))
As = A.to_scipy()
x = np.random.rand(A.shape[0])
y = A.matvec(torch.from_numpy(x).cuda()).cpu().numpy()
print('matvec bit-exact:', np.array_equal(y, As @ x))
np.random.seed(0)
jac = ('jacobi', {'omega': 4.0 / 3.0})
ml = amg.smoothed_aggregation_solver(A, presmoother=jac, postsmoother=jac, max_coarse=50)
res = []
ml.solve(np.random.rand(A.shape[0]), tol=1e-30, maxiter=2, residuals=res)
print('residuals', res)
