"""Generate tests/golden/ref_config2_<grid>.json by running the REAL reference (imported from /root/reference through
oracle/refimport.py) on BASELINE config 2 at full size: poisson((g,g,g)) + smoothed_aggregation_solver(Jacobi 4/3 pre/post,
max_coarse=500), seed protocol of SURVEY 8(d): np.random.seed(0) before the constructor, b = np.random.rand(n) after it,
x0 = 0, 10 V(1,1)-cycles.  Build container only (about 4 minutes and 40 GB at g = 256):

    python tests/golden/make_ref_config2.py [grid]

The file holds what the GPU arm of bench.py needs to pin itself to the reference at the BASELINE size without the
reference being present: the level table, the reference's rho(D^-1 A) per level (exact, as hex floats: the estimate of the
prolongation smoother and the one of the Jacobi relaxation, in call order), the residual
history (exact) and a few entries / the norm of the final x."""
import json
import os
import sys
import time
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
warnings.simplefilter('ignore')
from oracle import refimport  # noqa: E402

pyamg = refimport.import_pyamg()
import pyamg.aggregation.smooth as _smooth            # noqa: E402
import pyamg.relaxation.smoothing as _smoothing       # noqa: E402

# the reference keeps only the relaxation estimate (A.rho_D_inv); the one of the prolongation smoother is used and
# dropped (aggregation/smooth.py:164), so both call sites are wrapped to record what approximate_spectral_radius returned
RHO = {'smooth': [], 'relax': []}


def _recording(fn, key):
    def wrapped(*a, **k):
        r = fn(*a, **k)
        RHO[key].append(float(r))
        return r
    return wrapped


_smooth.approximate_spectral_radius = _recording(_smooth.approximate_spectral_radius, 'smooth')
_smoothing.approximate_spectral_radius = _recording(_smoothing.approximate_spectral_radius, 'relax')
JAC = ('jacobi', {'omega': 4.0 / 3.0})
g = int(sys.argv[1]) if len(sys.argv) > 1 else 256
A = pyamg.gallery.poisson((g, g, g), format='csr')
np.random.seed(0)
t0 = time.perf_counter()
ml = pyamg.smoothed_aggregation_solver(A, presmoother=JAC, postsmoother=JAC, max_coarse=500)
setup_s = time.perf_counter() - t0
b = np.random.rand(A.shape[0])
res = []
t0 = time.perf_counter()
x = ml.solve(b, tol=1e-30, maxiter=10, residuals=res)
solve_s = time.perf_counter() - t0
pick = [0, 1, g, g * g, A.shape[0] // 2, A.shape[0] - 1]
out = {'grid': g, 'levels': [[int(l.A.shape[0]), int(l.A.nnz)] for l in ml.levels],
       'rho_D_inv_A_hex': [float(l.A.rho_D_inv).hex() for l in ml.levels[:-1]],
       'rho_D_inv_A': [float(l.A.rho_D_inv) for l in ml.levels[:-1]],
       'rho_smooth_hex': [r.hex() for r in RHO['smooth']], 'rho_relax_hex': [r.hex() for r in RHO['relax']],
       'residuals_hex': [float(r).hex() for r in res], 'residuals': [float(r) for r in res],
       'x_pick_index': pick, 'x_pick_hex': [float(x[i]).hex() for i in pick],
       'x_norm_hex': float(np.linalg.norm(x)).hex(), 'b_sum_hex': float(b.sum()).hex(),
       'reference_setup_s_here': round(setup_s, 1), 'reference_solve10_s_here': round(solve_s, 2),
       'protocol': 'np.random.seed(0); ml = smoothed_aggregation_solver(poisson((g,g,g)), presmoother=JAC, postsmoother=JAC, '
                   'max_coarse=500); b = np.random.rand(n); ml.solve(b, tol=1e-30, maxiter=10, residuals=res)'}
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'ref_config2_%d.json' % g)
json.dump(out, open(path, 'w'), indent=1)
print(path, out['levels'], out['residuals'][:3], setup_s, solve_s)
