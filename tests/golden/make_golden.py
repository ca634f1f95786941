"""Generate tests/golden/*.npz by running the REAL reference (pyamg 3.0.2 fork, imported from
/root/reference through oracle/refimport.py, native kernels = oracle/_ref/_amg_core.so built
from the reference sources).  Run in the build container only:

    python tests/golden/make_golden.py

Each fixture = the reference-built hierarchy (A/P/R per level, exact formats and column order),
the resolved smoother parameters (omega/rho for Jacobi, MIS colour orderings for
gauss_seidel_indexed), b, the reference residual history and the final x.
"""
import os
import sys
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
warnings.simplefilter('ignore')

from oracle import refimport  # noqa: E402
import hier_io  # noqa: E402

pyamg = refimport.import_pyamg()
from pyamg.gallery import poisson, stencil_grid, linear_elasticity  # noqa: E402
from pyamg.gallery.diffusion import diffusion_stencil_2d  # noqa: E402
from pyamg.graph import vertex_coloring  # noqa: E402
from pyamg.relaxation.relaxation import gauss_seidel_indexed  # noqa: E402
import pyamg.amg_core as ref_core  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
JAC = ('jacobi', {'omega': 4.0 / 3.0})


def run(ml, n, cycles, cycle='V'):
    b = np.random.rand(n)
    res = []
    x = ml.solve(b, tol=1e-30, maxiter=cycles, residuals=res, cycle=cycle)
    return b, np.array(res), x


def jacobi_specs(ml):
    out = []
    for lvl in ml.levels[:-1]:
        om = (4.0 / 3.0) / lvl.A.rho_D_inv
        out.append((('jacobi', {'omega': float(om), 'iterations': 1}),) * 2)
    return out


def color_order(A):
    c = vertex_coloring(A.tocsr(), 'MIS')
    K = int(c.max()) + 1
    idx = np.concatenate([np.where(c == k)[0] for k in range(K)]).astype('intc')
    ptr = np.zeros(K + 1, 'intc')
    ptr[1:] = np.cumsum(np.bincount(c, minlength=K))
    return c, idx, ptr


def install_mcgs(ml):
    specs = []
    for lvl in ml.levels[:-1]:
        _, idx, ptr = color_order(lvl.A)

        def sm(A, x, b, idx=idx):
            gauss_seidel_indexed(A, x, b, idx, iterations=1, sweep='symmetric')
        lvl.presmoother = lvl.postsmoother = sm
        spec = ('gauss_seidel_indexed', {'indices': idx, 'color_ptr': ptr, 'sweep': 'symmetric',
                                         'iterations': 1})
        specs.append((spec, spec))
    return specs


def sa_poisson(name, grid, max_coarse, improve):
    A = poisson(grid, format='csr')
    np.random.seed(0)
    kw = {} if improve else {'improve_candidates': None}
    ml = pyamg.smoothed_aggregation_solver(A, presmoother=JAC, postsmoother=JAC,
                                           max_coarse=max_coarse, **kw)
    b, res, x = run(ml, A.shape[0], 10)
    rhos = [getattr(l, 'rho_smooth', 0.0) for l in ml.levels[:-1]]
    hier_io.save(os.path.join(OUT, name), ml.levels, jacobi_specs(ml),
                 {'b': b, 'residuals': res, 'x': x, 'rho_smooth': rhos})
    print(name, [l.A.shape[0] for l in ml.levels], res[0], res[-1])
    # W and F cycles on the same hierarchy
    for cyc in 'WF':
        r = []
        xx = ml.solve(b, tol=1e-30, maxiter=4, residuals=r, cycle=cyc)
        np.savez_compressed(os.path.join(OUT, name.replace('.npz', '_%s.npz' % cyc)),
                            residuals=np.array(r), x=xx)


def rs_mcgs(name, A, max_coarse):
    np.random.seed(0)
    ml = pyamg.ruge_stuben_solver(A, max_coarse=max_coarse)
    specs = install_mcgs(ml)
    b, res, x = run(ml, A.shape[0], 10)
    hier_io.save(os.path.join(OUT, name), ml.levels, specs, {'b': b, 'residuals': res, 'x': x})
    print(name, [l.A.shape[0] for l in ml.levels], [int(s[0][1]['color_ptr'].shape[0]) - 1 for s in specs],
          [bool(l.A.has_sorted_indices) for l in ml.levels], res[0], res[-1])


def tet_grid(nv):
    """nv^3 vertices of the unit cube, every cube cut into 6 tetrahedra around its main diagonal (fixed orientation)"""
    g = np.arange(nv) / float(nv - 1)
    V = np.array([[x, y, z] for z in g for y in g for x in g])
    vid = lambda i, j, k: i + nv * (j + nv * k)
    E = []
    for k in range(nv - 1):
        for j in range(nv - 1):
            for i in range(nv - 1):
                c = [vid(i + a, j + b, k + d) for d in (0, 1) for b in (0, 1) for a in (0, 1)]   # c[a + 2b + 4d]
                for p, q in ((1, 3), (3, 2), (2, 6), (6, 4), (4, 5), (5, 1)):                   # ring around c0-c7
                    E.append([c[0], c[p], c[q], c[7]])
    return V, np.array(E)


def sa_elasticity3d(name, nv, improve=True):
    """BASELINE config 4 in miniature: P1 tetrahedral elasticity (linear_elasticity_p1, E = 1e5, nu = 0.3), the face
    x = 0 clamped by eliminating its block rows / columns, SA with the 6 rigid-body modes and symmetric Gauss-Seidel:
    level 0 is BSR 3x3, the coarse levels BSR 6x6, P is 3x6 / 6x6 -- the block shapes of the full-size problem."""
    from pyamg.gallery import linear_elasticity_p1
    import scipy.sparse as sp
    V, E = tet_grid(nv)
    A, B = linear_elasticity_p1(V, E, format='bsr')
    free = np.where(V[:, 0] > 0)[0]
    dof = (3 * free[:, None] + np.arange(3)[None, :]).ravel()
    A = sp.csr_matrix(A)[dof][:, dof].tobsr(blocksize=(3, 3))
    B = B[dof]
    np.random.seed(0)
    gs = ('gauss_seidel', {'sweep': 'symmetric'})
    extra = {}
    if improve:
        ml = pyamg.smoothed_aggregation_solver(A, B, presmoother=gs, postsmoother=gs, max_coarse=10)
    else:
        # the SETUP fixture: candidates as given (improve_candidates=None, the option the device setup has), B kept
        ml = pyamg.smoothed_aggregation_solver(A, B, presmoother=gs, postsmoother=gs, max_coarse=10,
                                               improve_candidates=None)
        extra['B'] = B
    b, res, x = run(ml, A.shape[0], 8)
    specs = [((('gauss_seidel', {'sweep': 'symmetric', 'iterations': 1}),) * 2) for _ in ml.levels[:-1]]
    hier_io.save(os.path.join(OUT, name), ml.levels, specs, dict(extra, b=b, residuals=res, x=x))
    print(name, [(l.A.shape[0], l.A.format, getattr(l.A, 'blocksize', None)) for l in ml.levels],
          [getattr(l, 'P', None) is not None and l.P.blocksize for l in ml.levels], res[0], res[-1])


def sa_elasticity3d_block_smoothers(name, nv):
    """The reference's DEFAULT smoother on a block problem: symmetric block_gauss_seidel with the inverses of the
    diagonal blocks from get_block_diag (amg_core.pinv_array for 3x3 / 6x6 blocks), plus a block_jacobi(4/3) run on the
    same hierarchy.  The inverses and the resolved omega (4/3 / rho_block_D_inv_A) are stored, so that cycles replay
    with exactly the reference's numbers; candidates as given (improve_candidates=None)."""
    from pyamg.gallery import linear_elasticity_p1
    from pyamg.relaxation.smoothing import change_smoothers
    import scipy.sparse as sp
    V, E = tet_grid(nv)
    A, B = linear_elasticity_p1(V, E, format='bsr')
    free = np.where(V[:, 0] > 0)[0]
    dof = (3 * free[:, None] + np.arange(3)[None, :]).ravel()
    A = sp.csr_matrix(A)[dof][:, dof].tobsr(blocksize=(3, 3))
    B = B[dof]
    np.random.seed(0)
    ml = pyamg.smoothed_aggregation_solver(A, B, max_coarse=10, improve_candidates=None)     # default pre/post smoothers
    b, res, x = run(ml, A.shape[0], 8)
    specs = []
    for lvl in ml.levels[:-1]:
        Dinv = np.array(lvl.A.block_D_inv)
        sp_ = ('block_gauss_seidel', {'Dinv': Dinv, 'blocksize': int(Dinv.shape[1]), 'sweep': 'symmetric', 'iterations': 1})
        specs.append((sp_, sp_))
    extra = {'B': B, 'b': b, 'residuals': res, 'x': x}
    bj = ('block_jacobi', {'omega': 4.0 / 3.0})
    change_smoothers(ml, bj, bj)
    r2 = []
    x2 = ml.solve(b, tol=1e-30, maxiter=8, residuals=r2)
    extra.update({'bj_residuals': np.array(r2), 'bj_x': x2})
    for k, lvl in enumerate(ml.levels[:-1]):
        extra['bj_omega%d' % k] = (4.0 / 3.0) / lvl.A.rho_block_D_inv
    hier_io.save(os.path.join(OUT, name), ml.levels, specs, extra)
    print(name, [(l.A.shape[0], l.A.blocksize) for l in ml.levels], res[0], res[-1], r2[-1])


def sa_theta(name, A, B, theta, max_coarse):
    """SA with a symmetric strength threshold theta > 0 (entries dropped from the aggregation graph; BSR levels --
    every coarse level -- measured by the squared Frobenius norm of their blocks, strength.py:300-305), candidates as
    given, Jacobi(4/3) smoothers"""
    np.random.seed(0)
    ml = pyamg.smoothed_aggregation_solver(A, B, strength=('symmetric', {'theta': theta}), presmoother=JAC, postsmoother=JAC,
                                           max_coarse=max_coarse, improve_candidates=None)
    b, res, x = run(ml, A.shape[0], 8)
    extra = {'b': b, 'residuals': res, 'x': x, 'theta': theta}
    if B is not None:
        extra['B'] = B
    hier_io.save(os.path.join(OUT, name), ml.levels, jacobi_specs(ml), extra)
    print(name, [(l.A.shape[0], l.A.nnz, l.A.format) for l in ml.levels], res[0], res[-1])


def sa_theta_fixtures():
    st = diffusion_stencil_2d(epsilon=0.01, theta=np.pi / 6, type='FE')
    sa_theta('sa_aniso2d_40_theta005.npz', stencil_grid(st, (40, 40), format='csr'), None, 0.05, 10)
    A, B = linear_elasticity((16, 16), format='bsr')
    sa_theta('sa_elasticity2d_16_theta001.npz', A, B, 0.01, 10)


def sa_poisson_amli(name, grid, max_coarse):
    """cycle='AMLI' (multilevel.py:511-537: two A-orthogonalised recursive corrections per level) on the hierarchy and
    right-hand side of sa_poisson3d_12.npz (same seed, same call order)"""
    A = poisson(grid, format='csr')
    np.random.seed(0)
    ml = pyamg.smoothed_aggregation_solver(A, presmoother=JAC, postsmoother=JAC, max_coarse=max_coarse, improve_candidates=None)
    b = np.random.rand(A.shape[0])
    r = []
    xx = ml.solve(b, tol=1e-30, maxiter=4, residuals=r, cycle='AMLI')
    np.savez_compressed(os.path.join(OUT, name), residuals=np.array(r), x=xx, b=b)
    print(name, [l.A.shape[0] for l in ml.levels], r)


def sa_chebyshev(name, grid, max_coarse):
    """SA hierarchy with the reference's 'chebyshev' smoother (smoothing.py:438-449 -> relaxation.polynomial).  The
    coefficients the reference resolved (from its own random-start spectral-radius estimate) are read back out of the
    smoother closures and stored, so a cycle can be replayed with exactly those numbers."""
    A = poisson(grid, format='csr')
    np.random.seed(0)
    ch = ('chebyshev', {'degree': 3, 'iterations': 1})
    ml = pyamg.smoothed_aggregation_solver(A, presmoother=ch, postsmoother=ch, max_coarse=max_coarse,
                                           improve_candidates=None)
    specs = []
    for lvl in ml.levels[:-1]:
        pair = []
        for fn in (lvl.presmoother, lvl.postsmoother):
            cells = dict(zip(fn.__code__.co_freevars, [c.cell_contents for c in fn.__closure__]))
            pair.append(('polynomial', {'coefficients': np.asarray(cells['coefficients'], dtype=np.float64),
                                        'iterations': int(cells['iterations'])}))
        specs.append(tuple(pair))
    b, res, x = run(ml, A.shape[0], 8)
    hier_io.save(os.path.join(OUT, name), ml.levels, specs, {'b': b, 'residuals': res, 'x': x})
    print(name, [l.A.shape[0] for l in ml.levels], res[0], res[-1])


def sa_elasticity(name, grid):
    A, B = linear_elasticity(grid, format='bsr')
    np.random.seed(0)
    gs = ('gauss_seidel', {'sweep': 'symmetric'})
    ml = pyamg.smoothed_aggregation_solver(A, B, presmoother=gs, postsmoother=gs, max_coarse=10)
    b, res, x = run(ml, A.shape[0], 8)
    specs = [((('gauss_seidel', {'sweep': 'symmetric', 'iterations': 1}),) * 2) for _ in ml.levels[:-1]]
    extra = {'b': b, 'residuals': res, 'x': x}
    # block-multicolour variant: colour the BLOCK graph with the reference MIS colouring and
    # relax block rows colour by colour through the reference bsr_gauss_seidel kernel, one
    # block row per call (row_start=i,row_stop=i+-1) -- the sequential definition the GPU
    # block-colour kernel must reproduce.
    import scipy.sparse as sp
    orders = []
    for lvl in ml.levels[:-1]:
        Ab = lvl.A.tobsr()
        G = sp.csr_matrix((np.ones(len(Ab.indices)), Ab.indices, Ab.indptr),
                          shape=(Ab.shape[0] // Ab.blocksize[0],) * 2)
        _, idx, ptr = color_order(G)
        orders.append((idx, ptr))

        def sm(A, x, b, idx=idx):
            Ab = A.tobsr()
            bs = Ab.blocksize[0]
            data = np.ravel(Ab.data)
            for i in idx:
                ref_core.bsr_gauss_seidel(Ab.indptr, Ab.indices, data, x, b, int(i), int(i) + 1, 1, bs)
            for i in idx[::-1]:
                ref_core.bsr_gauss_seidel(Ab.indptr, Ab.indices, data, x, b, int(i), int(i) - 1, -1, bs)
        lvl.presmoother = lvl.postsmoother = sm
    r2 = []
    x2 = ml.solve(b, tol=1e-30, maxiter=8, residuals=r2)
    extra.update({'mc_residuals': np.array(r2), 'mc_x': x2})
    for k, (idx, ptr) in enumerate(orders):
        extra['mc_idx%d' % k] = idx
        extra['mc_ptr%d' % k] = ptr
    hier_io.save(os.path.join(OUT, name), ml.levels, specs, extra)
    print(name, [(l.A.shape[0], l.A.format, getattr(l.A, 'blocksize', None)) for l in ml.levels],
          res[0], res[-1], r2[-1])


if __name__ == '__main__':
    if sys.argv[1:] == ['elasticity3d_block_smoothers']:
        sa_elasticity3d_block_smoothers('sa_elasticity3d_tets_7_blockgs.npz', 7)
        sys.exit(0)
    if sys.argv[1:] == ['amli']:
        sa_poisson_amli('sa_poisson3d_12_AMLI.npz', (12, 12, 12), 20)
        sys.exit(0)
    if sys.argv[1:] == ['sa_theta']:
        sa_theta_fixtures()
        sys.exit(0)
    if sys.argv[1:] == ['elasticity3d_setup']:           # only the fixture added last (the others are unchanged)
        sa_elasticity3d('sa_elasticity3d_tets_7_setup.npz', 7, improve=False)
        sys.exit(0)
    sa_poisson('sa_poisson3d_12.npz', (12, 12, 12), 20, improve=False)
    sa_poisson('sa_poisson3d_16_default.npz', (16, 16, 16), 30, improve=True)
    rs_mcgs('rs_poisson2d_40_mcgs.npz', poisson((40, 40), format='csr'), 20)
    st = diffusion_stencil_2d(epsilon=0.001, theta=np.pi / 8, type='FE')
    rs_mcgs('rs_aniso_48_mcgs.npz', stencil_grid(st, (48, 48), format='csr'), 20)
    sa_elasticity('sa_elasticity2d_14.npz', (14, 14))
    sa_chebyshev('sa_poisson3d_12_chebyshev.npz', (12, 12, 12), 20)
    sa_elasticity3d('sa_elasticity3d_tets_7.npz', 7)
    sa_elasticity3d('sa_elasticity3d_tets_7_setup.npz', 7, improve=False)
    sa_elasticity3d_block_smoothers('sa_elasticity3d_tets_7_blockgs.npz', 7)
    sa_theta_fixtures()
    sa_poisson_amli('sa_poisson3d_12_AMLI.npz', (12, 12, 12), 20)
