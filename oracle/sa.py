"""TEST INFRASTRUCTURE ONLY -- restatement of the reference's smoothed-aggregation SETUP glue
so that hierarchies for the V-cycle parity tests / CPU baseline can be built where
/root/reference is absent (the GPU box).  Native pieces come from ``oracle.core`` (the
reference's compiled _amg_core when built, else the C port); sparse algebra is scipy's, with
the SAME expression sequence as the reference so formats (BSR(1,1) P/R/coarse A) and fp
orders coincide.

Follows: pyamg/aggregation/aggregation.py:30-435 (driver, extend_hierarchy),
pyamg/strength.py:213-318, pyamg/aggregation/aggregate.py:20-103,
pyamg/aggregation/tentative.py:19-166, pyamg/aggregation/smooth.py:67-208,
pyamg/util/linalg.py:155-380 (Arnoldi spectral-radius estimate, global NumPy RNG),
pyamg/relaxation/smoothing.py:172-200,326-332 (omega/rho(D^-1 A)),
pyamg/util/utils.py:133-200,526-590 (scale_rows, get_diagonal incl. its in-place sort).

Validated against the real Python reference by tests/golden/make_golden.py (run in the
build container, where /root/reference exists); fixtures in tests/golden/.
"""
import numpy as np
import scipy.linalg
import scipy.sparse as sparse

import oracle as _o
from .multilevel import Level


def _approximate_eigenvalues(A, maxiter, v0):
    """pyamg/util/linalg.py:155-254, symmetric=False branch (full Arnoldi)."""
    breakdown = np.finfo(float).eps * 1e6
    flag = False
    maxiter = min(A.shape[0], maxiter)
    v0 = v0 / np.sqrt(np.inner(v0.conj().ravel(), v0.ravel()).real)
    H = np.zeros((maxiter + 1, maxiter), dtype=np.result_type(v0.dtype, A.dtype))
    V = [v0]
    for j in range(maxiter):
        w = A * V[-1]
        for i, v in enumerate(V):
            H[i, j] = np.dot(np.conjugate(v.ravel()), w.ravel())
            w = w - H[i, j] * v
        H[j + 1, j] = np.sqrt(np.inner(w.conj().ravel(), w.ravel()).real)
        if H[j + 1, j] < breakdown:
            flag = True
            if H[j + 1, j] != 0.0:
                w = w / H[j + 1, j]
            V.append(w)
            break
        w = w / H[j + 1, j]
        V.append(w)
    ev, vec = scipy.linalg.eig(H[:j + 1, :j + 1], left=False, right=True)
    return vec, ev, H, V, flag


def approximate_spectral_radius(A, tol=0.01, maxiter=15, restart=5):
    """pyamg/util/linalg.py:282-380; start vector from the GLOBAL numpy RNG (:366-368)."""
    v0 = np.random.rand(A.shape[1], 1)
    for _ in range(restart + 1):
        evect, ev, H, V, flag = _approximate_eigenvalues(A, maxiter, v0)
        nvecs = ev.shape[0]
        k = np.abs(ev).argmax()
        error = H[nvecs, nvecs - 1] * evect[-1, k]
        v0 = np.dot(np.hstack(V[:-1]), evect[:, k].reshape(-1, 1))
        if (np.abs(error) / np.abs(ev[k]) < tol) or flag:
            break
    return float(np.abs(ev[k]))


def get_diagonal_inv(A):
    """utils.py:526-590 with inv=True; NB sorts A's indices IN PLACE (:571)."""
    A.sort_indices()
    D = A.diagonal()
    Dinv = np.zeros_like(D)
    m = D != 0.0
    Dinv[m] = 1.0 / D[m]
    return Dinv


def scale_rows(A, v):
    """utils.py:133-200, copy=True: data[jj] *= v[row] (csr_scale_rows / bsr_scale_rows)."""
    A = A.copy()
    if A.format == 'csr':
        A.data *= np.repeat(v, np.diff(A.indptr))
    else:
        R, C = A.blocksize
        rows = np.repeat(np.arange(A.shape[0] // R), np.diff(A.indptr))
        A.data *= v.reshape(-1, R)[rows][:, :, None]
    return A


def rho_D_inv_A(A):
    """smoothing.py:172-200 (cached on the matrix object)."""
    if not hasattr(A, 'rho_D_inv'):
        A.rho_D_inv = approximate_spectral_radius(scale_rows(A, get_diagonal_inv(A)))
    return A.rho_D_inv


def symmetric_strength_pattern(A, theta=0.0, core=None):
    """strength.py:213-318.  Only the PATTERN is returned as data=1 (standard_aggregation
    reads Ap/Aj only, aggregate.py:77-85), so the abs / row scaling of :311-316 is skipped."""
    core = core or _o.core
    if A.format == 'csr':
        Sp = np.empty_like(A.indptr)
        Sj = np.empty_like(A.indices)
        Sx = np.empty_like(A.data)
        core.symmetric_strength_of_connection(A.shape[0], float(theta), A.indptr, A.indices,
                                              A.data, Sp, Sj, Sx)
        nnz = Sp[-1]
        return sparse.csr_matrix((np.ones(nnz), Sj[:nnz], Sp), shape=A.shape)
    R, C = A.blocksize
    if R != C:
        raise ValueError('matrix must have square blocks')
    shape = (A.shape[0] // R, A.shape[1] // C)
    if theta == 0:
        return sparse.csr_matrix((np.ones(len(A.indices), A.dtype), A.indices.copy(),
                                  A.indptr.copy()), shape=shape)
    data = (np.conjugate(A.data) * A.data).reshape(-1, R * C).sum(axis=1)
    return symmetric_strength_pattern(sparse.csr_matrix((data, A.indices, A.indptr), shape=shape),
                                      theta, core)


def standard_aggregation(C, core=None):
    """aggregate.py:20-103 -> AggOp (csr int8, n x n_agg), Cpts."""
    core = core or _o.core
    n = C.shape[0]
    Tj = np.empty(n, dtype=C.indptr.dtype)
    Cpts = np.empty(n, dtype=C.indptr.dtype)
    nagg = core.standard_aggregation(n, C.indptr, C.indices, Tj, Cpts)
    Cpts = Cpts[:nagg]
    if nagg == 0:
        return sparse.csr_matrix((n, 1), dtype='int8'), Cpts, Tj
    if Tj.min() == -1:
        mask = Tj != -1
        row = np.arange(n, dtype=Tj.dtype)[mask]
        Agg = sparse.coo_matrix((np.ones(mask.sum(), 'int8'), (row, Tj[mask])),
                                shape=(n, nagg)).tocsr()
    else:
        Agg = sparse.csr_matrix((np.ones(n, 'int8'), Tj, np.arange(n + 1, dtype=Tj.dtype)),
                                shape=(n, nagg))
    return Agg, Cpts, Tj


def fit_candidates(AggOp, B, tol=1e-10, core=None):
    """tentative.py:19-166 -> (Q as BSR(K1,K2), coarse candidates R)."""
    core = core or _o.core
    B = np.asarray(B, dtype='float64')
    N_fine, N_coarse = AggOp.shape
    K1 = B.shape[0] // N_fine
    K2 = B.shape[1]
    R = np.empty((N_coarse, K2, K2), dtype=B.dtype)
    Qx = np.empty((AggOp.nnz, K1, K2), dtype=B.dtype)
    Acsc = AggOp.tocsc()
    core.fit_candidates(N_fine, N_coarse, K1, K2, Acsc.indptr, Acsc.indices, Qx.ravel(),
                        np.ascontiguousarray(B).ravel(), R.ravel(), tol)
    Q = sparse.bsr_matrix((Qx.swapaxes(1, 2).copy(), Acsc.indices, Acsc.indptr),
                          shape=(K2 * N_coarse, K1 * N_fine))
    Q = Q.T.tobsr()
    return Q, R.reshape(-1, K2)


def jacobi_prolongation_smoother(S, T, omega=4.0 / 3.0, degree=1):
    """smooth.py:67-208, weighting='diagonal', filter=False."""
    D_inv = get_diagonal_inv(S)
    D_inv_S = scale_rows(S, D_inv)
    rho = approximate_spectral_radius(D_inv_S)
    D_inv_S = (omega / rho) * D_inv_S
    P = T
    for _ in range(degree):
        P = P - (D_inv_S * P)
    return P, rho


def _levelize(spec, max_levels):
    """levelize_smooth_or_improve_candidates (util/utils.py): one spec for all levels, or a list padded with its last"""
    if isinstance(spec, list):
        out = list(spec)
        while len(out) < max_levels:
            out.append(out[-1])
        return out
    return [spec] * max_levels


def improve_B(A, B, spec, core=None):
    """aggregation.py:311-319: B <- relaxation_as_linear_operator(spec, A, 0) * B, every column relaxed on its own"""
    from . import relaxation as rx
    name, kw = spec if isinstance(spec, tuple) else (spec, {})
    if name is None:
        return B
    kw = dict(kw)
    iters, sweep = kw.pop('iterations', 1), kw.pop('sweep', 'forward')
    if kw:
        raise NotImplementedError('oracle.sa: improve_candidates options %r' % sorted(kw))
    bs = blocksize(A)
    b0 = np.zeros(A.shape[0], dtype=A.dtype)
    out = np.empty_like(B)
    Dinv = None
    for c in range(B.shape[1]):
        x = np.array(B[:, c], dtype=A.dtype)
        if name == 'gauss_seidel' or (name == 'block_gauss_seidel' and bs == 1):
            rx.gauss_seidel(A, x, b0, iterations=iters, sweep=sweep, core=core)
        elif name == 'block_gauss_seidel':
            if Dinv is None:
                Dinv = block_diag_inverse(A, core)
            rx.block_gauss_seidel(A, x, b0, Dinv, bs, iterations=iters, sweep=sweep, core=core)
        else:
            raise NotImplementedError('oracle.sa: improve_candidates method %r' % (name,))
        out[:, c] = x
    return out


def block_diag_inverse(A, core=None):
    """get_block_diag(A, blocksize, inv_flag=True) (util/utils.py:601-690) for blocks smaller than 7: the stored
    diagonal blocks (zeros where absent) through amg_core.pinv_array(..., 'T')"""
    core = core or _o.core
    bs = blocksize(A)
    if bs >= 7:
        raise NotImplementedError('oracle.sa: blocks of 7 and more go through LAPACK gelss in the reference')
    nb = A.shape[0] // bs
    D = np.zeros((nb, bs, bs), dtype=A.dtype)
    rows = np.repeat(np.arange(nb), np.diff(A.indptr))
    on_diag = np.where(A.indices == rows)[0]
    D[rows[on_diag]] = A.data[on_diag]
    core.pinv_array(D.ravel(), nb, bs, 'T')
    return D


def blocksize(A):
    return A.blocksize[0] if A.format == 'bsr' else 1


def smoothed_aggregation_hierarchy(A, B=None, theta=0.0, smooth_omega=4.0 / 3.0,
                                   presmoother=('jacobi', {'omega': 4.0 / 3.0}),
                                   postsmoother=('jacobi', {'omega': 4.0 / 3.0}),
                                   max_levels=10, max_coarse=500,
                                   improve_candidates=[('block_gauss_seidel', {'sweep': 'symmetric', 'iterations': 4}), None],
                                   core=None):
    """aggregation.py:30-290 for symmetry='hermitian', strength=('symmetric',{'theta':theta}),
    aggregate='standard', smooth=('jacobi',{'omega':smooth_omega}); returns a list of
    oracle.multilevel.Level with RESOLVED smoother specs (omega/rho for Jacobi).
    ``improve_candidates``: the reference's default (aggregation.py:38-41), None, or the reference's other forms (a spec
    or a per-level list) for the methods
    gauss_seidel / block_gauss_seidel: B <- relaxation on A B = 0, column by column (aggregation.py:311-319,
    util/utils.py:1140-1204)."""
    improve = _levelize(improve_candidates, max_levels)
    A = A.asfptype() if hasattr(A, 'asfptype') else A.astype(np.float64)
    if B is None:
        bs = blocksize(A)
        B = np.kron(np.ones((A.shape[0] // bs, 1), dtype=A.dtype), np.eye(bs))
    else:
        B = np.asarray(B, dtype=A.dtype)
        if B.ndim == 1:
            B = B.reshape(-1, 1)
    levels = [Level()]
    levels[-1].A = A
    levels[-1].B = B
    while len(levels) < max_levels and levels[-1].A.shape[0] // blocksize(levels[-1].A) > max_coarse:
        lvl = levels[-1]
        A, B = lvl.A, lvl.B
        B = lvl.B = improve_B(A, B, improve[len(levels) - 1], core)
        C = symmetric_strength_pattern(A, theta, core)
        AggOp, _, _ = standard_aggregation(C, core)
        T, Bc = fit_candidates(AggOp, B, core=core)
        P, rho = jacobi_prolongation_smoother(A, T, smooth_omega)
        R = P.conjugate().transpose()                 # P.H (aggregation.py:400-401)
        lvl.AggOp, lvl.T, lvl.P, lvl.R, lvl.rho_smooth = AggOp, T, P, R, rho
        levels.append(Level())
        levels[-1].A = R * A * P                      # aggregation.py:429
        levels[-1].B = Bc
    # change_smoothers (smoothing.py:24-169): one closure per non-coarsest level
    for lvl in levels[:-1]:
        lvl.pre = _resolve(presmoother, lvl)
        lvl.post = _resolve(postsmoother, lvl)
    return levels


def _resolve(spec, lvl):
    name, kw = spec if isinstance(spec, tuple) else (spec, {})
    kw = dict(kw)
    if name == 'jacobi':
        omega = kw.pop('omega', 1.0)
        if kw.pop('withrho', True):
            omega = omega / rho_D_inv_A(lvl.A)
        return ('jacobi', {'omega': omega, 'iterations': kw.pop('iterations', 1)})
    if name == 'gauss_seidel':
        return ('gauss_seidel', {'sweep': kw.pop('sweep', 'forward'),
                                 'iterations': kw.pop('iterations', 1)})
    if name == 'block_gauss_seidel':               # the reference's SA default (aggregation.py:34-37)
        sweep, iters = kw.pop('sweep', 'forward'), kw.pop('iterations', 1)
        bs = blocksize(lvl.A)
        if bs == 1:                                # smoothing.py:406-408
            return ('gauss_seidel', {'sweep': sweep, 'iterations': iters})
        return ('block_gauss_seidel', {'Dinv': block_diag_inverse(lvl.A), 'blocksize': bs, 'sweep': sweep, 'iterations': iters})
    raise ValueError('oracle.sa: unsupported smoother %r' % (name,))
