"""smoothed_aggregation_solver on the device (pyamg/aggregation/aggregation.py:30-435) for the
paths the BASELINE configs use: symmetric strength (theta = 0), standard aggregation, Jacobi-smoothed
prolongator, R = P^H, Galerkin A_c = R A P -- scalar CSR with one near-nullspace candidate (configs 2, 5)
and BSR with K1 x K1 blocks and K2 candidates (config 4: 3 x 3 elasticity blocks, 6 rigid-body modes;
the `_block` functions below, kernels in csrc/bsr_setup.cu).

Per level (extend_hierarchy, aggregation.py:293-435):
  C      strength pattern; theta = 0 keeps every stored entry (smoothed_aggregation.h:48-99)
  AggOp  amg_core.standard_aggregation -- sequential greedy, run on the HOST by
         b200amg_standard_aggregation_host (identical integers), pattern copied D2H once
  T, Bc  amg_core.fit_candidates, K1 = K2 = 1 (b200amg_tentative_scalar)
  P      T - (omega/rho(D^-1 A)) D^-1 A T  (csr_scale_rows + SpGEMM + sorted subtract; smooth.py:160-206)
  R      P^T (b200amg_csr_transpose)        A_c = (R A) P (b200amg_spgemm_*)
Every floating-point step keeps the reference's operation order, so given the same rho the
hierarchy equals the reference-built one to the bit; rho itself comes from a device Arnoldi run
started from the same numpy random vector (agreement ~1e-15 relative).
"""
import numpy as np
import torch

from . import _lib, amg_core
from .linalg import approximate_spectral_radius
from .multilevel import multilevel_solver
from .smoothing import change_smoothers
from .sparse import (DevMat, _dev, _ptr, _scan_i32, _stream, block_pattern, bsr_matmat, bsr_sort_indices, galerkin,
                     is_canonical_pattern, matmat, transpose)


def _unpack(v):
    return (v[0], dict(v[1])) if isinstance(v, tuple) else (v, {})


def sort_indices(A):
    """rows ascending in column (scipy sort_indices; get_diagonal sorts level matrices in place,
    pyamg/util/utils.py:571): a double transpose."""
    if A.has_sorted_indices:
        return A
    if not A.is_scalar:
        return bsr_sort_indices(A)
    S = transpose(transpose(A))
    return DevMat(S.indptr, S.indices, S.data, A.shape, A.blocksize, A.format, True)


def symmetric_strength_of_connection(A, theta=0.0):
    """pyamg/strength.py:213-318 -> the strength PATTERN as a scalar CSR DevMat over the (block) rows of A (all the
    aggregation reads, aggregate.py:77-85; the values kept are A's / the blocks' squared Frobenius norms, unscaled).
    theta == 0: every stored entry is strong -- A's own index arrays.  BSR levels (1 x 1 blocks included) are measured
    by the squared Frobenius norm of their blocks, like the reference (strength.py:300-305)."""
    if theta < 0:
        raise ValueError('expected a positive theta')
    if A.blocksize[0] != A.blocksize[1]:
        raise ValueError('matrix must have square blocks')
    if theta == 0:
        return block_pattern(A)
    st = _stream()
    dev = A.data.device
    n = A.n_brow
    if A.format == 'csr':
        vals = A.data
    else:
        vals = torch.empty(A.nblocks, dtype=torch.float64, device=dev)
        _lib.call('b200amg_block_sqnorm', A.nblocks, A.blocksize[0], A.blocksize[1], _ptr(A.data), _ptr(vals), st)
    diag = torch.empty(max(n, 1), dtype=torch.float64, device=dev)
    _lib.call('b200amg_abs_diag', n, _ptr(A.indptr), _ptr(A.indices), _ptr(vals), _ptr(diag), st)
    cnt = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    args = (n, float(theta), _ptr(A.indptr), _ptr(A.indices), _ptr(vals), _ptr(diag))
    _lib.call('b200amg_symmetric_strength', *args, 0, _ptr(cnt), 0, 0, 0, st)
    Sp = _scan_i32(cnt[:n])
    nnz = int(Sp[-1].item())
    Sj = torch.empty(nnz, dtype=torch.int32, device=dev)
    Sx = torch.empty(nnz, dtype=torch.float64, device=dev)
    _lib.call('b200amg_symmetric_strength', *args, 1, 0, _ptr(Sp), _ptr(Sj), _ptr(Sx), st)
    return DevMat(Sp, Sj, Sx, (n, n), (1, 1), 'csr', A.has_sorted_indices)


def standard_aggregation(A, method='standard'):
    """-> (agg int32 device tensor [n], n_agg, roots).  Pattern goes to the host once.  method: 'standard'
    (aggregate.py:20-103) or 'naive' (:106-190)."""
    n = A.shape[0]
    Ap, Aj = A.indptr.cpu().numpy(), A.indices.cpu().numpy()
    agg = np.empty(n, dtype=np.intc)
    roots = np.empty(n, dtype=np.intc)
    fn = amg_core.standard_aggregation if method == 'standard' else amg_core.naive_aggregation
    nagg = fn(n, Ap, Aj, agg, roots)
    return torch.from_numpy(agg).to(_dev()), nagg, roots[:nagg]


def _one_entry_rows(agg, vals, nagg):
    """CSR (n x nagg) with entry (agg[i], vals[i]) on every row with agg[i] >= 0"""
    n = agg.numel()
    st = _stream()
    flag = torch.empty(n, dtype=torch.int32, device=agg.device)
    _lib.call('b200amg_flag_nonneg', n, _ptr(agg), _ptr(flag), st)
    Tp = _scan_i32(flag)
    nnz = int(Tp[-1].item())
    Tj = torch.empty(nnz, dtype=torch.int32, device=agg.device)
    Tx = torch.empty(nnz, dtype=torch.float64, device=agg.device)
    _lib.call('b200amg_tentative_fill', n, _ptr(agg), _ptr(vals), _ptr(Tp), _ptr(Tj), _ptr(Tx), st)
    return DevMat(Tp, Tj, Tx, (n, nagg), (1, 1), 'bsr', True)


def fit_candidates_scalar(agg, nagg, B, tol=1e-10):
    """tentative.py:19-166 for one candidate: -> (T as BSR(1,1) DevMat, coarse candidate Bc)"""
    n = agg.numel()
    AggB = _one_entry_rows(agg, B, nagg)
    M = transpose(AggB)                                   # members of every aggregate, ascending
    Bc = torch.empty(nagg, dtype=torch.float64, device=B.device)
    node_val = torch.zeros(n, dtype=torch.float64, device=B.device)
    _lib.call('b200amg_tentative_scalar', nagg, _ptr(M.indptr), _ptr(M.indices), _ptr(M.data), float(tol),
              _ptr(Bc), _ptr(node_val), _stream())
    return _one_entry_rows(agg, node_val, nagg), Bc


def fit_candidates_block(agg, nagg, B, K1, K2, tol=1e-10):
    """tentative.py:19-166 for K1 unknowns per node and K2 candidates: B is the (n K1) x K2 candidate array, row-major
    on the device.  -> (T as BSR with K1 x K2 blocks, one per aggregated node; coarse candidates (nagg K2) x K2)"""
    n = agg.numel()
    dev = B.device
    AggM = _one_entry_rows(agg, torch.zeros(n, dtype=torch.float64, device=dev), nagg)     # pattern of AggOp
    M = transpose(AggM)                                   # AggOp in CSC form: members of every aggregate, ascending
    nmem = M.nnz
    Qx = torch.empty(nmem * K1 * K2, dtype=torch.float64, device=dev)
    Bc = torch.empty(nagg * K2 * K2, dtype=torch.float64, device=dev)
    _lib.call('b200amg_fit_candidates', nagg, K1, K2, _ptr(M.indptr), _ptr(M.indices), _ptr(B), float(tol), _ptr(Qx),
              _ptr(Bc), _stream())
    Tx = torch.empty_like(Qx)
    _lib.call('b200amg_tentative_scatter', nmem, K1, K2, _ptr(M.indices), _ptr(AggM.indptr), _ptr(Qx), _ptr(Tx), _stream())
    return DevMat(AggM.indptr, AggM.indices, Tx, (n * K1, nagg * K2), (K1, K2), 'bsr', True), Bc


def _scaled_block(A, v, c=None):
    out = torch.empty_like(A.data)
    R, C = A.blocksize
    _lib.call('b200amg_bsr_scale_rows', A.n_brow, R, C, _ptr(A.indptr), _ptr(A.data), _ptr(v),
              0.0 if c is None else float(c), 0 if c is None else 1, _ptr(out), _stream())
    return DevMat(A.indptr, A.indices, out, A.shape, A.blocksize, 'bsr', A.has_sorted_indices)


def bsr_subtract(A, B):
    """A - B on blocks with all-zero result blocks dropped: scipy bsr_binop_bsr, which merges when both operands are
    canonical (rows strictly ascending) and otherwise emits in reverse first-touch order -- what `T - D_inv_S*T` gets,
    because bsr_matmat returns its blocks in first-touch order"""
    general = not (is_canonical_pattern(A) and is_canonical_pattern(B))
    n = A.n_brow
    R, C = A.blocksize
    if B.blocksize != (R, C) or B.shape != A.shape:
        raise ValueError('inconsistent shapes')
    st = _stream()
    dev = A.data.device
    cnt = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    args = (n, R, C, _ptr(A.indptr), _ptr(A.indices), _ptr(A.data), _ptr(B.indptr), _ptr(B.indices), _ptr(B.data),
            int(general))
    _lib.call('b200amg_bsr_sub', *args, 0, _ptr(cnt), 0, 0, 0, st)
    Cp = _scan_i32(cnt[:n])
    nb = int(Cp[-1].item())
    Cj = torch.empty(nb, dtype=torch.int32, device=dev)
    Cx = torch.empty(nb * R * C, dtype=torch.float64, device=dev)
    _lib.call('b200amg_bsr_sub', *args, 1, 0, _ptr(Cp), _ptr(Cj), _ptr(Cx), st)
    return DevMat(Cp, Cj, Cx, A.shape, (R, C), 'bsr', not general)


def jacobi_prolongation_smoother_block(A, T, omega=4.0 / 3.0, degree=1, rho=None):
    """smooth.py:67-208 for BSR operands, weighting='diagonal', filter=False -> (P, rho)"""
    dinv = inverse_diagonal(A)
    if rho is None:
        rho = approximate_spectral_radius(_scaled_block(A, dinv))
    else:
        np.random.rand(A.shape[1], 1)       # a fed estimate still consumes the start vector's draw (util/linalg.py:367)
    S = _scaled_block(A, dinv, omega / rho)
    P = T
    for _ in range(degree):
        P = bsr_subtract(P, bsr_matmat(S, P))
    return P, rho


def _levelize_improve(spec, max_levels):
    """levelize_smooth_or_improve_candidates (pyamg/util/utils.py): one spec for every level, or a list padded with its
    last entry"""
    if isinstance(spec, list):
        out = list(spec)
        while len(out) < max_levels:
            out.append(out[-1])
        return out
    return [spec] * max_levels


def improve_candidates_on_level(A, Bv, K2, spec):
    """aggregation.py:311-319: B <- relaxation_as_linear_operator(spec, A, 0) * B (util/utils.py:1140-1204): every
    candidate (column of B) relaxed on A x = 0 on its own.  Methods: gauss_seidel, block_gauss_seidel (what the
    reference's default asks for) -- lexicographic sweeps, exact and sequential on the device."""
    from . import relaxation
    name, kw = _unpack(spec)
    if name is None:
        return Bv
    iters, sweep = kw.pop('iterations', 1), kw.pop('sweep', 'forward')
    if name not in ('gauss_seidel', 'block_gauss_seidel') or kw:
        raise NotImplementedError('improve_candidates: gauss_seidel / block_gauss_seidel with sweep and iterations only')
    n = A.shape[0]
    bs = A.blocksize[0]
    b0 = torch.zeros(n, dtype=torch.float64, device=Bv.device)
    Bm = Bv.view(n, K2)
    out = torch.empty_like(Bm)
    for c in range(K2):
        x = Bm[:, c].contiguous()
        if name == 'gauss_seidel' or bs == 1:            # block GS with 1 x 1 blocks is plain GS (smoothing.py:406-408)
            relaxation.gauss_seidel(A, x, b0, iterations=iters, sweep=sweep)
        else:
            relaxation.block_gauss_seidel(A, x, b0, iterations=iters, sweep=sweep, blocksize=bs)
        out[:, c] = x
    return out.reshape(-1).contiguous()


def _scaled(A, v, c=None):
    out = torch.empty_like(A.data)
    _lib.call('b200amg_csr_scale_rows', A.shape[0], _ptr(A.indptr), _ptr(A.data), _ptr(v),
              0.0 if c is None else float(c), 0 if c is None else 1, _ptr(out), _stream())
    return DevMat(A.indptr, A.indices, out, A.shape, (1, 1), 'csr', A.has_sorted_indices)


def csr_subtract(A, B, general=True):
    """A - B with zero results dropped: scipy csr_binop_csr, canonical (sorted operands, sorted
    result) or general (any order in, reverse first-touch order out -- what the reference's
    `T - D_inv_S*T` produces, because csr_matmat returns unsorted rows)"""
    fn = 'b200amg_csr_sub_general' if general else 'b200amg_csr_sub_sorted'
    n = A.shape[0]
    st = _stream()
    cnt = torch.empty(max(n, 1), dtype=torch.int32, device=A.data.device)
    args = (n, _ptr(A.indptr), _ptr(A.indices), _ptr(A.data), _ptr(B.indptr), _ptr(B.indices), _ptr(B.data))
    _lib.call(fn, *args, 0, _ptr(cnt), 0, 0, 0, st)
    Cp = _scan_i32(cnt[:n])
    nnz = int(Cp[-1].item())
    Cj = torch.empty(nnz, dtype=torch.int32, device=A.data.device)
    Cx = torch.empty(nnz, dtype=torch.float64, device=A.data.device)
    _lib.call(fn, *args, 1, 0, _ptr(Cp), _ptr(Cj), _ptr(Cx), st)
    return DevMat(Cp, Cj, Cx, A.shape, (1, 1), 'bsr', not general)


def inverse_diagonal(A):
    d = A.diagonal()
    return torch.where(d != 0, 1.0 / d, torch.zeros_like(d))


def jacobi_prolongation_smoother(A, T, omega=4.0 / 3.0, degree=1, rho=None):
    """smooth.py:67-208, weighting='diagonal', filter=False -> (P, rho)"""
    dinv = inverse_diagonal(A)
    if rho is None:
        rho = approximate_spectral_radius(_scaled(A, dinv))
    else:
        np.random.rand(A.shape[1], 1)       # a fed estimate still consumes the start vector's draw (util/linalg.py:367)
    S = _scaled(A, dinv, omega / rho)
    P = T
    for _ in range(degree):
        P = csr_subtract(P, matmat(S, P, scipy_order=True), general=True)
    return P, rho


def smoothed_aggregation_solver(A, B=None, BH=None, symmetry='hermitian', strength='symmetric',
                                aggregate='standard', smooth=('jacobi', {'omega': 4.0 / 3.0}),
                                presmoother=('jacobi', {'omega': 4.0 / 3.0}),
                                postsmoother=('jacobi', {'omega': 4.0 / 3.0}),
                                improve_candidates=[('block_gauss_seidel', {'sweep': 'symmetric', 'iterations': 4}), None],
                                max_levels=10, max_coarse=500,
                                diagonal_dominance=False, keep=False, rho_override=None, **kwargs):
    """Same call as the reference's smoothed_aggregation_solver (pyamg/aggregation/aggregation.py:28-42) for the options
    that have a device path; unsupported options raise (no CPU fallback).  `improve_candidates` defaults to the
    reference's (4 symmetric block Gauss-Seidel sweeps on A B = 0 on the finest level): those sweeps are lexicographic,
    and run level-parallel over their dependency wavefronts with the reference's bits (relaxation.wavefront_ordering).
    One deviation from the reference DEFAULTS: `presmoother/postsmoother` default to Jacobi(4/3) instead of the
    lexicographic block Gauss-Seidel (available by asking for it:
    presmoother=postsmoother=('block_gauss_seidel', {'sweep': 'symmetric'})).  `rho_override` (list of per-level (rho_smooth, rho_relax) or None) lets the parity
    tests feed the reference's spectral-radius estimates; a fed estimate still draws (and drops) the start vector the
    estimate would have drawn, so that a seeded setup leaves the global numpy RNG where the reference leaves it."""
    if symmetry not in ('hermitian', 'symmetric'):
        raise NotImplementedError('only symmetric / hermitian problems have a device setup path')
    improve = _levelize_improve(improve_candidates, max_levels)
    if diagonal_dominance:
        raise NotImplementedError('diagonal_dominance filtering is not available on the device')
    sname, skw = _unpack(strength)
    if sname not in ('symmetric', None) or set(skw) - {'theta'}:
        raise NotImplementedError("strength must be ('symmetric', {'theta': t}) or None")
    theta = float(skw.get('theta', 0.0)) if sname is not None else 0.0
    if theta < 0:
        raise ValueError('expected a positive theta')
    aname = _unpack(aggregate)[0]
    if aname not in ('standard', 'naive'):
        raise NotImplementedError("aggregate must be 'standard' or 'naive'")
    mname, mkw = _unpack(smooth)
    if mname not in ('jacobi', None) or set(mkw) - {'omega', 'degree'}:
        raise NotImplementedError("smooth must be ('jacobi', {'omega', 'degree'}) or None")
    if not isinstance(A, DevMat):
        A = DevMat.from_scipy(A)
    if A.shape[0] != A.shape[1]:
        raise ValueError('expected square matrix')
    if A.blocksize[0] != A.blocksize[1]:
        raise ValueError('expected square blocks')
    dev = A.data.device
    bs = A.blocksize[0]
    if B is None:                       # one constant per unknown of a node (aggregation.py:234-236)
        Bn = np.kron(np.ones((A.shape[0] // bs, 1)), np.eye(bs))
    else:
        Bn = np.asarray(B.cpu().numpy() if isinstance(B, torch.Tensor) else B, dtype=np.float64)
        if Bn.ndim == 1:
            Bn = Bn.reshape(-1, 1)
        if Bn.shape[0] != A.shape[0]:
            raise ValueError('The near null-space modes B have incorrect dimensions for matrix A')
    K2 = Bn.shape[1]
    block = not A.is_scalar
    if block and sname is None:
        raise NotImplementedError("strength=None on a BSR matrix aggregates scalar unknowns in the reference; use 'symmetric'")
    if not block and K2 != 1:
        raise NotImplementedError('several candidates need a BSR matrix (the reference mixes CSR and BSR operands there)')
    Bv = torch.from_numpy(np.ascontiguousarray(Bn.ravel())).to(dev)

    levels = [multilevel_solver.level()]
    levels[-1].A = A
    levels[-1].B = Bv
    while len(levels) < max_levels and levels[-1].A.n_brow > max_coarse:
        lvl = levels[-1]
        k = len(levels) - 1
        lvl.B = improve_candidates_on_level(lvl.A, lvl.B, K2, improve[k])
        if block:
            # the same steps on blocks: aggregation on the block graph AS STORED (first-touch order out of bsr_matmat
            # on coarse levels), then the in-place block sort of get_diagonal
            K1 = lvl.A.blocksize[0]
            agg, nagg, roots = standard_aggregation(symmetric_strength_of_connection(lvl.A, theta), aname)
            T, Bc = fit_candidates_block(agg, nagg, lvl.B, K1, K2)
            A = lvl.A = sort_indices(lvl.A) if mname == 'jacobi' else lvl.A
            if mname == 'jacobi':
                rho = rho_override[k][0] if rho_override is not None and k < len(rho_override) else None
                P, lvl.rho_smooth = jacobi_prolongation_smoother_block(A, T, mkw.get('omega', 4.0 / 3.0),
                                                                       mkw.get('degree', 1), rho)
            else:
                P = T
            R = transpose(P)
            if keep:
                lvl.AggOp, lvl.T, lvl.Cpts = agg, T, roots
            lvl.P, lvl.R = P, R
            levels.append(multilevel_solver.level())
            levels[-1].A = galerkin(R, A, P)
            levels[-1].B = Bc
            continue
        # the reference aggregates on the matrix AS STORED at that moment (unsorted rows straight
        # out of csr_matmat on coarse levels; pass 2 of the aggregation picks the FIRST aggregated
        # neighbour in storage order) and only afterwards sorts it in place inside
        # jacobi_prolongation_smoother -> get_diagonal (pyamg/util/utils.py:571)
        agg, nagg, roots = standard_aggregation(lvl.A if theta == 0 else symmetric_strength_of_connection(lvl.A, theta), aname)
        T, Bc = fit_candidates_scalar(agg, nagg, lvl.B)
        A = lvl.A = sort_indices(lvl.A) if mname == 'jacobi' else lvl.A
        if mname == 'jacobi':
            rho = rho_override[k][0] if rho_override is not None and k < len(rho_override) else None
            P, lvl.rho_smooth = jacobi_prolongation_smoother(A, T, mkw.get('omega', 4.0 / 3.0), mkw.get('degree', 1), rho)
        else:
            P = T
        R = transpose(P)
        if keep:
            lvl.AggOp, lvl.T, lvl.Cpts = agg, T, roots
        lvl.P, lvl.R = P, R
        Ac = galerkin(R, A, P, scipy_order=True)
        levels.append(multilevel_solver.level())
        levels[-1].A = DevMat(Ac.indptr, Ac.indices, Ac.data, Ac.shape, (1, 1), 'bsr', False)
        levels[-1].B = Bc
    if rho_override is not None:
        for k, lvl in enumerate(levels[:-1]):
            if k < len(rho_override) and rho_override[k][1] is not None:
                lvl.A.rho_D_inv = rho_override[k][1]
                np.random.rand(lvl.A.shape[1], 1)       # keep the seeded random stream where the reference's is
    ml = multilevel_solver(levels, **kwargs)
    change_smoothers(ml, presmoother, postsmoother)
    return ml
