"""ruge_stuben_solver on the device (pyamg/classical/classical.py:22-190): classical strength of
connection, first-pass Ruge-Stuben C/F splitting, direct interpolation, R = P^T, Galerkin A_c = (R A) P.

Per level (extend_hierarchy, classical.py:120-190):
  C          b200amg_rs_strength -- amg_core.classical_strength_of_connection (ruge_stuben.h:45-93), row-parallel
  splitting  amg_core.rs_cf_splitting (ruge_stuben.h:157-297) -- sequential bucket algorithm, run on the HOST by
             b200amg_rs_cf_splitting_host (identical integers) on S = C without diagonal (rows sorted, like
             remove_diagonal's coo -> csr round trip, util/utils.py:1815-1827) and T = S^T, both made on the device
  P          b200amg_rs_direct_interpolation_pass1/2 (ruge_stuben.h:496-595) on C = pattern(S) .* A in the row
             order scipy's csr_elmul_csr gives it (interpolate.py:55-57): storage order when A is canonical,
             REVERSE storage order on coarse levels, whose rows come unsorted out of csr_matmat
  R          P^T (b200amg_csr_transpose)          A_c = (R A) P (b200amg_spgemm_*, scipy's row order)
Every floating-point step keeps the reference's operation order: the hierarchy equals the reference-built one
to the bit (tests/test_gpu_cycle.py::test_rs_setup_matches_reference).
"""
import numpy as np
import torch

from . import _lib, amg_core
from .multilevel import multilevel_solver
from .smoothing import change_smoothers
from .sparse import DevMat, _ptr, _scan_i32, _stream, galerkin, transpose

_DROP_DIAG, _DROP_ZERO, _REVERSE = 1, 2, 4


def _unpack(v):
    return (v[0], dict(v[1])) if isinstance(v, tuple) else (v, {})


def _strength(A, theta, flags):
    n = A.shape[0]
    st = _stream()
    dev = A.data.device
    cnt = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    args = (n, float(theta), _ptr(A.indptr), _ptr(A.indices), _ptr(A.data), int(flags))
    _lib.call('b200amg_rs_strength', *args, 0, _ptr(cnt), 0, 0, 0, st)
    Cp = _scan_i32(cnt[:n])
    nnz = int(Cp[-1].item())
    Cj = torch.empty(nnz, dtype=torch.int32, device=dev)
    Cx = torch.empty(nnz, dtype=torch.float64, device=dev)
    _lib.call('b200amg_rs_strength', *args, 1, 0, _ptr(Cp), _ptr(Cj), _ptr(Cx), st)
    return DevMat(Cp, Cj, Cx, A.shape, (1, 1), 'csr', None)


def is_canonical(A):
    """scipy's csr_has_canonical_format: every row strictly increasing in column"""
    flag = torch.zeros(1, dtype=torch.int32, device=A.data.device)
    _lib.call('b200amg_csr_not_canonical', A.shape[0], _ptr(A.indptr), _ptr(A.indices), _ptr(flag), _stream())
    return int(flag.item()) == 0


def classical_strength_of_connection(A, theta=0.0):
    """pyamg/strength.py:139-210: S keeps A's strong entries (and every stored diagonal entry) in A's row order.
    The reference then stores |a_ij| / max_k |a_ik| in S; everything downstream (RS, direct_interpolation) uses
    the PATTERN of S only, so the entries of A are kept as the values here."""
    if not A.is_scalar:
        raise NotImplementedError('classical strength: scalar (CSR) matrices only')
    if theta < 0 or theta > 1:
        raise ValueError('expected theta in [0,1]')
    return _strength(A, theta, 0)


def RS(S, graph=None):
    """pyamg/classical/split.py:110-156 -> int32 device tensor, 1 = C point, 0 = F point.  `S` is a strength matrix
    (DevMat); the diagonal is removed and rows sorted on the device, the sequential selection runs on the host.
    `graph`: S without its diagonal when the caller has it already."""
    n = S.shape[0]
    Snd = _strength(S, 0.0, _DROP_DIAG) if graph is None else graph
    T = transpose(Snd)                       # rows ascending in the original row index (csr_tocsc order)
    Ss = transpose(T)                        # = S without diagonal, rows sorted
    Sp, Sj = Ss.indptr.cpu().numpy(), Ss.indices.cpu().numpy()
    Tp, Tj = T.indptr.cpu().numpy(), T.indices.cpu().numpy()
    splitting = np.empty(n, dtype=np.intc)
    amg_core.rs_cf_splitting(n, Sp, Sj, Tp, Tj, splitting)
    return torch.from_numpy(splitting).to(S.data.device)


def direct_interpolation(A, C, splitting):
    """pyamg/classical/interpolate.py:13-83.  `C` must hold A's entries on the strength pattern in the row order
    scipy's `C.multiply(A)` produces (see _interp_pattern)."""
    n = A.shape[0]
    st = _stream()
    dev = A.data.device
    cnt = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    _lib.call('b200amg_rs_direct_interpolation_pass1', n, _ptr(C.indptr), _ptr(C.indices), _ptr(splitting), _ptr(cnt), st)
    Pp = _scan_i32(cnt[:n])
    cmap = _scan_i32(splitting)
    nc = int(cmap[-1].item())
    nnz = int(Pp[-1].item())
    Pj = torch.empty(nnz, dtype=torch.int32, device=dev)
    Px = torch.empty(nnz, dtype=torch.float64, device=dev)
    _lib.call('b200amg_rs_direct_interpolation_pass2', n, _ptr(A.indptr), _ptr(A.indices), _ptr(A.data), _ptr(C.indptr),
              _ptr(C.indices), _ptr(C.data), _ptr(splitting), _ptr(cmap), _ptr(Pp), _ptr(Pj), _ptr(Px), st)
    return DevMat(Pp, Pj, Px, (n, nc), (1, 1), 'csr', None)


def _interp_pattern(A, theta):
    """C = pattern(S) .* A as scipy's csr_elmul_csr returns it: exact zeros not stored; rows in storage order for a
    canonical A, reversed otherwise (csr_binop_csr_general walks its linked list backwards)"""
    return _strength(A, theta, _DROP_ZERO | (0 if is_canonical(A) else _REVERSE))


def ruge_stuben_solver(A, strength=('classical', {'theta': 0.25}), CF='RS',
                       presmoother=('gauss_seidel', {'sweep': 'symmetric'}),
                       postsmoother=('gauss_seidel', {'sweep': 'symmetric'}),
                       max_levels=10, max_coarse=500, keep=False, **kwargs):
    """Same call as the reference's ruge_stuben_solver for the options that have a device path
    (strength 'classical' or None, CF 'RS'); anything else raises (no CPU fallback).  The default smoother is the
    reference's lexicographic symmetric Gauss-Seidel -- exact but sequential on a GPU; pass
    ('multicolor_gauss_seidel', {'sweep': 'symmetric'}) or ('jacobi', ...) for a parallel one."""
    sname, skw = _unpack(strength)
    if sname not in ('classical', None):
        raise NotImplementedError("strength must be 'classical' or None on the device")
    if set(skw) - {'theta'}:
        raise TypeError('classical strength: unexpected options %r' % sorted(set(skw) - {'theta'}))
    if _unpack(CF)[0] != 'RS':
        raise NotImplementedError("CF must be 'RS' on the device")
    # strength=None makes every stored entry strong (C = A, classical.py:148): the same pattern as theta = 0
    theta = skw.get('theta', 0.0) if sname is not None else 0.0
    if theta < 0 or theta > 1:
        raise ValueError('expected theta in [0,1]')
    if not isinstance(A, DevMat):
        import scipy.sparse as sp
        if sp.issparse(A) and A.format != 'csr':
            A = A.tocsr()
        A = DevMat.from_scipy(A)
    if not A.is_scalar or A.format != 'csr':
        raise TypeError('Argument A must have type csr_matrix, or be convertible to csr_matrix')
    if A.shape[0] != A.shape[1]:
        raise ValueError('expected square matrix')
    levels = [multilevel_solver.level()]
    levels[-1].A = A
    while len(levels) < max_levels and levels[-1].A.shape[0] > max_coarse:
        lvl = levels[-1]
        A = lvl.A
        # graph of the splitting: strong entries without the diagonal (explicit zeros stay: they are entries of S)
        splitting = RS(A, graph=_strength(A, theta, _DROP_DIAG))
        C = _interp_pattern(A, theta)
        P = direct_interpolation(A, C, splitting)
        R = transpose(P)
        if keep:
            lvl.C, lvl.splitting = C, splitting
        lvl.P, lvl.R = P, R
        Ac = galerkin(R, A, P, scipy_order=True)
        levels.append(multilevel_solver.level())
        levels[-1].A = DevMat(Ac.indptr, Ac.indices, Ac.data, Ac.shape, (1, 1), 'csr', False)
    ml = multilevel_solver(levels, **kwargs)
    change_smoothers(ml, presmoother, postsmoother)
    return ml
