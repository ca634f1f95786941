"""TEST INFRASTRUCTURE ONLY -- restatement of the reference's Ruge-Stuben SETUP glue, so that
classical-AMG hierarchies for the parity tests / CPU baseline can be built where /root/reference
is absent (the GPU box).  Native pieces come from ``oracle.core`` (the reference's compiled
_amg_core when built, else the C port); sparse algebra is scipy's, with the SAME expression
sequence as the reference so storage orders and fp orders coincide.

Follows: pyamg/classical/classical.py:22-190 (driver, extend_hierarchy),
pyamg/strength.py:139-210 (classical_strength_of_connection; the |.| and row scaling of :205-208
change values only and the values of S are never used below, so they are left out),
pyamg/classical/split.py:110-156 (RS) with pyamg/util/utils.py:1815-1827 (remove_diagonal),
pyamg/classical/interpolate.py:13-83 (direct_interpolation).

Validated against the real Python reference by tests/golden/make_golden.py / the RS fixtures in
tests/golden/ (tests/test_oracle.py::test_rs_setup_reproduces_reference_hierarchy).
"""
import numpy as np
import scipy.sparse as sparse

import oracle as _o
from .multilevel import Level


def classical_strength_of_connection(A, theta=0.0, core=None):
    core = core or _o.core
    Sp = np.empty_like(A.indptr)
    Sj = np.empty_like(A.indices)
    Sx = np.empty_like(A.data)
    core.classical_strength_of_connection(A.shape[0], theta, A.indptr, A.indices, A.data, Sp, Sj, Sx)
    return sparse.csr_matrix((Sx, Sj, Sp), shape=A.shape)


def remove_diagonal(S):
    S = sparse.coo_matrix(S)
    mask = S.row != S.col
    S.row, S.col, S.data = S.row[mask], S.col[mask], S.data[mask]
    return S.tocsr()


def RS(S, core=None):
    core = core or _o.core
    S = remove_diagonal(S)
    T = S.T.tocsr()
    splitting = np.empty(S.shape[0], dtype='intc')
    core.rs_cf_splitting(S.shape[0], S.indptr, S.indices, T.indptr, T.indices, splitting)
    return splitting


def direct_interpolation(A, C, splitting, core=None):
    core = core or _o.core
    C = C.copy()
    C.data[:] = 1.0
    C = C.multiply(A)
    if not sparse.isspmatrix_csr(C):
        C = sparse.csr_matrix(C)
    Pp = np.empty_like(A.indptr)
    core.rs_direct_interpolation_pass1(A.shape[0], C.indptr, C.indices, splitting, Pp)
    nnz = Pp[-1]
    Pj = np.empty(nnz, dtype=Pp.dtype)
    Px = np.empty(nnz, dtype=A.dtype)
    core.rs_direct_interpolation_pass2(A.shape[0], A.indptr, A.indices, A.data, C.indptr, C.indices, C.data,
                                       splitting, Pp, Pj, Px)
    return sparse.csr_matrix((Px, Pj, Pp))


def ruge_stuben_hierarchy(A, theta=0.25, max_levels=10, max_coarse=500, core=None, keep=False):
    """-> list of oracle.multilevel.Level with A, P, R (smoothers are attached by the caller)"""
    A = sparse.csr_matrix(A).astype(np.float64)
    levels = [Level()]
    levels[-1].A = A
    while len(levels) < max_levels and levels[-1].A.shape[0] > max_coarse:
        A = levels[-1].A
        C = classical_strength_of_connection(A, theta, core)
        splitting = RS(C, core)
        P = direct_interpolation(A, C, splitting, core)
        R = P.T.tocsr()
        if keep:
            levels[-1].C, levels[-1].splitting = C, splitting
        levels[-1].P, levels[-1].R = P, R
        levels.append(Level())
        levels[-1].A = R * A * P
    return levels
