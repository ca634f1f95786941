"""Pre-hardware check of the block (BSR) setup kernels: pyamg_b200/csrc/bsr_setup.cu is compiled UNCHANGED for the host
with tests/emu/cuda_runtime.h (every CUDA thread an OS thread, __syncwarp a barrier) and every entry point is compared
bit for bit with scipy / the reference's amg_core on small block matrices, then a whole block smoothed-aggregation
hierarchy is assembled from the emulated kernels and compared with the fixture the REAL reference produced
(tests/golden/sa_elasticity3d_tets_7_setup.npz).

This is test infrastructure: the emulated library takes host pointers and is never loaded by pyamg_b200.  It checks the
kernel SOURCE (indexing, operation order, storage order) -- launch geometry and device behaviour are what the
`-m gpu` tests check on hardware."""
import ctypes
import os
import shutil
import subprocess

import numpy as np
import pytest
import scipy.sparse as sp

import hier_io
import oracle
from oracle import sa as osa

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
EMU = os.path.join(HERE, 'emu')
c_int, c_ll, c_dbl, vp = ctypes.c_int, ctypes.c_longlong, ctypes.c_double, ctypes.c_void_p


def build_emu(stem, extra=()):
    """pyamg_b200/csrc/<stem>.cu (+ extra stems) compiled for the host with the stand-in CUDA header -> ctypes library"""
    if shutil.which('g++') is None:
        pytest.skip('g++ not available')
    out = os.path.join(EMU, '_build', 'libemu_%s.so' % stem)
    srcs = [os.path.join(ROOT, 'pyamg_b200', 'csrc', st + '.cu') for st in (stem,) + tuple(extra)]
    src = srcs[0]
    deps = srcs + [os.path.join(EMU, 'cuda_runtime.h'), os.path.join(EMU, 'emu_globals.cpp'),
            os.path.join(ROOT, 'pyamg_b200', 'csrc', 'common.cuh'), os.path.join(ROOT, 'pyamg_b200', 'csrc', 'internal.cuh'),
            os.path.join(ROOT, 'include', 'b200amg.h')]
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(d) for d in deps):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.run(['g++', '-std=c++20', '-O1', '-ffp-contract=off', '-fPIC', '-shared', '-pthread', '-I', EMU,
                        '-x', 'c++'] + srcs + [os.path.join(EMU, 'emu_globals.cpp'), '-o', out], check=True)
    return ctypes.CDLL(out)


@pytest.fixture(scope='module')
def emu():
    lib = build_emu('bsr_setup')
    sigs = {
        'b200amg_bsr_scale_rows': [c_int, c_int, c_int, vp, vp, vp, c_dbl, c_int, vp, vp],
        'b200amg_bsr_gather_blocks': [c_ll, c_int, c_int, vp, vp, vp, c_int, vp],
        'b200amg_fit_candidates': [c_int, c_int, c_int, vp, vp, vp, c_dbl, vp, vp, vp],
        'b200amg_tentative_scatter': [c_ll, c_int, c_int, vp, vp, vp, vp, vp],
        'b200amg_bsr_sub': [c_int, c_int, c_int, vp, vp, vp, vp, vp, vp, c_int, c_int, vp, vp, vp, vp, vp],
        'b200amg_bsr_slot_map': [c_int, vp, vp, vp, vp, vp, vp],
        'b200amg_bsr_matmat_numeric': [c_int, c_int, c_int, c_int, vp, vp, vp, vp, vp, vp, vp, vp, vp, c_ll, vp, vp],
    }
    for name, args in sigs.items():
        f = getattr(lib, name)
        f.restype, f.argtypes = c_int, args
    # the signatures used here are the ones the product binds (pyamg_b200/_lib.py)
    from pyamg_b200 import _lib
    for name, args in sigs.items():
        assert len(_lib._SIGS[name][1]) == len(args), name
    return lib


def P(a):
    return a.ctypes.data if a is not None else 0


def i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


# ---- host mirrors of the product's Python orchestration (pyamg_b200/sparse.py, aggregation.py) -----------------
def pattern(M):
    return sp.csr_matrix((np.ones(len(M.indices)), M.indices, M.indptr), shape=(M.shape[0] // M.blocksize[0],
                                                                                 M.shape[1] // M.blocksize[1]))


def emu_bsr_matmat(lib, A, B):
    """sparse.bsr_matmat: pattern = scalar SpGEMM of the block patterns (scipy's csr_matmat here; the device's own
    SpGEMM is pinned to it bit for bit by tests/test_gpu_kernels.py), blocks by slot_map + numeric"""
    R, N = A.blocksize
    _, C = B.blocksize
    L = pattern(A) * pattern(B)                          # csr_matmat order = reverse first touch
    Ssorted = L.copy()
    Ssorted.sort_indices()
    Cp, Sj, Lj = i32(L.indptr), i32(Ssorted.indices), i32(L.indices)
    nb = len(Lj)
    slot, Cj = np.full(nb, -1, np.int32), np.full(nb, -1, np.int32)
    Cx = np.full(nb * R * C, np.nan)
    nbrow = A.shape[0] // R
    assert lib.b200amg_bsr_slot_map(nbrow, P(Cp), P(Sj), P(Lj), P(slot), P(Cj), None) == 0
    Ap, Aj, Ax = i32(A.indptr), i32(A.indices), np.ascontiguousarray(A.data).ravel()
    Bp, Bj, Bx = i32(B.indptr), i32(B.indices), np.ascontiguousarray(B.data).ravel()
    assert lib.b200amg_bsr_matmat_numeric(nbrow, R, N, C, P(Ap), P(Aj), P(Ax), P(Bp), P(Bj), P(Bx), P(Cp), P(Sj), P(slot),
                                          nb, P(Cx), None) == 0
    return sp.bsr_matrix((Cx.reshape(-1, R, C), Cj, Cp), shape=(A.shape[0], B.shape[1]), blocksize=(R, C))


def canonical(M):
    return all(np.all(np.diff(M.indices[M.indptr[i]:M.indptr[i + 1]]) > 0) for i in range(len(M.indptr) - 1))


def emu_bsr_sub(lib, A, B):
    R, C = A.blocksize
    n = A.shape[0] // R
    general = int(not (canonical(A) and canonical(B)))
    Ap, Aj, Ax = i32(A.indptr), i32(A.indices), np.ascontiguousarray(A.data).ravel()
    Bp, Bj, Bx = i32(B.indptr), i32(B.indices), np.ascontiguousarray(B.data).ravel()
    cnt = np.zeros(n, np.int32)
    assert lib.b200amg_bsr_sub(n, R, C, P(Ap), P(Aj), P(Ax), P(Bp), P(Bj), P(Bx), general, 0, P(cnt), None, None, None, None) == 0
    Cp = i32(np.concatenate([[0], np.cumsum(cnt)]))
    Cj = np.full(Cp[-1], -1, np.int32)
    Cx = np.full(Cp[-1] * R * C, np.nan)
    assert lib.b200amg_bsr_sub(n, R, C, P(Ap), P(Aj), P(Ax), P(Bp), P(Bj), P(Bx), general, 1, None, P(Cp), P(Cj), P(Cx), None) == 0
    return sp.bsr_matrix((Cx.reshape(-1, R, C), Cj, Cp), shape=A.shape, blocksize=(R, C))


def emu_gather(lib, A, perm, tr):
    R, C = A.blocksize
    data = np.ascontiguousarray(A.data).ravel()
    out = np.full_like(data, np.nan)
    permd = np.ascontiguousarray(perm, dtype=np.float64)
    assert lib.b200amg_bsr_gather_blocks(len(perm), R, C, P(permd), P(data), P(out), int(tr), None) == 0
    return out


def emu_bsr_transpose(lib, A):
    """sparse.bsr_transpose: csr_tocsc over block numbers (scipy here), then the block gather + transpose"""
    R, C = A.blocksize
    ids = sp.csr_matrix((np.arange(len(A.indices), dtype=np.float64), A.indices, A.indptr),
                        shape=(A.shape[0] // R, A.shape[1] // C))
    T = ids.tocsc()                                      # (indptr, indices, data) of the CSC = CSR of the transpose
    out = emu_gather(lib, A, T.data, True)
    return sp.bsr_matrix((out.reshape(-1, C, R), i32(T.indices), i32(T.indptr)), shape=(A.shape[1], A.shape[0]), blocksize=(C, R))


def emu_bsr_sort(lib, A):
    R, C = A.blocksize
    ids = sp.csr_matrix((np.arange(len(A.indices), dtype=np.float64), A.indices.copy(), A.indptr.copy()),
                        shape=(A.shape[0] // R, A.shape[1] // C))
    ids.sort_indices()
    out = emu_gather(lib, A, ids.data, False)
    return sp.bsr_matrix((out.reshape(-1, R, C), i32(ids.indices), i32(ids.indptr)), shape=A.shape, blocksize=(R, C))


def emu_scale_rows(lib, A, v, c=None):
    R, C = A.blocksize
    data = np.ascontiguousarray(A.data).ravel()
    out = np.full_like(data, np.nan)
    Ap = i32(A.indptr)
    v = np.ascontiguousarray(v, dtype=np.float64)
    assert lib.b200amg_bsr_scale_rows(A.shape[0] // R, R, C, P(Ap), P(data), P(v), 0.0 if c is None else float(c),
                                      0 if c is None else 1, P(out), None) == 0
    return sp.bsr_matrix((out.reshape(-1, R, C), A.indices, A.indptr), shape=A.shape, blocksize=(R, C))


def emu_fit_candidates(lib, agg, nagg, B, K1, K2, tol=1e-10):
    """aggregation.fit_candidates_block"""
    n = len(agg)
    keep = agg >= 0
    AggOp = sp.csr_matrix((np.ones(keep.sum()), agg[keep], np.concatenate([[0], np.cumsum(keep)])), shape=(n, nagg))
    M = AggOp.tocsc()
    Mp, Mi = i32(M.indptr), i32(M.indices)
    nmem = len(Mi)
    Bf = np.ascontiguousarray(B, dtype=np.float64).ravel()
    Qx = np.full(nmem * K1 * K2, np.nan)
    Rx = np.full(nagg * K2 * K2, np.nan)
    assert lib.b200amg_fit_candidates(nagg, K1, K2, P(Mp), P(Mi), P(Bf), tol, P(Qx), P(Rx), None) == 0
    Tp = i32(AggOp.indptr)
    Tx = np.full(nmem * K1 * K2, np.nan)
    assert lib.b200amg_tentative_scatter(nmem, K1, K2, P(Mi), P(Tp), P(Qx), P(Tx), None) == 0
    T = sp.bsr_matrix((Tx.reshape(-1, K1, K2), i32(AggOp.indices), Tp), shape=(n * K1, nagg * K2), blocksize=(K1, K2))
    return T, Rx.reshape(-1, K2)


def same(M, Rf):
    return (M.format == Rf.format and getattr(M, 'blocksize', (1, 1)) == getattr(Rf, 'blocksize', (1, 1)) and M.shape == Rf.shape
            and np.array_equal(M.indptr, Rf.indptr) and np.array_equal(M.indices, Rf.indices)
            and np.array_equal(np.ravel(M.data), np.ravel(Rf.data)))


def random_bsr(nbr, nbc, R, C, density, seed, unsorted=False):
    rng = np.random.RandomState(seed)
    pat = sp.random(nbr, nbc, density=density, format='csr', random_state=rng)
    pat.sort_indices()
    indptr, indices = pat.indptr.astype(np.int32), pat.indices.astype(np.int32)
    if unsorted:
        for i in range(nbr):
            seg = indices[indptr[i]:indptr[i + 1]]
            indices[indptr[i]:indptr[i + 1]] = seg[rng.permutation(len(seg))]
    data = rng.randn(len(indices), R, C)
    data[rng.rand(*data.shape) < 0.2] = 0.0                 # exact zeros inside blocks
    return sp.bsr_matrix((data, indices, indptr), shape=(nbr * R, nbc * C), blocksize=(R, C))


# ---- entry points against scipy ----------------------------------------------------------------------------------
@pytest.mark.parametrize('R,N,C', [(3, 3, 6), (6, 3, 3), (6, 3, 6), (2, 2, 3), (6, 6, 6), (1, 2, 1)])
def test_bsr_matmat_equals_scipy(emu, R, N, C):
    """first-touch block order, zero blocks kept, visiting-order sums: identical arrays to scipy's bsr_matmat, for
    sorted and unsorted operands"""
    for seed, unsorted in ((1, False), (2, True)):
        A = random_bsr(37, 29, R, N, 0.2, seed, unsorted)
        B = random_bsr(29, 41, N, C, 0.15, seed + 10, unsorted)
        ref = A * B
        assert ref.format == 'bsr' and ref.blocksize == (R, C)
        assert same(emu_bsr_matmat(emu, A, B), ref)


def test_bsr_matmat_keeps_zero_blocks_and_empty_rows(emu):
    A = random_bsr(12, 9, 2, 2, 0.3, 5)
    B = random_bsr(9, 7, 2, 2, 0.3, 6)
    B.data[:] = 0.0                                         # every product block is exactly zero: scipy keeps them
    A.data[A.indptr[3]:A.indptr[4]] = 0.0
    ref = A * B
    got = emu_bsr_matmat(emu, A, B)
    assert same(got, ref) and len(ref.indices) > 0
    E = sp.bsr_matrix((12 * 2, 9 * 2), blocksize=(2, 2))
    assert same(emu_bsr_matmat(emu, E, B), E * B)


@pytest.mark.parametrize('R,C', [(3, 6), (6, 6), (2, 3)])
def test_bsr_sub_equals_scipy(emu, R, C):
    """general (reverse first-touch) and canonical (merge) forms of scipy's bsr_binop_bsr, dropped zero blocks included"""
    for seed, unsorted in ((3, True), (4, False)):
        A = random_bsr(40, 23, R, C, 0.08, seed, False)
        B = random_bsr(40, 23, R, C, 0.2, seed + 20, unsorted)
        # make some blocks cancel exactly (same column, same values) and some rows empty
        Bl = B.tolil() if False else B
        for i in range(0, 40, 5):
            a0, a1 = A.indptr[i], A.indptr[i + 1]
            b0, b1 = B.indptr[i], B.indptr[i + 1]
            common = np.intersect1d(A.indices[a0:a1], B.indices[b0:b1])
            for col in common:
                pa = a0 + int(np.where(A.indices[a0:a1] == col)[0][0])
                pb = b0 + int(np.where(B.indices[b0:b1] == col)[0][0])
                B.data[pb] = A.data[pa]
        ref = A - B
        assert same(emu_bsr_sub(emu, A, B), ref), (seed, unsorted)
    assert canonical(random_bsr(40, 23, R, C, 0.2, 24, False)) and not canonical(random_bsr(40, 23, R, C, 0.2, 23, True))


def test_bsr_transpose_sort_scale_equal_scipy(emu):
    A = random_bsr(31, 17, 3, 6, 0.25, 7, unsorted=True)
    T = A.T
    assert T.format == 'bsr' and same(emu_bsr_transpose(emu, A), T)
    Ssrt = A.copy()
    Ssrt.sort_indices()
    assert same(emu_bsr_sort(emu, A), Ssrt)
    Sq = random_bsr(20, 20, 3, 3, 0.3, 8)
    v = np.random.RandomState(0).rand(60)
    assert same(emu_scale_rows(emu, Sq, v), osa.scale_rows(Sq, v))
    ref = 0.37 * osa.scale_rows(Sq, v)
    assert same(emu_scale_rows(emu, Sq, v, 0.37), ref)


@pytest.mark.parametrize('K1,K2', [(3, 6), (6, 6), (2, 3), (1, 1), (1, 2)])
def test_fit_candidates_equals_reference_kernel(emu, K1, K2):
    """amg_core.fit_candidates (the reference's compiled kernel when built, else the C port) incl. unaggregated nodes,
    dependent candidates (dropped by the tolerance) and single-node aggregates"""
    rng = np.random.RandomState(K1 * 10 + K2)
    n, nagg = 57, 9
    agg = rng.randint(0, nagg, n).astype(np.int32)
    agg[[4, 19]] = -1                                       # not aggregated
    agg[agg == 8] = 7
    agg[5] = 8                                              # aggregate 8 = one node
    B = rng.randn(n * K1, K2)
    if K2 > 1:
        B[:, 1] = 2.0 * B[:, 0]                             # numerically dependent second candidate
    keep = agg >= 0
    AggOp = sp.csr_matrix((np.ones(keep.sum(), 'int8'), agg[keep], np.concatenate([[0], np.cumsum(keep)]).astype(np.int32)),
                          shape=(n, nagg))
    for core in ([oracle.ref] if oracle.ref is not None else []) + [oracle.port]:
        Tref, Rref = osa.fit_candidates(AggOp, B, core=core)
        T, Rc = emu_fit_candidates(emu, agg, nagg, B, K1, K2)
        assert same(T, Tref)
        assert np.array_equal(Rc, Rref)


# ---- the whole block hierarchy from the emulated kernels ------------------------------------------------------------
def test_block_sa_hierarchy_equals_reference_fixture(emu, golden_dir):
    """pyamg_b200.aggregation's block path, step for step, on the kernels' source: A, P, R of every level equal the arrays
    the REAL reference built for P1 tetrahedral elasticity (BSR 3x3 -> 6x6, six rigid-body modes), given the reference's
    spectral-radius estimates (an unconverged Krylov number, fed like in test_sa_setup_matches_reference)."""
    levels, ex = hier_io.load(os.path.join(golden_dir, 'sa_elasticity3d_tets_7_setup.npz'))
    A, B = levels[0].A, ex['B']
    gs = ('gauss_seidel', {'sweep': 'symmetric'})
    np.random.seed(0)
    ref = osa.smoothed_aggregation_hierarchy(A.copy(), B, presmoother=gs, postsmoother=gs, max_coarse=10, improve_candidates=None)
    assert len(ref) == len(levels) == 3
    K2 = B.shape[1]
    cur_A, cur_B = A.copy(), B
    omega = 4.0 / 3.0
    for k in range(len(levels) - 1):
        K1 = cur_A.blocksize[0]
        nb = cur_A.shape[0] // K1
        agg = np.empty(nb, np.int32)
        roots = np.empty(nb, np.int32)
        nagg = oracle.core.standard_aggregation(nb, i32(cur_A.indptr), i32(cur_A.indices), agg, roots)   # host in the product too
        T, Bc = emu_fit_candidates(emu, agg, nagg, cur_B, K1, K2)
        assert same(T, ref[k].T)
        cur_A = emu_bsr_sort(emu, cur_A)                   # get_diagonal's in-place sort: what the reference keeps as A
        assert same(cur_A, levels[k].A), k
        d = cur_A.diagonal()
        dinv = np.where(d != 0, 1.0 / np.where(d != 0, d, 1.0), 0.0)
        S = emu_scale_rows(emu, cur_A, dinv, omega / ref[k].rho_smooth)
        U = emu_bsr_matmat(emu, S, T)
        Pk = emu_bsr_sub(emu, T, U)
        Rk = emu_bsr_transpose(emu, Pk)
        Ac = emu_bsr_matmat(emu, emu_bsr_matmat(emu, Rk, cur_A), Pk)
        for got, name in ((Pk, 'P'), (Rk, 'R')):
            assert same(got, getattr(levels[k], name)), (k, name)
        assert np.array_equal(Bc, ref[k + 1].B)
        cur_A, cur_B = Ac, Bc
    assert same(cur_A, levels[-1].A)                       # the coarsest operator stays in bsr_matmat's first-touch order


# ---- block relaxation (csrc/block_relax.cu) against the reference's compiled kernels -----------------------------------
@pytest.fixture(scope='module')
def emu_relax():
    lib = build_emu('pinv', extra=('block_relax',))
    from pyamg_b200 import _lib
    for name in ('b200amg_block_jacobi', 'b200amg_block_gauss_seidel', 'b200amg_block_diag_inverse',
                 'b200amg_block_diag_inverse_gj', 'b200amg_pinv_array'):
        f = getattr(lib, name)
        f.restype, f.argtypes = _lib._SIGS[name]
    return lib


def block_problem(nb, bs, seed, diag=True):
    rng = np.random.RandomState(seed)
    pat = sp.csr_matrix(sp.random(nb, nb, 0.15, random_state=rng, format='csr') + (sp.identity(nb) if diag else 0))
    pat = sp.csr_matrix(pat + pat.T)
    pat.sort_indices()
    data = rng.randn(len(pat.indices), bs, bs)
    A = sp.bsr_matrix((data, pat.indices.astype(np.int32), pat.indptr.astype(np.int32)), shape=(nb * bs, nb * bs))
    return A, rng.randn(nb, bs, bs), rng.randn(nb * bs), rng.randn(nb * bs)


CORE = oracle.ref if oracle.ref is not None else oracle.port          # the reference's own block kernels when built


@pytest.mark.parametrize('bs', [1, 2, 3, 5, 6, 8])
def test_block_gauss_seidel_sequential_equals_reference(emu_relax, bs):
    A, Dinv, b, x0 = block_problem(45, bs, bs)
    Ap, Aj, Ax, Tx = i32(A.indptr), i32(A.indices), np.ravel(A.data).copy(), np.ravel(Dinv).copy()
    for r0, r1, dr in ((0, 45, 1), (44, -1, -1), (3, 33, 3), (7, 7, 1)):
        want, got = x0.copy(), x0.copy()
        if r0 != r1:
            CORE.block_gauss_seidel(Ap, Aj, Ax, want, b, Tx, r0, r1, dr, bs)
        assert emu_relax.b200amg_block_gauss_seidel(P(Ap), P(Aj), P(Ax), P(got), P(b), P(Tx), r0, r1, dr, bs, None, None, 0, None) == 0
        assert np.array_equal(got, want), (bs, r0, r1, dr)
    # an index list with repeats, walked backwards: one reference call per listed block row
    Id = i32(np.random.RandomState(1).randint(0, 45, 60))
    want, got = x0.copy(), x0.copy()
    for p_ in range(59, -1, -1):
        CORE.block_gauss_seidel(Ap, Aj, Ax, want, b, Tx, int(Id[p_]), int(Id[p_]) - 1, -1, bs)
    assert emu_relax.b200amg_block_gauss_seidel(P(Ap), P(Aj), P(Ax), P(got), P(b), P(Tx), 59, -1, -1, bs, P(Id), None, 0, None) == 0
    assert np.array_equal(got, want)
    assert emu_relax.b200amg_block_gauss_seidel(P(Ap), P(Aj), P(Ax), P(got), P(b), P(Tx), 0, 10, -1, bs, None, None, 0, None) != 0


@pytest.mark.parametrize('bs', [2, 3, 6])
def test_block_gauss_seidel_colour_ordered_equals_row_by_row_reference(emu_relax, bs):
    """the parallel form: block rows grouped by MIS colour of the block graph, one launch per colour, forward and
    backward, against the reference kernel driven one block row at a time in the same order"""
    from oracle import relaxation as orx
    A, Dinv, b, x0 = block_problem(150, bs, 10 + bs)
    Ap, Aj, Ax, Tx = i32(A.indptr), i32(A.indices), np.ravel(A.data).copy(), np.ravel(Dinv).copy()
    G = sp.csr_matrix((np.ones(len(Aj)), Aj, Ap), shape=(150, 150))
    idx, ptr = orx.color_ordering(orx.vertex_coloring_mis(G, oracle.port))
    idx, ptr = i32(idx), np.ascontiguousarray(ptr, dtype=np.intc)
    K = len(ptr) - 1
    assert K >= 3
    cp = ptr.ctypes.data_as(ctypes.POINTER(ctypes.c_int))
    want, got = x0.copy(), x0.copy()
    for i in idx:
        CORE.block_gauss_seidel(Ap, Aj, Ax, want, b, Tx, int(i), int(i) + 1, 1, bs)
    for i in idx[::-1]:
        CORE.block_gauss_seidel(Ap, Aj, Ax, want, b, Tx, int(i), int(i) - 1, -1, bs)
    n = len(idx)
    assert emu_relax.b200amg_block_gauss_seidel(P(Ap), P(Aj), P(Ax), P(got), P(b), P(Tx), 0, n, 1, bs, P(idx), cp, K, None) == 0
    assert emu_relax.b200amg_block_gauss_seidel(P(Ap), P(Aj), P(Ax), P(got), P(b), P(Tx), n - 1, -1, -1, bs, P(idx), cp, K, None) == 0
    assert np.array_equal(got, want)
    # a range that cuts a colour is refused
    assert emu_relax.b200amg_block_gauss_seidel(P(Ap), P(Aj), P(Ax), P(got), P(b), P(Tx), 0, n - 1, 1, bs, P(idx), cp, K, None) != 0


@pytest.mark.parametrize('bs', [1, 2, 3, 6, 7])
def test_block_jacobi_equals_reference(emu_relax, bs):
    A, Dinv, b, x0 = block_problem(70, bs, 20 + bs)
    Ap, Aj, Ax, Tx = i32(A.indptr), i32(A.indices), np.ravel(A.data).copy(), np.ravel(Dinv).copy()
    for (r0, r1, dr), omega in (((0, 70, 1), 0.7), ((0, 70, 1), 1.0), ((4, 64, 4), 1.3), ((69, -1, -1), 0.5)):
        want, got = x0.copy(), x0.copy()
        tw, tg = np.full_like(x0, 7.0), np.full_like(x0, 7.0)
        CORE.block_jacobi(Ap, Aj, Ax, want, b, Tx, tw, r0, r1, dr, np.array([omega]), bs)
        assert emu_relax.b200amg_block_jacobi(P(Ap), P(Aj), P(Ax), P(got), P(b), P(Tx), P(tg), r0, r1, dr, omega, bs, None) == 0
        # partial sweeps read temp entries outside the sweep: both sides see the same caller-filled values
        assert np.array_equal(got, want) and np.array_equal(tg, tw), (bs, r0, r1, dr)


@pytest.mark.parametrize('bs', [1, 2, 3, 6, 8])
def test_block_diag_inverse_gauss_jordan(emu_relax, bs):
    """the Gauss-Jordan path (used for 7x7 / 8x8 blocks, where the reference calls LAPACK gelss: no bit contract): to
    rounding against numpy; no diagonal block or a singular one -> zero block"""
    rng = np.random.RandomState(bs)
    nb = 60
    A, _, _, _ = block_problem(nb, bs, 30 + bs)
    A = A.copy()
    for i in range(nb):                                     # SPD-ish diagonal blocks like a stiffness matrix
        q = A.indptr[i] + int(np.where(A.indices[A.indptr[i]:A.indptr[i + 1]] == i)[0][0])
        M = rng.randn(bs, bs)
        A.data[q] = M @ M.T + bs * np.eye(bs)
    q5 = A.indptr[5] + int(np.where(A.indices[A.indptr[5]:A.indptr[6]] == 5)[0][0])
    A.data[q5] = 0.0
    if bs > 1:
        q9 = A.indptr[9] + int(np.where(A.indices[A.indptr[9]:A.indptr[10]] == 9)[0][0])
        A.data[q9] = np.ones((bs, bs))                      # rank one
    Ap, Aj, Ax = i32(A.indptr), i32(A.indices), np.ravel(A.data).copy()
    Dinv = np.full(nb * bs * bs, np.nan)
    assert emu_relax.b200amg_block_diag_inverse_gj(nb, bs, P(Ap), P(Aj), P(Ax), P(Dinv), None) == 0
    Dinv = Dinv.reshape(nb, bs, bs)
    for i in range(nb):
        q = A.indptr[i] + int(np.where(A.indices[A.indptr[i]:A.indptr[i + 1]] == i)[0][0])
        if i == 5 or (i == 9 and bs > 1):
            assert np.all(Dinv[i] == 0.0)
        else:
            ref = np.linalg.inv(A.data[q])
            assert np.abs(Dinv[i] - ref).max() <= 1e-12 * np.abs(ref).max()


@pytest.mark.parametrize('n', [1, 2, 3, 4, 5, 6])
def test_pinv_array_equals_reference_kernel(emu_relax, n):
    """amg_core.pinv_array (one-sided Jacobi SVD): SPD, general, zero and rank-deficient blocks, both storage orders --
    the bits of the reference's compiled kernel"""
    rng = np.random.RandomState(n)
    a = rng.randn(150, n, n)
    for i in range(0, 150, 4):
        M = rng.randn(n, n)
        a[i] = M @ M.T + n * np.eye(n)
    a[7] = 0.0
    if n > 1:
        a[9] = np.ones((n, n))
        a[11][:, 0] = a[11][:, 1]
    for trans in ('T', 'F'):
        want, got = a.copy(), a.copy()
        CORE.pinv_array(want.ravel(), 150, n, trans)
        assert emu_relax.b200amg_pinv_array(P(got), 150, n, 1 if trans == 'T' else 0, None) == 0
        assert np.array_equal(got, want), (n, trans)
    assert emu_relax.b200amg_pinv_array(P(a), 150, 7, 1, None) != 0


@pytest.mark.parametrize('bs', [2, 3, 6])
def test_block_diag_inverse_equals_reference_get_block_diag(emu_relax, bs):
    """the Dinv the block smoothers get by default: diagonal blocks (zeros where a block row stores none) through
    pinv_array(..., 'T') -- pyamg/util/utils.py:655-677"""
    A, _, _, _ = block_problem(80, bs, 40 + bs)
    # drop the diagonal block of a few rows
    keep = np.ones(len(A.indices), bool)
    for i in (3, 17):
        q = A.indptr[i] + int(np.where(A.indices[A.indptr[i]:A.indptr[i + 1]] == i)[0][0])
        keep[q] = False
    cnt = np.add.reduceat(keep.astype(np.int64), A.indptr[:-1])
    A2 = sp.bsr_matrix((A.data[keep], A.indices[keep], np.concatenate([[0], np.cumsum(cnt)]).astype(np.int32)), shape=A.shape)
    Ap, Aj, Ax = i32(A2.indptr), i32(A2.indices), np.ravel(A2.data).copy()
    want = np.zeros((80, bs, bs))
    for i in range(80):
        for q in range(Ap[i], Ap[i + 1]):
            if Aj[q] == i:
                want[i] = A2.data[q]
    CORE.pinv_array(want.ravel(), 80, bs, 'T')
    got = np.full(80 * bs * bs, np.nan)
    assert emu_relax.b200amg_block_diag_inverse(80, bs, P(Ap), P(Aj), P(Ax), P(got), None) == 0
    assert np.array_equal(got.reshape(80, bs, bs), want)
    assert np.all(want[3] == 0.0)


# ---- symmetric strength of connection, theta > 0 (csrc/strength.cu) ------------------------------------------------
@pytest.fixture(scope='module')
def emu_strength():
    lib = build_emu('strength')
    from pyamg_b200 import _lib
    for name in ('b200amg_block_sqnorm', 'b200amg_abs_diag', 'b200amg_symmetric_strength'):
        f = getattr(lib, name)
        f.restype, f.argtypes = _lib._SIGS[name]
    return lib


def emu_symmetric_strength(lib, A, theta):
    """aggregation.symmetric_strength_of_connection for theta > 0 -> (Sp, Sj)"""
    if A.format == 'csr':
        n, vals = A.shape[0], np.ascontiguousarray(A.data, dtype=np.float64)
    else:
        R, C = A.blocksize
        n = A.shape[0] // R
        data = np.ascontiguousarray(A.data, dtype=np.float64).ravel()
        vals = np.full(len(A.indices), np.nan)
        assert lib.b200amg_block_sqnorm(len(A.indices), R, C, P(data), P(vals), None) == 0
    Ap, Aj = i32(A.indptr), i32(A.indices)
    diag = np.full(n, np.nan)
    assert lib.b200amg_abs_diag(n, P(Ap), P(Aj), P(vals), P(diag), None) == 0
    cnt = np.zeros(n, np.int32)
    assert lib.b200amg_symmetric_strength(n, theta, P(Ap), P(Aj), P(vals), P(diag), 0, P(cnt), None, None, None, None) == 0
    Sp = i32(np.concatenate([[0], np.cumsum(cnt)]))
    Sj, Sx = np.full(Sp[-1], -1, np.int32), np.full(Sp[-1], np.nan)
    assert lib.b200amg_symmetric_strength(n, theta, P(Ap), P(Aj), P(vals), P(diag), 1, None, P(Sp), P(Sj), P(Sx), None) == 0
    return Sp, Sj, vals


@pytest.mark.parametrize('bs', [0, 1, 2, 3, 6, 8])
def test_symmetric_strength_equals_reference(emu_strength, bs):
    """the strength pattern for theta > 0 against oracle/sa.py's restatement of pyamg/strength.py:213-318 (numpy block
    norms + the reference's compiled amg_core.symmetric_strength_of_connection); bs = 0: a CSR matrix.  Ties included:
    a constant-coefficient stencil puts many entries exactly ON the threshold."""
    rng = np.random.RandomState(bs)
    if bs == 0:
        T = sp.diags([-1.0, 2.0, -1.0], [-1, 0, 1], shape=(12, 12))
        A = sp.csr_matrix(sp.kron(T, sp.identity(12)) + 0.25 * sp.kron(sp.identity(12), T))
        A.sort_indices()
        mats = [A, sp.csr_matrix(A + sp.random(144, 144, 0.05, random_state=rng))]
    else:
        mats = [random_bsr(60, 60, bs, bs, 0.15, 70 + bs, unsorted=True) + sp.identity(60 * bs, format='bsr')]
        mats = [sp.bsr_matrix(m, blocksize=(bs, bs)) for m in mats]
    for A in mats:
        for theta in (0.25, 0.5, 0.01):
            Sp, Sj, vals = emu_symmetric_strength(emu_strength, A, theta)
            ref = osa.symmetric_strength_pattern(A, theta, core=CORE)
            assert np.array_equal(Sp, ref.indptr) and np.array_equal(Sj, ref.indices[:ref.indptr[-1]]), (bs, theta)
            if bs:
                assert np.array_equal(vals, (A.data * A.data).reshape(-1, bs * bs).sum(axis=1))     # numpy's own summation order
    assert emu_strength.b200amg_symmetric_strength(4, -0.1, None, None, None, None, 0, None, None, None, None, None) != 0


# ---- the scalar setup kernels (already pinned on the GPU in round 1), compiled for the host for the CPU harness -----
@pytest.fixture(scope='module')
def emu_setup():
    lib = build_emu('setup_kernels', extra=('rs_setup',))
    from pyamg_b200 import _lib
    for name in ('b200amg_csr_scale_rows', 'b200amg_tentative_scalar', 'b200amg_flag_nonneg', 'b200amg_tentative_fill',
                 'b200amg_csr_sub_sorted', 'b200amg_csr_sub_general', 'b200amg_rs_strength', 'b200amg_csr_not_canonical',
                 'b200amg_rs_direct_interpolation_pass1', 'b200amg_rs_direct_interpolation_pass2'):
        f = getattr(lib, name)
        f.restype, f.argtypes = _lib._SIGS[name]
    return lib


# ---- the BSR solve-phase kernels (csrc/bsr.cu; pinned on the GPU in round 1) for the CPU harness -------------------
@pytest.fixture(scope='module')
def emu_bsr():
    if shutil.which('g++') is None:
        pytest.skip('g++ not available')
    csrc = os.path.join(ROOT, 'pyamg_b200', 'csrc')
    srcs = [os.path.join(csrc, 'bsr.cu'), os.path.join(EMU, 'emu_bsr_api.cpp'), os.path.join(EMU, 'emu_globals.cpp')]
    out = os.path.join(EMU, '_build', 'libemu_bsr.so')
    deps = srcs + [os.path.join(EMU, 'cuda_runtime.h'), os.path.join(csrc, 'internal.cuh'), os.path.join(csrc, 'common.cuh')]
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(d) for d in deps):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.run(['g++', '-std=c++20', '-O1', '-ffp-contract=off', '-fPIC', '-shared', '-pthread', '-I', EMU, '-I', csrc,
                        '-x', 'c++'] + srcs + ['-o', out], check=True)
    lib = ctypes.CDLL(out)
    lib.emu_bsr_matvec.restype = c_int
    lib.emu_bsr_matvec.argtypes = [c_int, c_int, c_int, vp, vp, vp, vp, vp, vp, c_int]
    lib.emu_bsr_matvec_warp.restype = c_int
    lib.emu_bsr_matvec_warp.argtypes = [c_int, c_int, c_int, vp, vp, vp, vp, vp, vp, c_int]
    lib.emu_bsr_jacobi.restype = c_int
    lib.emu_bsr_jacobi.argtypes = [vp, vp, vp, vp, vp, vp, c_int, c_int, c_int, c_dbl]
    lib.emu_bsr_gauss_seidel.restype = c_int
    lib.emu_bsr_gauss_seidel.argtypes = [vp, vp, vp, vp, vp, c_int, c_int, c_int, c_int, vp, ctypes.POINTER(c_int), c_int]
    lib.emu_bsr_gauss_seidel_warp.restype = c_int
    lib.emu_bsr_gauss_seidel_warp.argtypes = [vp, vp, vp, vp, vp, c_int, c_int, c_int, c_int, vp, ctypes.POINTER(c_int), c_int]
    return lib


@pytest.mark.parametrize('bs', [2, 3, 6])
def test_bsr_solve_kernels_equal_reference(emu_bsr, bs):
    """bsr_matvec (3 modes, rectangular blocks too), bsr_jacobi, bsr_gauss_seidel lexicographic and colour-ordered: the
    kernel source of csrc/bsr.cu on the host against scipy's bsr_matvec and the reference's compiled kernels"""
    from oracle import relaxation as orx
    A, _, b, x0 = block_problem(90, bs, 60 + bs)
    Ap, Aj, Ax = i32(A.indptr), i32(A.indices), np.ravel(A.data).copy()
    n = 90 * bs
    for mode, want in ((0, A * x0), (1, b - A * x0), (2, b + A * x0)):
        y = np.full(n, np.nan)
        assert emu_bsr.emu_bsr_matvec(90, bs, bs, P(Ap), P(Aj), P(Ax), P(x0), P(b), P(y), mode) == 0
        assert np.array_equal(y, want), mode
    Rm = random_bsr(50, 70, 3, 6, 0.2, 5)
    xr = np.random.RandomState(1).rand(70 * 6)
    y = np.full(150, np.nan)
    assert emu_bsr.emu_bsr_matvec(50, 3, 6, P(i32(Rm.indptr)), P(i32(Rm.indices)), P(np.ravel(Rm.data).copy()), P(xr), None, P(y), 0) == 0
    assert np.array_equal(y, Rm * xr)
    want, tw, got, tg = x0.copy(), np.zeros(n), x0.copy(), np.zeros(n)
    CORE.bsr_jacobi(Ap, Aj, Ax, want, b, tw, 0, 90, 1, bs, np.array([0.8]))
    assert emu_bsr.emu_bsr_jacobi(P(Ap), P(Aj), P(Ax), P(got), P(b), P(tg), 0, 90, bs, 0.8) == 0
    assert np.array_equal(got, want)
    for r0, r1, dr in ((0, 90, 1), (89, -1, -1)):
        want, got = x0.copy(), x0.copy()
        CORE.bsr_gauss_seidel(Ap, Aj, Ax, want, b, r0, r1, dr, bs)
        assert emu_bsr.emu_bsr_gauss_seidel(P(Ap), P(Aj), P(Ax), P(got), P(b), r0, r1, dr, bs, None, None, 0) == 0
        assert np.array_equal(got, want)
    G = sp.csr_matrix((np.ones(len(Aj)), Aj, Ap), shape=(90, 90))
    idx, ptr = orx.color_ordering(orx.vertex_coloring_mis(G, oracle.port))
    idx, ptr = i32(idx), np.ascontiguousarray(ptr, dtype=np.intc)
    cp = ptr.ctypes.data_as(ctypes.POINTER(c_int))
    want, got = x0.copy(), x0.copy()
    oracle.port.bsr_gauss_seidel_indexed(Ap, Aj, Ax, want, b, idx, 0, 90, 1, bs)
    oracle.port.bsr_gauss_seidel_indexed(Ap, Aj, Ax, want, b, idx, 89, -1, -1, bs)
    assert emu_bsr.emu_bsr_gauss_seidel(P(Ap), P(Aj), P(Ax), P(got), P(b), 0, 90, 1, bs, P(idx), cp, len(ptr) - 1) == 0
    assert emu_bsr.emu_bsr_gauss_seidel(P(Ap), P(Aj), P(Ax), P(got), P(b), 89, -1, -1, bs, P(idx), cp, len(ptr) - 1) == 0
    assert np.array_equal(got, want)
    # the warp-per-block-row schedules (long rows in the product): same bits
    got = x0.copy()
    assert emu_bsr.emu_bsr_gauss_seidel_warp(P(Ap), P(Aj), P(Ax), P(got), P(b), 0, 90, 1, bs, P(idx), cp, len(ptr) - 1) == 0
    assert emu_bsr.emu_bsr_gauss_seidel_warp(P(Ap), P(Aj), P(Ax), P(got), P(b), 89, -1, -1, bs, P(idx), cp, len(ptr) - 1) == 0
    assert np.array_equal(got, want)
    for mode, wantv in ((0, A * x0), (1, b - A * x0), (2, b + A * x0)):
        y = np.full(n, np.nan)
        assert emu_bsr.emu_bsr_matvec_warp(90, bs, bs, P(Ap), P(Aj), P(Ax), P(x0), P(b), P(y), mode) == 0
        assert np.array_equal(y, wantv), mode
    for (nr, nc, R, C) in ((50, 70, 3, 6), (40, 90, 6, 3), (7, 300, 6, 6)):
        Rm = random_bsr(nr, nc, R, C, 0.3, 5)
        xr = np.random.RandomState(1).rand(nc * C)
        y = np.full(nr * R, np.nan)
        assert emu_bsr.emu_bsr_matvec_warp(nr, R, C, P(i32(Rm.indptr)), P(i32(Rm.indices)), P(np.ravel(Rm.data).copy()), P(xr),
                                           None, P(y), 0) == 0
        assert np.array_equal(y, Rm * xr), (R, C)


# ---- whole smoothing steps of small levels in one launch (csrc/small_level.cu) -------------------------------------------
@pytest.fixture(scope='module')
def emu_small():
    if shutil.which('g++') is None:
        pytest.skip('g++ not available')
    csrc = os.path.join(ROOT, 'pyamg_b200', 'csrc')
    srcs = [os.path.join(csrc, 'small_level.cu'), os.path.join(EMU, 'emu_small_level.cpp'), os.path.join(EMU, 'emu_globals.cpp')]
    out = os.path.join(EMU, '_build', 'libemu_small.so')
    deps = srcs + [os.path.join(EMU, 'cuda_runtime.h'), os.path.join(csrc, 'internal.cuh'), os.path.join(csrc, 'common.cuh')]
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(d) for d in deps):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.run(['g++', '-std=c++20', '-O1', '-ffp-contract=off', '-fPIC', '-shared', '-pthread', '-I', EMU, '-I', csrc,
                        '-x', 'c++'] + srcs + ['-o', out], check=True)
    lib = ctypes.CDLL(out)
    lib.emu_gs_level_cta.restype = c_int
    lib.emu_gs_level_cta.argtypes = [c_int, c_int, c_int, ctypes.c_longlong, vp, vp, vp, vp, vp, c_int, vp, vp, c_int, c_int, c_int]
    return lib


@pytest.mark.parametrize('n,per_row,sweep,iters,zero', [(700, 9, 2, 1, False), (300, 25, 0, 2, True), (1500, 5, 1, 1, False),
                                                        (40, 3, 2, 3, False)])
@pytest.mark.parametrize('bsrf', [False, True])
def test_small_level_step_scalar(emu_small, monkeypatch, n, per_row, sweep, iters, zero, bsrf):
    """gs_level_cta_kernel (iterate in shared memory, the NEXT colour's row requested before the current one is folded;
    rows longer than the 16 prefetched entries and colours wider than the 256 threads included) against the oracle's
    sequential gauss_seidel_indexed / bsr_gauss_seidel_indexed(blocksize 1) sweeps over the same MIS colour ordering"""
    from oracle import relaxation as orx
    monkeypatch.setenv('B200AMG_SMALL_LEVEL', '2')          # the scalar form is opt-in (per-colour launches are faster on a B200)
    rng = np.random.RandomState(n + per_row)
    rows = np.repeat(np.arange(n), per_row)
    cols = (rows + rng.randint(-4 * per_row, 4 * per_row, rows.size)) % n
    A = sp.csr_matrix((rng.rand(rows.size) - 0.4, (rows, cols)), shape=(n, n))
    A = sp.csr_matrix(A + A.T + sp.diags(rng.rand(n) + 2.0 * per_row))
    A.sort_indices()
    if n == 700:                                    # a row without a diagonal entry is left alone, like the reference
        A = A.tolil(); A[5, 5] = 0.0; A = sp.csr_matrix(A); A.eliminate_zeros()
    Ap, Aj, Ax = i32(A.indptr), i32(A.indices), A.data.copy()
    idx, ptr = orx.color_ordering(orx.vertex_coloring_mis(A, oracle.port))
    idx, ptr = i32(idx), i32(ptr)
    K = len(ptr) - 1
    if n == 1500:
        assert np.max(np.diff(ptr)) > 256           # a colour wider than the block
    if per_row == 25:
        assert np.max(np.diff(Ap)) > 16             # rows longer than the prefetch depth
    b, x0 = rng.rand(n), (np.zeros(n) if zero else rng.rand(n))
    want = x0.copy()
    for _ in range(iters):
        for (r0, r1, dr) in (((0, n, 1),) if sweep == 0 else ((n - 1, -1, -1),) if sweep == 1 else ((0, n, 1), (n - 1, -1, -1))):
            if bsrf:
                oracle.port.bsr_gauss_seidel_indexed(Ap, Aj, Ax, want, b, idx, r0, r1, dr, 1)
            else:
                oracle.port.gauss_seidel_indexed(Ap, Aj, Ax, want, b, idx, r0, r1, dr)
    got = np.full(n, np.nan) if zero else x0.copy()
    assert emu_small.emu_gs_level_cta(1, int(bsrf), n, len(Ax), P(Ap), P(Aj), P(Ax), P(idx), P(ptr), K, P(got), P(b), iters, sweep,
                                      int(zero)) == 0
    assert np.array_equal(got, want)


@pytest.mark.parametrize('bs,sweep', [(2, 2), (3, 0), (6, 1)])
def test_small_level_step_block(emu_small, bs, sweep):
    """bsr_gs_level_cta_kernel against the oracle's bsr_gauss_seidel_indexed sweeps"""
    from oracle import relaxation as orx
    nb = 60
    A, _, b, x0 = block_problem(nb, bs, 40 + bs)
    Ap, Aj, Ax = i32(A.indptr), i32(A.indices), np.ravel(A.data).copy()
    G = sp.csr_matrix((np.ones(len(Aj)), Aj, Ap), shape=(nb, nb))
    idx, ptr = orx.color_ordering(orx.vertex_coloring_mis(G, oracle.port))
    idx, ptr = i32(idx), i32(ptr)
    want = x0.copy()
    for (r0, r1, dr) in (((0, nb, 1),) if sweep == 0 else ((nb - 1, -1, -1),) if sweep == 1 else ((0, nb, 1), (nb - 1, -1, -1))):
        oracle.port.bsr_gauss_seidel_indexed(Ap, Aj, Ax, want, b, idx, r0, r1, dr, bs)
    got = x0.copy()
    assert emu_small.emu_gs_level_cta(bs, 1, nb * bs, len(Ax), P(Ap), P(Aj), P(Ax), P(idx), P(ptr), len(ptr) - 1, P(got), P(b), 1,
                                      sweep, 0) == 0
    assert np.array_equal(got, want)


@pytest.mark.parametrize('bs,nb,density', [(2, 190, 0.7), (3, 130, 0.7), (6, 80, 0.7)])
def test_bsr_warp_gauss_seidel_long_rows(emu_bsr, bs, nb, density):
    """the batched warp-per-block-row sweep on rows that need SEVERAL batches (4 or 8 passes of 32 / bs blocks each) and a
    ragged last one, against the reference's bsr_gauss_seidel_indexed on the same colour ordering"""
    from oracle import relaxation as orx
    rng = np.random.RandomState(bs)
    pat = sp.csr_matrix(sp.random(nb, nb, density, random_state=rng, format='csr') + sp.identity(nb))
    pat = sp.csr_matrix(pat + pat.T)
    pat.sort_indices()
    assert np.max(np.diff(pat.indptr)) > (32 // bs) * (8 if bs <= 3 else 4)
    data = 0.01 * rng.randn(len(pat.indices), bs, bs)
    rows_of = np.repeat(np.arange(nb), np.diff(pat.indptr))
    data[rows_of == pat.indices] += 3.0 * np.eye(bs)        # diagonally dominant: the sweeps stay finite
    A = sp.bsr_matrix((data, pat.indices.astype(np.int32), pat.indptr.astype(np.int32)), shape=(nb * bs, nb * bs))
    Ap, Aj, Ax = i32(A.indptr), i32(A.indices), np.ravel(A.data).copy()
    b, x0 = rng.randn(nb * bs), rng.randn(nb * bs)
    # any partition into "colours" is a valid index list for the indexed sweep; rows of a dense pattern all conflict, so
    # the parallel kernel is only the sequential sweep if every colour is ONE row
    idx = i32(rng.permutation(nb))
    ptr = np.ascontiguousarray(np.arange(nb + 1), dtype=np.intc)
    cp = ptr.ctypes.data_as(ctypes.POINTER(c_int))
    want, got = x0.copy(), x0.copy()
    oracle.port.bsr_gauss_seidel_indexed(Ap, Aj, Ax, want, b, idx, 0, nb, 1, bs)
    oracle.port.bsr_gauss_seidel_indexed(Ap, Aj, Ax, want, b, idx, nb - 1, -1, -1, bs)
    assert emu_bsr.emu_bsr_gauss_seidel_warp(P(Ap), P(Aj), P(Ax), P(got), P(b), 0, nb, 1, bs, P(idx), cp, nb) == 0
    assert emu_bsr.emu_bsr_gauss_seidel_warp(P(Ap), P(Aj), P(Ax), P(got), P(b), nb - 1, -1, -1, bs, P(idx), cp, nb) == 0
    assert np.all(np.isfinite(want)) and np.array_equal(got, want)
    y = np.full(nb * bs, np.nan)
    assert emu_bsr.emu_bsr_matvec_warp(nb, bs, bs, P(Ap), P(Aj), P(Ax), P(x0), P(b), P(y), 1) == 0
    assert np.array_equal(y, b - A * x0)


# ---- the sliced-ELL layout (csrc/sell.cu): copies with 32-bit columns / 16-bit column steps / without diagonal entries ---
@pytest.fixture(scope='module')
def emu_sell():
    if shutil.which('g++') is None:
        pytest.skip('g++ not available')
    csrc = os.path.join(ROOT, 'pyamg_b200', 'csrc')
    srcs = [os.path.join(csrc, 'sell.cu'), os.path.join(EMU, 'emu_sell_api.cpp'), os.path.join(EMU, 'emu_globals.cpp')]
    out = os.path.join(EMU, '_build', 'libemu_sell.so')
    deps = srcs + [os.path.join(EMU, 'cuda_runtime.h'), os.path.join(csrc, 'internal.cuh'), os.path.join(csrc, 'common.cuh'),
                   os.path.join(csrc, 'setup.cuh')]
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(d) for d in deps):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.run(['g++', '-std=c++20', '-O1', '-ffp-contract=off', '-fPIC', '-shared', '-pthread', '-I', EMU, '-I', csrc,
                        '-x', 'c++'] + srcs + ['-o', out], check=True)
    lib = ctypes.CDLL(out)
    ip = ctypes.POINTER(c_int)
    lib.emu_sell_apply.restype = c_int
    lib.emu_sell_apply.argtypes = [c_int, ctypes.c_longlong, vp, vp, vp, c_int, c_int, vp, vp, vp, c_dbl, ip, ip]
    lib.emu_sell_gs.restype = c_int
    lib.emu_sell_gs.argtypes = [c_int, ctypes.c_longlong, vp, vp, vp, vp, vp, c_int, c_int, c_int, c_int, vp, vp, ip]
    return lib


def _long_row_matrix(n, ncols, per_row, span, seed, square_diag=True):
    rng = np.random.RandomState(seed)
    rows = np.repeat(np.arange(n), per_row)
    cols = (rows * (ncols // n) + rng.randint(-span, span, rows.size)) % ncols
    A = sp.csr_matrix((rng.rand(rows.size) - 0.3, (rows, cols)), shape=(n, ncols))       # duplicates summed: ragged rows
    if square_diag:
        A = sp.csr_matrix(A + sp.diags(rng.rand(n) + 1.0 + per_row))
    A.sort_indices()
    return A, rng


@pytest.mark.parametrize('steps', [False, True])
@pytest.mark.parametrize('drop_diag', [False, True])
def test_sliced_ell_ops_on_the_host(emu_sell, monkeypatch, steps, drop_diag):
    """sell_build + sell_kernel compiled for the host: y = A x, b - A x, y + A x and both Jacobi forms from copies with
    32-bit columns or 16-bit column steps, with or without the diagonal entries, against scipy's csr_matvec and the
    reference's jacobi / bsr_jacobi loops (oracle port): same bits.  Rows without a diagonal entry, a row that holds ONLY
    its diagonal and an empty row included."""
    monkeypatch.setenv('B200AMG_SELL_COL16', '2' if steps else '0')
    n = 900
    A, rng = _long_row_matrix(n, n, 40, 120, 5 + steps)
    A = A.tolil()
    A[11, 11] = 0.0                                         # no diagonal entry
    A[12, :] = 0.0; A[12, 12] = 3.0                         # only the diagonal
    A[13, :] = 0.0                                          # empty row
    A = sp.csr_matrix(A); A.eliminate_zeros(); A.sort_indices()
    Ap, Aj, Ax = i32(A.indptr), i32(A.indices), A.data.copy()
    x, b = rng.rand(n), rng.rand(n)
    lay, esc = c_int(0), c_int(0)
    core = oracle.port
    Axv = np.zeros(n); core.csr_matvec(n, n, Ap, Aj, Ax, x, Axv)
    ops = ([] if drop_diag else [(0, np.zeros(n), Axv), (2, np.zeros(n), b - Axv), (3, b.copy(), b + Axv)])
    xo = x.copy(); core.jacobi(Ap, Aj, Ax, xo, b, np.empty(n), 0, n, 1, np.array([0.7]))
    xb = x.copy(); core.bsr_jacobi(Ap, Aj, Ax, xb, b, np.empty(n), 0, n, 1, 1, np.array([0.7]))
    ops += [(4, np.zeros(n), xo), (5, np.zeros(n), xb)]
    for op, y0, want in ops:
        y = y0.copy()
        bb = y if op == 3 else b
        assert emu_sell.emu_sell_apply(n, len(Ax), P(Ap), P(Aj), P(Ax), int(drop_diag), op, P(x), P(bb), P(y), 0.7,
                                       ctypes.byref(lay), ctypes.byref(esc)) == 0
        assert lay.value == (7 if steps else 2)
        assert np.array_equal(y, want), op


def test_sliced_ell_escape_slots_on_the_host(emu_sell, monkeypatch):
    """16-bit column steps of 65 535 columns or more take escape slots: a 300 x 200 000 operator with 14 entries per row
    (a few steps that wide per hundred rows) keeps layout 7 and the bits of csr_matvec; with 4 entries per row most steps
    are that wide and the copy keeps 32-bit columns"""
    monkeypatch.setenv('B200AMG_SELL_COL16', '2')
    lay, esc = c_int(0), c_int(0)
    for per_row, want_layout in ((14, 7), (4, 2)):
        A, rng = _long_row_matrix(300, 200000, per_row, 100000, 9, square_diag=False)
        Ap, Aj, Ax = i32(A.indptr), i32(A.indices), A.data.copy()
        st = np.diff(A.indices.astype(np.int64))
        inside = np.ones(len(st), bool); inside[A.indptr[1:-1] - 1] = False
        assert np.count_nonzero(st[inside] >= 65535) > 0
        x = rng.rand(200000)
        want = np.zeros(300); oracle.port.csr_matvec(300, 200000, Ap, Aj, Ax, x, want)
        y = np.full(300, np.nan)
        assert emu_sell.emu_sell_apply(300, len(Ax), P(Ap), P(Aj), P(Ax), 0, 0, P(x), None, P(y), 0.0, ctypes.byref(lay),
                                       ctypes.byref(esc)) == 0
        assert lay.value == want_layout and (esc.value > 0) == (want_layout == 7)
        assert np.array_equal(y, want)


@pytest.mark.parametrize('bsrf', [False, True])
@pytest.mark.parametrize('steps,drop_diag', [(False, False), (True, True), (False, True)])
def test_sliced_ell_colour_sweeps_on_the_host(emu_sell, monkeypatch, bsrf, steps, drop_diag):
    """sell_gs_kernel on a colour-grouped copy (32-slot rounds with tied loads, with / without diagonal entries among the
    slots, 32-bit columns / 16-bit steps) against the reference's sequential gauss_seidel_indexed / bsr_gauss_seidel over
    the same MIS colour ordering, forward then backward"""
    from oracle import relaxation as orx
    monkeypatch.setenv('B200AMG_SELL_COL16', '2' if steps else '0')
    n = 700
    A, rng = _long_row_matrix(n, n, 9, 40, 21)
    A = sp.csr_matrix(A + A.T); A.sort_indices()
    Ap, Aj, Ax = i32(A.indptr), i32(A.indices), A.data.copy()
    idx, ptr = orx.color_ordering(orx.vertex_coloring_mis(A, oracle.port))
    idx, ptr = i32(idx), i32(ptr)
    b, x0 = rng.rand(n), rng.rand(n)
    want = x0.copy()
    for (r0, r1, dr) in ((0, n, 1), (n - 1, -1, -1)):
        if bsrf:
            oracle.port.bsr_gauss_seidel_indexed(Ap, Aj, Ax, want, b, idx, r0, r1, dr, 1)
        else:
            oracle.port.gauss_seidel_indexed(Ap, Aj, Ax, want, b, idx, r0, r1, dr)
    got, lay = x0.copy(), c_int(0)
    assert emu_sell.emu_sell_gs(n, len(Ax), P(Ap), P(Aj), P(Ax), P(idx), P(ptr), len(ptr) - 1, int(bsrf), int(drop_diag), 2,
                                P(got), P(b), ctypes.byref(lay)) == 0
    assert lay.value == (7 if steps else 2)
    assert np.array_equal(got, want)


# ---- Galerkin-product rows without a sort (csrc/spgemm_hash.cu) --------------------------------------------------------
@pytest.fixture(scope='module')
def emu_spgemm():
    if shutil.which('g++') is None:
        pytest.skip('g++ not available')
    csrc = os.path.join(ROOT, 'pyamg_b200', 'csrc')
    srcs = [os.path.join(csrc, 'spgemm_hash.cu'), os.path.join(EMU, 'emu_spgemm_api.cpp'), os.path.join(EMU, 'emu_globals.cpp')]
    out = os.path.join(EMU, '_build', 'libemu_spgemm.so')
    deps = srcs + [os.path.join(EMU, 'cuda_runtime.h'), os.path.join(csrc, 'internal.cuh'), os.path.join(csrc, 'common.cuh'),
                   os.path.join(csrc, 'setup.cuh')]
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(d) for d in deps):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.run(['g++', '-std=c++20', '-O1', '-ffp-contract=off', '-fPIC', '-shared', '-pthread', '-I', EMU, '-I', csrc,
                        '-x', 'c++'] + srcs + ['-o', out], check=True)
    lib = ctypes.CDLL(out)
    lib.emu_spgemm_hash_bin.restype = c_int
    lib.emu_spgemm_hash_bin.argtypes = [c_int, c_int, vp, c_int] + [vp] * 10
    return lib


def _sort_free_product(lib, A, B, mode):
    """C = A B through spgemm_hash_bin the way b200amg_spgemm_numeric drives it: rows binned by their number of products
    (32 / 128 / 512 / 2048 / 8192), one call per bin.  -> (indptr, indices, data) in the kernel's output order, bins it took"""
    Ap, Aj, Ax = i32(A.indptr), i32(A.indices), np.ascontiguousarray(A.data, np.float64)
    Bp, Bj, Bx = i32(B.indptr), i32(B.indices), np.ascontiguousarray(B.data, np.float64)
    n = A.shape[0]
    blen = np.diff(Bp).astype(np.int64)
    ub = np.array([blen[Aj[Ap[i]:Ap[i + 1]]].sum() for i in range(n)], dtype=np.int64)
    ubp = np.concatenate([[0], np.cumsum(ub)]).astype(np.int64)
    Tj = np.full(max(int(ubp[-1]), 1), -7, dtype=np.int32)
    Tx = np.full(max(int(ubp[-1]), 1), np.nan)
    cnt = np.zeros(max(n, 1), dtype=np.int32)
    limits = [0, 32, 128, 512, 2048, 8192]
    took = []
    cnt[:n][ub > limits[-1]] = -1                               # rows beyond the last bin stay with the sorting kernels
    for b in range(5):
        rows = i32(np.nonzero((ub > limits[b]) & (ub <= limits[b + 1]))[0][::-1])        # (any order inside a bin)
        if rows.size == 0:
            continue
        rc = lib.emu_spgemm_hash_bin(b, mode, P(rows), rows.size, P(Ap), P(Aj), P(Ax), P(Bp), P(Bj), P(Bx), P(ubp), P(Tj), P(Tx),
                                     P(cnt))
        assert rc in (0, 1)
        if rc == 1:
            took.append(b)
        else:
            cnt[rows] = -1
    return ubp, Tj, Tx, cnt, took


@pytest.mark.parametrize('case', ['stencil_RA', 'random_wide', 'cancelling', 'unsorted_operands', 'up_to_2048', 'up_to_8192',
                                  'full_table'])
def test_sort_free_galerkin_rows_on_the_host(emu_spgemm, case):
    """every row of C = A B equals scipy's csr_matmat row: the same columns IN THE SAME (reverse first touch) ORDER, the
    same bits, exact zeros dropped -- for tentative-prolongator-like operands, wide random rows (probe chains, groups of
    equal columns inside a 32-product step), sums that cancel to 0.0, operands with unsorted rows and duplicate entries"""
    rng = np.random.RandomState(5)
    mode = 1
    if case == 'stencil_RA':
        d1 = lambda k: sp.diags([-1.0, 2.0, -1.0], [-1, 0, 1], shape=(k, k))           # noqa: E731
        eye = sp.identity
        A = sp.csr_matrix(sp.kron(sp.kron(d1(12), eye(11)), eye(10)) + sp.kron(sp.kron(eye(12), d1(11)), eye(10))
                          + sp.kron(sp.kron(eye(12), eye(11)), d1(10)))
        A.sort_indices()
        agg = (np.arange(A.shape[0]) // 27).astype(np.int32)
        T = sp.csr_matrix((rng.rand(A.shape[0]) + 0.5, (np.arange(A.shape[0]), agg)))
        L, R_ = sp.csr_matrix(T.T), A
        L.sort_indices()
    elif case == 'random_wide':
        L = sp.random(60, 300, 0.12, random_state=rng, format='csr')
        R_ = sp.random(300, 90, 0.08, random_state=rng, format='csr')
    elif case == 'cancelling':
        # rows whose products cancel exactly: column j gets +v and -v (dropped, scipy drops it too), and -0.0 sums
        L = sp.csr_matrix(np.array([[1.0, 1.0, 0.0], [2.0, 0.0, 3.0], [0.0, 0.0, 0.0], [1.0, -1.0, 1.0]]))
        R_ = sp.csr_matrix(np.array([[4.0, -2.5, 0.0, 1.0], [-4.0, 2.5, 7.0, 0.0], [0.0, 5.0, 0.0, -1.0]]))
    elif case == 'unsorted_operands':
        L = sp.random(50, 80, 0.3, random_state=rng, format='csr')
        R_ = sp.random(80, 70, 0.2, random_state=rng, format='csr')
        for M in (L, R_):                                                  # shuffle every row's storage order
            for i in range(M.shape[0]):
                s, e = M.indptr[i], M.indptr[i + 1]
                q = rng.permutation(e - s)
                M.indices[s:e], M.data[s:e] = M.indices[s:e][q], M.data[s:e][q]
        # duplicate entries inside rows of the left operand (non-canonical CSR)
        L = sp.csr_matrix((np.concatenate([L.data, L.data[:40]]), np.concatenate([L.indices, L.indices[:40]]),
                           np.concatenate([L.indptr[:-1], [L.indptr[-1] + 40]])), shape=L.shape)
    elif case == 'up_to_2048':
        mode = 2
        L = sp.random(12, 400, 0.3, random_state=rng, format='csr')
        R_ = sp.random(400, 1500, 0.01, random_state=rng, format='csr')
    elif case == 'up_to_8192':
        # products staged in the row's global scratch (mode 3): rows of 600 .. 8000 products, few and many distinct columns
        mode = 3
        L = sp.vstack([sp.random(6, 500, 0.25, random_state=rng, format='csr'), sp.random(6, 500, 0.6, random_state=rng, format='csr'),
                       sp.random(3, 500, 0.02, random_state=rng, format='csr'), sp.random(5, 500, 0.09, random_state=rng, format='csr')]).tocsr()
        R_ = sp.hstack([sp.random(500, 300, 0.06, random_state=rng, format='csr'),
                        sp.random(500, 6000, 0.001, random_state=rng, format='csr')]).tocsr()
        L.data -= 0.5
    else:
        # a row whose 8192 products all have different columns: the 8192-slot table fills to the last slot
        mode = 3
        L = sp.csr_matrix(np.ones((1, 2)))
        R_ = sp.csr_matrix((rng.rand(8192) + 0.1, (np.repeat([0, 1], 4096), rng.permutation(20000)[:8192])), shape=(2, 20000))
    ubp, Tj, Tx, cnt, took = _sort_free_product(emu_spgemm, L, R_, mode)
    import scipy.sparse._sparsetools as st
    n, nc = L.shape[0], R_.shape[1]
    cap = st.csr_matmat_maxnnz(n, nc, i32(L.indptr), i32(L.indices), i32(R_.indptr), i32(R_.indices))
    Cp, Cj, Cx = np.zeros(n + 1, np.int32), np.zeros(max(cap, 1), np.int32), np.zeros(max(cap, 1))
    st.csr_matmat(n, nc, i32(L.indptr), i32(L.indices), np.ascontiguousarray(L.data, np.float64), i32(R_.indptr), i32(R_.indices),
                  np.ascontiguousarray(R_.data, np.float64), Cp, Cj, Cx)
    assert took and (mode == 1 or 3 in took or 4 in took) and (case != 'up_to_8192' or (3 in took and 4 in took))
    checked = 0
    for i in range(n):
        if cnt[i] < 0:
            continue
        o = int(ubp[i])
        assert cnt[i] == Cp[i + 1] - Cp[i], i
        assert np.array_equal(Tj[o:o + cnt[i]], Cj[Cp[i]:Cp[i + 1]]), i
        assert np.array_equal(Tx[o:o + cnt[i]], Cx[Cp[i]:Cp[i + 1]]), i
        checked += 1
    assert checked >= n - 2 or case in ('up_to_2048', 'up_to_8192')
