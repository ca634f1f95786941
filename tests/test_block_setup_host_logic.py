"""pyamg_b200.smoothed_aggregation_solver's BLOCK path run end to end in this GPU-less container, to check the Python
orchestration (argument order of every new C-ABI call, shapes, level loop) before hardware sees it:

  * the new entry points (csrc/bsr_setup.cu) are served by that source compiled for the host (tests/emu);
  * the scalar building blocks already pinned on the GPU (SpGEMM, transpose, scan, diagonal, ...) are replaced by
    scipy / numpy doubles;
  * tensors are torch CPU tensors, so every `data_ptr()` is a host pointer.

All of this is monkeypatching inside the test: the product itself still raises without a CUDA device
(tests/test_cabi.py::test_no_cpu_fallback_without_cuda)."""
import ctypes
import os

import numpy as np
import pytest
import scipy.sparse as sp
import torch

import hier_io
from oracle import sa as osa
from test_emulated_kernels import emu, emu_bsr, emu_relax, emu_setup, emu_strength, same  # noqa: F401  (the host-compiled kernel sources)


def _arr(ptr, n, ctype):
    return np.ctypeslib.as_array((ctype * n).from_address(ptr)) if n else np.zeros(0, dtype=ctype)


@pytest.fixture()
def host_amg(emu, emu_relax, emu_strength, emu_setup, emu_bsr, monkeypatch):  # noqa: F811
    import pyamg_b200 as amg
    from pyamg_b200 import _lib, aggregation, classical, linalg, multilevel, smoothing, sparse
    DevMat = sparse.DevMat
    cpu = torch.device('cpu')
    for mod in (sparse, aggregation, classical, linalg, multilevel, smoothing):
        if hasattr(mod, '_dev'):
            monkeypatch.setattr(mod, '_dev', lambda: cpu)
        if hasattr(mod, '_stream'):
            monkeypatch.setattr(mod, '_stream', lambda: 0)
    emu_names = ('b200amg_bsr_scale_rows', 'b200amg_bsr_gather_blocks', 'b200amg_fit_candidates', 'b200amg_tentative_scatter',
                 'b200amg_bsr_sub', 'b200amg_bsr_slot_map', 'b200amg_bsr_matmat_numeric')
    relax_names = ('b200amg_block_gauss_seidel', 'b200amg_block_jacobi', 'b200amg_block_diag_inverse', 'b200amg_pinv_array')
    for name in emu_names:                                  # bind exactly the product's signature table
        f = getattr(emu, name)
        f.restype, f.argtypes = _lib._SIGS[name]
    for name in relax_names:
        f = getattr(emu_relax, name)
        f.restype, f.argtypes = _lib._SIGS[name]
    strength_names = ('b200amg_block_sqnorm', 'b200amg_abs_diag', 'b200amg_symmetric_strength')
    setup_names = ('b200amg_csr_scale_rows', 'b200amg_tentative_scalar', 'b200amg_flag_nonneg', 'b200amg_tentative_fill',
                   'b200amg_csr_sub_sorted', 'b200amg_csr_sub_general', 'b200amg_rs_strength', 'b200amg_csr_not_canonical',
                   'b200amg_rs_direct_interpolation_pass1', 'b200amg_rs_direct_interpolation_pass2')
    real_call = _lib.call
    seen = set()

    def call(name, *args):
        seen.add(name)
        if name in emu_names or name in relax_names or name in strength_names or name in setup_names:
            lib = (emu if name in emu_names else emu_relax if name in relax_names else emu_strength if name in strength_names
                   else emu_setup)
            rc = getattr(lib, name)(*args)
            assert rc == 0, name
        elif name == 'b200amg_bsr_matvec' and (args[2], args[3]) != (1, 1):      # blocks: the kernel source of csrc/bsr.cu
            nbr, nbc, R, C, Ap, Aj, Ax, x, bb, y, mode, _ = args
            assert emu_bsr.emu_bsr_matvec(nbr, R, C, Ap, Aj, Ax, x, bb, y, mode) == 0
        elif name == 'b200amg_bsr_gauss_seidel' and args[8] > 1:
            Ap, Aj, Ax, x, bb, r0, r1, dr, bs, Id, cp, K, _ = args
            assert emu_bsr.emu_bsr_gauss_seidel(Ap, Aj, Ax, x, bb, r0, r1, dr, bs, Id, cp, K) == 0
        elif name == 'b200amg_bsr_matvec':                  # 1 x 1: scipy's csr matvec (what the CSR kernels are pinned to)
            nbr, nbc, R, C, Ap, Aj, Ax, x, bb, y, mode, _ = args
            ap = _arr(Ap, nbr + 1, ctypes.c_int)
            nb_ = int(ap[-1])
            M = sp.bsr_matrix((_arr(Ax, nb_ * R * C, ctypes.c_double).reshape(-1, R, C), _arr(Aj, nb_, ctypes.c_int), ap),
                              shape=(nbr * R, nbc * C))
            if R == 1 and C == 1:
                M = sp.csr_matrix((_arr(Ax, nb_, ctypes.c_double), _arr(Aj, nb_, ctypes.c_int), ap), shape=(nbr, nbc))
            Ax_ = M * _arr(x, nbc * C, ctypes.c_double)
            yv = _arr(y, nbr * R, ctypes.c_double)
            if mode == 0:
                yv[:] = Ax_
            elif mode == 1:
                yv[:] = _arr(bb, nbr * R, ctypes.c_double) - Ax_
            else:
                yv[:] = (_arr(bb, nbr * R, ctypes.c_double) if bb else yv) + Ax_
        elif name == 'b200amg_dense_matvec':
            n_, M_, x_, y_, _ = args
            _arr(y_, n_, ctypes.c_double)[:] = np.dot(_arr(M_, n_ * n_, ctypes.c_double).reshape(n_, n_), _arr(x_, n_, ctypes.c_double))
        elif name == 'b200amg_dot':
            n_, x_, y_, out_, _, _ = args
            _arr(out_, 1, ctypes.c_double)[0] = np.inner(_arr(x_, n_, ctypes.c_double), _arr(y_, n_, ctypes.c_double))
        elif name == 'b200amg_extract_diag':
            nb, bs, Ap, Aj, Ax, d, _ = args
            ap = _arr(Ap, nb + 1, ctypes.c_int)
            M = sp.bsr_matrix((_arr(Ax, int(ap[-1]) * bs * bs, ctypes.c_double).reshape(-1, bs, bs),
                               _arr(Aj, int(ap[-1]), ctypes.c_int), ap), shape=(nb * bs, nb * bs))
            _arr(d, nb * bs, ctypes.c_double)[:] = M.diagonal()
        else:
            real_call(name, *args)                          # host entry points of the real library (aggregation, colouring)

    monkeypatch.setattr(_lib, 'call', call)

    from pyamg_b200 import amg_core

    def chk(t, dtype, name):                                # the product's check minus "must be a CUDA tensor"
        assert isinstance(t, torch.Tensor) and t.dtype == dtype and t.dim() == 1 and t.is_contiguous(), name
        return t.data_ptr()
    monkeypatch.setattr(amg_core, '_chk', chk)
    monkeypatch.setattr(amg_core, '_stream', lambda: 0)

    class _S:
        cuda_stream = 0
    monkeypatch.setattr(torch.cuda, 'current_stream', lambda *a, **k: _S())

    def scan(v):
        out = torch.zeros(v.numel() + 1, dtype=torch.int32)
        out[1:] = torch.cumsum(v.to(torch.int64), 0).to(torch.int32)
        return out
    monkeypatch.setattr(sparse, '_scan_i32', scan)
    monkeypatch.setattr(aggregation, '_scan_i32', scan)

    def from_sp(M, fmt, sorted_flag):
        return DevMat(torch.from_numpy(np.ascontiguousarray(M.indptr, np.int32)), torch.from_numpy(np.ascontiguousarray(M.indices, np.int32)),
                      torch.from_numpy(np.ascontiguousarray(M.data, np.float64)), M.shape, (1, 1), fmt, sorted_flag)

    real_bsr_matmat, real_bsr_transpose = sparse.bsr_matmat, sparse.bsr_transpose

    def matmat(A, B, scipy_order=False, batch_products=None):
        if not (A.is_scalar and B.is_scalar):
            return real_bsr_matmat(A, B)
        C = sp.csr_matrix(A.to_scipy()) * sp.csr_matrix(B.to_scipy())        # csr_matmat: reverse first-touch rows
        if not scipy_order:
            C.sort_indices()
        return from_sp(C, 'csr', not scipy_order)

    def transpose(A):
        if not A.is_scalar:
            return real_bsr_transpose(A)
        T = sp.csr_matrix(A.to_scipy()).tocsc()
        return DevMat(torch.from_numpy(np.ascontiguousarray(T.indptr, np.int32)), torch.from_numpy(np.ascontiguousarray(T.indices, np.int32)),
                      torch.from_numpy(np.ascontiguousarray(T.data, np.float64)), (A.shape[1], A.shape[0]), (1, 1), A.format, True)

    for mod in (sparse, aggregation, classical):
        if hasattr(mod, 'matmat'):
            monkeypatch.setattr(mod, 'matmat', matmat)
        monkeypatch.setattr(mod, 'transpose', transpose)
        if hasattr(mod, '_scan_i32'):
            monkeypatch.setattr(mod, '_scan_i32', scan)
    amg._seen_calls = seen
    return amg


def test_block_path_builds_the_reference_hierarchy(host_amg, golden_dir):
    amg = host_amg
    levels, ex = hier_io.load(os.path.join(golden_dir, 'sa_elasticity3d_tets_7_setup.npz'))
    A, B = levels[0].A, ex['B']
    gs = ('gauss_seidel', {'sweep': 'symmetric'})
    np.random.seed(0)
    ref = osa.smoothed_aggregation_hierarchy(A.copy(), B, presmoother=gs, postsmoother=gs, max_coarse=10, improve_candidates=None)
    rhos = [(l.rho_smooth, None) for l in ref[:-1]]
    ml = amg.smoothed_aggregation_solver(A, B, presmoother=gs, postsmoother=gs, max_coarse=10, rho_override=rhos, improve_candidates=None)
    assert len(ml.levels) == len(levels) == 3
    for k, (m, r) in enumerate(zip(ml.levels, levels)):
        for nm in ('A', 'P', 'R'):
            if hasattr(r, nm):
                assert same(getattr(m, nm).to_scipy(), getattr(r, nm)), (k, nm)
    assert [l.A.blocksize for l in ml.levels] == [(3, 3), (6, 6), (6, 6)]
    assert {'b200amg_fit_candidates', 'b200amg_tentative_scatter', 'b200amg_bsr_scale_rows', 'b200amg_bsr_sub',
            'b200amg_bsr_slot_map', 'b200amg_bsr_matmat_numeric', 'b200amg_bsr_gather_blocks'} <= amg._seen_calls


def test_block_path_default_candidates_and_argument_checks(host_amg, golden_dir):
    """B=None -> one constant per unknown of a node (K2 = blocksize); the options without a device path raise"""
    amg = host_amg
    levels, ex = hier_io.load(os.path.join(golden_dir, 'sa_elasticity2d_14.npz'))
    A = levels[0].A
    assert A.format == 'bsr' and A.blocksize == (2, 2)
    gs = ('gauss_seidel', {'sweep': 'symmetric'})
    np.random.seed(0)
    ref = osa.smoothed_aggregation_hierarchy(A.copy(), None, presmoother=gs, postsmoother=gs, max_coarse=10, improve_candidates=None)
    rhos = [(l.rho_smooth, None) for l in ref[:-1]]
    ml = amg.smoothed_aggregation_solver(A, presmoother=gs, postsmoother=gs, max_coarse=10, rho_override=rhos, improve_candidates=None)
    assert len(ml.levels) == len(ref)
    for k, (m, r) in enumerate(zip(ml.levels, ref)):
        for nm in ('A', 'P', 'R'):
            if getattr(r, nm, None) is not None:
                assert same(getattr(m, nm).to_scipy(), getattr(r, nm)), (k, nm)
    with pytest.raises(NotImplementedError):
        amg.smoothed_aggregation_solver(A, strength=None, improve_candidates=None)
    with pytest.raises(NotImplementedError):
        amg.smoothed_aggregation_solver(sp.csr_matrix(A), np.ones((A.shape[0], 2)), improve_candidates=None)
    with pytest.raises(ValueError):
        amg.smoothed_aggregation_solver(A, np.ones((A.shape[0] + 2, 3)), improve_candidates=None)


def test_bench_config4_host_pieces(host_amg, golden_dir):
    """bench.py's config-4 plumbing on a block hierarchy: the generator, block-multicolour smoother specs (MIS colouring
    of the BLOCK graph), their oracle mirror (block rows relaxed in colour order) and the block byte model"""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    import oracle
    from oracle import multilevel as oml
    amg = host_amg
    A, B = amg.gallery.linear_elasticity_tets(7, device=torch.device('cpu'))
    gs = ('gauss_seidel', {'sweep': 'symmetric'})
    np.random.seed(0)             # spectral radii from the oracle's setup: the device Arnoldi needs the GPU's dot / axpy
    ref = osa.smoothed_aggregation_hierarchy(A.to_scipy(), B.numpy(), presmoother=gs, postsmoother=gs, max_coarse=10, improve_candidates=None)
    mc = ('multicolor_gauss_seidel', {'sweep': 'symmetric'})
    ml = amg.smoothed_aggregation_solver(A, B, presmoother=mc, postsmoother=mc, max_coarse=10,
                                         rho_override=[(l.rho_smooth, None) for l in ref[:-1]], improve_candidates=None)
    assert [l.A.blocksize for l in ml.levels] == [(3, 3), (6, 6), (6, 6)]
    sm = ml.levels[0].presmoother
    assert sm.kind == 2 and sm.indices.numel() == ml.levels[0].A.n_brow          # block rows, not scalar rows
    levels = bench.host_levels(ml)
    assert levels[0].pre[0] == 'bsr_gauss_seidel_indexed' and levels[0].A.blocksize == (3, 3)
    b = np.random.rand(A.shape[0])
    res = []
    oml.Solver(levels, core=oracle.port).solve(b, tol=1e-30, maxiter=6, residuals=res)
    assert res[-1] < 0.1 * res[0]
    lex = bench.lexicographic(levels)
    assert lex[0].pre == ('gauss_seidel', {'sweep': 'symmetric', 'iterations': 1}) and lex[0].A is levels[0].A
    A0 = ml.levels[0].A
    assert bench.block_table(ml)[0] == [A0.n_brow, A0.nblocks, 3]
    vb = bench.vcycle_bytes(ml, 2)
    assert vb > 4 * bench.bsr_gs_sweep_bytes(A0.n_brow, A0.nblocks, 3) + bench.block_bytes(A0.nblocks, 3, 3)


def test_block_path_with_the_reference_default_improve_candidates(host_amg, golden_dir):
    """improve_candidates = the reference's default (4 symmetric block_gauss_seidel sweeps on A B = 0, inverses of the
    diagonal blocks by pinv_array): the orchestration + the kernel sources rebuild the hierarchy the REAL reference built
    with its defaults (tests/golden/sa_elasticity3d_tets_7.npz)"""
    amg = host_amg
    levels, ex = hier_io.load(os.path.join(golden_dir, 'sa_elasticity3d_tets_7.npz'))
    B = hier_io.load(os.path.join(golden_dir, 'sa_elasticity3d_tets_7_setup.npz'))[1]['B']
    A = levels[0].A
    gs = ('gauss_seidel', {'sweep': 'symmetric'})
    default = [('block_gauss_seidel', {'sweep': 'symmetric', 'iterations': 4}), None]
    np.random.seed(0)
    ref = osa.smoothed_aggregation_hierarchy(A.copy(), B, presmoother=gs, postsmoother=gs, max_coarse=10, improve_candidates=default)
    ml = amg.smoothed_aggregation_solver(A, B, presmoother=gs, postsmoother=gs, max_coarse=10, improve_candidates=default,
                                         rho_override=[(l.rho_smooth, None) for l in ref[:-1]])
    assert len(ml.levels) == len(levels)
    for k, (m, r) in enumerate(zip(ml.levels, levels)):
        for nm in ('A', 'P', 'R'):
            if hasattr(r, nm):
                assert same(getattr(m, nm).to_scipy(), getattr(r, nm)), (k, nm)
    assert 'b200amg_block_gauss_seidel' in amg._seen_calls and 'b200amg_block_diag_inverse' in amg._seen_calls
    # block smoother names resolve against the same inverses
    amg.change_smoothers(ml, ('multicolor_block_gauss_seidel', {'sweep': 'symmetric'}), ('block_jacobi', {'omega': 1.0, 'withrho': False}))
    pre, post = ml.levels[0].presmoother, ml.levels[0].postsmoother
    assert (pre.kind, post.kind) == (6, 5) and pre.Dinv is post.Dinv and tuple(pre.Dinv.shape) == (294, 3, 3)
    assert pre.indices.numel() == 294


def test_bench_helpers_on_a_scalar_hierarchy(host_amg, golden_dir):
    """the helpers bench.py's DEFAULT (config 2) path calls on the solver object -- byte model, level tables, oracle
    specs -- on a scalar SA hierarchy held in DevMats (the attribute surface they rely on)"""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    levels, ex = hier_io.load(os.path.join(golden_dir, 'sa_poisson3d_12.npz'))
    ml = host_amg.import_hierarchy(levels)
    want = 0
    for l in levels[:-1]:
        n, nnz, nc = l.A.shape[0], l.A.nnz, l.P.shape[1]
        want += 3 * (12 * nnz + 4 * n + 24 * n) + (12 * l.R.nnz + 4 * nc + 8 * n + 8 * nc) + (12 * l.P.nnz + 4 * n + 8 * nc + 16 * n)
    want += 8 * levels[-1].A.shape[0] ** 2
    assert bench.vcycle_bytes(ml) == want
    assert bench.block_table(ml) is None
    assert bench.level_table(ml) == [[l.A.shape[0], l.A.nnz] for l in levels]
    hl = bench.host_levels(ml)
    assert hl[0].pre[0] == 'jacobi' and abs(hl[0].pre[1]['omega'] - levels[0].pre[1]['omega']) == 0.0
    assert bench.block_bytes(ml.levels[0].A.nblocks, *ml.levels[0].A.blocksize) == 12 * levels[0].A.nnz


def test_block_path_with_a_strength_threshold(host_amg, golden_dir):
    """strength=('symmetric', {'theta': 0.01}) on a BSR 2x2 problem: block rows measured by the squared Frobenius norm of
    their blocks, entries below the threshold leave the aggregation graph -- the hierarchy the REAL reference built"""
    amg = host_amg
    levels, ex = hier_io.load(os.path.join(golden_dir, 'sa_elasticity2d_16_theta001.npz'))
    A, B, theta = levels[0].A, ex['B'], float(ex['theta'])
    jac = ('jacobi', {'omega': 4.0 / 3.0})
    np.random.seed(0)
    ref = osa.smoothed_aggregation_hierarchy(A.copy(), B, theta=theta, presmoother=jac, postsmoother=jac, max_coarse=10, improve_candidates=None)
    gs = ('gauss_seidel', {'sweep': 'symmetric'})      # smoothers are not the subject here (Jacobi needs the device Arnoldi)
    ml = amg.smoothed_aggregation_solver(A, B, strength=('symmetric', {'theta': theta}), presmoother=gs, postsmoother=gs,
                                         max_coarse=10, rho_override=[(l.rho_smooth, None) for l in ref[:-1]], improve_candidates=None)
    assert len(ml.levels) == len(levels) == 4
    for k, (m, r) in enumerate(zip(ml.levels, levels)):
        for nm in ('A', 'P', 'R'):
            if hasattr(r, nm):
                assert same(getattr(m, nm).to_scipy(), getattr(r, nm)), (k, nm)
    assert {'b200amg_block_sqnorm', 'b200amg_abs_diag', 'b200amg_symmetric_strength'} <= amg._seen_calls
    # fewer strong connections than stored blocks on some level: the threshold did something
    dropped = [l.A.nblocks - amg.aggregation.symmetric_strength_of_connection(l.A, theta).nnz for l in ml.levels[:-1]]
    assert min(dropped) >= 0 and max(dropped) > 0, dropped
    with pytest.raises(ValueError):
        amg.smoothed_aggregation_solver(A, B, strength=('symmetric', {'theta': -1.0}), improve_candidates=None)


@pytest.mark.parametrize('name,kw', [('sa_poisson3d_12', {}), ('sa_aniso2d_40_theta005', {'strength': ('symmetric', {'theta': 0.05})})])
def test_scalar_sa_path_builds_the_reference_hierarchy(host_amg, golden_dir, name, kw):
    """the scalar smoothed-aggregation setup (pinned on the GPU in round 1: tests/test_gpu_cycle.py) through the same CPU
    harness -- setup_kernels.cu compiled for the host, SpGEMM / transpose by scipy -- against fixtures from the REAL
    reference, with and without a strength threshold"""
    amg = host_amg
    levels, ex = hier_io.load(os.path.join(golden_dir, name + '.npz'))
    jac = ('jacobi', {'omega': 4.0 / 3.0})
    mc = {'sa_poisson3d_12': 20}.get(name, 10)
    np.random.seed(0)
    ref = osa.smoothed_aggregation_hierarchy(levels[0].A.copy(), theta=kw.get('strength', (0, {'theta': 0.0}))[1]['theta'],
                                             presmoother=jac, postsmoother=jac, max_coarse=mc, improve_candidates=None)
    rhos = [(l.rho_smooth, l.A.rho_D_inv) for l in ref[:-1]]
    ml = amg.smoothed_aggregation_solver(levels[0].A, presmoother=jac, postsmoother=jac, max_coarse=mc, rho_override=rhos, **kw, improve_candidates=None)
    assert len(ml.levels) == len(levels)
    for k, (m, r) in enumerate(zip(ml.levels, levels)):
        for nm in ('A', 'P', 'R'):
            if hasattr(r, nm):
                assert same(getattr(m, nm).to_scipy(), getattr(r, nm)), (k, nm)
    assert abs(ml.levels[0].presmoother.omega - levels[0].pre[1]['omega']) == 0.0


@pytest.mark.parametrize('name', ['rs_poisson2d_40_mcgs', 'rs_aniso_48_mcgs'])
def test_ruge_stuben_path_builds_the_reference_hierarchy(host_amg, golden_dir, name):
    """classical Ruge-Stuben setup (pyamg_b200/classical.py; rs_setup.cu compiled for the host, the C/F splitting by the
    real library's host function) against the fixtures from the REAL reference, unsorted coarse rows included"""
    amg = host_amg
    levels, ex = hier_io.load(os.path.join(golden_dir, name + '.npz'))
    ml = amg.ruge_stuben_solver(levels[0].A, max_coarse=20, presmoother=None, postsmoother=None)
    assert len(ml.levels) == len(levels)
    for k, (m, r) in enumerate(zip(ml.levels, levels)):
        for nm in ('A', 'P', 'R'):
            if hasattr(r, nm):
                assert same(getattr(m, nm).to_scipy(), getattr(r, nm)), (k, nm)


def test_host_driven_cycles_on_a_block_hierarchy(host_amg, golden_dir):
    """multilevel_solver.cycle_host (the host-driven cycle behind cycle='AMLI'; V / W / F as the engine's cross-check) on
    the hierarchy the REAL reference built with its default block_gauss_seidel smoothers: V-cycles against the history
    the reference recorded, W and AMLI against the oracle's restatement (itself pinned to reference fixtures).  Block
    smoother and pinv kernels from the host-compiled source, SpMV / dense solve / dots by scipy and numpy doubles."""
    import oracle
    from oracle import multilevel as oml
    amg = host_amg
    levels, ex = hier_io.load(os.path.join(golden_dir, 'sa_elasticity3d_tets_7_blockgs.npz'))
    bgs = ('block_gauss_seidel', {'sweep': 'symmetric'})
    ml = amg.import_hierarchy(levels, presmoother=bgs, postsmoother=bgs)
    for m, r in zip(ml.levels[:-1], levels[:-1]):
        assert m.presmoother.kind == 7 and np.array_equal(m.presmoother.Dinv.numpy(), r.pre[1]['Dinv'])   # pinv_array bits
    b = torch.from_numpy(ex['b'].copy())
    A0 = levels[0].A

    def history(cycle, ncyc):
        x = torch.zeros_like(b)
        res = [float(np.sqrt(np.inner(ex['b'], ex['b'])))]
        for _ in range(ncyc):
            ml.cycle_host(0, x, b, cycle)
            r = ex['b'] - A0 * x.numpy()
            res.append(float(np.sqrt(np.inner(r, r))))
        return np.array(res), x.numpy()
    res, x = history('V', 4)                                  # (the emulated sequential sweep is slow: 4 of the 8 recorded cycles)
    assert np.array_equal(res, ex['residuals'][:5])
    for cyc in ('W', 'AMLI'):
        ro = []
        xo = oml.Solver(levels, core=oracle.port).solve(ex['b'], tol=1e-30, maxiter=1, residuals=ro, cycle=cyc)
        res, x = history(cyc, 1)
        assert np.all(np.abs(res - np.array(ro)) <= 1e-12 * ro[0]), cyc
        assert np.abs(x - xo).max() <= 1e-12 * np.abs(xo).max(), cyc
    with pytest.raises(TypeError):
        ml.cycle_host(0, torch.zeros_like(b), b, 'Z')


def test_config4_in_miniature_end_to_end(host_amg, golden_dir):
    """BASELINE config 4 end to end on the kernel SOURCES, without a GPU: generator -> block SA setup with the six
    rigid-body modes -> symmetric block-multicolour bsr_gauss_seidel V-cycles (host-driven), every block kernel from the
    host-compiled csrc files -- against the oracle's cycle on the same hierarchy with the same block rows relaxed one by
    one in the same colour order by the reference row update (what bench.py's parity_in_run does at full size)"""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    import oracle
    from oracle import multilevel as oml
    amg = host_amg
    A, B = amg.gallery.linear_elasticity_tets(7, device=torch.device('cpu'))
    gs = ('gauss_seidel', {'sweep': 'symmetric'})
    np.random.seed(0)
    ref = osa.smoothed_aggregation_hierarchy(A.to_scipy(), B.numpy(), presmoother=gs, postsmoother=gs, max_coarse=10, improve_candidates=None)
    mc = ('multicolor_gauss_seidel', {'sweep': 'symmetric'})
    ml = amg.smoothed_aggregation_solver(A, B, presmoother=mc, postsmoother=mc, max_coarse=10,
                                         rho_override=[(l.rho_smooth, None) for l in ref[:-1]], improve_candidates=None)
    rng = np.random.RandomState(1)
    bh = rng.rand(A.shape[0])
    levels = bench.host_levels(ml)
    ro = []
    xo = oml.Solver(levels, core=oracle.port).solve(bh, tol=1e-30, maxiter=3, residuals=ro)
    b, x = torch.from_numpy(bh.copy()), torch.zeros(A.shape[0], dtype=torch.float64)
    for _ in range(3):
        ml.cycle_host(0, x, b, 'V')
    assert np.array_equal(x.numpy(), xo)                        # same arithmetic, same order: the bits
    assert ro[-1] < 0.5 * ro[0]
    assert {'b200amg_bsr_gauss_seidel', 'b200amg_bsr_matvec', 'b200amg_fit_candidates', 'b200amg_bsr_matmat_numeric'} <= amg._seen_calls


def test_coarse_solver_names(host_amg, golden_dir):
    """coarse_solver = 'pinv2' (default), 'pinv', 'lu', 'cholesky', 'splu' (multilevel.py:608-650): the inverse formed on
    the host once; all of them solve the SPD coarse system of the fixture"""
    amg = host_amg
    levels, _ = hier_io.load(os.path.join(golden_dir, 'sa_poisson3d_12.npz'))
    Ac = levels[-1].A
    rhs = np.random.RandomState(0).rand(Ac.shape[0])
    want = np.linalg.solve(Ac.toarray(), rhs)
    for name in ('pinv2', 'pinv', 'lu', 'cholesky', 'splu'):
        cs = amg.multilevel.coarse_grid_solver(name)
        x = cs(amg.DevMat.from_scipy(Ac), torch.from_numpy(rhs.copy()))
        assert np.abs(x.numpy() - want).max() <= 1e-12 * np.abs(want).max(), name
    with pytest.raises(ValueError):
        amg.multilevel.coarse_grid_solver('gmres')
    # 'splu' with an empty row / column: zero there
    D = Ac.toarray()
    D[3, :] = 0.0
    D[:, 3] = 0.0
    cs = amg.multilevel.coarse_grid_solver('splu')
    x = cs(amg.DevMat.from_scipy(sp.csr_matrix(D)), torch.from_numpy(rhs.copy())).numpy()
    keep = np.array([i for i in range(D.shape[0]) if i != 3])
    assert x[3] == 0.0 and np.allclose(x[keep], np.linalg.solve(D[np.ix_(keep, keep)], rhs[keep]), rtol=1e-11)
