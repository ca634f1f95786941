"""CPU tests (-m "not gpu"): the C-ABI library loads and exports exactly what include/b200amg.h
declares; the host-side (sequential) setup functions equal the oracle; the product refuses to
run without a CUDA device instead of falling back."""
import ctypes
import os
import re

import numpy as np
import pytest
import scipy.sparse as sp

import oracle
from oracle import relaxation as orx

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, 'include', 'b200amg.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(b200amg_[a-z0-9_]+)\s*\(', text)))


def test_library_exports_every_declared_symbol():
    from pyamg_b200 import _lib
    lib = _lib.load()
    names = declared_symbols()
    assert len(names) >= 35
    for n in names:
        assert hasattr(lib, n), 'missing export ' + n
    assert set(_lib.EXPORTS) <= set(names), set(_lib.EXPORTS) - set(names)
    assert lib.b200amg_version() >= 100
    assert ctypes.sizeof(_lib.LevelDesc) > 0


def test_no_torch_types_in_header():
    text = open(os.path.join(ROOT, 'include', 'b200amg.h')).read()
    assert 'torch' not in text and 'at::' not in text and 'extern "C"' in text


def test_host_setup_functions_equal_oracle():
    from pyamg_b200 import amg_core
    rng = np.random.RandomState(0)
    for trial in range(60):
        n = rng.randint(1, 80)
        G = sp.random(n, n, rng.uniform(0.01, 0.2), random_state=rng, format='csr')
        G = sp.csr_matrix(G + G.T + (sp.eye(n) if trial % 2 else 0 * sp.eye(n)))
        Ap, Aj = G.indptr.astype(np.intc), G.indices.astype(np.intc)
        c = np.empty(n, np.intc)
        K = amg_core.vertex_coloring_mis(n, Ap, Aj, c)
        c2 = np.empty(n, np.intc)
        K2 = oracle.port.vertex_coloring_mis(n, Ap, Aj, c2)
        assert K == K2 and np.array_equal(c, c2)
        a, r = np.empty(n, np.intc), np.empty(n, np.intc)
        a2, r2 = np.empty(n, np.intc), np.empty(n, np.intc)
        k = amg_core.standard_aggregation(n, Ap, Aj, a, r)
        k2 = oracle.port.standard_aggregation(n, Ap, Aj, a2, r2)
        assert k == k2 and np.array_equal(a, a2) and np.array_equal(r[:k], r2[:k2])
    # grids: the BASELINE problem class
    from helpers_mats import poisson3d
    G = poisson3d(11)
    Ap, Aj = G.indptr.astype(np.intc), G.indices.astype(np.intc)
    a, r = np.empty(G.shape[0], np.intc), np.empty(G.shape[0], np.intc)
    a2, r2 = a.copy(), r.copy()
    assert amg_core.standard_aggregation(G.shape[0], Ap, Aj, a, r) == oracle.port.standard_aggregation(G.shape[0], Ap, Aj, a2, r2)
    assert np.array_equal(a, a2)


def test_host_rs_splitting_equals_oracle():
    """b200amg_rs_cf_splitting_host == the C restatement (itself pinned against the reference .so) on random
    graphs, grids and directed (non-symmetric) strength graphs"""
    from pyamg_b200 import amg_core
    from oracle import rs as ors
    rng = np.random.RandomState(4)
    mats = []
    for trial in range(50):
        n = rng.randint(1, 120)
        G = sp.random(n, n, rng.uniform(0.01, 0.25), random_state=rng, format='csr')
        mats.append(sp.csr_matrix(G + G.T) if trial % 3 else sp.csr_matrix(G))
    from helpers_mats import poisson3d
    mats.append(poisson3d(9))
    g = 31
    T = sp.diags([-1.0, 2.0, -1.0], [-1, 0, 1], shape=(g, g))
    mats.append(sp.csr_matrix(sp.kron(T, sp.eye(g)) + sp.kron(sp.eye(g), T)))
    for G in mats:
        S = ors.remove_diagonal(sp.csr_matrix(G))
        Tt = S.T.tocsr()
        n = S.shape[0]
        a, b = np.empty(n, np.intc), np.empty(n, np.intc)
        args = [np.ascontiguousarray(v, np.intc) for v in (S.indptr, S.indices, Tt.indptr, Tt.indices)]
        amg_core.rs_cf_splitting(n, *args, a)
        oracle.port.rs_cf_splitting(n, *args, b)
        assert np.array_equal(a, b)


def test_product_never_imports_oracle():
    bad = []
    for d, _, files in os.walk(os.path.join(ROOT, 'pyamg_b200')):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                src = open(os.path.join(d, f)).read()
                if re.search(r'^\s*(import|from)\s+oracle\b', src, flags=re.M) or 'oracle/' in src:
                    bad.append(f)
    assert not bad, bad


def test_no_cpu_fallback_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip('CUDA present')
    import pyamg_b200
    with pytest.raises(pyamg_b200._lib.B200AmgError):
        pyamg_b200.DevMat.from_scipy(sp.eye(4, format='csr'))
    with pytest.raises(pyamg_b200._lib.B200AmgError):
        pyamg_b200.gallery.poisson((4, 4))


def test_color_ordering_and_stencil_helpers():
    from pyamg_b200 import relaxation, gallery
    c = np.array([1, 0, 2, 0, 1], dtype=np.intc)
    idx, ptr = relaxation.color_ordering(c)
    assert idx.tolist() == [1, 3, 0, 4, 2] and ptr.tolist() == [0, 2, 4, 5]
    i2, p2 = orx.color_ordering(c)
    assert np.array_equal(idx, i2) and np.array_equal(ptr, p2)
    st = gallery.diffusion_stencil_2d(0.001, np.pi / 8, 'FE')
    assert st.shape == (3, 3) and abs(st.sum()) < 1e-12


def test_chebyshev_coefficients_host_formula():
    """pyamg/relaxation/chebyshev.py:9-56 (docstring example :37-38) and the sign / truncation setup_chebyshev applies
    (smoothing.py:438-449); pure host arithmetic, no GPU needed"""
    from pyamg_b200.smoothing import chebyshev_polynomial_coefficients
    c = chebyshev_polynomial_coefficients(1.0, 2.0, 3)
    assert np.allclose(c, [-0.32323232, 1.45454545, -2.12121212, 1.0], atol=5e-9)
    assert abs(np.polyval(c, 0.0) - 1.0) < 1e-15
    roots = np.sort(np.roots(c).real)
    assert np.all(roots > 1.0) and np.all(roots < 2.0)
    import pytest
    with pytest.raises(ValueError):
        chebyshev_polynomial_coefficients(2.0, 1.0, 3)


def test_tet_elasticity_generator_equals_reference_gallery(golden_dir):
    """pyamg_b200.gallery.linear_elasticity_tets (stencil-wise assembly, BASELINE config 4's workload) against the
    matrix the REAL reference's linear_elasticity_p1 assembled element by element on the same mesh: identical block
    pattern, values to the rounding of the duplicate sums, rigid-body modes identical.  Index arithmetic only (torch on
    the CPU device passed explicitly): no kernel of the product runs here."""
    import torch
    import hier_io
    from pyamg_b200 import gallery
    levels, ex = hier_io.load(os.path.join(golden_dir, 'sa_elasticity3d_tets_7_setup.npz'))
    Aref = levels[0].A
    A, B = gallery.linear_elasticity_tets(7, device=torch.device('cpu'))
    As = A.to_scipy()
    assert As.blocksize == (3, 3) and As.shape == Aref.shape
    assert np.array_equal(As.indptr, Aref.indptr) and np.array_equal(As.indices, Aref.indices)
    assert np.abs(As.data - Aref.data).max() <= 1e-13 * np.abs(Aref.data).max()
    assert np.array_equal(B.numpy(), ex['B'])
    # the operator is symmetric and annihilates the rigid-body modes away from the clamped face
    D = (As - As.T)
    assert abs(D).max() <= 1e-10 * np.abs(Aref.data).max()


def test_level_descriptor_mirror_matches_the_compiled_struct():
    """the ctypes mirror of b200amg_level_desc (pyamg_b200/_lib.py) has the size the library was compiled with"""
    from pyamg_b200 import _lib
    assert ctypes.sizeof(_lib.LevelDesc) == _lib.load().b200amg_level_desc_size()
    assert _lib.LevelDesc.gs_csr.offset + 8 == ctypes.sizeof(_lib.LevelDesc)          # last field + tail padding


def test_host_jp_ldf_colourings():
    """vertex_coloring_jones_plassmann / vertex_coloring_LDF on the host: the reference's known-answer vectors
    (pyamg/tests/test_graph.py:172-227), equality with the reference's compiled kernels / the C port on random graphs
    with tied weights and self loops, and validity (no edge inside a colour)"""
    from pyamg_b200 import amg_core
    G0 = sp.csr_matrix(np.array([[0, 1, 0, 1, 0], [1, 0, 1, 1, 1], [0, 1, 0, 0, 1], [1, 1, 0, 0, 1], [0, 1, 1, 1, 0]]))
    G1 = sp.csr_matrix(np.array([[0, 1, 1, 0, 0, 0], [1, 0, 1, 1, 0, 0], [1, 1, 0, 0, 0, 0], [0, 1, 0, 0, 1, 1],
                                 [0, 0, 0, 1, 0, 1], [0, 0, 0, 1, 1, 0]]))
    w0, w1 = np.array([0.8, 0.1, 0.9, 0.7, 0.6]), np.array([0.1, 0.2, 0.3, 0.1, 0.2, 0.3])
    gold = {('vertex_coloring_jones_plassmann', 0): [2, 0, 1, 1, 2], ('vertex_coloring_jones_plassmann', 1): [2, 0, 1, 1, 2, 0],
            ('vertex_coloring_LDF', 0): [2, 0, 1, 1, 2], ('vertex_coloring_LDF', 1): [2, 0, 1, 2, 1, 0]}
    cores = [amg_core, oracle.port] + ([oracle.ref] if oracle.ref is not None else [])
    for (fn, which), want in gold.items():
        G, w = ((G0, w0), (G1, w1))[which]
        for core in cores:
            c = np.empty(G.shape[0], np.intc)
            getattr(core, fn)(G.shape[0], G.indptr.astype(np.intc), G.indices.astype(np.intc), c, w.copy())
            assert list(c) == want, (fn, which, core)
    rng = np.random.RandomState(3)
    for trial in range(40):
        n = rng.randint(1, 300)
        G = sp.random(n, n, rng.uniform(0.01, 0.1), random_state=rng, format='csr')
        G = sp.csr_matrix(G + G.T + (sp.eye(n) if trial % 2 else 0 * sp.eye(n)))
        Ap, Aj = G.indptr.astype(np.intc), G.indices.astype(np.intc)
        w = rng.rand(n)
        w[::5] = w[0]
        for fn in ('vertex_coloring_jones_plassmann', 'vertex_coloring_LDF'):
            a, b = np.empty(n, np.intc), np.empty(n, np.intc)
            K = getattr(amg_core, fn)(n, Ap, Aj, a, w.copy())
            getattr(cores[-1], fn)(n, Ap, Aj, b, w.copy())
            assert np.array_equal(a, b) and K == a.max() + 1
            rows = np.repeat(np.arange(n), np.diff(Ap))
            off = rows != Aj
            assert not np.any(a[rows[off]] == a[Aj[off]])


def test_host_naive_aggregation_equals_reference():
    """b200amg_naive_aggregation_host (aggregate='naive') against the reference's compiled amg_core.naive_aggregation
    (1-based there, 0-based here like the reference's Python wrapper makes them) and its docstring example"""
    from pyamg_b200 import amg_core
    G = sp.csr_matrix(np.array([[2., -1, 0, 0], [-1, 2, -1, 0], [0, -1, 2, -1], [0, 0, -1, 2]]))
    a, r = np.empty(4, np.intc), np.empty(4, np.intc)
    assert amg_core.naive_aggregation(4, G.indptr, G.indices, a, r) == 2 and list(a) == [0, 0, 1, 1] and list(r[:2]) == [0, 2]
    if oracle.ref is None:
        pytest.skip('reference kernels not built')
    rng = np.random.RandomState(0)
    for t in range(40):
        n = rng.randint(1, 200)
        M = sp.random(n, n, rng.uniform(0.01, 0.1), random_state=rng, format='csr')
        M = sp.csr_matrix(M + M.T + (sp.eye(n) if t % 2 else 0 * sp.eye(n)))
        Ap, Aj = M.indptr.astype(np.intc), M.indices.astype(np.intc)
        a, r, a2, r2 = (np.empty(n, np.intc) for _ in range(4))
        k = amg_core.naive_aggregation(n, Ap, Aj, a, r)
        k2 = oracle.ref.naive_aggregation(n, Ap, Aj, a2, r2)
        assert k == k2 and np.array_equal(a, a2 - 1) and np.array_equal(r[:k], r2[:k2])


def test_gs_wavefronts_host_gives_the_sequential_sweep():
    """b200amg_gs_wavefronts_host: rows grouped by level, levels in ascending (descending) order with the rows of a
    level updated TOGETHER from the values before the level, reproduce the lexicographic forward (backward)
    Gauss-Seidel sweep of the reference kernel bit for bit -- symmetric and non-symmetric patterns, empty rows"""
    from pyamg_b200 import amg_core
    rng = np.random.RandomState(8)
    n = 700
    rows = np.repeat(np.arange(n), 3)
    cols = np.clip(rows + rng.randint(-25, 26, rows.size), 0, n - 1)
    N = sp.coo_matrix((rng.rand(rows.size) - 0.5, (rows, cols)), shape=(n, n)).tocsr()
    N = sp.csr_matrix(N + sp.diags(np.full(n, 3.0)))
    from helpers_mats import poisson3d
    for A in (N, sp.csr_matrix(poisson3d(7))):
        A.sort_indices()
        m = A.shape[0]
        lvl = np.empty(m, np.intc)
        K = amg_core.gs_wavefronts(m, A.indptr.astype(np.intc), A.indices.astype(np.intc), lvl)
        assert lvl.min() == 0 and lvl.max() == K - 1
        C = sp.coo_matrix(A + A.T)
        lt = C.row < C.col
        assert np.all(lvl[C.row[lt]] < lvl[C.col[lt]])
        x0, b = rng.rand(m), rng.rand(m)
        for sweep, order in (('forward', range(K)), ('backward', range(K - 1, -1, -1))):
            want = x0.copy()
            orx.gauss_seidel(A, want, b, sweep=sweep, core=oracle.port)
            x = x0.copy()
            for L in order:
                new = x.copy()
                for i in np.where(lvl == L)[0]:
                    one = x.copy()
                    orx.gauss_seidel_indexed(A, one, b, np.array([i], dtype=np.intc), core=oracle.port)
                    new[i] = one[i]
                x = new
            assert np.array_equal(x, want), sweep
