"""GPU parity tests (-m gpu): every kernel of libb200amg, called through the C-ABI bindings,
against the oracle (C port / reference .so / scipy) on the same seeded inputs.
fp64 results are required to be BIT-IDENTICAL (np.array_equal) unless a tolerance is stated."""
import numpy as np
import pytest
import scipy.sparse as sp

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu

import oracle
from oracle import relaxation as orx


@pytest.fixture(scope='module')
def amg():
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    import pyamg_b200
    return pyamg_b200


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def rand_matrix(n, density, rng, diag=3.0, sym=True):
    A = sp.random(n, n, density, random_state=rng, format='csr')
    if sym:
        A = A + A.T
    A = sp.csr_matrix(A + diag * sp.eye(n))
    A.sort_indices()
    return A


def lap1d(n):
    return sp.spdiags([-np.ones(n), 2 * np.ones(n), -np.ones(n)], [-1, 0, 1], n, n, format='csr')


def poisson3d(m):
    I = sp.eye(m); T = lap1d(m)
    return sp.csr_matrix(sp.kron(sp.kron(T, I), I) + sp.kron(sp.kron(I, T), I) + sp.kron(sp.kron(I, I), T))


CASES = [('rand300', lambda rng: rand_matrix(300, 0.03, rng)),
         ('rand5000_long', lambda rng: rand_matrix(5000, 0.02, rng)),      # ~200 nnz/row: multi-chunk CTAs
         ('poisson3d_17', lambda rng: poisson3d(17)),
         ('empty_rows', lambda rng: sp.csr_matrix(sp.random(700, 700, 0.002, random_state=rng))),
         ('one', lambda rng: sp.csr_matrix(np.array([[2.0]])))]


@pytest.mark.parametrize('name,make', CASES)
def test_spmv_family_bit_exact(amg, name, make):
    rng = np.random.RandomState(0)
    A = make(rng)
    n = A.shape[0]
    x, b = rng.rand(n), rng.rand(n)
    dA = amg.DevMat.from_scipy(A)
    y = dA.matvec(dev(x)).cpu().numpy()
    assert np.array_equal(y, A * x)
    r = dA.residual(dev(x), dev(b)).cpu().numpy()
    assert np.array_equal(r, b - A * x)
    xx, dx = dev(b.copy()), dev(x)
    amg._lib.call('b200amg_csr_matvec_add', n, dA.indptr.data_ptr(), dA.indices.data_ptr(), dA.data.data_ptr(),
                  dx.data_ptr(), xx.data_ptr(), 0)
    torch.cuda.synchronize()
    assert np.array_equal(xx.cpu().numpy(), b + A * x)


def test_spmv_unaligned_and_rectangular(amg):
    rng = np.random.RandomState(1)
    A = sp.random(1000, 333, 0.05, random_state=rng, format='csr')
    x = rng.rand(333)
    dA = amg.DevMat.from_scipy(A)
    assert np.array_equal(dA.matvec(dev(x)).cpu().numpy(), A * x)
    # pointers offset by one element: the scalar (non-128-bit) path
    pad_j = torch.cat([torch.zeros(1, dtype=torch.int32, device='cuda'), dA.indices])[1:]
    pad_x = torch.cat([torch.zeros(1, dtype=torch.float64, device='cuda'), dA.data])[1:]
    y = torch.empty(1000, dtype=torch.float64, device='cuda')
    dx = dev(x)
    amg._lib.call('b200amg_csr_matvec', 1000, 333, dA.indptr.data_ptr(), pad_j.data_ptr(), pad_x.data_ptr(),
                  dx.data_ptr(), y.data_ptr(), 0)
    assert np.array_equal(y.cpu().numpy(), A * x)


@pytest.mark.parametrize('name,make', CASES[:4])
def test_jacobi_bit_exact(amg, name, make):
    rng = np.random.RandomState(2)
    A = make(rng)
    n = A.shape[0]
    x0, b = rng.rand(n), rng.rand(n)
    for omega, iters in ((1.0, 1), (4.0 / 3.0 / 1.9, 3)):
        xo = x0.copy()
        orx.jacobi(A, xo, b, iterations=iters, omega=omega, core=oracle.port)
        xd = dev(x0.copy())
        amg.relaxation.jacobi(amg.DevMat.from_scipy(A), xd, dev(b), iterations=iters, omega=omega)
        assert np.array_equal(xd.cpu().numpy(), xo)
    # BSR(1,1) container -> amg_core.bsr_jacobi arithmetic (different association)
    Ab = A.tobsr(blocksize=(1, 1))
    xo = x0.copy(); orx.jacobi(Ab, xo, b, omega=0.7, core=oracle.port)
    xd = dev(x0.copy()); amg.relaxation.jacobi(amg.DevMat.from_scipy(Ab), xd, dev(b), omega=0.7)
    assert np.array_equal(xd.cpu().numpy(), xo)


@pytest.mark.parametrize('bs', [2, 3, 6])
def test_bsr_kernels_bit_exact(amg, bs):
    rng = np.random.RandomState(3)
    n = 60 * bs
    A = rand_matrix(n, 0.02, rng).tobsr(blocksize=(bs, bs))
    x0, b = rng.rand(n), rng.rand(n)
    dA = amg.DevMat.from_scipy(A)
    assert np.array_equal(dA.matvec(dev(x0)).cpu().numpy(), A * x0)
    assert np.array_equal(dA.residual(dev(x0), dev(b)).cpu().numpy(), b - A * x0)
    xo = x0.copy(); orx.jacobi(A, xo, b, iterations=2, omega=0.8, core=oracle.port)
    xd = dev(x0.copy()); amg.relaxation.jacobi(dA, xd, dev(b), iterations=2, omega=0.8)
    assert np.array_equal(xd.cpu().numpy(), xo)
    for sweep in ('forward', 'backward', 'symmetric'):
        xo = x0.copy(); orx.gauss_seidel(A, xo, b, sweep=sweep, core=oracle.port)
        xd = dev(x0.copy()); amg.relaxation.gauss_seidel(dA, xd, dev(b), sweep=sweep)     # sequential on device
        assert np.array_equal(xd.cpu().numpy(), xo), sweep
    # block-multicolour sweep == sequential sweep over the colour-grouped block rows
    G = sp.csr_matrix((np.ones(len(A.indices)), A.indices, A.indptr), shape=(n // bs, n // bs))
    colors = amg.relaxation.vertex_coloring(G)
    assert np.array_equal(colors, orx.vertex_coloring_mis(G, oracle.port))
    idx, ptr = amg.relaxation.color_ordering(colors)
    xo = x0.copy()
    data = np.ravel(A.data)
    for i in idx:
        oracle.port.bsr_gauss_seidel(A.indptr, A.indices, data, xo, b, int(i), int(i) + 1, 1, bs)
    for i in idx[::-1]:
        oracle.port.bsr_gauss_seidel(A.indptr, A.indices, data, xo, b, int(i), int(i) - 1, -1, bs)
    xd = dev(x0.copy())
    amg.relaxation.gauss_seidel(dA, xd, dev(b), sweep='symmetric', indices=dev(idx), color_ptr=ptr)
    assert np.array_equal(xd.cpu().numpy(), xo)
    # rectangular blocks (P of an elasticity hierarchy: bs x K)
    P = sp.random(n, 40 * 3, 0.05, random_state=rng, format='csr').tobsr(blocksize=(bs, 3))
    xc = rng.rand(P.shape[1])
    assert np.array_equal(amg.DevMat.from_scipy(P).matvec(dev(xc)).cpu().numpy(), P * xc)


@pytest.mark.parametrize('name,make', CASES[:3])
def test_multicolor_gauss_seidel_bit_exact(amg, name, make):
    rng = np.random.RandomState(4)
    A = make(rng)
    n = A.shape[0]
    x0, b = rng.rand(n), rng.rand(n)
    colors = amg.relaxation.vertex_coloring(A)
    assert np.array_equal(colors, orx.vertex_coloring_mis(A, oracle.port))
    idx, ptr = amg.relaxation.color_ordering(colors)
    dA = amg.DevMat.from_scipy(A)
    for sweep in ('forward', 'backward', 'symmetric'):
        xo = x0.copy(); orx.gauss_seidel_indexed(A, xo, b, idx, iterations=2, sweep=sweep, core=oracle.port)
        xd = dev(x0.copy())
        amg.relaxation.gauss_seidel_indexed(dA, xd, dev(b), dev(idx), iterations=2, sweep=sweep, color_ptr=ptr)
        assert np.array_equal(xd.cpu().numpy(), xo), sweep
    # without colour information: strictly sequential on the device, any list (repeats, subsets)
    lst = rng.randint(0, n, 3 * n // 2).astype('intc')
    xo = x0.copy(); orx.gauss_seidel_indexed(A, xo, b, lst, sweep='symmetric', core=oracle.port)
    xd = dev(x0.copy()); amg.relaxation.gauss_seidel_indexed(dA, xd, dev(b), lst, sweep='symmetric')
    assert np.array_equal(xd.cpu().numpy(), xo)
    # lexicographic gauss_seidel (sequential on the device)
    xo = x0.copy(); orx.gauss_seidel(A, xo, b, sweep='symmetric', core=oracle.port)
    xd = dev(x0.copy()); amg.relaxation.gauss_seidel(dA, xd, dev(b), sweep='symmetric')
    assert np.array_equal(xd.cpu().numpy(), xo)


def nonsymmetric_band(rng, n=6000):
    """rows reference 2-3 columns within +-40 of the diagonal, pattern NOT symmetric: a row may be read by an earlier
    row that it does not reference itself (the anti-dependence of the wavefront levels)"""
    rows = np.repeat(np.arange(n), 3)
    cols = np.clip(rows + rng.randint(-40, 41, rows.size), 0, n - 1)
    A = sp.coo_matrix((rng.rand(rows.size) - 0.5, (rows, cols)), shape=(n, n)).tocsr()
    A = sp.csr_matrix(A + sp.diags(np.full(n, 3.0)))
    A.sort_indices()
    return A


@pytest.mark.parametrize('name,make', [('poisson3d_17', lambda rng: poisson3d(17)), ('lap2d_90', lambda rng: lap2d(90)),
                                       ('nonsymmetric_band', nonsymmetric_band)])
def test_lexicographic_gauss_seidel_over_wavefronts_bit_exact(amg, monkeypatch, name, make):
    """amg_core.gauss_seidel (relaxation.h:33-62) run level-parallel over the dependency wavefronts of the sweep
    (relaxation.wavefront_ordering): the bits of the sequential sweep, forward, backward and symmetric"""
    monkeypatch.setattr(amg.relaxation, 'WAVEFRONT_MIN_WIDTH', 1)
    rng = np.random.RandomState(8)
    A = make(rng)
    n = A.shape[0]
    x0, b = rng.rand(n), rng.rand(n)
    dA = amg.DevMat.from_scipy(A)
    wf = amg.relaxation.wavefront_ordering(dA)
    assert wf is not None and int(wf[1][-1]) == n and len(wf[1]) - 1 < n
    lvl = np.empty(n, dtype=np.intc)
    lvl[wf[0].cpu().numpy()] = np.repeat(np.arange(len(wf[1]) - 1), np.diff(wf[1]))
    # definition of a valid schedule: every coupled pair (i, j), i < j, in either direction, has level[i] < level[j]
    C = sp.coo_matrix(A + A.T)
    m = C.row < C.col
    assert np.all(lvl[C.row[m]] < lvl[C.col[m]])
    for sweep in ('forward', 'backward', 'symmetric'):
        xo = x0.copy(); orx.gauss_seidel(A, xo, b, iterations=2, sweep=sweep, core=oracle.port)
        xd = dev(x0.copy())
        amg.relaxation.gauss_seidel(dA, xd, dev(b), iterations=2, sweep=sweep)
        assert np.array_equal(xd.cpu().numpy(), xo), sweep


@pytest.mark.parametrize('bs', [3, 6])
def test_lexicographic_block_sweeps_over_wavefronts_bit_exact(amg, bs):
    """bsr_gauss_seidel (relaxation.h:89-173) and block_gauss_seidel (:753-806) in their lexicographic order through the
    wavefront schedule of the block graph"""
    rng = np.random.RandomState(9)
    G = poisson3d(12)                                                     # 1728 block rows, 34 wavefronts
    Bk = rng.rand(bs, bs); Bk = Bk @ Bk.T + bs * np.eye(bs)
    A = sp.kron(G, Bk).tobsr(blocksize=(bs, bs)); A.sort_indices()
    n = A.shape[0]
    x0, b = rng.rand(n), rng.rand(n)
    dA = amg.DevMat.from_scipy(A)
    assert amg.relaxation.wavefront_ordering(dA) is not None
    for sweep in ('forward', 'backward', 'symmetric'):
        xo = x0.copy(); orx.gauss_seidel(A, xo, b, sweep=sweep, core=oracle.port)
        xd = dev(x0.copy()); amg.relaxation.gauss_seidel(dA, xd, dev(b), sweep=sweep)
        assert np.array_equal(xd.cpu().numpy(), xo), sweep
    Dinv = amg.relaxation.get_block_diag_inverse(dA, bs)
    xo = x0.copy()
    orx.block_gauss_seidel(A, xo, b, Dinv=Dinv.cpu().numpy(), blocksize=bs, sweep='symmetric', core=oracle.port)
    xd = dev(x0.copy()); amg.relaxation.block_gauss_seidel(dA, xd, dev(b), blocksize=bs, sweep='symmetric', Dinv=Dinv)
    assert np.array_equal(xd.cpu().numpy(), xo)


def test_reference_golden_vectors_on_device(amg):
    """pyamg/relaxation/tests/test_relaxation.py:139-188, :291-402 through the device path"""
    rx, D = amg.relaxation, amg.DevMat.from_scipy
    A = D(lap1d(3))
    x = dev(np.arange(3, dtype=float)); rx.gauss_seidel(A, x, dev(np.zeros(3)))
    assert np.allclose(x.cpu().numpy(), [1.0 / 2, 5.0 / 4, 5.0 / 8])
    x = dev(np.arange(3, dtype=float)); rx.gauss_seidel(A, x, dev(np.zeros(3)), sweep='backward')
    assert np.allclose(x.cpu().numpy(), [1.0 / 8, 1.0 / 4, 1.0 / 2])
    x = dev(np.arange(3, dtype=float)); rx.gauss_seidel(A, x, dev(np.array([10., 20, 30])))
    assert np.allclose(x.cpu().numpy(), [11.0 / 2, 55.0 / 4, 175.0 / 8])
    x = dev(np.zeros(3)); rx.jacobi(A, x, dev(np.arange(3, dtype=float)))
    assert np.allclose(x.cpu().numpy(), [0.0, 0.5, 1.0])
    x = dev(np.arange(3, dtype=float)); rx.jacobi(A, x, dev(np.array([10., 20, 30])), omega=1.0 / 3.0)
    assert np.allclose(x.cpu().numpy(), 2.0 / 3.0 * np.arange(3) + 1.0 / 3.0 * np.array([5.5, 11.0, 15.5]))
    x = dev(np.arange(3, dtype=float)); rx.gauss_seidel_indexed(A, x, dev(np.zeros(3)), [2, 1, 0], sweep='backward')
    assert np.allclose(x.cpu().numpy(), [1.0 / 2, 5.0 / 4, 5.0 / 8])
    A4 = D(lap1d(4))
    x = dev(np.ones(4)); rx.gauss_seidel_indexed(A4, x, dev(np.zeros(4)), [0, 3])
    assert np.allclose(x.cpu().numpy(), [0.5, 1.0, 1.0, 0.5])
    x = dev(np.ones(4)); rx.gauss_seidel_indexed(A4, x, dev(np.zeros(4)), [0, 0])
    assert np.allclose(x.cpu().numpy(), [0.5, 1.0, 1.0, 1.0])


def test_argument_validation(amg):
    """TestCommonRelaxation, test_relaxation.py:23-102: TypeError for dtype mismatch, ValueError
    for bad sizes / non-contiguous x / non-square A"""
    rx, D = amg.relaxation, amg.DevMat.from_scipy
    A = D(lap1d(6))
    b = dev(np.zeros(6))
    with pytest.raises(TypeError):
        rx.jacobi(A, torch.zeros(6, dtype=torch.float32, device='cuda'), b)
    with pytest.raises(ValueError):
        rx.jacobi(A, dev(np.zeros(5)), b)
    with pytest.raises(ValueError):
        rx.gauss_seidel(A, dev(np.zeros(12))[::2], b)
    with pytest.raises(ValueError):
        rx.jacobi(D(sp.random(6, 5, 0.5, format='csr')), dev(np.zeros(6)), b)
    with pytest.raises(ValueError):
        rx.gauss_seidel(A, dev(np.zeros(6)), b, sweep='sideways')
    with pytest.raises(TypeError):
        amg.amg_core.jacobi(A.indptr.long(), A.indices, A.data, dev(np.zeros(6)), b, dev(np.zeros(6)), 0, 6, 1, [1.0])


def test_dot_norm_dense(amg):
    rng = np.random.RandomState(5)
    for n in (1, 7, 1000, 1 << 20):
        x, y = rng.rand(n), rng.rand(n)
        assert abs(amg.linalg.dot(dev(x), dev(y)) - np.dot(x, y)) <= 1e-13 * np.dot(x, y)   # reduction order differs
    M = rng.rand(37, 37); v = rng.rand(37)
    out = torch.empty(37, dtype=torch.float64, device='cuda')
    dM, dv = dev(M), dev(v)                       # keep the device copies alive across the call
    amg._lib.call('b200amg_dense_matvec', 37, dM.data_ptr(), dv.data_ptr(), out.data_ptr(), 0)
    torch.cuda.synchronize()
    assert np.allclose(out.cpu().numpy(), M @ v, rtol=1e-13)


# ---- schedules of the CSR stream family: persistent TMA pipeline (short rows), sliced-ELL (engine
#      layout for long rows), chunked stream kernel (everything else) -- same bits everywhere -------
def lap2d(m):
    I = sp.eye(m); T = lap1d(m)
    return sp.csr_matrix(sp.kron(T, I) + sp.kron(I, T))


def ragged_matrix(rng, n=20000):
    """tridiagonal + three rows with 5000 entries (tiles larger than a pipeline stage) + a band of
    300 rows with ~40 entries (entry-parallel tiles inside a short-row matrix) + empty rows"""
    A = sp.lil_matrix((n, n))
    A.setdiag(rng.rand(n) + 2.0)
    A.setdiag(-rng.rand(n - 1), 1)
    A.setdiag(-rng.rand(n - 1), -1)
    for r in (17, 9000, n - 1):
        cols = rng.choice(n, 5000, replace=False)
        A[r, cols] = rng.rand(5000)
    for r in range(4000, 4300):
        cols = rng.choice(n, 40, replace=False)
        A[r, cols] = rng.rand(40)
    A = sp.csr_matrix(A)
    A.sort_indices()
    keep = np.ones(n, bool); keep[[100, 101, 5000, n - 2]] = False          # empty rows
    D = sp.diags(keep.astype(float))
    A = sp.csr_matrix(D @ A); A.eliminate_zeros(); A.sort_indices()
    return A


PIPE_CASES = [('lap2d_800_ring_wrap', lambda rng: lap2d(800)),          # 2500 tiles: >5 per persistent CTA
              ('lap2d_333_tail_tile', lambda rng: lap2d(333)),          # last tile partial, n % 4 != 0
              ('ragged', ragged_matrix)]


@pytest.mark.parametrize('name,make', PIPE_CASES)
def test_pipeline_schedule_bit_exact(amg, name, make):
    rng = np.random.RandomState(7)
    A = make(rng)
    n = A.shape[0]
    x, b = rng.rand(n), rng.rand(n)
    dA = amg.DevMat.from_scipy(A)
    assert np.array_equal(dA.matvec(dev(x)).cpu().numpy(), A * x)
    assert np.array_equal(dA.residual(dev(x), dev(b)).cpu().numpy(), b - A * x)
    for it in range(2):                                                  # second call: cached nnz, warm rings
        xo = x.copy(); orx.jacobi(A, xo, b, iterations=1, omega=0.8, core=oracle.port)
        xd = dev(x)
        amg.relaxation.jacobi(dA, xd, dev(b), iterations=1, omega=0.8)
        assert np.array_equal(xd.cpu().numpy(), xo)


def few_valued_irregular(rng):
    """short rows, 6 distinct values, irregular columns: the value dictionary applies, the pattern one does not"""
    n = 70000
    rows = np.repeat(np.arange(n), 4)
    cols = (rows + rng.randint(-40, 40, rows.size)) % n
    vals = rng.choice(np.array([-1.0, 0.5, 2.0, -0.0, 1e-300, 3.25]), rows.size)
    A = sp.coo_matrix((vals, (rows, cols)), shape=(n, n)).tocsr()        # duplicates summed: a few more values
    A = sp.csr_matrix(A + sp.diags(np.full(n, 4.0))); A.sort_indices()
    return A


def stencil_with_oddities(rng):
    """5-point stencil with an empty row, a row without a diagonal and a stored zero on a diagonal"""
    A = lap2d(400)
    A.sort_indices()
    A.data[A.indptr[5000]:A.indptr[5001]] = 0.0                          # becomes an empty row
    k = A.indptr[123] + int(np.where(A.indices[A.indptr[123]:A.indptr[124]] == 123)[0][0])
    A.data[k] = 0.0                                                      # row 123 loses its diagonal entry
    A.eliminate_zeros()
    k = A.indptr[17] + int(np.where(A.indices[A.indptr[17]:A.indptr[18]] == 17)[0][0])
    A.data[k] = 0.0                                                      # explicit zero stays stored
    return A


def stencil_many_row_patterns(rng):
    """5-point stencil whose diagonal takes 400 different values: few (offset, value) pairs per offset would
    still fit... but not 255 row patterns -> the row-pattern layout declines, so does the pattern dictionary"""
    A = lap2d(300)
    A.sort_indices()
    A = sp.csr_matrix(A + sp.diags((np.arange(A.shape[0]) % 400).astype(float)))
    A.sort_indices()
    return A


def poisson3d_box(nx, ny, nz):
    """7-point Laplacian on an nx x ny x nz box, x fastest: column offsets +-1, +-nx, +-nx*ny"""
    Tx, Ty, Tz = lap1d(nx), lap1d(ny), lap1d(nz)
    Ix, Iy, Iz = sp.eye(nx), sp.eye(ny), sp.eye(nz)
    A = sp.kron(sp.kron(Iz, Iy), Tx) + sp.kron(sp.kron(Iz, Ty), Ix) + sp.kron(sp.kron(Tz, Iy), Ix)
    A = sp.csr_matrix(A)
    A.sort_indices()
    return A


def lap2d_periodic(m):
    """5-point stencil on a torus: the wrap-around rows have column offsets no interior row has, so in the row-pattern
    layout they are NOT the common pattern with entries missing (per-row global path of the TMA-staged sweep)"""
    S = sp.diags([np.ones(m - 1), np.ones(m - 1), [1.0], [1.0]], [-1, 1, m - 1, -(m - 1)], shape=(m, m), format='csr')
    I = sp.eye(m)
    A = sp.csr_matrix(4.0 * sp.eye(m * m) - sp.kron(S, I) - sp.kron(I, S))
    A.sort_indices()
    return A


# lap2d_301 / poisson3d_41 / lap1d: odd offsets, an odd number of rows (the last element of a staged segment travels
# on its own), 5-, 7- and 3-entry common patterns of the TMA-staged row-pattern sweep (csrc/rowpat_tile.cu)
DICT_CASES = [('lap2d_800', lambda rng: lap2d(800), {'0': 1, '1': 3, '2': 4, '3': 5}),
              ('lap2d_301', lambda rng: lap2d(301), {'0': 1, '1': 3, '2': 4, '3': 5}),
              ('poisson3d_41', lambda rng: poisson3d(41), {'0': 1, '1': 3, '2': 4, '3': 5}),
              ('lap1d_100001', lambda rng: lap1d(100001), {'0': 1, '1': 3, '2': 4, '3': 5}),
              ('lap2d_periodic_300', lambda rng: lap2d_periodic(300), {'0': 1, '1': 3, '2': 4, '3': 5}),
              # grid lines of 600 points: the merged middle segment is wide, the stage ring drops to 2 stages (the 512^3
              # problem of 8 GPUs); lines of 1100 points: five separate segments
              ('poisson3d_600x5x5', lambda rng: poisson3d_box(600, 5, 5), {'0': 1, '1': 3, '2': 4, '3': 5}),
              ('poisson3d_1100x5x4', lambda rng: poisson3d_box(1100, 5, 4), {'0': 1, '1': 3, '2': 4, '3': 5}),
              ('few_valued_irregular', few_valued_irregular, {'0': 1, '1': 3, '2': 3, '3': 3}),
              ('stencil_with_oddities', stencil_with_oddities, {'0': 1, '1': 3, '2': 4, '3': 5}),
              ('stencil_many_row_patterns', stencil_many_row_patterns, {'0': 1, '1': 1, '2': 1, '3': 1}),
              ('ragged_many_values', ragged_matrix, {'0': 1, '1': 1, '2': 1, '3': 1})]


@pytest.mark.parametrize('level', ['0', '1', '2', '3'])
@pytest.mark.parametrize('name,make,layouts', DICT_CASES)
def test_dictionary_layouts_bit_exact(amg, monkeypatch, name, make, layouts, level):
    """B200AMG_VALUE_DICT = 0 / 1 / 2 / 3: as stored, 8-bit value codes, 8-bit (offset, value) pattern codes,
    one row-pattern byte per row -- the operator handle must pick the expected layout (falling back when the
    operator does not qualify) and give the same bits in every one"""
    import ctypes
    monkeypatch.setenv('B200AMG_VALUE_DICT', level)
    rng = np.random.RandomState(5)
    A = make(rng)
    n = A.shape[0]
    x, b = rng.rand(n), rng.rand(n)
    dA = amg.DevMat.from_scipy(A)
    P = lambda t: t.data_ptr()
    h = ctypes.c_void_p()
    amg._lib.call('b200amg_op_create', n, P(dA.indptr), P(dA.indices), P(dA.data), 0, ctypes.byref(h), 0)
    try:
        assert amg._lib.load().b200amg_op_layout(h) == layouts[level]

        def run(op, y0):
            y = dev(y0)
            dx, db = dev(x), dev(b)
            amg._lib.call('b200amg_op_apply', h, op, P(dx), P(db), P(y), 0.7, 0)
            torch.cuda.synchronize()
            return y.cpu().numpy()
        assert np.array_equal(run(0, np.zeros(n)), A * x)
        assert np.array_equal(run(2, np.zeros(n)), b - A * x)
        assert np.array_equal(run(3, b), b + A * x)
        xo = x.copy(); orx.jacobi(A, xo, b, iterations=1, omega=0.7, core=oracle.port)
        assert np.array_equal(run(4, np.zeros(n)), xo)
        Ab = sp.bsr_matrix(A, blocksize=(1, 1))
        xo = x.copy(); orx.jacobi(Ab, xo, b, iterations=1, omega=0.7, core=oracle.port)
        assert np.array_equal(run(5, np.zeros(n)), xo)
    finally:
        amg._lib.call('b200amg_op_destroy', h)


@pytest.mark.parametrize('name,n,per_row,few_values', [('coarse_like', 40000, 31, False), ('restriction_like', 9000, 240, False),
                                                        ('small_pipe_form', 3000, 60, False),
                                                        ('coded_values', 40000, 30, True)])
def test_sliced_ell_layout_bit_exact(amg, name, n, per_row, few_values):
    """the engine's SELL-32 copy must give the bits of the CSR kernels / the reference loops"""
    rng = np.random.RandomState(11)
    rows = np.repeat(np.arange(n), per_row)
    cols = (rows + rng.randint(-3 * per_row, 3 * per_row, rows.size)) % n
    if few_values:      # restriction-like: a handful of distinct values -> 8-bit codes in the value slots
        vals = rng.choice(np.array([0.25, -0.125, 1.0, 3.5, -0.0, 1e-310]), rows.size)
        A = sp.coo_matrix((vals, (rows, cols)), shape=(n, n)).tocsr()
        A = sp.csr_matrix(A + sp.diags(np.full(n, 2.0))); A.sort_indices()
    else:
        A = sp.csr_matrix((rng.rand(rows.size) - 0.3, (rows, cols)), shape=(n, n))  # duplicates summed, ragged
        A = sp.csr_matrix(A + sp.diags(rng.rand(n) + 1.0)); A.sort_indices()
    x, b = rng.rand(n), rng.rand(n)
    dA = amg.DevMat.from_scipy(A)
    P = lambda t: t.data_ptr()

    import ctypes
    h = ctypes.c_void_p()
    amg._lib.call('b200amg_op_create', n, P(dA.indptr), P(dA.indices), P(dA.data), 2, ctypes.byref(h), 0)
    assert amg._lib.load().b200amg_op_layout(h) == (6 if few_values else 2)

    def run(op, y0):
        y = dev(y0)
        dx, db = dev(x), dev(b)
        amg._lib.call('b200amg_op_apply', h, op, P(dx), P(db), P(y), 0.7, 0)
        torch.cuda.synchronize()
        return y.cpu().numpy()
    assert np.array_equal(run(0, np.zeros(n)), A * x)
    assert np.array_equal(run(2, np.zeros(n)), b - A * x)
    assert np.array_equal(run(3, b), b + A * x)
    xo = x.copy(); orx.jacobi(A, xo, b, iterations=1, omega=0.7, core=oracle.port)
    assert np.array_equal(run(4, np.zeros(n)), xo)
    Ab = sp.bsr_matrix(A, blocksize=(1, 1))
    xo = x.copy(); orx.jacobi(Ab, xo, b, iterations=1, omega=0.7, core=oracle.port)
    assert np.array_equal(run(5, np.zeros(n)), xo)
    amg._lib.load().b200amg_op_destroy(h)


@pytest.mark.parametrize('name,n,per_row,span,sort,expect', [('steps_fit', 40000, 31, 90, True, 7), ('long_rows', 9000, 240, 700, True, 7),
                                                             ('some_wide_steps', 300000, 40, 140000, True, 7),
                                                             ('mostly_wide_steps', 300000, 4, 150000, True, 2),
                                                             ('unsorted_rows', 40000, 31, 90, False, 2)])
def test_sliced_ell_16bit_column_steps(amg, monkeypatch, name, n, per_row, span, sort, expect):
    """layout 7: columns stored as 16-bit steps from the previous stored column (10 bytes per slot).  Needs rows
    stored by ascending column; a step of 65 535 columns or more takes escape slots ("65 535 on, no entry"), and when
    those would exceed 1/32 of the entries the copy keeps 32-bit columns.  Same bits either way."""
    monkeypatch.setenv('B200AMG_SELL_COL16', '2')           # the engine uses it from 400 000 rows; here for every size
    rng = np.random.RandomState(12)
    rows = np.repeat(np.arange(n), per_row)
    cols = (rows + rng.randint(-span, span, rows.size)) % n
    A = sp.csr_matrix((rng.rand(rows.size) - 0.3, (rows, cols)), shape=(n, n))      # duplicates summed: ragged rows
    A = sp.csr_matrix(A + sp.diags(rng.rand(n) + 1.0)); A.sort_indices()
    if name == 'steps_fit':
        A = sp.csr_matrix(A)
        A.data[A.indptr[7]:A.indptr[8]] = 0.0; A.eliminate_zeros()                  # an empty row
    if not sort:            # reverse the storage order of every 5th row
        for i in range(0, n, 5):
            lo, hi = A.indptr[i], A.indptr[i + 1]
            A.indices[lo:hi] = A.indices[lo:hi][::-1].copy(); A.data[lo:hi] = A.data[lo:hi][::-1].copy()
        A.has_sorted_indices = False
    if name == 'some_wide_steps':       # rows of 40 columns within +-140 000: a few steps per hundred rows need an escape
        st = np.diff(A.indices.astype(np.int64))
        assert 0 < np.count_nonzero(st >= 65535) * 32 < A.nnz
    x, b = rng.rand(n), rng.rand(n)
    dA = amg.DevMat.from_scipy(A)                           # storage order kept as given
    P = lambda t: t.data_ptr()
    import ctypes
    h = ctypes.c_void_p()
    amg._lib.call('b200amg_op_create', n, P(dA.indptr), P(dA.indices), P(dA.data), 2, ctypes.byref(h), 0)
    try:
        assert amg._lib.load().b200amg_op_layout(h) == expect

        def run(op, y0):
            y = dev(y0)
            dx, db = dev(x), dev(b)
            amg._lib.call('b200amg_op_apply', h, op, P(dx), P(db), P(y), 0.7, 0)
            torch.cuda.synchronize()
            return y.cpu().numpy()
        core = oracle.port
        Ax = np.zeros(n); core.csr_matvec(n, n, A.indptr, A.indices, A.data, x, Ax)
        assert np.array_equal(run(0, np.zeros(n)), Ax)
        assert np.array_equal(run(2, np.zeros(n)), b - Ax)
        assert np.array_equal(run(3, b), b + Ax)
        xo = x.copy(); tmp = np.empty(n)
        core.jacobi(A.indptr, A.indices, A.data, xo, b, tmp, 0, n, 1, np.array([0.7]))
        assert np.array_equal(run(4, np.zeros(n)), xo)
    finally:
        amg._lib.call('b200amg_op_destroy', h)
