"""Smoother front-ends with the reference's names, arguments and error behaviour
(pyamg/relaxation/relaxation.py: make_system :21-105, gauss_seidel :280-354, jacobi :357-427,
gauss_seidel_indexed :671-741), operating on device data: `A` is a pyamg_b200.sparse.DevMat,
`x` and `b` are float64 torch CUDA tensors; `x` is updated in place, nothing is returned.

Multicolour Gauss-Seidel = vertex_coloring(A) + color_ordering + gauss_seidel_indexed
with `color_ptr`; without `color_ptr` the indexed sweep runs strictly sequentially on the GPU.
"""
import numpy as np
import torch

from . import _lib, amg_core
from .sparse import DevMat


def make_system(A, x, b, formats=None):
    """relaxation.py:21-105 on device containers."""
    if not isinstance(A, DevMat):
        raise TypeError('expected a pyamg_b200 device matrix (DevMat.from_scipy(A))')
    if formats == ['csr'] and A.format == 'bsr':
        if not A.is_scalar:
            raise NotImplementedError('BSR -> CSR expansion on the device is only done for 1x1 blocks')
        A = A.tocsr_scalar()
    if not isinstance(x, torch.Tensor):
        raise ValueError('expected a torch tensor for argument x')
    if not isinstance(b, torch.Tensor):
        raise ValueError('expected a torch tensor for argument b')
    M, N = A.shape
    if M != N:
        raise ValueError('expected square matrix')
    if tuple(x.shape) not in [(M,), (M, 1)]:
        raise ValueError('x has invalid dimensions')
    if tuple(b.shape) not in [(M,), (M, 1)]:
        raise ValueError('b has invalid dimensions')
    if x.dtype != torch.float64 or b.dtype != torch.float64:
        raise TypeError('arguments A, x, and b must have the same dtype')
    if not x.is_contiguous():
        raise ValueError('x must be contiguous in memory')
    return A, x.view(-1), b.reshape(-1).contiguous()


def _range(sweep, n):
    if sweep == 'forward':
        return 0, n, 1
    if sweep == 'backward':
        return n - 1, -1, -1
    raise ValueError("valid sweep directions are 'forward', 'backward', and 'symmetric'")


def jacobi(A, x, b, iterations=1, omega=1.0):
    A, x, b = make_system(A, x, b, formats=['csr', 'bsr'])
    n = A.shape[0]
    if n <= 0:
        return
    temp = torch.empty_like(x)
    for _ in range(iterations):
        if A.format == 'csr':
            amg_core.jacobi(A.indptr, A.indices, A.data, x, b, temp, 0, n, 1, [omega])
        else:
            R, C = A.blocksize
            if R != C:
                raise ValueError('BSR blocks must be square')
            amg_core.bsr_jacobi(A.indptr, A.indices, A.data, x, b, temp, 0, n // R, 1, R, [omega])


WAVEFRONT_MIN_ROWS = 1024        # below this a lexicographic sweep stays on one thread block
WAVEFRONT_MIN_WIDTH = 32         # average rows per dependency level for the level-parallel form to pay


def wavefront_ordering(A):
    """Rows (block rows) of the LEXICOGRAPHIC Gauss-Seidel sweep grouped by dependency level (amg_core.gs_wavefronts),
    cached on the matrix: -> (indices int32 device tensor, level_ptr host int32) or None when the sweep is too narrow
    (small or densely coupled matrices: one level per few rows) to beat the sequential kernel.  Feeding the pair to the
    indexed kernels as a colour ordering runs the reference's sweep order, bit for bit, one launch per level."""
    cached = getattr(A, '_gs_wavefront', False)
    if cached is not False:
        return cached
    n = A.n_brow
    out = None
    if n >= WAVEFRONT_MIN_ROWS:
        level = np.empty(n, dtype=np.intc)
        nlev = amg_core.gs_wavefronts(n, A.indptr.cpu().numpy(), A.indices.cpu().numpy(), level)
        if n >= WAVEFRONT_MIN_WIDTH * nlev:
            idx, ptr = color_ordering(level)
            out = (torch.from_numpy(idx).to(A.data.device), ptr)
    A._gs_wavefront = out
    return out


def gauss_seidel(A, x, b, iterations=1, sweep='forward', indices=None, color_ptr=None):
    """Lexicographic sweep unless a colour ordering (`indices`, `color_ptr` over rows / block rows) is given.  The
    lexicographic sweep runs level-parallel over its dependency wavefronts when the matrix is large and sparse enough
    (wavefront_ordering; same bits), else sequentially on one thread block."""
    A, x, b = make_system(A, x, b, formats=['csr', 'bsr'])
    if A.format == 'csr':
        bs = 1
    else:
        R, C = A.blocksize
        if R != C:
            raise ValueError('BSR blocks must be square')
        bs = R
    if indices is None and color_ptr is None:
        wf = wavefront_ordering(A)
        if wf is not None:
            indices, color_ptr = wf
    if sweep == 'symmetric':
        for _ in range(iterations):
            gauss_seidel(A, x, b, 1, 'forward', indices, color_ptr)
            gauss_seidel(A, x, b, 1, 'backward', indices, color_ptr)
        return
    nrows = len(x) // bs if indices is None else int(indices.numel())
    r0, r1, dr = _range(sweep, nrows)
    for _ in range(iterations):
        if A.format == 'csr' and indices is None:
            amg_core.gauss_seidel(A.indptr, A.indices, A.data, x, b, r0, r1, dr)
        elif A.format == 'csr':
            amg_core.gauss_seidel_indexed(A.indptr, A.indices, A.data, x, b, indices, r0, r1, dr, color_ptr)
        else:
            amg_core.bsr_gauss_seidel(A.indptr, A.indices, A.data, x, b, r0, r1, dr, bs, indices, color_ptr)


def gauss_seidel_indexed(A, x, b, indices, iterations=1, sweep='forward', color_ptr=None):
    A, x, b = make_system(A, x, b, formats=['csr'])
    if not isinstance(indices, torch.Tensor):
        indices = torch.from_numpy(np.ascontiguousarray(indices, dtype='intc')).to(x.device)
    if sweep == 'symmetric':
        for _ in range(iterations):
            gauss_seidel_indexed(A, x, b, indices, 1, 'forward', color_ptr)
            gauss_seidel_indexed(A, x, b, indices, 1, 'backward', color_ptr)
        return
    r0, r1, dr = _range(sweep, int(indices.numel()))
    for _ in range(iterations):
        amg_core.gauss_seidel_indexed(A.indptr, A.indices, A.data, x, b, indices, r0, r1, dr, color_ptr)


def get_block_diag_inverse(A, blocksize=None):
    """inverses of the diagonal blocks, [n_brow, bs, bs] on the device (pyamg/util/utils.py:601-690 with
    inv_flag=True; cached on the matrix like the reference's A.block_D_inv).  Blocks up to 6 x 6 go through the device
    restatement of amg_core.pinv_array like in the reference (its bits, pseudo-inverse of singular blocks included);
    7 x 7 and 8 x 8 -- LAPACK gelss in the reference -- through Gauss-Jordan (equal to rounding)."""
    bs = A.blocksize[0] if blocksize is None else int(blocksize)
    if A.blocksize != (bs, bs):
        raise NotImplementedError('re-blocking on the device is not available: A must be BSR with the requested blocksize')
    cached = getattr(A, 'block_D_inv', None)
    if cached is not None and tuple(cached.shape) == (A.n_brow, bs, bs):
        return cached
    Dinv = torch.empty((A.n_brow, bs, bs), dtype=torch.float64, device=A.data.device)
    _lib.call('b200amg_block_diag_inverse', A.n_brow, bs, A.indptr.data_ptr(), A.indices.data_ptr(), A.data.data_ptr(),
              Dinv.data_ptr(), torch.cuda.current_stream().cuda_stream)
    A.block_D_inv = Dinv
    return Dinv


def _block_args(A, x, b, blocksize, Dinv):
    A, x, b = make_system(A, x, b, formats=['csr', 'bsr'])
    blocksize = int(blocksize)
    if A.blocksize != (blocksize, blocksize):
        raise NotImplementedError('re-blocking on the device is not available: A must be BSR with the requested blocksize')
    if Dinv is None:
        Dinv = get_block_diag_inverse(A, blocksize)
    elif Dinv.shape[0] != A.shape[0] // blocksize:
        raise ValueError('Dinv and A have incompatible dimensions')
    elif Dinv.shape[1] != blocksize or Dinv.shape[2] != blocksize:
        raise ValueError('Dinv and blocksize are incompatible')
    return A, x, b, blocksize, Dinv.contiguous().view(-1)


def block_jacobi(A, x, b, Dinv=None, blocksize=1, iterations=1, omega=1.0):
    """pyamg/relaxation/relaxation.py:430-492"""
    A, x, b, bs, Tx = _block_args(A, x, b, blocksize, Dinv)
    nb = A.shape[0] // bs
    if nb <= 0:
        return
    temp = torch.empty_like(x)
    for _ in range(iterations):
        amg_core.block_jacobi(A.indptr, A.indices, A.data, x, b, Tx, temp, 0, nb, 1, [omega], bs)


def block_gauss_seidel(A, x, b, iterations=1, sweep='forward', blocksize=1, Dinv=None, indices=None, color_ptr=None):
    """pyamg/relaxation/relaxation.py:495-592.  Without `indices`: the reference's lexicographic sweep (sequential on the
    device).  With `indices` (block rows grouped by colour) and `color_ptr`: the colours one after the other, the rows
    of a colour in parallel."""
    A, x, b, bs, Tx = _block_args(A, x, b, blocksize, Dinv)
    if indices is not None and not isinstance(indices, torch.Tensor):
        indices = torch.from_numpy(np.ascontiguousarray(indices, dtype='intc')).to(x.device)
    if indices is None and color_ptr is None:
        wf = wavefront_ordering(A)                      # the lexicographic sweep, level-parallel (same bits)
        if wf is not None:
            indices, color_ptr = wf
    if sweep == 'symmetric':
        for _ in range(iterations):
            block_gauss_seidel(A, x, b, 1, 'forward', bs, Dinv, indices, color_ptr)
            block_gauss_seidel(A, x, b, 1, 'backward', bs, Dinv, indices, color_ptr)
        return
    r0, r1, dr = _range(sweep, A.shape[0] // bs if indices is None else int(indices.numel()))
    for _ in range(iterations):
        amg_core.block_gauss_seidel(A.indptr, A.indices, A.data, x, b, Tx, r0, r1, dr, bs, indices, color_ptr)


def polynomial(A, x, b, coefficients, iterations=1):
    """pyamg/relaxation/relaxation.py:593-668: x += p(A)(b - A x) with p given by its coefficients (highest power
    first), Horner form: h = c0 r; h = c_k r + A h.  The reference's "norm(x) == 0 -> residual = b" shortcut gives
    the bits of b - A*0, so the residual is always formed."""
    A, x, b = make_system(A, x, b, formats=None)
    n = A.shape[0]
    st = torch.cuda.current_stream().cuda_stream
    h, t = torch.empty_like(x), torch.empty_like(x)
    for _ in range(iterations):
        r = A.residual(x, b)
        _lib.call('b200amg_scal', n, float(coefficients[0]), r.data_ptr(), h.data_ptr(), st)
        for c in coefficients[1:]:
            _lib.call('b200amg_scal', n, float(c), r.data_ptr(), t.data_ptr(), st)
            R, C = A.blocksize
            _lib.call('b200amg_bsr_matvec', A.n_brow, A.n_bcol, R, C, A.indptr.data_ptr(), A.indices.data_ptr(),
                      A.data.data_ptr(), h.data_ptr(), 0, t.data_ptr(), 2, st)              # t += A h
            h, t = t, h
        _lib.call('b200amg_axpy', n, 1.0, h.data_ptr(), x.data_ptr(), st)


def vertex_coloring(G, method='MIS'):
    """pyamg/graph.py:84-125: 'MIS' (graph.h:171-188), 'JP' (Jones-Plassmann, :242-287), 'LDF' (largest degree first,
    :290-335); host, setup-time.  'JP' and 'LDF' draw one random weight per vertex from the global numpy RNG like the
    reference's sp.rand(N).  G: DevMat or scipy matrix (pattern of block rows for BSR)."""
    if method not in ('MIS', 'JP', 'LDF'):
        raise ValueError('unknown method (%s)' % method)
    if isinstance(G, DevMat):
        Ap, Aj = G.indptr.cpu().numpy(), G.indices.cpu().numpy()
        n = G.n_brow
    else:
        Ap, Aj = np.asarray(G.indptr, np.intc), np.asarray(G.indices, np.intc)
        n = len(Ap) - 1
    colors = np.empty(n, dtype=np.intc)
    if method == 'MIS':
        amg_core.vertex_coloring_mis(n, Ap, Aj, colors)
    elif method == 'JP':
        amg_core.vertex_coloring_jones_plassmann(n, Ap, Aj, colors, np.random.rand(n))
    else:
        amg_core.vertex_coloring_LDF(n, Ap, Aj, colors, np.random.rand(n))
    return colors


def color_ordering(colors):
    """Rows grouped by colour, ascending inside each colour -> (indices int32[n], color_ptr int32[K+1])."""
    colors = np.asarray(colors)
    K = int(colors.max()) + 1 if len(colors) else 0
    key = colors.astype(np.uint16) if 0 < K <= 65536 else colors      # 16-bit keys: numpy's stable sort is a radix sort
    idx = np.argsort(key, kind='stable').astype(np.intc)
    ptr = np.zeros(K + 1, dtype=np.intc)
    np.cumsum(np.bincount(colors, minlength=K), out=ptr[1:])
    return idx, ptr
