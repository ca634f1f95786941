"""Device-resident CSR/BSR matrices and the sparse products of the hot path.

`DevMat` mirrors the few scipy.sparse members the reference's cycle touches (`shape`, `nnz`,
`format`, `blocksize`, `indptr/indices/data`, `A * x`), with the arrays living in HBM as torch
tensors (int32 indices, fp64 values).  All arithmetic is done by libb200amg kernels:
  A * x          -> b200amg_csr_matvec / b200amg_bsr_matvec   (scipy csr_matvec / bsr_matvec)
  matmat(A, B)   -> b200amg_spgemm_*                          (scipy csr_matmat, pyamg aggregation.py:429)
  transpose(A)   -> b200amg_csr_transpose                     (scipy csr_tocsc,  pyamg classical.py:174)
Block operands (BSR with R x C blocks, BASELINE config 4) take the same two calls: the block PATTERN goes through
the scalar kernels, the blocks follow through b200amg_bsr_slot_map / _bsr_matmat_numeric / _bsr_gather_blocks
(scipy bsr_matmat / bsr_transpose).
"""
import numpy as np
import torch

from . import _lib


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _dev():
    if not torch.cuda.is_available():
        raise _lib.B200AmgError('pyamg_b200 needs a CUDA device (no CPU fallback)')
    return torch.device('cuda', torch.cuda.current_device())


def _ptr(t):
    return 0 if t is None else t.data_ptr()


def as_dev_vector(v, dtype=torch.float64):
    """numpy / torch -> contiguous 1-D device tensor (copy only when needed)."""
    if isinstance(v, torch.Tensor):
        t = v
        if t.device.type != 'cuda':
            t = t.to(_dev())
        if t.dtype != dtype:
            raise TypeError('expected dtype %s, got %s' % (dtype, t.dtype))
        return t.reshape(-1) if t.is_contiguous() else t.contiguous().reshape(-1)
    a = np.ascontiguousarray(np.ravel(v))
    return torch.from_numpy(a).to(_dev()).to(dtype)


class DevMat:
    """CSR (blocksize (1,1), format 'csr') or BSR (format 'bsr', R x C row-major blocks) in HBM."""

    def __init__(self, indptr, indices, data, shape, blocksize=(1, 1), fmt=None, sorted_indices=None):
        self.indptr, self.indices, self.data = indptr, indices, data
        self.shape = (int(shape[0]), int(shape[1]))
        self.blocksize = (int(blocksize[0]), int(blocksize[1]))
        self.format = fmt or ('csr' if self.blocksize == (1, 1) else 'bsr')
        self.dtype = np.dtype('float64')
        self.has_sorted_indices = sorted_indices
        assert indptr.dtype == torch.int32 and indices.dtype == torch.int32 and data.dtype == torch.float64
        assert indptr.is_contiguous() and indices.is_contiguous() and data.is_contiguous()

    # -- scipy-like surface ------------------------------------------------------------------
    @property
    def nnz(self):
        return int(self.data.numel())

    @property
    def nblocks(self):
        return int(self.indices.numel())

    @property
    def n_brow(self):
        return self.shape[0] // self.blocksize[0]

    @property
    def n_bcol(self):
        return self.shape[1] // self.blocksize[1]

    @property
    def is_scalar(self):
        return self.blocksize == (1, 1)

    @classmethod
    def from_scipy(cls, M):
        import scipy.sparse as sp
        if not sp.issparse(M):
            M = sp.csr_matrix(M)
        if M.format not in ('csr', 'bsr'):
            M = M.tocsr()
        if M.dtype != np.float64:
            M = M.astype(np.float64)
        dev = _dev()
        bs = M.blocksize if M.format == 'bsr' else (1, 1)
        return cls(torch.from_numpy(np.ascontiguousarray(M.indptr, np.int32)).to(dev),
                   torch.from_numpy(np.ascontiguousarray(M.indices, np.int32)).to(dev),
                   torch.from_numpy(np.ascontiguousarray(M.data, np.float64).ravel()).to(dev),
                   M.shape, bs, M.format, bool(M.has_sorted_indices))

    def to_scipy(self):
        import scipy.sparse as sp
        ip, ix, dx = self.indptr.cpu().numpy(), self.indices.cpu().numpy(), self.data.cpu().numpy()
        if self.format == 'bsr':
            R, C = self.blocksize
            return sp.bsr_matrix((dx.reshape(-1, R, C), ix, ip), shape=self.shape, blocksize=(R, C))
        return sp.csr_matrix((dx, ix, ip), shape=self.shape)

    def tocsr_scalar(self):
        """Same arrays viewed as CSR when the blocks are 1x1 (scipy does the same, bsr_matvec)."""
        if not self.is_scalar:
            raise ValueError('only 1x1 blocks can be viewed as CSR')
        return DevMat(self.indptr, self.indices, self.data, self.shape, (1, 1), 'csr', self.has_sorted_indices)

    # -- products ------------------------------------------------------------------------------
    def matvec(self, x, out=None):
        x = as_dev_vector(x)
        if x.numel() != self.shape[1]:
            raise ValueError('dimension mismatch')
        y = torch.empty(self.shape[0], dtype=torch.float64, device=x.device) if out is None else out
        R, C = self.blocksize
        _lib.call('b200amg_bsr_matvec', self.n_brow, self.n_bcol, R, C, _ptr(self.indptr), _ptr(self.indices),
                  _ptr(self.data), _ptr(x), 0, _ptr(y), 0, _stream())
        return y

    def __mul__(self, other):
        if isinstance(other, DevMat):
            return matmat(self, other)
        return self.matvec(other)

    __matmul__ = __mul__

    def residual(self, x, b, out=None):
        """b - A x in one pass (pyamg/multilevel.py:496)."""
        x, b = as_dev_vector(x), as_dev_vector(b)
        r = torch.empty_like(b) if out is None else out
        R, C = self.blocksize
        _lib.call('b200amg_bsr_matvec', self.n_brow, self.n_bcol, R, C, _ptr(self.indptr), _ptr(self.indices),
                  _ptr(self.data), _ptr(x), _ptr(b), _ptr(r), 1, _stream())
        return r

    def diagonal(self):
        if self.blocksize[0] != self.blocksize[1]:
            raise ValueError('square blocks required')
        d = torch.empty(self.shape[0], dtype=torch.float64, device=self.data.device)
        _lib.call('b200amg_extract_diag', self.n_brow, self.blocksize[0], _ptr(self.indptr), _ptr(self.indices),
                  _ptr(self.data), _ptr(d), _stream())
        return d

    @property
    def T(self):
        return transpose(self)


def _scan_i32(v):
    n = v.numel()
    out = torch.empty(n + 1, dtype=torch.int32, device=v.device)
    scr = torch.empty(_lib.load().b200amg_scan_scratch_elems(n), dtype=torch.int32, device=v.device)
    _lib.call('b200amg_exclusive_scan_i32', n, _ptr(v), _ptr(out), _ptr(scr), _stream())
    return out


def _scan_i64(v):
    n = v.numel()
    out = torch.empty(n + 1, dtype=torch.int64, device=v.device)
    scr = torch.empty(_lib.load().b200amg_scan_scratch_elems(n), dtype=torch.int64, device=v.device)
    _lib.call('b200amg_exclusive_scan_i64', n, _ptr(v), _ptr(out), _ptr(scr), _stream())
    return out


MATMAT_BATCH_PRODUCTS = 400_000_000      # scratch bound of one SpGEMM pass: 12 bytes per product -> 4.8 GB


def matmat(A, B, scipy_order=False, batch_products=None):
    """C = A B for scalar CSR/BSR(1,1) operands: exact zeros dropped, values bit-identical to
    scipy's csr_matmat (see csrc/spgemm.cu).  Rows sorted by column, or (scipy_order=True) in
    scipy's own reverse-first-touch order.  Rows are independent, so large products run in row
    batches whose expand-sort-compress scratch stays below `batch_products` entries (same result)."""
    if not (A.is_scalar and B.is_scalar):
        return bsr_matmat(A, B)
    if A.shape[1] != B.shape[0]:
        raise ValueError('dimension mismatch')
    n = A.shape[0]
    dev = A.data.device
    st = _stream()
    limit = int(batch_products or MATMAT_BATCH_PRODUCTS)
    ub = torch.empty(max(n, 1), dtype=torch.int64, device=dev)
    _lib.call('b200amg_spgemm_count', n, _ptr(A.indptr), _ptr(A.indices), _ptr(B.indptr), _ptr(ub), st)
    ub = ub[:n]
    ubp = _scan_i64(ub)
    total = int(ubp[-1].item())
    cnt = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    # row batches [r0, r1) with at most `limit` products each (a single row may exceed it)
    cuts = [0]
    if total > limit:
        targets = torch.arange(limit, total, limit, dtype=torch.int64, device=dev)
        for c in torch.searchsorted(ubp, targets, right=True).tolist():
            c = min(max(int(c) - 1, cuts[-1] + 1), n)
            if c > cuts[-1]:
                cuts.append(c)
    if cuts[-1] != n:
        cuts.append(n)
    pieces = []
    for r0, r1 in zip(cuts[:-1], cuts[1:]):
        m = r1 - r0
        if len(cuts) == 2:
            ubp_b, tot_b = ubp, total
        else:
            ubp_b = (ubp[r0:r1 + 1] - ubp[r0]).contiguous()
            tot_b = int(ubp_b[-1].item())
        ub_b = ub[r0:r1]
        big = bool((ub_b.max() > 8192).item()) if m else False
        Tj = torch.empty(max(tot_b, 1), dtype=torch.int32, device=dev)
        Tx = torch.empty(max(tot_b, 1), dtype=torch.float64, device=dev)
        Tk = torch.empty(max(tot_b, 1), dtype=torch.int64, device=dev) if big else None
        work = torch.empty(m + 256, dtype=torch.int32, device=dev)
        cnt_b = cnt[r0:r1]
        _lib.call('b200amg_spgemm_numeric', m, _ptr(A.indptr) + 4 * r0, _ptr(A.indices), _ptr(A.data), _ptr(B.indptr),
                  _ptr(B.indices), _ptr(B.data), _ptr(ub_b), _ptr(ubp_b), _ptr(Tk), _ptr(Tj), _ptr(Tx), _ptr(cnt_b),
                  _ptr(work), int(bool(scipy_order)), st)
        Cp_b = _scan_i32(cnt_b)
        nnz_b = int(Cp_b[-1].item())
        Cj = torch.empty(nnz_b, dtype=torch.int32, device=dev)
        Cx = torch.empty(nnz_b, dtype=torch.float64, device=dev)
        _lib.call('b200amg_spgemm_compact', m, _ptr(ubp_b), _ptr(Tj), _ptr(Tx), _ptr(Cp_b), _ptr(Cj), _ptr(Cx), st)
        pieces.append((Cp_b, Cj, Cx))
        del Tj, Tx, Tk, work
    if len(pieces) == 1:
        Cp, Cj, Cx = pieces[0]
    else:
        Cp = _scan_i32(cnt[:n])
        if int(Cp[-1].item()) != sum(int(pc[1].numel()) for pc in pieces):
            raise _lib.B200AmgError('matmat: result exceeds int32 indexing')
        Cj = torch.cat([pc[1] for pc in pieces])
        Cx = torch.cat([pc[2] for pc in pieces])
    return DevMat(Cp, Cj, Cx, (A.shape[0], B.shape[1]), (1, 1), 'csr', not scipy_order)


def transpose(A):
    """A^T as CSR with rows ascending in the original row index (scipy csr_tocsc order)."""
    if not A.is_scalar:
        return bsr_transpose(A)
    n_row, n_col = A.shape
    nnz = A.nnz
    dev = A.data.device
    Bp = torch.empty(n_col + 1, dtype=torch.int32, device=dev)
    Bi = torch.empty(nnz, dtype=torch.int32, device=dev)
    Bx = torch.empty(nnz, dtype=torch.float64, device=dev)
    nbytes = _lib.load().b200amg_csr_transpose_work_bytes(n_row, n_col, nnz)
    work = torch.empty(nbytes // 8 + 1, dtype=torch.int64, device=dev)
    _lib.call('b200amg_csr_transpose', n_row, n_col, nnz, _ptr(A.indptr), _ptr(A.indices), _ptr(A.data),
              _ptr(Bp), _ptr(Bi), _ptr(Bx), _ptr(work), _stream())
    return DevMat(Bp, Bi, Bx, (n_col, n_row), (1, 1), A.format, True)


GALERKIN_BATCH_NNZ = 48_000_000          # entries of R per pass of the triple product


def row_slice(M, r0, r1):
    """rows [r0, r1) of a scalar CSR matrix as a DevMat over views of the same arrays"""
    s, e = int(M.indptr[r0].item()), int(M.indptr[r1].item())
    ip = (M.indptr[r0:r1 + 1] - M.indptr[r0]).contiguous()
    return DevMat(ip, M.indices[s:e], M.data[s:e], (r1 - r0, M.shape[1]), (1, 1), M.format, M.has_sorted_indices)


def galerkin(R, A, P, scipy_order=True, batch_nnz=None):
    """A_c = (R A) P, evaluated left to right like the reference (aggregation.py:429).  With
    scipy_order (default) the intermediate R A and the result keep scipy's row-internal order, so
    values AND storage order equal the reference's; otherwise both are column-sorted.
    Output rows depend on the matching rows of R only, so the product runs in batches of rows of R
    (about `batch_nnz` entries each): the intermediate R A never exists as a whole (at 512^3 it would
    not fit int32 indexing), the result is identical."""
    if not (R.is_scalar and A.is_scalar and P.is_scalar):
        return bsr_matmat(bsr_matmat(R, A), P)
    limit = int(batch_nnz or GALERKIN_BATCH_NNZ)
    n = R.shape[0]
    if R.nnz <= limit or n <= 1:
        return matmat(matmat(R, A, scipy_order), P, scipy_order)
    targets = torch.arange(limit, R.nnz, limit, dtype=torch.int64, device=R.indptr.device)
    cuts = [0]
    for c in torch.searchsorted(R.indptr.to(torch.int64), targets, right=True).tolist():
        c = min(max(int(c) - 1, cuts[-1] + 1), n)
        if c > cuts[-1]:
            cuts.append(c)
    if cuts[-1] != n:
        cuts.append(n)
    cps, cjs, cxs, base = [], [], [], 0
    for r0, r1 in zip(cuts[:-1], cuts[1:]):
        C = matmat(matmat(row_slice(R, r0, r1), A, scipy_order), P, scipy_order)
        cps.append(C.indptr[:-1].to(torch.int64) + base)
        base += C.nnz
        cjs.append(C.indices)
        cxs.append(C.data)
        del C
    if base >= 2 ** 31:
        raise _lib.B200AmgError('galerkin: result exceeds int32 indexing')
    Cp = torch.cat(cps + [torch.tensor([base], dtype=torch.int64, device=R.indptr.device)]).to(torch.int32)
    return DevMat(Cp, torch.cat(cjs), torch.cat(cxs), (R.shape[0], P.shape[1]), (1, 1), 'csr', not scipy_order)


# ---- block (BSR) operands ------------------------------------------------------------------------------------
def block_pattern(A, values=None):
    """the block graph of a BSR matrix as a scalar CSR DevMat over the SAME index arrays (unit values, or `values`
    per block)"""
    v = torch.ones(A.nblocks, dtype=torch.float64, device=A.data.device) if values is None else values
    return DevMat(A.indptr, A.indices, v, (A.n_brow, A.n_bcol), (1, 1), 'csr', A.has_sorted_indices)


def is_canonical_pattern(A):
    """scipy's csr_has_canonical_format on the (block) index arrays: every row strictly increasing in column"""
    flag = torch.zeros(1, dtype=torch.int32, device=A.data.device)
    _lib.call('b200amg_csr_not_canonical', A.n_brow, _ptr(A.indptr), _ptr(A.indices), _ptr(flag), _stream())
    return int(flag.item()) == 0


def bsr_matmat(A, B):
    """C = A B on blocks (R x N) (N x C) with scipy bsr_matmat's result: blocks of a row in FIRST-TOUCH order, zero
    blocks kept, every block summed in visiting order with separately rounded multiplies and adds.  The block pattern
    is the scalar SpGEMM of the two block patterns (sorted for the lookups, csr_matmat order for the slots)."""
    R, N = A.blocksize
    N2, C = B.blocksize
    if N != N2 or A.shape[1] != B.shape[0]:
        raise ValueError('dimension mismatch')
    pa, pb = block_pattern(A), block_pattern(B)
    Ssorted = matmat(pa, pb, scipy_order=False)
    Lorder = matmat(pa, pb, scipy_order=True)
    if not torch.equal(Ssorted.indptr, Lorder.indptr):
        raise _lib.B200AmgError('bsr_matmat: the two pattern products disagree')
    nb = Ssorted.nnz
    dev = A.data.device
    st = _stream()
    slot = torch.empty(max(nb, 1), dtype=torch.int32, device=dev)
    Cj = torch.empty(nb, dtype=torch.int32, device=dev)
    Cx = torch.empty(nb * R * C, dtype=torch.float64, device=dev)
    _lib.call('b200amg_bsr_slot_map', A.n_brow, _ptr(Ssorted.indptr), _ptr(Ssorted.indices), _ptr(Lorder.indices),
              _ptr(slot), _ptr(Cj), st)
    _lib.call('b200amg_bsr_matmat_numeric', A.n_brow, R, N, C, _ptr(A.indptr), _ptr(A.indices), _ptr(A.data),
              _ptr(B.indptr), _ptr(B.indices), _ptr(B.data), _ptr(Ssorted.indptr), _ptr(Ssorted.indices), _ptr(slot),
              nb, _ptr(Cx), st)
    return DevMat(Ssorted.indptr, Cj, Cx, (A.shape[0], B.shape[1]), (R, C), 'bsr', False)


def _gather_blocks(A, perm, transpose_blocks):
    out = torch.empty_like(A.data)
    R, C = A.blocksize
    _lib.call('b200amg_bsr_gather_blocks', A.nblocks, R, C, _ptr(perm), _ptr(A.data), _ptr(out),
              int(bool(transpose_blocks)), _stream())
    return out


def bsr_transpose(A):
    """scipy bsr_transpose: csr_tocsc on the block numbers, every block transposed"""
    ids = torch.arange(A.nblocks, dtype=torch.float64, device=A.data.device)
    T = transpose(block_pattern(A, ids))
    R, C = A.blocksize
    return DevMat(T.indptr, T.indices, _gather_blocks(A, T.data, True), (A.shape[1], A.shape[0]), (C, R), 'bsr', True)


def bsr_sort_indices(A):
    """blocks of every row ascending in column (scipy bsr sort_indices; stable, rows hold no duplicates)"""
    if A.has_sorted_indices:
        return A
    ids = torch.arange(A.nblocks, dtype=torch.float64, device=A.data.device)
    S = transpose(transpose(block_pattern(A, ids)))
    return DevMat(S.indptr, S.indices, _gather_blocks(A, S.data, False), A.shape, A.blocksize, 'bsr', True)
