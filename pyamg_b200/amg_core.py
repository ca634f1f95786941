"""`amg_core`-named entry points with the reference's POSITIONAL signatures
(pyamg/amg_core/amg_core.i:163-194; C++ in pyamg/amg_core/relaxation.h, graph.h), taking
device arrays (torch CUDA tensors: int32 indices, float64 values).  Like the SWIG
INPLACE_ARRAY1 typemaps (amg_core.i:27-92) they refuse wrong dtypes (TypeError) and
non-contiguous arrays (ValueError) instead of converting, mutate `x` / `temp` in place and
return None.  Two optional trailing arguments extend gauss_seidel_indexed / bsr_gauss_seidel /
block_gauss_seidel: `color_ptr` (host ints) declares colour groups inside `Id`, enabling the parallel sweep.
"""
import ctypes

import numpy as np
import torch

from . import _lib
from .sparse import _stream


def _chk(t, dtype, name):
    if not isinstance(t, torch.Tensor):
        raise TypeError('%s: expected a torch CUDA tensor' % name)
    if t.dtype != dtype:
        raise TypeError('%s: expected dtype %s, got %s' % (name, dtype, t.dtype))
    if t.device.type != 'cuda':
        raise TypeError('%s: expected a CUDA tensor (no CPU fallback)' % name)
    if t.dim() != 1:
        raise ValueError('%s: expected a 1-D array' % name)
    if not t.is_contiguous():
        raise ValueError('%s: array must be contiguous' % name)
    return t.data_ptr()


def _csr(Ap, Aj, Ax):
    return (_chk(Ap, torch.int32, 'Ap'), _chk(Aj, torch.int32, 'Aj'), _chk(Ax, torch.float64, 'Ax'))


def _omega(omega):
    if isinstance(omega, torch.Tensor):
        return float(omega.reshape(-1)[0].item())
    return float(np.ravel(omega)[0])


def _cptr(color_ptr):
    if color_ptr is None:
        return None, 0
    cp = np.ascontiguousarray(color_ptr, dtype=np.intc)
    return cp.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), len(cp) - 1


def jacobi(Ap, Aj, Ax, x, b, temp, row_start, row_stop, row_step, omega):
    """relaxation.h:201-239"""
    p = _csr(Ap, Aj, Ax)
    _lib.call('b200amg_jacobi', *p, _chk(x, torch.float64, 'x'), _chk(b, torch.float64, 'b'),
              _chk(temp, torch.float64, 'temp'), int(row_start), int(row_stop), int(row_step), _omega(omega),
              _stream())


def bsr_jacobi(Ap, Aj, Ax, x, b, temp, row_start, row_stop, row_step, blocksize, omega):
    """relaxation.h:267-360"""
    p = _csr(Ap, Aj, Ax)
    _lib.call('b200amg_bsr_jacobi', *p, _chk(x, torch.float64, 'x'), _chk(b, torch.float64, 'b'),
              _chk(temp, torch.float64, 'temp'), int(row_start), int(row_stop), int(row_step), int(blocksize),
              _omega(omega), _stream())


def gauss_seidel(Ap, Aj, Ax, x, b, row_start, row_stop, row_step):
    """relaxation.h:33-62 (sequential sweep on the device)"""
    p = _csr(Ap, Aj, Ax)
    _lib.call('b200amg_gauss_seidel', *p, _chk(x, torch.float64, 'x'), _chk(b, torch.float64, 'b'),
              int(row_start), int(row_stop), int(row_step), _stream())


def gauss_seidel_indexed(Ap, Aj, Ax, x, b, Id, row_start, row_stop, row_step, color_ptr=None):
    """relaxation.h:394-426"""
    p = _csr(Ap, Aj, Ax)
    cp, K = _cptr(color_ptr)
    _lib.call('b200amg_gauss_seidel_indexed', *p, _chk(x, torch.float64, 'x'), _chk(b, torch.float64, 'b'),
              _chk(Id, torch.int32, 'Id'), int(row_start), int(row_stop), int(row_step), cp, K, _stream())


def bsr_gauss_seidel(Ap, Aj, Ax, x, b, row_start, row_stop, row_step, blocksize, Id=None, color_ptr=None):
    """relaxation.h:89-173"""
    p = _csr(Ap, Aj, Ax)
    cp, K = _cptr(color_ptr)
    _lib.call('b200amg_bsr_gauss_seidel', *p, _chk(x, torch.float64, 'x'), _chk(b, torch.float64, 'b'),
              int(row_start), int(row_stop), int(row_step), int(blocksize),
              0 if Id is None else _chk(Id, torch.int32, 'Id'), cp, K, _stream())


def vertex_coloring_mis(num_rows, Ap, Aj, x):
    """graph.h:171-188.  Setup-time, inherently sequential: HOST int32 numpy arrays; returns K."""
    Ap = np.ascontiguousarray(Ap, np.intc)
    Aj = np.ascontiguousarray(Aj, np.intc)
    if x.dtype != np.intc or not x.flags.c_contiguous:
        raise TypeError('x must be a contiguous int32 numpy array')
    lib = _lib.load()
    K = lib.b200amg_vertex_coloring_mis_host(int(num_rows), Ap.ctypes.data, Aj.ctypes.data, x.ctypes.data)
    if K < 0:
        raise _lib.B200AmgError(lib.b200amg_last_error().decode())
    return K


def gs_wavefronts(num_rows, Ap, Aj, level):
    """dependency levels of the lexicographic Gauss-Seidel sweep over the pattern (host int32 arrays); returns the
    number of levels.  No reference counterpart: it is what lets relaxation.h:33-62 run in parallel with its bits."""
    Ap = np.ascontiguousarray(Ap, np.intc)
    Aj = np.ascontiguousarray(Aj, np.intc)
    if level.dtype != np.intc or not level.flags.c_contiguous:
        raise TypeError('level must be a contiguous int32 numpy array')
    lib = _lib.load()
    K = lib.b200amg_gs_wavefronts_host(int(num_rows), Ap.ctypes.data, Aj.ctypes.data, level.ctypes.data)
    if K < 0:
        raise _lib.B200AmgError(lib.b200amg_last_error().decode())
    return K


def naive_aggregation(n_row, Ap, Aj, x, y):
    """smoothed_aggregation.h:244-277 (host arrays); x receives 0-based aggregate numbers (the reference's are 1-based
    and its Python wrapper subtracts 1, aggregate.py:174-176); returns the number of aggregates"""
    Ap = np.ascontiguousarray(Ap, np.intc)
    Aj = np.ascontiguousarray(Aj, np.intc)
    if x.dtype != np.intc or y.dtype != np.intc or not x.flags.c_contiguous or not y.flags.c_contiguous:
        raise TypeError('x, y must be contiguous int32 numpy arrays')
    lib = _lib.load()
    k = lib.b200amg_naive_aggregation_host(int(n_row), Ap.ctypes.data, Aj.ctypes.data, x.ctypes.data, y.ctypes.data)
    if k < 0:
        raise _lib.B200AmgError(lib.b200amg_last_error().decode())
    return k


def _weighted_coloring(entry, num_rows, Ap, Aj, x, w):
    Ap = np.ascontiguousarray(Ap, np.intc)
    Aj = np.ascontiguousarray(Aj, np.intc)
    w = np.ascontiguousarray(w, np.float64)
    if x.dtype != np.intc or not x.flags.c_contiguous:
        raise TypeError('x must be a contiguous int32 numpy array')
    if w.shape[0] < num_rows:
        raise ValueError('one weight per vertex expected')
    lib = _lib.load()
    K = getattr(lib, entry)(int(num_rows), Ap.ctypes.data, Aj.ctypes.data, x.ctypes.data, w.ctypes.data)
    if K < 0:
        raise _lib.B200AmgError(lib.b200amg_last_error().decode())
    return K


def vertex_coloring_jones_plassmann(num_rows, Ap, Aj, x, z):
    """graph.h:242-287 (host arrays; z = random weights, left unmodified here); returns the number of colours"""
    return _weighted_coloring('b200amg_vertex_coloring_jones_plassmann_host', num_rows, Ap, Aj, x, z)


def vertex_coloring_LDF(num_rows, Ap, Aj, x, y):
    """graph.h:290-335 (host arrays; y = random weights); returns the number of colours"""
    return _weighted_coloring('b200amg_vertex_coloring_ldf_host', num_rows, Ap, Aj, x, y)


def standard_aggregation(n_row, Ap, Aj, x, y):
    """smoothed_aggregation.h:121-221 (host arrays); returns the number of aggregates."""
    Ap = np.ascontiguousarray(Ap, np.intc)
    Aj = np.ascontiguousarray(Aj, np.intc)
    lib = _lib.load()
    n = lib.b200amg_standard_aggregation_host(int(n_row), Ap.ctypes.data, Aj.ctypes.data, x.ctypes.data,
                                              y.ctypes.data)
    if n < 0:
        raise _lib.B200AmgError(lib.b200amg_last_error().decode())
    return n


def rs_cf_splitting(n_nodes, Sp, Sj, Tp, Tj, splitting):
    """ruge_stuben.h:157-297 (host int32 numpy arrays; sequential by definition): splitting[i] = 1 C / 0 F."""
    Sp, Sj, Tp, Tj = (np.ascontiguousarray(v, np.intc) for v in (Sp, Sj, Tp, Tj))
    if splitting.dtype != np.intc or not splitting.flags.c_contiguous:
        raise TypeError('splitting must be a contiguous int32 numpy array')
    lib = _lib.load()
    rc = lib.b200amg_rs_cf_splitting_host(int(n_nodes), Sp.ctypes.data, Sj.ctypes.data, Tp.ctypes.data, Tj.ctypes.data,
                                          splitting.ctypes.data)
    if rc != 0:
        raise _lib.B200AmgError(lib.b200amg_last_error().decode())


def block_jacobi(Ap, Aj, Ax, x, b, Tx, temp, row_start, row_stop, row_step, omega, blocksize):
    """relaxation.h:661-727 (Tx = inverses of the diagonal blocks, [n_brow * bs * bs])"""
    p = _csr(Ap, Aj, Ax)
    _lib.call('b200amg_block_jacobi', *p, _chk(x, torch.float64, 'x'), _chk(b, torch.float64, 'b'),
              _chk(Tx, torch.float64, 'Tx'), _chk(temp, torch.float64, 'temp'), int(row_start), int(row_stop),
              int(row_step), _omega(omega), int(blocksize), _stream())


def block_gauss_seidel(Ap, Aj, Ax, x, b, Tx, row_start, row_stop, row_step, blocksize, Id=None, color_ptr=None):
    """relaxation.h:753-806; with (Id, color_ptr): block rows grouped by colour, one parallel launch per colour"""
    p = _csr(Ap, Aj, Ax)
    cp, K = _cptr(color_ptr)
    _lib.call('b200amg_block_gauss_seidel', *p, _chk(x, torch.float64, 'x'), _chk(b, torch.float64, 'b'),
              _chk(Tx, torch.float64, 'Tx'), int(row_start), int(row_stop), int(row_step), int(blocksize),
              0 if Id is None else _chk(Id, torch.int32, 'Id'), cp, K, _stream())
