"""Device restatement of the few util.linalg routines the hot path and its setup need
(pyamg/util/linalg.py): norm :17-54, residual_norm :109-112, approximate_spectral_radius
:282-380 with its Arnoldi helper :155-254.  Vectors are torch CUDA tensors; SpMV, dot and
axpy run in libb200amg."""
import numpy as np
import scipy.linalg
import torch

from . import _lib
from .sparse import _ptr, _stream, _dev

_work = {}


def _dotwork(dev):
    if dev not in _work:
        _work[dev] = (torch.empty(1024, dtype=torch.float64, device=dev), torch.empty(1, dtype=torch.float64, device=dev))
    return _work[dev]


def dot(x, y):
    w, out = _dotwork(x.device)
    _lib.call('b200amg_dot', x.numel(), _ptr(x), _ptr(y), _ptr(out), _ptr(w), _stream())
    return float(out.item())


def norm(x):
    return float(np.sqrt(dot(x, x)))


def axpy(a, x, y):
    """y += a x"""
    _lib.call('b200amg_axpy', x.numel(), float(a), _ptr(x), _ptr(y), _stream())


def residual_norm(A, x, b):
    r = A.residual(x, b)
    return norm(r)


def _apply(A, v, row_scale):
    w = A.matvec(v)
    if row_scale is not None:
        w.mul_(row_scale)       # (D^-1 A) v == D^-1 (A v): the scaled copy of A is never formed
    return w


def _arnoldi(A, v0, maxiter, row_scale):
    """util/linalg.py:155-254, symmetric=False branch, real arithmetic."""
    breakdown = np.finfo(float).eps * 1e6
    flag = False
    n = A.shape[0]
    maxiter = min(n, maxiter)
    v0 = v0 / norm(v0)
    H = np.zeros((maxiter + 1, maxiter))
    V = [v0]
    j = 0
    for j in range(maxiter):
        w = _apply(A, V[-1], row_scale)
        for i, v in enumerate(V):
            H[i, j] = dot(v, w)
            axpy(-H[i, j], v, w)
        H[j + 1, j] = norm(w)
        if H[j + 1, j] < breakdown:
            flag = True
            if H[j + 1, j] != 0.0:
                w = w / H[j + 1, j]
            V.append(w)
            break
        w = w / H[j + 1, j]
        V.append(w)
    ev, vec = scipy.linalg.eig(H[:j + 1, :j + 1], left=False, right=True)
    return vec, ev, H, V, flag


def approximate_spectral_radius(A, tol=0.01, maxiter=15, restart=5, row_scale=None):
    """rho(diag(row_scale) A) by restarted Arnoldi; the start vector comes from the GLOBAL
    numpy RNG (np.random.rand(n, 1)) exactly like util/linalg.py:366-368, so that a seeded
    setup consumes the random stream like the reference."""
    n = A.shape[1]
    v0 = torch.from_numpy(np.random.rand(n, 1).ravel()).to(_dev())
    ev = None
    for _ in range(restart + 1):
        evect, ev, H, V, flag = _arnoldi(A, v0, maxiter, row_scale)
        nvecs = ev.shape[0]
        k = int(np.abs(ev).argmax())
        error = H[nvecs, nvecs - 1] * evect[-1, k]
        coef = np.real(evect[:, k])
        v0 = torch.zeros_like(V[0])
        for c, v in zip(coef, V[:-1]):
            axpy(float(c), v, v0)
        if (np.abs(error) / np.abs(ev[k]) < tol) or flag:
            break
    return float(np.abs(ev[k]))
