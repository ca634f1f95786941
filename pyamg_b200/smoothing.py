"""change_smoothers / rho_D_inv_A on the device (pyamg/relaxation/smoothing.py:24-200,320-332).

A smoother spec is `name`, `(name, {opts})`, or a per-level list of those -- the same forms the
reference accepts (smoothing.py:37-57,126-169).  Names with a device kernel:
jacobi, gauss_seidel, gauss_seidel_indexed, multicolor_gauss_seidel, chebyshev / polynomial, block_jacobi,
block_gauss_seidel, multicolor_block_gauss_seidel, None.
"""
import numpy as np
import torch

from .linalg import approximate_spectral_radius
from .multilevel import _Smoother
from .sparse import DevMat, _dev
from . import relaxation


def rho_D_inv_A(A):
    """spectral radius of D^-1 A, cached on the matrix like the reference (smoothing.py:195-200).
    The estimate draws its start vector from the global numpy RNG exactly like the reference."""
    if not hasattr(A, 'rho_D_inv'):
        d = A.diagonal()
        dinv = torch.where(d != 0, 1.0 / d, torch.zeros_like(d))
        A.rho_D_inv = approximate_spectral_radius(A, row_scale=dinv)
    return A.rho_D_inv


class _BlockDinvA:
    """x -> Dinv (A x) without forming the product (smoothing.py:243-247)"""

    def __init__(self, A, Dinv):
        n = A.n_brow
        bs = A.blocksize[0]
        ar = torch.arange(n + 1, dtype=torch.int32, device=A.data.device)
        self.D = DevMat(ar, ar[:-1].contiguous(), Dinv.contiguous().view(-1), A.shape, (bs, bs), 'bsr', True)
        self.A = A
        self.shape = A.shape

    def matvec(self, v):
        return self.D.matvec(self.A.matvec(v))


def rho_block_D_inv_A(A, Dinv):
    """spectral radius of (block D)^-1 A, cached on the matrix (smoothing.py:203-250)"""
    if not hasattr(A, 'rho_block_D_inv'):
        A.rho_block_D_inv = approximate_spectral_radius(_BlockDinvA(A, Dinv))
    return A.rho_block_D_inv


def _block_setup(A, kw):
    """blocksize / Dinv resolution of setup_block_jacobi / setup_block_gauss_seidel (smoothing.py:366-420)"""
    Dinv, blocksize = kw.pop('Dinv', None), kw.pop('blocksize', None)
    if blocksize is None:
        blocksize = A.blocksize[0] if Dinv is None else int(Dinv.shape[1])
    if blocksize != 1 and Dinv is None:
        Dinv = relaxation.get_block_diag_inverse(A, blocksize)
    if Dinv is not None and not isinstance(Dinv, torch.Tensor):
        Dinv = torch.from_numpy(np.ascontiguousarray(Dinv, dtype=np.float64)).to(_dev())
    if blocksize != 1 and A.blocksize != (blocksize, blocksize):
        raise NotImplementedError('re-blocking on the device is not available: A must be BSR with the requested blocksize')
    return int(blocksize), Dinv


def chebyshev_polynomial_coefficients(a, b, degree):
    """pyamg/relaxation/chebyshev.py:9-56: coefficients (highest power first) of the polynomial with the Chebyshev
    roots of [a, b], scaled so that C(0) = 1.  A handful of host flops at setup time."""
    if a >= b or a <= 0:
        raise ValueError('invalid interval [%s,%s]' % (a, b))
    std_roots = np.cos(np.pi * (np.arange(degree) + 0.5) / degree)
    scaled_roots = 0.5 * (b - a) * (1 + std_roots) + a
    scaled_poly = np.poly(scaled_roots)
    scaled_poly /= np.polyval(scaled_poly, 0)
    return scaled_poly


def _unpack(spec):
    if isinstance(spec, tuple):
        return spec[0], dict(spec[1])
    return spec, {}


def _levelize(spec, nlev):
    if isinstance(spec, list):
        out = list(spec)
        while len(out) < nlev:
            out.append(out[-1])
        return out[:nlev]
    return [spec] * nlev


def _resolve(spec, lvl):
    name, kw = _unpack(spec)
    A = lvl.A
    if name is None:
        return _Smoother(0, name=None)
    if name == 'jacobi':
        omega = kw.pop('omega', 1.0)
        iters = kw.pop('iterations', 1)
        if kw.pop('withrho', True):
            omega = omega / rho_D_inv_A(A)
        if kw:
            raise TypeError('jacobi: unexpected options %r' % sorted(kw))
        return _Smoother(1, omega, iters, name='jacobi')
    if name == 'gauss_seidel':
        iters, sweep = kw.pop('iterations', 1), kw.pop('sweep', 'forward')
        if kw:
            raise TypeError('gauss_seidel: unexpected options %r' % sorted(kw))
        wf = relaxation.wavefront_ordering(A)           # the lexicographic sweep over its dependency levels: same bits
        if wf is not None:
            return _Smoother(2, 1.0, iters, sweep, wf[0], np.ascontiguousarray(wf[1], dtype=np.intc), name='gauss_seidel')
        return _Smoother(3, 1.0, iters, sweep, name='gauss_seidel')
    if name in ('gauss_seidel_indexed', 'multicolor_gauss_seidel'):
        iters, sweep = kw.pop('iterations', 1), kw.pop('sweep', 'forward')
        if name == 'multicolor_gauss_seidel':
            method = kw.pop('coloring', 'MIS')                      # 'MIS' (default), 'JP', 'LDF': pyamg/graph.py:84-125
            if getattr(lvl, '_mc_method', method) != method or not hasattr(lvl, '_mc_order'):
                idx, ptr = relaxation.color_ordering(relaxation.vertex_coloring(A, method))
                lvl._mc_order = (torch.from_numpy(idx).to(_dev()), ptr)
                lvl._mc_method = method
            idx, ptr = lvl._mc_order
        else:
            idx, ptr = kw.pop('indices'), kw.pop('color_ptr', None)
            if ptr is None:
                raise ValueError('gauss_seidel_indexed inside a cycle needs color_ptr (colour groups of `indices`)')
            if not isinstance(idx, torch.Tensor):
                idx = torch.from_numpy(np.ascontiguousarray(idx, dtype=np.intc)).to(_dev())
        return _Smoother(2, 1.0, iters, sweep, idx, np.ascontiguousarray(ptr, dtype=np.intc), name=name)
    if name == 'block_jacobi':
        bs, Dinv = _block_setup(A, kw)
        if bs == 1:                                     # block Jacobi is equivalent to normal Jacobi (smoothing.py:377-380)
            return _resolve(('jacobi', kw), lvl)
        omega, iters = kw.pop('omega', 1.0), kw.pop('iterations', 1)
        if kw.pop('withrho', True):
            omega = omega / rho_block_D_inv_A(A, Dinv)
        if kw:
            raise TypeError('block_jacobi: unexpected options %r' % sorted(kw))
        return _Smoother(5, omega, iters, name=name, Dinv=Dinv)
    if name in ('block_gauss_seidel', 'multicolor_block_gauss_seidel'):
        bs, Dinv = _block_setup(A, kw)
        if bs == 1:
            return _resolve(('gauss_seidel' if name == 'block_gauss_seidel' else 'multicolor_gauss_seidel', kw), lvl)
        iters, sweep = kw.pop('iterations', 1), kw.pop('sweep', 'forward')
        if kw:
            raise TypeError('%s: unexpected options %r' % (name, sorted(kw)))
        if name == 'block_gauss_seidel':                # the reference's lexicographic order
            wf = relaxation.wavefront_ordering(A)       # ... level-parallel where the matrix allows it (same bits)
            if wf is not None:
                return _Smoother(6, 1.0, iters, sweep, wf[0], np.ascontiguousarray(wf[1], dtype=np.intc), name=name, Dinv=Dinv)
            return _Smoother(7, 1.0, iters, sweep, name=name, Dinv=Dinv)
        if not hasattr(lvl, '_mc_order'):
            idx, ptr = relaxation.color_ordering(relaxation.vertex_coloring(A, 'MIS'))
            lvl._mc_order = (torch.from_numpy(idx).to(_dev()), ptr)
        idx, ptr = lvl._mc_order
        return _Smoother(6, 1.0, iters, sweep, idx, np.ascontiguousarray(ptr, dtype=np.intc), name=name, Dinv=Dinv)
    if name in ('polynomial', 'chebyshev'):
        iters = kw.pop('iterations', 1)
        if name == 'polynomial':
            coef = kw.pop('coefficients')
        else:                       # setup_chebyshev, smoothing.py:438-449
            lower, upper, degree = kw.pop('lower_bound', 1.0 / 30.0), kw.pop('upper_bound', 1.1), kw.pop('degree', 3)
            rho = kw.pop('rho', None)
            if rho is None:
                rho = approximate_spectral_radius(A)
            coef = -chebyshev_polynomial_coefficients(rho * lower, rho * upper, degree)[:-1]
        if kw:
            raise TypeError('%s: unexpected options %r' % (name, sorted(kw)))
        return _Smoother(4, 1.0, iters, name=name, coefficients=np.asarray(coef, dtype=np.float64))
    raise ValueError('smoother %r has no device kernel (available: jacobi, gauss_seidel, '
                     'gauss_seidel_indexed, multicolor_gauss_seidel, chebyshev, polynomial, block_jacobi, block_gauss_seidel, '
                     'multicolor_block_gauss_seidel, None)' % (name,))


def change_smoothers(ml, presmoother, postsmoother):
    """Attach resolved device smoothers to every non-coarsest level (smoothing.py:24-169)."""
    nlev = len(ml.levels) - 1
    pre, post = _levelize(presmoother, nlev), _levelize(postsmoother, nlev)
    for k, lvl in enumerate(ml.levels[:-1]):
        lvl.presmoother = _resolve(pre[k], lvl)
    for k, lvl in enumerate(ml.levels[:-1]):
        lvl.postsmoother = _resolve(post[k], lvl)
    ml._destroy_engine()
