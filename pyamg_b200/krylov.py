"""Krylov acceleration around the device cycle (pyamg/multilevel.py:381-422 `accel=`;
pyamg/krylov/_cg.py).  Preconditioned CG runs entirely on the device: SpMV, dots and axpys in
libb200amg, the preconditioner is one cycle of the engine; one scalar per iteration returns to
the host for the stopping test, like the reference."""
import numpy as np
import torch

from . import linalg


def cg(ml, b, x0=None, tol=1e-5, maxiter=None, cycle='V', callback=None, residuals=None):
    """pyamg/krylov/_cg.py semantics: stop when ||r|| <= tol*||b||; `residuals` receives the
    PRECONDITIONED residual norms sqrt(<r, z>) first entry included, like the reference's cg."""
    A = ml.levels[0].A
    n = A.shape[0]
    dev = A.data.device
    as_numpy = not isinstance(b, torch.Tensor)
    bt = (torch.from_numpy(np.ascontiguousarray(np.ravel(b), np.float64)).to(dev) if as_numpy else b.reshape(-1))
    x = torch.zeros_like(bt) if x0 is None else (
        torch.from_numpy(np.array(np.ravel(x0), np.float64)).to(dev) if not isinstance(x0, torch.Tensor) else x0.clone())
    if maxiter is None:
        maxiter = int(1.3 * n) + 2
    normb = linalg.norm(bt)
    if normb == 0.0:
        normb = 1.0
    r = A.residual(x, bt)
    normr = linalg.norm(r)
    if residuals is not None:
        residuals[:] = [normr]
    if normr < tol * normb:
        return (x.cpu().numpy() if as_numpy else x), 0
    z = ml.solve(r, maxiter=1, cycle=cycle, tol=1e-12, _norms=False)
    p = z.clone()
    rz = linalg.dot(r, z)
    it = 0
    while True:
        Ap = A.matvec(p)
        alpha = rz / linalg.dot(Ap, p)
        linalg.axpy(alpha, p, x)
        linalg.axpy(-alpha, Ap, r)
        z = ml.solve(r, maxiter=1, cycle=cycle, tol=1e-12, _norms=False)
        rz_old, rz = rz, linalg.dot(r, z)
        beta = rz / rz_old
        p.mul_(beta).add_(z)
        it += 1
        normr = linalg.norm(r)
        if residuals is not None:
            residuals.append(normr)
        if callback is not None:
            callback(x.cpu().numpy() if as_numpy else x)
        if normr < tol * normb:
            return (x.cpu().numpy() if as_numpy else x), 0
        if it == maxiter:
            return (x.cpu().numpy() if as_numpy else x), it


def accelerated_solve(ml, accel, b, x0, tol, maxiter, cycle, callback, residuals):
    if accel == 'cg' or accel is cg:
        return cg(ml, b, x0=x0, tol=tol, maxiter=maxiter, cycle=cycle, callback=callback, residuals=residuals)[0]
    if callable(accel):
        # scipy-style solver on the host driving the device preconditioner
        M = ml.aspreconditioner(cycle=cycle)
        A = ml.levels[0].A.to_scipy()
        return accel(A, b, x0=x0, M=M, maxiter=maxiter, callback=callback)[0]
    raise NotImplementedError("accel=%r: only 'cg' runs on the device" % (accel,))
