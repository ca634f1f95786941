"""TEST INFRASTRUCTURE ONLY -- restatement of the reference cycle engine
(pyamg/multilevel.py: solve :316-471, __solve :473-548, coarse_grid_solver('pinv2') :608-612,
:694-720; residual_norm pyamg/util/linalg.py:109-112) on scipy.sparse matrices, with the
smoothers of oracle.relaxation.  SpMV is scipy's own (`A * x`), as in the reference.

A hierarchy is a list of ``Level`` objects (A, P, R scipy matrices; ``pre``/``post`` smoother
specs with RESOLVED parameters, e.g. ('jacobi', {'omega': 4/3/rho, 'iterations': 1}) or
('gauss_seidel_indexed', {'indices': idx, 'sweep': 'symmetric', 'iterations': 1})).
"""
import numpy as np
import scipy.linalg

from . import relaxation as _rx


class Level:
    pass


def _smoother(spec, core=None):
    name, kw = spec if isinstance(spec, tuple) else (spec, {})
    kw = dict(kw)
    kw.pop('color_ptr', None)
    if name is None:
        return lambda A, x, b: None
    fn = {'jacobi': _rx.jacobi, 'gauss_seidel': _rx.gauss_seidel,
          'gauss_seidel_indexed': _rx.gauss_seidel_indexed, 'bsr_gauss_seidel_indexed': _rx.bsr_gauss_seidel_indexed,
          'block_jacobi': _rx.block_jacobi, 'block_gauss_seidel': _rx.block_gauss_seidel,
          'polynomial': _rx.polynomial}[name]
    return lambda A, x, b: fn(A, x, b, core=core, **kw)


def residual_norm(A, x, b):
    r = b - A * x
    return float(np.sqrt(np.inner(r.conj(), r).real))


class Solver:
    """Mirror of multilevel_solver for V/W/F cycles with the default 'pinv2' coarse solve."""

    def __init__(self, levels, core=None, wasted_prolongation=True):
        self.levels = levels
        self.core = core
        # multilevel.py:548 -- the fork returns P*coarse_x (a second product) from __solve;
        # results are unaffected, only the CPU time.
        self.wasted_prolongation = wasted_prolongation
        for lvl in levels[:-1]:
            lvl._pre = _smoother(lvl.pre, core)
            lvl._post = _smoother(lvl.post, core)
        self._pinv = None

    def coarse_solve(self, A, b):
        if A.nnz == 0:                                   # multilevel.py:699-701
            return np.zeros(b.shape)
        if self._pinv is None:
            self._pinv = scipy.linalg.pinv(A.toarray())  # pinv2's successor
        return np.dot(self._pinv, b)

    def coarse_matrix(self):
        A = self.levels[-1].A
        if self._pinv is None and A.nnz:
            self._pinv = scipy.linalg.pinv(A.toarray())
        return self._pinv

    def solve(self, b, x0=None, tol=1e-5, maxiter=100, cycle='V', callback=None, residuals=None):
        x = np.zeros_like(b) if x0 is None else np.array(x0)
        cycle = str(cycle).upper()
        normb = float(np.sqrt(np.inner(b.conj(), b).real))
        if normb != 0:
            tol = tol * normb
        if residuals is None:
            residuals = []
        else:
            residuals[:] = []
        b = np.ravel(b)
        x = np.ravel(x)
        A = self.levels[0].A
        residuals.append(residual_norm(A, x, b))
        while len(residuals) <= maxiter and residuals[-1] > tol:
            if len(self.levels) == 1:
                x = self.coarse_solve(A, b)
            else:
                self._cycle(0, x, b, cycle)
            residuals.append(residual_norm(A, x, b))
            if callback is not None:
                callback(x)
        return x

    def _cycle(self, k, x, b, cycle):
        lvl = self.levels[k]
        A = lvl.A
        lvl._pre(A, x, b)
        residual = b - A * x
        coarse_b = lvl.R * residual
        coarse_x = np.zeros_like(coarse_b)
        if k == len(self.levels) - 2:
            coarse_x[:] = self.coarse_solve(self.levels[-1].A, coarse_b)
        elif cycle == 'V':
            self._cycle(k + 1, coarse_x, coarse_b, 'V')
        elif cycle == 'W':
            self._cycle(k + 1, coarse_x, coarse_b, cycle)
            self._cycle(k + 1, coarse_x, coarse_b, cycle)
        elif cycle == 'F':
            self._cycle(k + 1, coarse_x, coarse_b, cycle)
            self._cycle(k + 1, coarse_x, coarse_b, 'V')
        elif cycle == 'AMLI':
            # multilevel.py:511-537: nAMLI recursive corrections, A-orthogonalised against each other
            nAMLI = 2
            Ac = self.levels[k + 1].A
            p = np.zeros((nAMLI, coarse_b.shape[0]), dtype=coarse_b.dtype)
            beta = np.zeros((nAMLI, nAMLI), dtype=coarse_b.dtype)
            for q in range(nAMLI):
                p[q, :] = 1
                self._cycle(k + 1, p[q, :].reshape(coarse_b.shape), coarse_b, cycle)
                for j in range(q):
                    beta[q, j] = np.inner(p[j, :].conj(), Ac * p[q, :]) / np.inner(p[j, :].conj(), Ac * p[j, :])
                    p[q, :] -= beta[q, j] * p[j, :]
                Ap = Ac * p[q, :]
                alpha = np.inner(p[q, :].conj(), np.ravel(coarse_b)) / np.inner(p[q, :].conj(), Ap)
                coarse_x += alpha * p[q, :].reshape(coarse_x.shape)
                coarse_b -= alpha * Ap.reshape(coarse_b.shape)
        else:
            raise TypeError('Unrecognized cycle type (%s)' % cycle)
        x += lvl.P * coarse_x
        lvl._post(A, x, b)
        if self.wasted_prolongation:
            return lvl.P * coarse_x
        return None
