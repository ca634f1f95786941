"""multilevel_solver on the B200: same constructor, `solve()`, `aspreconditioner()`,
complexities and repr as the reference's (pyamg/multilevel.py:14-548); levels hold device
matrices (pyamg_b200.sparse.DevMat) and the cycle runs inside libb200amg
(b200amg_vcycle_*: a CUDA-graph replay of the whole V/W/F cycle).

Smoothers are attached with `change_smoothers(ml, presmoother, postsmoother)` like the
reference (pyamg/relaxation/smoothing.py:24-169) using specs
    ('jacobi', {'omega': w, 'iterations': k, 'withrho': True})
    ('gauss_seidel', {'sweep': 'forward'|'backward'|'symmetric', 'iterations': k})      lexicographic (exact, sequential)
    ('gauss_seidel_indexed', {'indices': Id, 'color_ptr': ptr, 'sweep': ..., 'iterations': k})
    ('multicolor_gauss_seidel', {'sweep': ..., 'iterations': k})    MIS colouring + indexed sweep
Anything else raises: there is no CPU fallback.
"""
import ctypes

import numpy as np
import torch

from . import _lib
from .sparse import DevMat, _dev, _ptr

_SWEEP = {'forward': 0, 'backward': 1, 'symmetric': 2}
_CYCLE = {'V': 0, 'W': 1, 'F': 2}


class _Smoother:
    """resolved smoother of one level: kind, omega, iterations, sweep (+ colour ordering)"""

    def __init__(self, kind=0, omega=1.0, iterations=1, sweep='forward', indices=None, color_ptr=None, name=None,
                 coefficients=None, Dinv=None):
        self.kind, self.omega, self.iterations, self.sweep = kind, float(omega), int(iterations), sweep
        self.indices, self.color_ptr, self.name = indices, color_ptr, name
        self.Dinv = Dinv                              # kinds 5-7: inverses of the diagonal blocks, device [n_brow, bs, bs]
        self.coefficients = None if coefficients is None else [float(c) for c in coefficients]   # kind 4: polynomial


def coarse_grid_solver(solver):
    """pyamg/multilevel.py:554-720 for the direct solvers: 'pinv2' (default) / 'pinv' -- the pseudo-inverse of the
    (<= max_coarse) coarsest matrix is formed ONCE on the host with LAPACK, exactly as the reference does
    (sp.linalg.pinv2(A.todense())), and applied on the device every cycle (b200amg_dense_matvec) -- and 'lu' /
    'cholesky' / 'splu', whose factorisation is likewise done once on the host and kept as the explicit inverse (for
    'splu' of the matrix without its empty rows / columns, zero there, like the reference's Map), so that the per-cycle
    work is the same dense product; equal to the reference's triangular solves up to rounding (not to the bit)."""
    name, kwargs = solver if isinstance(solver, tuple) else (solver, {})
    if name not in ('pinv', 'pinv2', 'lu', 'cholesky', 'splu'):
        raise ValueError('coarse solver %r is not available on the device (pinv, pinv2, lu, cholesky, splu)' % (name,))

    def host_inverse(dense):
        import scipy.linalg
        if name in ('pinv', 'pinv2'):
            return scipy.linalg.pinv(dense, **kwargs)
        if name == 'lu':
            return scipy.linalg.lu_solve(scipy.linalg.lu_factor(dense, **kwargs), np.eye(dense.shape[0]))
        if name == 'cholesky':
            return scipy.linalg.cho_solve(scipy.linalg.cho_factor(dense, **kwargs), np.eye(dense.shape[0]))
        keep = np.where(np.abs(dense).sum(axis=0) != 0)[0]          # 'splu': drop empty columns (and the same rows)
        inv = np.zeros_like(dense)
        if len(keep):
            sub = dense[np.ix_(keep, keep)]
            inv[np.ix_(keep, keep)] = scipy.linalg.lu_solve(scipy.linalg.lu_factor(sub), np.eye(len(keep)))
        return inv

    class generic_solver:
        def __init__(self):
            self.P = None

        def matrix(self, A):
            """device pseudo-inverse (n x n row-major) or None when A has no entries"""
            if A.nnz == 0:
                return None
            if self.P is None:
                dense = A.to_scipy().toarray()
                self.P = torch.from_numpy(np.ascontiguousarray(host_inverse(dense))).to(_dev())
            return self.P

        def __call__(self, A, b):
            P = self.matrix(A)
            x = torch.zeros_like(b)
            if P is not None:
                _lib.call('b200amg_dense_matvec', A.shape[0], _ptr(P), _ptr(b), _ptr(x),
                          torch.cuda.current_stream().cuda_stream)
            return x

        def __repr__(self):
            return 'coarse_grid_solver(' + repr(name) + ')'

        def name(self):
            return repr(name)

    return generic_solver()


class multilevel_solver:
    """Stores the multigrid hierarchy and runs the cycle (pyamg/multilevel.py:14-44)."""

    class level:
        """struct: A (and P, R, presmoother, postsmoother on all but the coarsest level)"""

        def __init__(self):
            pass

    def __init__(self, levels, coarse_solver='pinv2'):
        self.levels = levels
        self.coarse_solver_spec = coarse_solver
        self.coarse_solver = coarse_grid_solver(coarse_solver)
        for lvl in levels:
            for nm in ('A', 'P', 'R'):
                M = getattr(lvl, nm, None)
                if M is not None and not isinstance(M, DevMat):
                    setattr(lvl, nm, DevMat.from_scipy(M))
        for lvl in levels[:-1]:
            if not hasattr(lvl, 'R'):
                from .sparse import transpose
                lvl.R = transpose(lvl.P)                  # level.R = level.P.H (multilevel.py:154-156)
        self._engine = None
        self._stream = None
        self._keep = []

    # ---- statistics (multilevel.py:158-269) -----------------------------------------------
    def __repr__(self):
        out = 'multilevel_solver\n'
        out += 'Number of Levels:     %d\n' % len(self.levels)
        out += 'Operator Complexity: %6.3f\n' % self.operator_complexity()
        out += 'Grid Complexity:     %6.3f\n' % self.grid_complexity()
        out += 'Coarse Solver:        %s\n' % self.coarse_solver.name()
        total = sum(l.A.nnz for l in self.levels)
        out += '  level   unknowns     nonzeros\n'
        for n, l in enumerate(self.levels):
            out += '   %2d   %10d   %10d [%5.2f%%]\n' % (n, l.A.shape[1], l.A.nnz, 100 * float(l.A.nnz) / float(total))
        return out

    def cycle_complexity(self, cycle='V'):
        cycle = str(cycle).upper()
        nnz = [l.A.nnz for l in self.levels]
        last = len(self.levels) - 2

        def cost(k, kind):
            if len(self.levels) == 1:
                return nnz[0]
            if k == last:
                return 2 * nnz[k] + nnz[k + 1]
            if kind == 'V':
                return 2 * nnz[k] + cost(k + 1, 'V')
            if kind == 'W':
                return 2 * nnz[k] + 2 * cost(k + 1, 'W')
            return 2 * nnz[k] + cost(k + 1, 'F') + cost(k + 1, 'V')
        if cycle == 'AMLI':
            cycle = 'W'
        if cycle not in 'VWF':
            raise TypeError('Unrecognized cycle type (%s)' % cycle)
        return float(cost(0, cycle)) / float(nnz[0])

    def operator_complexity(self):
        return sum(l.A.nnz for l in self.levels) / float(self.levels[0].A.nnz)

    def grid_complexity(self):
        return sum(l.A.shape[0] for l in self.levels) / float(self.levels[0].A.shape[0])

    # ---- cycle engine ------------------------------------------------------------------------
    def _destroy_engine(self):
        if self._engine is not None:
            _lib.load().b200amg_vcycle_destroy(self._engine)
            self._engine = None
            self._keep = []

    def __del__(self):
        try:
            self._destroy_engine()
        except Exception:
            pass

    def _build_engine(self):
        lv = self.levels
        nlev = len(lv)
        dev = _dev()
        descs = (_lib.LevelDesc * (nlev - 1))()
        keep = []
        for k in range(nlev - 1):
            L, d = lv[k], descs[k]
            A, P, R = L.A, L.P, L.R
            if A.blocksize[0] != A.blocksize[1]:
                raise ValueError('BSR blocks must be square')
            if R.blocksize != (P.blocksize[1], P.blocksize[0]) or R.shape != (P.shape[1], P.shape[0]):
                raise ValueError('R must have the transposed shape / block shape of P')
            pre = getattr(L, 'presmoother', None) or _Smoother()
            post = getattr(L, 'postsmoother', None) or _Smoother()
            if not isinstance(pre, _Smoother) or not isinstance(post, _Smoother):
                raise TypeError('smoothers must be set with pyamg_b200.change_smoothers (device kernels only)')
            d.n, d.bs, d.flavor = A.shape[0], A.blocksize[0], 0 if A.format == 'csr' else 1
            d.Ap, d.Aj, d.Ax = _ptr(A.indptr), _ptr(A.indices), _ptr(A.data)
            d.nc, d.p_br, d.p_bc = P.shape[1], P.blocksize[0], P.blocksize[1]
            d.Pp, d.Pj, d.Px = _ptr(P.indptr), _ptr(P.indices), _ptr(P.data)
            d.Rp, d.Rj, d.Rx = _ptr(R.indptr), _ptr(R.indices), _ptr(R.data)
            d.pre_kind, d.post_kind = pre.kind, post.kind
            d.pre_omega, d.post_omega = pre.omega, post.omega
            d.pre_iters, d.post_iters = pre.iterations, post.iterations
            d.pre_sweep, d.post_sweep = _SWEEP[pre.sweep], _SWEEP[post.sweep]
            d.Id, d.K = 0, 0
            for side, sm in (('pre', pre), ('post', post)):
                if sm.kind == 4:
                    if not 1 <= len(sm.coefficients) <= 8:
                        raise ValueError('polynomial smoother: 1..8 coefficients')
                    setattr(d, side + '_ncoef', len(sm.coefficients))
                    for q, c in enumerate(sm.coefficients):
                        getattr(d, side + '_coef')[q] = c
            blocked = [s for s in (pre, post) if s.kind in (5, 6, 7)]
            if blocked:
                Dinv = blocked[0].Dinv
                if any(s.Dinv is not Dinv and not torch.equal(s.Dinv, Dinv) for s in blocked[1:]):
                    raise ValueError('pre- and post-smoother must share one set of diagonal-block inverses per level')
                if tuple(Dinv.shape) != (A.n_brow, A.blocksize[0], A.blocksize[0]):
                    raise ValueError('Dinv and A have incompatible dimensions')
                Dinv = Dinv.contiguous()
                keep.append(Dinv)
                d.Dinv = _ptr(Dinv)
            ordered = [s for s in (pre, post) if s.kind in (2, 6)]
            if ordered:
                s0 = ordered[0]
                for s in ordered[1:]:
                    if s.indices is not s0.indices and not (
                            np.array_equal(np.asarray(s.color_ptr), np.asarray(s0.color_ptr)) and
                            (torch.equal(s.indices, s0.indices) if isinstance(s.indices, torch.Tensor) and isinstance(s0.indices, torch.Tensor)
                             else np.array_equal(np.asarray(s.indices.cpu() if isinstance(s.indices, torch.Tensor) else s.indices),
                                                 np.asarray(s0.indices.cpu() if isinstance(s0.indices, torch.Tensor) else s0.indices)))):
                        raise ValueError('pre- and post-smoother must share one colour ordering per level')
                Id = s0.indices
                if not isinstance(Id, torch.Tensor):
                    Id = torch.from_numpy(np.ascontiguousarray(Id, dtype=np.intc)).to(dev)
                cp = np.ascontiguousarray(s0.color_ptr, dtype=np.intc)
                keep += [Id, cp]
                d.Id, d.K = _ptr(Id), len(cp) - 1
                # gauss_seidel_indexed (and the multicolour construct over it) works on the CSR form of the matrix in
                # the reference (relaxation.py:713); the lexicographic gauss_seidel keeps the level's format
                d.gs_csr = int(s0.name in ('gauss_seidel_indexed', 'multicolor_gauss_seidel'))
                d.color_ptr_host = cp.ctypes.data_as(ctypes.POINTER(ctypes.c_int))
        lib = _lib.load()
        nbytes = lib.b200amg_vcycle_work_bytes(nlev, descs)
        work = torch.empty(nbytes // 8 + 8, dtype=torch.float64, device=dev)
        cinv = self.coarse_solver.matrix(lv[-1].A)
        handle = ctypes.c_void_p()
        _lib.call('b200amg_vcycle_create', nlev, descs, _ptr(cinv), _ptr(work), nbytes, ctypes.byref(handle))
        n = lv[0].A.shape[0]
        self._x = torch.zeros(n, dtype=torch.float64, device=dev)
        self._b = torch.zeros(n, dtype=torch.float64, device=dev)
        self._nrm = torch.zeros(2, dtype=torch.float64, device=dev)
        self._keep = keep + [work, cinv, descs]
        self._engine = handle
        if self._stream is None:
            self._stream = torch.cuda.Stream(device=dev)

    def _residual_norm(self):
        _lib.call('b200amg_vcycle_residual_norm2', self._engine, _ptr(self._x), _ptr(self._b), _ptr(self._nrm),
                  self._stream.cuda_stream)
        return float(np.sqrt(self._nrm[0].item()))

    def cycle_device(self, cycle='V', use_graph=True, from_zero=False):
        """one cycle on the engine's resident (x, b): x <- cycle(x, b).  Asynchronous.  V / W / F replay the engine's
        CUDA graph; 'AMLI' needs inner products between its recursive corrections and is driven from the host
        (cycle_host).  from_zero: x is known to be zero (and need not be cleared): the first pre-smoothing sweep
        collapses to its bit-identical shortcut (b200amg_vcycle_run_zero)."""
        if cycle == 'AMLI':
            if from_zero:
                self._x.zero_()
            self.cycle_host(0, self._x, self._b, 'AMLI')
            return
        _lib.call('b200amg_vcycle_run_zero' if from_zero else 'b200amg_vcycle_run', self._engine, _ptr(self._x), _ptr(self._b),
                  _CYCLE[cycle], int(use_graph), self._stream.cuda_stream)

    def _apply_smoother(self, sm, A, x, b):
        """one resolved smoother through the stand-alone entry points (same kernels' arithmetic as the engine)"""
        from . import relaxation as rx
        if sm is None or sm.kind == 0:
            return
        if sm.kind == 1:
            rx.jacobi(A, x, b, iterations=sm.iterations, omega=sm.omega)
        elif sm.kind == 2:
            rx.gauss_seidel(A, x, b, iterations=sm.iterations, sweep=sm.sweep, indices=sm.indices, color_ptr=sm.color_ptr)
        elif sm.kind == 3:
            rx.gauss_seidel(A, x, b, iterations=sm.iterations, sweep=sm.sweep)
        elif sm.kind == 4:
            rx.polynomial(A, x, b, sm.coefficients, iterations=sm.iterations)
        elif sm.kind == 5:
            rx.block_jacobi(A, x, b, Dinv=sm.Dinv, blocksize=A.blocksize[0], iterations=sm.iterations, omega=sm.omega)
        elif sm.kind in (6, 7):
            rx.block_gauss_seidel(A, x, b, iterations=sm.iterations, sweep=sm.sweep, blocksize=A.blocksize[0], Dinv=sm.Dinv,
                                  indices=sm.indices if sm.kind == 6 else None, color_ptr=sm.color_ptr if sm.kind == 6 else None)
        else:
            raise ValueError('unknown smoother kind %r' % (sm.kind,))

    def cycle_host(self, k, x, b, cycle='V'):
        """multilevel_solver.__solve (pyamg/multilevel.py:473-548) driven from the host, one kernel call per operation,
        on device vectors x (updated in place) and b: every cycle type of the reference, 'AMLI' included (:511-537: two
        recursive corrections per level, the second A-orthogonalised against the first, step lengths from inner
        products that return to the host).  V / W / F give the engine's numbers and serve as its cross-check."""
        from . import linalg
        lvl = self.levels[k]
        A = lvl.A
        self._apply_smoother(getattr(lvl, 'presmoother', None), A, x, b)
        r = A.residual(x, b)
        cb = lvl.R.matvec(r)
        cx = torch.zeros_like(cb)
        if k == len(self.levels) - 2:
            cx.copy_(self.coarse_solver(self.levels[-1].A, cb))
        elif cycle == 'V':
            self.cycle_host(k + 1, cx, cb, 'V')
        elif cycle == 'W':
            self.cycle_host(k + 1, cx, cb, cycle)
            self.cycle_host(k + 1, cx, cb, cycle)
        elif cycle == 'F':
            self.cycle_host(k + 1, cx, cb, cycle)
            self.cycle_host(k + 1, cx, cb, 'V')
        elif cycle == 'AMLI':
            Ac = self.levels[k + 1].A
            p = []
            for q in range(2):                                        # nAMLI = 2
                pq = torch.ones_like(cb)                              # the reference starts every direction from ones
                self.cycle_host(k + 1, pq, cb, cycle)
                for pj in p:
                    beta = linalg.dot(pj, Ac.matvec(pq)) / linalg.dot(pj, Ac.matvec(pj))
                    pq.sub_(pj * beta)
                Ap = Ac.matvec(pq)
                alpha = linalg.dot(pq, cb) / linalg.dot(pq, Ap)
                cx.add_(pq * alpha)
                cb.sub_(Ap * alpha)
                p.append(pq)
        else:
            raise TypeError('Unrecognized cycle type (%s)' % cycle)
        x.add_(lvl.P.matvec(cx))
        self._apply_smoother(getattr(lvl, 'postsmoother', None), A, x, b)

    def profile_cycle(self, cycle='V'):
        """one un-graphed cycle on the resident (x, b) with a CUDA event after every operation:
        -> list of (op, level, ms), op in presmooth/residual/restrict/coarse/prolong/postsmooth/copyback"""
        names = ('presmooth', 'residual', 'restrict', 'coarse', 'prolong', 'postsmooth', 'copyback')
        cap = 64 * len(self.levels) * (4 if cycle != 'V' else 1)
        code, lev = (ctypes.c_int * cap)(), (ctypes.c_int * cap)()
        ms, n = (ctypes.c_float * cap)(), ctypes.c_int(0)
        _lib.call('b200amg_vcycle_profile', self._engine, _ptr(self._x), _ptr(self._b), _CYCLE[cycle], cap,
                  code, lev, ms, ctypes.byref(n), self._stream.cuda_stream)
        return [(names[code[i]], lev[i], float(ms[i])) for i in range(n.value)]

    def layouts(self):
        """[{'A': code, 'P': code, 'R': code, 'GS': code} per level below the coarsest]: how the engine streams each
        operator (1 CSR, 2 sliced-ELL, 3 CSR + value codes, 4 CSR rows + pattern codes, 5 row-pattern bytes, 6 sliced-ELL
        + value codes, 7 sliced-ELL + 16-bit column steps, 0 block operator); GS: the colour-grouped copy of A that
        colour-ordered Gauss-Seidel sweeps, J: the diagonal-free copy of A for the Jacobi sweeps of a big level (0: none)"""
        if self._engine is None:
            self._build_engine()
        f = _lib.load().b200amg_vcycle_layout
        return [{k: int(f(self._engine, lev, w)) for w, k in enumerate(('A', 'P', 'R', 'GS', 'J'))} for lev in range(len(self.levels) - 1)]

    def cycle_from_zero(self, b_dev, cycle='V'):
        """x = one cycle from a zero guess for a DEVICE right-hand side: no norms, no host sync, the
        result stays in the engine's x buffer (valid until the next call).  Used for the agglomerated
        coarse levels of the row-partitioned cycle (distributed.py)."""
        if len(self.levels) == 1:
            return self.coarse_solver(self.levels[0].A, b_dev)
        if self._engine is None:
            self._build_engine()
        if torch.cuda.is_current_stream_capturing():
            # inside an outer CUDA-graph capture (distributed cycle): launch the level kernels directly
            # into the capturing stream instead of replaying the engine's own graph
            self._b.copy_(b_dev, non_blocking=True)
            self._x.zero_()
            _lib.call('b200amg_vcycle_run', self._engine, _ptr(self._x), _ptr(self._b), _CYCLE[cycle], 0,
                      torch.cuda.current_stream().cuda_stream)
            return self._x
        cur = torch.cuda.current_stream()
        self._stream.wait_stream(cur)
        with torch.cuda.stream(self._stream):
            self._b.copy_(b_dev, non_blocking=True)
            self._x.zero_()
            self.cycle_device(cycle)
        cur.wait_stream(self._stream)
        return self._x

    # ---- public API (multilevel.py:271-471) -------------------------------------------------
    def psolve(self, b):
        return self.solve(b, maxiter=1)

    def aspreconditioner(self, cycle='V'):
        """LinearOperator whose matvec is ONE cycle from a zero guess (multilevel.py:274-314).
        The reference computes two residual norms around that cycle only to discard them; they
        are skipped here (same result)."""
        from scipy.sparse.linalg import LinearOperator
        shape = self.levels[0].A.shape

        def matvec(b):
            return self.solve(b, maxiter=1, cycle=cycle, tol=1e-12, _norms=False)
        return LinearOperator(shape, matvec, dtype=np.float64)

    def distributed_solver(self, **kw):
        """the row-partitioned solver over the ranks of torch.distributed for this hierarchy (collective; cached):
        the other ranks hold pyamg_b200.distributed.participant() in place of this object"""
        if getattr(self, '_dist', None) is None:
            from .distributed import DistributedSolver
            self._dist = DistributedSolver.from_ml(self, **kw)
        return self._dist

    def solve(self, b, x0=None, tol=1e-5, maxiter=100, cycle='V', accel=None, callback=None, residuals=None,
              return_residuals=False, _norms=True, _alias_out=False, distributed=False):
        """Same contract as the reference's solve (multilevel.py:316-471): relative-residual
        stopping test `||r_k|| <= tol * ||b||`, residual history appended to `residuals`,
        `callback(x)` after every cycle.  numpy in -> numpy out, torch CUDA in -> torch CUDA out, torch CPU (pinned) in
        -> torch CPU out (a fresh tensor; `_alias_out=True` hands out the solver's persistent pinned buffer instead,
        which the next call overwrites).  `_norms=False` skips the residual norms (fixed `maxiter` cycles): what
        aspreconditioner's matvec needs.  `distributed=True` (one process per GPU, torch.distributed initialised): the
        same call on the row-partitioned hierarchy; collective -- the other ranks call
        pyamg_b200.distributed.participant().solve(distributed=True); x comes back on this rank."""
        cycle = str(cycle).upper()
        if cycle not in _CYCLE and cycle != 'AMLI':
            raise TypeError('Unrecognized cycle type (%s)' % cycle)
        if distributed:
            # one process per GPU (torch.distributed initialised): the fine levels are cut into row slabs over the
            # ranks, halos refreshed over NVLink, coarse levels agglomerated here (distributed.py); same stopping rule
            if cycle != 'V' or accel is not None or callback is not None:
                raise NotImplementedError('distributed solve: plain V-cycles (no accel / callback)')
            res = [] if residuals is None else residuals
            x = self.distributed_solver().solve(b, x0, tol=tol, maxiter=maxiter, residuals=res)
            out = x if isinstance(b, torch.Tensor) else x.cpu().numpy().reshape(np.shape(b))
            return (out, res) if return_residuals else out
        if accel is not None:
            from . import krylov
            return krylov.accelerated_solve(self, accel, b, x0, tol, maxiter, cycle, callback, residuals)
        n = self.levels[0].A.shape[0]
        as_numpy = not isinstance(b, torch.Tensor)
        host_tensor = isinstance(b, torch.Tensor) and b.device.type == 'cpu'
        if self._engine is None and len(self.levels) > 1:
            self._build_engine()
        if len(self.levels) == 1:
            return self._solve_one_level(b, x0, tol, maxiter, callback, residuals, as_numpy)
        cur = torch.cuda.current_stream()
        self._stream.wait_stream(cur)
        with torch.cuda.stream(self._stream):
            bt = b if isinstance(b, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(np.ravel(b), np.float64))
            if bt.numel() != n:
                raise ValueError('b has invalid dimensions')
            self._b.copy_(bt.reshape(-1), non_blocking=True)
            zero_guess = x0 is None
            if zero_guess and _norms:
                self._x.zero_()                       # the residual norm before the first cycle reads x
            elif not zero_guess:
                xt = x0 if isinstance(x0, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(np.ravel(x0), np.float64))
                self._x.copy_(xt.reshape(-1), non_blocking=True)
            if residuals is None:
                residuals = []
            else:
                residuals[:] = []
            if _norms:
                _lib.call('b200amg_dot', n, _ptr(self._b), _ptr(self._b), _ptr(self._nrm) + 8, self._keep_dotwork(),
                          self._stream.cuda_stream)
                normb = float(np.sqrt(self._nrm[1].item()))
                if normb != 0:
                    tol = tol * normb
                residuals.append(self._residual_norm())
                while len(residuals) <= maxiter and residuals[-1] > tol:
                    self.cycle_device(cycle, from_zero=zero_guess and len(residuals) == 1)
                    residuals.append(self._residual_norm())
                    if callback is not None:
                        callback(self._x.cpu().numpy() if as_numpy else self._x)
            else:
                if maxiter == 0 and zero_guess:
                    self._x.zero_()
                for k in range(maxiter):
                    self.cycle_device(cycle, from_zero=zero_guess and k == 0)
            if host_tensor:
                # host tensor in -> host tensor out through a persistent pinned buffer (one D2H copy)
                if getattr(self, '_xhost', None) is None:
                    self._xhost = torch.empty(n, dtype=torch.float64, pin_memory=True)
                self._xhost.copy_(self._x, non_blocking=True)
                self._stream.synchronize()
                out = self._xhost if _alias_out else self._xhost.clone()
            else:
                out = self._x.cpu().numpy() if as_numpy else self._x.clone()
        cur.wait_stream(self._stream)
        if as_numpy and isinstance(b, np.ndarray):
            out = out.reshape(b.shape)
        if return_residuals:
            return out, residuals
        return out

    def solve_many(self, bs, maxiter=1, cycle='V', out=None):
        """x_k = `maxiter` cycles from a zero initial guess for EVERY right-hand side of `bs` (host vectors: pinned
        torch CPU tensors for full overlap, numpy arrays accepted) -- the multiple-right-hand-side form of
        `aspreconditioner().matvec` / `solve(b, maxiter=k)` without residual norms.  Three streams pipeline the work of
        consecutive right-hand sides: while the cycle of b_k runs, b_{k+1} crosses PCIe host->device and x_{k-1}
        device->host (full duplex), over two device buffer pairs.  Every x_k is the same bits as
        solve(b_k, maxiter=maxiter, _norms=False).  -> list of host tensors (`out`: preallocated pinned tensors to fill;
        default: fresh pinned tensors)."""
        cycle = str(cycle).upper()
        if cycle not in _CYCLE:
            raise TypeError('solve_many: cycle must be V, W or F (got %s)' % cycle)
        if len(self.levels) == 1:
            raise NotImplementedError('solve_many needs a multilevel hierarchy')
        if self._engine is None:
            self._build_engine()
        n = self.levels[0].A.shape[0]
        dev = self._x.device
        if getattr(self, '_pipe', None) is None:
            self._pipe = {'x': [self._x, torch.zeros_like(self._x)], 'b': [self._b, torch.zeros_like(self._b)],
                          'h2d': torch.cuda.Stream(device=dev), 'd2h': torch.cuda.Stream(device=dev)}
        P = self._pipe
        comp, h2d, d2h = self._stream, P['h2d'], P['d2h']
        cur = torch.cuda.current_stream()
        for s_ in (comp, h2d, d2h):
            s_.wait_stream(cur)
        b_ready = [torch.cuda.Event(), torch.cuda.Event()]
        x_ready = [None, None]
        x_copied = [None, None]
        outs = []
        for k, bk in enumerate(bs):
            slot = k & 1
            bt = bk if isinstance(bk, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(np.ravel(bk), np.float64))
            if bt.numel() != n:
                raise ValueError('b has invalid dimensions')
            xo = out[k] if out is not None else torch.empty(n, dtype=torch.float64, pin_memory=True)
            xd, bd = P['x'][slot], P['b'][slot]
            if x_ready[slot] is not None:
                h2d.wait_event(x_ready[slot])             # the cycle two right-hand sides ago has finished reading b
            with torch.cuda.stream(h2d):
                bd.copy_(bt.reshape(-1), non_blocking=True)
                b_ready[slot].record(h2d)
            comp.wait_event(b_ready[slot])
            if x_copied[slot] is not None:
                comp.wait_event(x_copied[slot])           # ... and its x has left for the host
            with torch.cuda.stream(comp):
                for it in range(maxiter):
                    _lib.call('b200amg_vcycle_run_zero' if it == 0 else 'b200amg_vcycle_run', self._engine, _ptr(xd), _ptr(bd),
                              _CYCLE[cycle], 1, comp.cuda_stream)
                if maxiter == 0:
                    xd.zero_()
                x_ready[slot] = torch.cuda.Event()
                x_ready[slot].record(comp)
            d2h.wait_event(x_ready[slot])
            with torch.cuda.stream(d2h):
                xo.copy_(xd, non_blocking=True)
                x_copied[slot] = torch.cuda.Event()
                x_copied[slot].record(d2h)
            b_ready[slot] = torch.cuda.Event()
            outs.append(xo)
        d2h.synchronize()
        comp.synchronize()
        cur.wait_stream(comp)
        return outs

    def _keep_dotwork(self):
        if not hasattr(self, '_dotwork'):
            self._dotwork = torch.empty(1024, dtype=torch.float64, device=self._x.device)
        return _ptr(self._dotwork)

    def _solve_one_level(self, b, x0, tol, maxiter, callback, residuals, as_numpy):
        A = self.levels[0].A
        bt = b if isinstance(b, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(np.ravel(b), np.float64)).to(_dev())
        bt = bt.reshape(-1)
        x = torch.zeros_like(bt) if x0 is None else (x0.clone() if isinstance(x0, torch.Tensor) else
                                                      torch.from_numpy(np.array(np.ravel(x0), np.float64)).to(_dev()))
        normb = float(torch.linalg.vector_norm(bt).item())
        if normb != 0:
            tol = tol * normb
        if residuals is None:
            residuals = []
        else:
            residuals[:] = []
        rn = lambda: float(torch.linalg.vector_norm(A.residual(x, bt)).item())
        residuals.append(rn())
        while len(residuals) <= maxiter and residuals[-1] > tol:
            x = self.coarse_solver(A, bt)
            residuals.append(rn())
            if callback is not None:
                callback(x.cpu().numpy() if as_numpy else x)
        return x.cpu().numpy() if as_numpy else x
