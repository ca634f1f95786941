"""Krylov acceleration around the device cycle (pyamg/multilevel.py:381-422 `accel=`; pyamg/krylov/_cg.py).

Preconditioned CG runs on the device WITHOUT the host in the loop: alpha, beta, the curvature checks and the stopping
test are computed by device kernels from device scalars (b200amg_pcg_*), A p and b - A x run on the engine's layout of
the finest operator (b200amg_vcycle_apply_op), the preconditioner is one cycle from a zero guess
(b200amg_vcycle_run_zero).  One whole ITERATION -- A p, both dots, the scalar kernels, the vector updates and the
cycle's ~35 kernels -- is captured once as a CUDA graph over vectors and a state array that live on the solver, and
replayed (B200AMG_PCG_GRAPH=0: ten Python-driven launches per iteration instead).  The host enqueues `check_every`
iterations ahead and then reads one small state array; after convergence the already-enqueued iterations are no-ops, so x is the iterate the reference
returns.  Semantics follow _cg.py:109-170: preconditioned norm sqrt(<r, z>), tolerance scaled by the initial one, r
recomputed as b - A x every 8th iteration, abort on pAp < 0 / rz < 0 / rz == 0."""
import numpy as np
import torch

from . import _lib
from .sparse import _ptr

_HIST0 = 16
_HIST_MAX = 1 << 20
RECOMPUTE_R = 8
_ITERATION_GRAPHS = __import__('os').environ.get('B200AMG_PCG_GRAPH', '1') != '0'     # A/B switch


def cg(ml, b, x0=None, tol=1e-5, maxiter=None, cycle='V', callback=None, residuals=None, check_every=4):
    """pyamg/krylov/_cg.py on the device.  -> (x, info): info 0 converged, -1 aborted (indefinite matrix or
    preconditioner, singular preconditioner), maxiter when the iteration limit was reached -- the reference's codes.
    `residuals` receives the PRECONDITIONED residual norms sqrt(<r, M r>), the initial one included."""
    cycle = str(cycle).upper()
    if cycle == 'AMLI':
        raise ValueError('AMLI cycles require acceleration (accel) to be fgmres, or no acceleration')
    A = ml.levels[0].A
    n = A.shape[0]
    dev = A.data.device
    as_numpy = not isinstance(b, torch.Tensor)
    if maxiter is None:
        maxiter = int(1.3 * n) + 2
    elif maxiter < 1:
        raise ValueError('Number of iterations must be positive')
    if ml._engine is None and len(ml.levels) > 1:
        ml._build_engine()
    if len(ml.levels) == 1:
        raise NotImplementedError('cg around a one-level hierarchy: call the coarse solver directly')
    cyc = {'V': 0, 'W': 1, 'F': 2}[cycle]
    cur = torch.cuda.current_stream()
    st = ml._stream
    st.wait_stream(cur)
    with torch.cuda.stream(st):
        s = st.cuda_stream
        # every vector, the right-hand side and the state array live on the solver: the preconditioner's CUDA graph is
        # keyed by the (z, r) addresses, and one ITERATION is replayed as a CUDA graph of its own (below)
        # history slots: fixed (the iteration graphs hold the state array's address); a run of more than a million
        # iterations keeps iterating, its residual list stops growing there
        cap = min(int(maxiter) + 2, _HIST_MAX)
        pc = getattr(ml, '_pcg', None)
        if pc is None or pc['n'] != n:
            pc = ml._pcg = {'n': n, 'vec': tuple(torch.empty(n, dtype=torch.float64, device=dev) for _ in range(6)),
                            'state': torch.zeros(_HIST0 + _HIST_MAX, dtype=torch.float64, device=dev), 'graphs': {}}
        r, z, Ap, p, x, bt = pc['vec']
        state = pc['state']
        state.zero_()
        if (np.size(b) if as_numpy else b.numel()) != n:
            raise ValueError('b has invalid dimensions')
        bt.copy_(torch.from_numpy(np.ascontiguousarray(np.ravel(b), np.float64)) if as_numpy else b.reshape(-1), non_blocking=True)
        if x0 is None:
            x.zero_()
        elif isinstance(x0, torch.Tensor):
            x.copy_(x0.reshape(-1))
        else:
            x.copy_(torch.from_numpy(np.array(np.ravel(x0), np.float64)))
        work = ml._keep_dotwork()
        E = ml._engine

        def apply_A(op, xin, bin_, out):
            _lib.call('b200amg_vcycle_apply_op', E, 0, op, _ptr(xin), _ptr(bin_) if bin_ is not None else 0, _ptr(out), s)

        def precondition(rin, zout, graph=1):
            _lib.call('b200amg_vcycle_run_zero', E, _ptr(zout), _ptr(rin), cyc, graph, s)

        def dot_to(u, v, slot):
            _lib.call('b200amg_dot', n, _ptr(u), _ptr(v), _ptr(state) + 8 * slot, work, s)

        # setup (_cg.py:109-131)
        apply_A(2, x, bt, r)
        precondition(r, z)
        p.copy_(z)
        dot_to(r, z, 0)
        dot_to(bt, bt, 2)
        head = state[:3].cpu().numpy()
        rz, normb = float(head[0]), float(np.sqrt(head[2]))
        normr = float(np.sqrt(rz)) if rz >= 0 else float('nan')
        if residuals is not None:
            residuals[:] = [normr]
        if normb == 0.0:
            normb = 1.0
        out = lambda: (x.cpu().numpy().reshape(np.shape(b)) if as_numpy else x.clone())   # noqa: E731
        if normr < tol * normb:
            cur.wait_stream(st)
            return out(), 0
        tol_abs = tol * normr if normr != 0.0 else tol
        init = torch.zeros(_HIST0 + 1, dtype=torch.float64)
        init[0], init[5], init[8], init[9], init[_HIST0] = rz, tol_abs, float(maxiter), float(cap), normr
        state[:_HIST0 + 1].copy_(init.to(dev))

        def body(recompute, graph):
            apply_A(0, p, None, Ap)
            dot_to(Ap, p, 1)
            _lib.call('b200amg_pcg_scalars', _ptr(state), 0, s)
            _lib.call('b200amg_pcg_update_xr', n, _ptr(state), _ptr(p), _ptr(Ap), _ptr(x), _ptr(r), 0 if recompute else 1, s)
            if recompute:
                apply_A(2, x, bt, r)
            precondition(r, z, graph)
            dot_to(r, z, 2)
            _lib.call('b200amg_pcg_scalars', _ptr(state), 1, s)
            _lib.call('b200amg_pcg_update_p', n, _ptr(state), _ptr(z), _ptr(p), s)

        def iteration_graph(recompute):
            """one whole iteration (about 40 kernels with the cycle) as ONE CUDA graph: a replay costs one launch
            instead of ten Python-driven ones plus a nested graph launch"""
            key = (getattr(E, 'value', E), cyc, bool(recompute))       # (a rebuilt engine gets graphs of its own)
            g = pc['graphs'].get(key)
            if g is None:
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=st):    # capture records, it does not execute: no state is touched
                    body(recompute, 0)                  # the cycle's kernels go straight into this capture
                pc['graphs'][key] = g
            return g

        use_graphs = _ITERATION_GRAPHS

        def enqueue(it):
            recompute = not (it % RECOMPUTE_R and it > 0)
            if use_graphs:
                iteration_graph(recompute).replay()
            else:
                body(recompute, 1)

        step = 1 if callback is not None else max(1, int(check_every))
        enq, seen, flag = 0, 0, 0.0
        while flag == 0.0:
            for _ in range(min(step, maxiter - enq)):
                enqueue(enq)
                enq += 1
            hs = state[:_HIST0 + min(enq, cap - 1) + 1].cpu().numpy()        # one D2H read per batch
            flag, done = float(hs[6]), int(hs[7])
            if residuals is not None:
                residuals.extend(float(v) for v in hs[_HIST0 + seen + 1:_HIST0 + done + 1])
            if callback is not None:
                for _ in range(done - seen):
                    callback(x.cpu().numpy() if as_numpy else x)
            seen = done
            if enq >= maxiter and flag == 0.0:
                flag = 2.0
        if flag < 0:
            import warnings
            warnings.warn({-1.0: 'Indefinite matrix detected in CG, aborting', -2.0: 'Indefinite preconditioner detected in CG, aborting',
                           -3.0: 'Singular preconditioner detected in CG, ceasing iterations'}[flag])
        info = 0 if flag == 1.0 else (-1 if flag < 0 else seen)
    cur.wait_stream(st)
    return out(), info


def accelerated_solve(ml, accel, b, x0, tol, maxiter, cycle, callback, residuals):
    """multilevel.py:381-422: accel names a pyamg-style solver (here only 'cg' runs on the device) or is a callable;
    tol, maxiter, callback and the residual list are forwarded like the reference forwards them."""
    if str(cycle).upper() == 'AMLI' and accel != 'fgmres':
        raise ValueError('AMLI cycles require acceleration (accel) to be fgmres, or no acceleration')
    if accel == 'cg' or accel is cg:
        return cg(ml, b, x0=x0, tol=tol, maxiter=maxiter, cycle=cycle, callback=callback, residuals=residuals)[0]
    if callable(accel):
        # a host solver (scipy.sparse.linalg style, or pyamg.krylov style) driving the device preconditioner
        M = ml.aspreconditioner(cycle=cycle)
        A = ml.levels[0].A.to_scipy()
        try:                        # PyAMG style interface, which has a residuals parameter (multilevel.py:399-401)
            return accel(A, b, x0=x0, tol=tol, maxiter=maxiter, M=M, callback=callback, residuals=residuals)[0]
        except TypeError:
            pass
        cb = callback
        if residuals is not None:   # scipy style: the history comes from a callback (multilevel.py:402-420)
            xs = np.zeros(A.shape[0]) if x0 is None else np.ravel(x0)
            residuals[:] = [float(np.linalg.norm(np.ravel(b) - A * xs))]

            def callback(xk):       # noqa: F811
                residuals.append(float(xk) if np.isscalar(xk) else float(np.linalg.norm(np.ravel(b) - A * np.ravel(xk))))
                if cb is not None:
                    cb(xk)
        try:
            return accel(A, b, x0=x0, tol=tol, maxiter=maxiter, M=M, callback=callback)[0]
        except TypeError:           # scipy >= 1.12 renamed tol to rtol
            return accel(A, b, x0=x0, rtol=tol, maxiter=maxiter, M=M, callback=callback)[0]
    raise NotImplementedError("accel=%r: only 'cg' runs on the device" % (accel,))
