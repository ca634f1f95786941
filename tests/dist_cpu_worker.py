"""worker of tests/test_distributed_cpu.py: world_size-2 gloo run of the row-partitioned cycle with a
CPU checker backend (oracle kernels on numpy views).  TEST INFRASTRUCTURE: exercises the host logic of
pyamg_b200/partition.py + distributed.py (slab cutting, column renumbering, halo plans, all-to-all
refresh, agglomeration gather/scatter); the product backend is the CUDA one."""
import os
import sys

import numpy as np
import scipy.sparse as sp
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))

import oracle                                       # noqa: E402
from oracle import multilevel as oml, sa as osa     # noqa: E402
from helpers_mats import poisson3d                  # noqa: E402
import importlib.util                               # noqa: E402


def _load(name):
    """import pyamg_b200.<name> WITHOUT importing the package (which needs the CUDA library loaded lazily only)"""
    import pyamg_b200.partition  # noqa: F401  (pure torch)
    return importlib.import_module('pyamg_b200.' + name)


class CpuOp:
    def __init__(self, ip, ix, dx, n_row, R=1, C=1):
        self.ip, self.ix, self.dx, self.n = ip.numpy(), ix.numpy(), dx.numpy(), int(n_row)
        self.ncol = int(self.ix.max()) + 1 if self.ix.size else 0
        self.R, self.C = int(R), int(C)

    def _apply_block(self, op, x, b, y, omega):
        """block rows: scipy bsr_matvec / amg_core.bsr_jacobi (the oracle's restatements)"""
        xn, yn = x.numpy(), y.numpy()
        nb, R, C = self.n, self.R, self.C
        n = nb * R
        if op in (0, 2, 3):
            t = np.zeros(n)
            oracle.port.bsr_matvec(nb, self.ncol, R, C, self.ip, self.ix, self.dx, np.ascontiguousarray(xn), t)
            yn[:n] = t if op == 0 else (b.numpy()[:n] - t if op == 2 else yn[:n] + t)
        else:
            xc, temp = xn.copy(), xn.copy()
            oracle.port.bsr_jacobi(self.ip, self.ix, self.dx, xc, np.ascontiguousarray(b.numpy()[:n]), temp, 0, nb, 1, R,
                                   np.array([omega]))
            yn[:n] = xc[:n]

    def apply(self, op, x, b, y, omega=0.0):
        if self.R * self.C > 1:
            return self._apply_block(op, x, b, y, omega)
        xn, yn = x.numpy(), y.numpy()
        n = self.n
        if op in (0, 2, 3):
            t = np.zeros(n)
            oracle.port.csr_matvec(n, self.ncol, self.ip, self.ix, self.dx, np.ascontiguousarray(xn), t)
            if op == 0:
                yn[:n] = t
            elif op == 2:
                yn[:n] = b.numpy()[:n] - t
            else:
                yn[:n] = yn[:n] + t
        else:
            xc, temp = xn.copy(), xn.copy()
            bb = np.ascontiguousarray(b.numpy()[:n])
            if op == 4:
                oracle.port.jacobi(self.ip, self.ix, self.dx, xc, bb, temp, 0, n, 1, np.array([omega]))
            else:
                oracle.port.bsr_jacobi(self.ip, self.ix, self.dx, xc, bb, temp, 0, n, 1, 1, np.array([omega]))
            yn[:n] = xc[:n]

    def diagonal(self):
        R = self.R
        d = np.zeros(self.n * R)
        blocks = self.dx.reshape(-1, R, self.C)
        for i in range(self.n):
            for k in range(self.ip[i], self.ip[i + 1]):
                if self.ix[k] == i:
                    d[i * R:(i + 1) * R] = np.diag(blocks[k])
        return torch.from_numpy(d)


class CpuBackend:
    device = torch.device('cpu')

    def make_op(self, ip, ix, dx, n_row, R=1, C=1):
        return CpuOp(ip, ix, dx, n_row, R, C)

    def scale_diag(self, omega, b, diag, x, flavor):
        n = diag.numel()
        bn, dn = b.numpy()[:n], diag.numpy()
        with np.errstate(divide='ignore', invalid='ignore'):
            v = (omega * bn) / dn if flavor else omega * (bn / dn)
        x.numpy()[:n] = np.where(dn != 0, v, 0.0)

    def gs_color(self, op, x, b, Id, p0, p1, flavor, backward=False):
        if p1 <= p0:
            return
        xn, bn = x.numpy(), np.ascontiguousarray(b.numpy()[:op.n * op.R])
        idx = np.ascontiguousarray(Id.numpy(), dtype=np.intc)
        xc = np.ascontiguousarray(xn)
        if flavor or op.R > 1:
            r0, r1, dr = (int(p1) - 1, int(p0) - 1, -1) if backward else (int(p0), int(p1), 1)
            oracle.port.bsr_gauss_seidel_indexed(op.ip, op.ix, op.dx, xc, bn, idx, r0, r1, dr, op.R)
        else:
            oracle.port.gauss_seidel_indexed(op.ip, op.ix, op.dx, xc, bn, idx, int(p0), int(p1), 1)
        xn[:] = xc

    def dot(self, x, y, n):
        return torch.tensor([float(np.dot(x.numpy()[:n], y.numpy()[:n]))], dtype=torch.float64)


class CoarseOracle:
    def __init__(self, levels):
        self.S = oml.Solver(levels, core=oracle.port, wasted_prolongation=False)

    def cycle_from_zero(self, b):
        bn = b.numpy()
        if len(self.S.levels) == 1:
            return torch.from_numpy(np.asarray(self.S.coarse_solve(self.S.levels[0].A, bn)).copy())
        x = np.zeros_like(bn)
        self.S._cycle(0, x, bn, 'V')
        return torch.from_numpy(x)


def t3(M):
    M = sp.csr_matrix(M) if M.format != 'bsr' else M
    return (torch.from_numpy(np.ascontiguousarray(M.indptr, np.int32)), torch.from_numpy(np.ascontiguousarray(M.indices, np.int32)),
            torch.from_numpy(np.ascontiguousarray(np.ravel(M.data), np.float64)))


def main():
    """argv: one or more scenarios "grid:thresh:cycles[:mode]" run one after the other in the same process group (a
    torchrun start costs more than a scenario); exit code 0 only if every scenario is bit-identical"""
    dist.init_process_group('gloo')
    bad = 0
    for spec in sys.argv[1:]:
        f = spec.split(':')
        ok = scenario(int(f[0]), int(f[1]), int(f[2]), f[3] if len(f) > 3 else '')
        if dist.get_rank() == 0:
            print('scenario', spec, 'ok' if ok else 'FAILED', flush=True)
        bad += 0 if ok else 1
    dist.destroy_process_group()
    sys.exit(0 if bad == 0 else 1)


def scenario(grid, thresh, cycles, mode):
    mcgs = mode in ('mcgs', 'bsr_mcgs')
    block = mode in ('bsr_jacobi', 'bsr_mcgs')
    rank = dist.get_rank()
    D = _load('distributed')
    levels_desc, coarse, host, b = None, None, None, None
    if rank == 0:
        A = poisson3d(grid); A.sort_indices()
        np.random.seed(5)
        jac = ('jacobi', {'omega': 4.0 / 3.0})
        if block:
            # a vector problem: 3 unknowns per grid point (BSR 3x3), 6 candidates -> coarse levels with 6 x 6 blocks,
            # P with 3 x 6 / 6 x 6 blocks (the block shapes of the elasticity hierarchy of BASELINE config 4)
            rng = np.random.RandomState(3)
            M = rng.rand(3, 3); M = M @ M.T + 3 * np.eye(3)
            A = sp.bsr_matrix(sp.kron(A, M), blocksize=(3, 3)); A.sort_indices()
            B = np.hstack([np.kron(np.ones((A.shape[0] // 3, 1)), np.eye(3)), rng.rand(A.shape[0], 3)])
            jac = ('jacobi', {'omega': 0.9, 'withrho': False})
            host = osa.smoothed_aggregation_hierarchy(A, B, presmoother=jac, postsmoother=jac, max_coarse=6, core=oracle.port,
                                                      improve_candidates=None)
        else:
            host = osa.smoothed_aggregation_hierarchy(A, presmoother=jac, postsmoother=jac, max_coarse=12, core=oracle.port,
                                                      improve_candidates=None)
        b = np.random.rand(A.shape[0])
        levels_desc = []
        for L in host[:-1]:
            d = {'n': L.A.shape[0], 'nc': L.P.shape[1], 'flavor': 0 if L.A.format == 'csr' else 1,
                 'pre': (1, L.pre[1]['omega'], 1), 'post': (1, L.post[1]['omega'], 1),
                 'A': t3(L.A), 'P': t3(L.P), 'R': t3(L.R)}
            if block:
                d['bs'], d['bsc'] = int(L.A.blocksize[0]), int(L.P.blocksize[1])
                assert L.P.blocksize[0] == d['bs'] and L.R.blocksize == (d['bsc'], d['bs'])
            if mcgs and block:      # block rows relaxed colour by colour (MIS colours of the BLOCK graph): bsr_gauss_seidel
                from oracle import relaxation as orx
                G = sp.csr_matrix((np.ones(len(L.A.indices)), L.A.indices, L.A.indptr), shape=(d['n'] // d['bs'],) * 2)
                idx, ptr = orx.color_ordering(orx.vertex_coloring_mis(G, oracle.port))
                L.pre = L.post = ('bsr_gauss_seidel_indexed', {'indices': idx, 'sweep': 'symmetric', 'iterations': 1})
                d['pre'] = d['post'] = (2, 1.0, 1, 'symmetric')
                d['Id'], d['color_ptr'], d['gs_csr'] = torch.from_numpy(np.ascontiguousarray(idx, np.int32)), [int(v) for v in ptr], 0
            elif mcgs:        # symmetric multicolour Gauss-Seidel over the MIS colours of the GLOBAL matrix graph
                from oracle import relaxation as orx
                idx, ptr = orx.color_ordering(orx.vertex_coloring_mis(sp.csr_matrix(L.A), oracle.port))
                L.pre = L.post = ('gauss_seidel_indexed', {'indices': idx, 'sweep': 'symmetric', 'iterations': 1})
                d['pre'] = d['post'] = (2, 1.0, 1, 'symmetric')
                d['Id'], d['color_ptr'] = torch.from_numpy(np.ascontiguousarray(idx, np.int32)), [int(v) for v in ptr]
            levels_desc.append(d)
        coarse = lambda ka: CoarseOracle(host[ka:])
    S = D.DistributedSolver(levels_desc, coarse, agglomerate_below=thresh, backend=CpuBackend())
    res = []
    x = S.solve(b, tol=1e-30, maxiter=cycles, residuals=res)
    ok = True
    if rank == 0:
        ref_res = []
        xr = oml.Solver(host, core=oracle.port, wasted_prolongation=False).solve(b, tol=1e-30, maxiter=cycles, residuals=ref_res)
        ok = bool(np.array_equal(x.numpy(), xr)) and np.allclose(res, ref_res, rtol=1e-12, atol=0)
        print('partitioned levels', len(S.levels), 'of', S.nlev, 'exchanges', S.exchanges, 'bit-identical x', np.array_equal(x.numpy(), xr),
              'res', res[-1], ref_res[-1], flush=True)
    # slab-local entry points: every rank loads its own slab of b and reads its own slab of x (no gather through rank 0)
    S.load(b)
    lo, hi = S.slab()
    b_loc = S.levels[0].b[:hi - lo].clone()
    S.load_local(b_loc)
    for _ in range(cycles):
        S.cycle()
    x_loc = S.store_local(torch.empty(hi - lo, dtype=torch.float64))
    xg = S.gather_x()
    S.check_halo()
    ok_loc = torch.tensor([1])
    if rank == 0:
        ok_loc[0] = int(np.array_equal(xg.numpy(), xr) and np.array_equal(x_loc.numpy(), xr[lo:hi]))
    flag = torch.tensor([1 if ok and int(ok_loc.item()) == 1 else 0])
    dist.broadcast(flag, 0)
    return int(flag.item()) == 1


if __name__ == '__main__':
    main()
