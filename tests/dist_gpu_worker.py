"""worker of tests/test_gpu_distributed.py: one process per GPU (NCCL), row-partitioned cycle through
libb200amg vs the single-GPU engine on rank 0."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    grid, thresh, cycles = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    mcgs = len(sys.argv) > 4 and sys.argv[4] == 'mcgs'
    block = len(sys.argv) > 4 and sys.argv[4] == 'bsr'
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    rank, world = dist.get_rank(), dist.get_world_size()
    import pyamg_b200 as amg
    from pyamg_b200.distributed import DistributedSolver
    ml, b = None, None
    if rank == 0 and block:
        # P1 elasticity on grid^3 vertices: BSR 3 x 3 on the finest level, 6 x 6 below (6 rigid-body modes), symmetric
        # colour-ordered block Gauss-Seidel -- the hierarchy shape of BASELINE config 4
        A, B = amg.gallery.linear_elasticity_tets(grid)
        np.random.seed(2)
        mc = ('multicolor_gauss_seidel', {'sweep': 'symmetric'})
        ml = amg.smoothed_aggregation_solver(A, B, presmoother=mc, postsmoother=mc, max_coarse=10)
        b = np.random.rand(A.shape[0])
    elif rank == 0:
        A = amg.gallery.poisson((grid * world, grid, grid))
        np.random.seed(2)
        jac = ('jacobi', {'omega': 4.0 / 3.0})
        ml = amg.smoothed_aggregation_solver(A, presmoother=jac, postsmoother=jac, max_coarse=40, improve_candidates=None)
        b = np.random.rand(A.shape[0])
        if mcgs:            # symmetric multicolour Gauss-Seidel over GLOBAL MIS colours on every level
            mc = ('multicolor_gauss_seidel', {'sweep': 'symmetric'})
            amg.change_smoothers(ml, mc, mc)
    # the public call: rank 0 holds the hierarchy, the others a participant; solve(distributed=True) is collective
    from pyamg_b200.distributed import participant
    res = []
    if rank == 0:
        x = ml.solve(torch.from_numpy(b).cuda(), tol=1e-30, maxiter=cycles, residuals=res, distributed=True)
        S = ml.distributed_solver()
    else:
        me = participant()
        x = me.solve(tol=1e-30, maxiter=cycles, residuals=res, distributed=True)
        S = me.solver()
    assert isinstance(S, DistributedSolver)
    if thresh != 200:                                     # explicit agglomeration threshold: a second solver
        S = DistributedSolver.from_ml(ml, agglomerate_below=thresh)
        x = S.solve(b, tol=1e-30, maxiter=cycles, residuals=res)
    ok = True
    if rank == 0:
        ref = []
        xr = ml.solve(torch.from_numpy(b).cuda(), tol=1e-30, maxiter=cycles, residuals=ref)
        same = bool(torch.equal(x, xr))
        ok = same and np.allclose(res, ref, rtol=1e-12, atol=0)
        print('world', world, 'partitioned levels', len(S.levels), 'of', S.nlev, 'exchanges', S.exchanges,
              'bit-identical x', same, 'res', res[-1], ref[-1], flush=True)
    # slab-local I/O: every rank loads its own slab of b / stores its own slab of x through pinned host memory
    S.load(b)
    lo, hi = S.slab()
    b_loc = torch.empty(hi - lo, dtype=torch.float64).pin_memory()
    b_loc.copy_(S.levels[0].b[:hi - lo])
    torch.cuda.synchronize()
    S.load_local(b_loc)
    for _ in range(cycles):
        S.cycle()
    x_loc = S.store_local(torch.empty(hi - lo, dtype=torch.float64).pin_memory())
    torch.cuda.synchronize()
    S.check_halo()
    xg = S.gather_x()
    if rank == 0:
        ok = ok and bool(torch.equal(xg, xr)) and bool(torch.equal(x_loc, xr[lo:hi].cpu()))
        print('slab-local x bit-identical', bool(torch.equal(x_loc, xr[lo:hi].cpu())), flush=True)
    flag = torch.tensor([1 if ok else 0], device='cuda')
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == '__main__':
    main()
