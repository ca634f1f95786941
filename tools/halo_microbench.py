"""What does ONE halo refresh cost when nothing else runs?  torchrun --nproc-per-node 2 tools/halo_microbench.py [grid]
Builds the row-partitioned solver of a (2g, g, g) Poisson problem, captures 200 back-to-back refreshes of the level-0
(and level-1) x halo into a CUDA graph and times replays with CUDA events: the lock-step floor of the peer-memory push /
signal / wait kernel (csrc/halo.cu), to set against the ~23 us per refresh that a distributed cycle pays."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    g = int(sys.argv[1]) if len(sys.argv) > 1 else 128
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    rank, world = dist.get_rank(), dist.get_world_size()
    import pyamg_b200 as amg
    from pyamg_b200.distributed import DistributedSolver
    ml = None
    if rank == 0:
        A = amg.gallery.poisson((g * world, g, g))
        np.random.seed(1)
        jac = ('jacobi', {'omega': 4.0 / 3.0})
        ml = amg.smoothed_aggregation_solver(A, presmoother=jac, postsmoother=jac, max_coarse=500, improve_candidates=None)
    S = DistributedSolver.from_ml(ml)
    out = {}
    for k, L in enumerate(S.levels):
        for name, plan, vec in (('A', L.planA, L.x), ('R', L.planR, L.r)):
            for _ in range(3):
                S.exchange(plan, vec)
            torch.cuda.synchronize(); dist.barrier()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                for _ in range(200):
                    S.exchange(plan, vec)
            torch.cuda.synchronize(); dist.barrier()
            best = 1e9
            for _ in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                dist.barrier(); torch.cuda.synchronize()
                e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize()
                best = min(best, e0.elapsed_time(e1) * 1e3 / 200)
            out['L%d plan%s (%d sent, %d halo)' % (k, name, int(plan.send_idx.numel()), plan.n_halo)] = round(best, 2)
            gr.reset()
    S.check_halo()
    if rank == 0:
        print('us per refresh, back to back, graph replay:', out, flush=True)
    S.release_graph()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
