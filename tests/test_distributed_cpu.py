"""world_size-2 (and 3) gloo runs of the row-partitioned V-cycle's HOST logic on CPU (no GPU needed):
partition.py + distributed.py with a checker backend; the gathered x must equal the single-process
oracle's x bit for bit (slab arithmetic is the global arithmetic), residual histories to 1e-12."""
import os
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def run(world, port, *scenarios):
    env = dict(os.environ, MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), OMP_NUM_THREADS='1')
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(world),
           '--master-addr', '127.0.0.1', '--master-port', str(port), os.path.join(HERE, 'dist_cpu_worker.py')] + list(scenarios)
    return subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900)


# scenario = "grid:thresh:cycles[:mode]"; the scenarios of one world size share a torchrun start
JACOBI = {2: ['10:50:4',        # levels 0 and 1 partitioned, rest agglomerated
              '9:100000:4',     # only level 0 partitioned
              '14:5:4'],        # every level but the coarsest partitioned (3 of 4): the shape of the 8-GPU 512^3 run, only
                                # the dense coarse solve stays on rank 0
          3: ['8:30:4']}        # uneven slabs
MCGS = {2: ['10:50:3:mcgs'], 3: ['9:30:3:mcgs']}
BLOCK = {2: ['8:100:3:bsr_jacobi', '8:100:3:bsr_mcgs'], 3: ['7:100:3:bsr_mcgs']}


@pytest.mark.parametrize('world', [2, 3])
def test_partitioned_cycles_equal_global(world):
    """Jacobi cycles on slabs; colour-ordered Gauss-Seidel with GLOBAL colours (one halo refresh per colour);
    BSR levels (3 x 3 blocks on the finest level, 6 x 6 below, 3 x 6 / 6 x 6 transfer blocks) cut between block rows with
    bsr_matvec / bsr_jacobi / block-row colour-ordered bsr_gauss_seidel -- every scenario bit-identical to the
    single-process cycle, slab-local load / store included"""
    scen = JACOBI[world] + MCGS[world] + BLOCK[world]
    r = run(world, 29500 + world * 10, *scen)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    for sc in scen:
        assert 'scenario %s ok' % sc in r.stdout, r.stdout[-3000:]
    assert r.stdout.count('bit-identical x True') == len(scen)
    if world == 2:
        assert 'partitioned levels 3 of 4' in r.stdout          # the 14:5:4 scenario
        assert r.stdout.count('partitioned levels 2 of 3') >= 2  # both block scenarios (and small scalar ones)
