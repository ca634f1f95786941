"""world_size-2 (and 3) gloo runs of the row-partitioned V-cycle's HOST logic on CPU (no GPU needed):
partition.py + distributed.py with a checker backend; the gathered x must equal the single-process
oracle's x bit for bit (slab arithmetic is the global arithmetic), residual histories to 1e-12."""
import os
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def run(world, grid, thresh, cycles, port, *extra):
    env = dict(os.environ, MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), OMP_NUM_THREADS='1')
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(world),
           '--master-addr', '127.0.0.1', '--master-port', str(port), os.path.join(HERE, 'dist_cpu_worker.py'),
           str(grid), str(thresh), str(cycles)] + list(extra)
    return subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)


@pytest.mark.parametrize('world,grid,thresh', [(2, 10, 50),       # levels 0 and 1 partitioned, rest agglomerated
                                               (2, 9, 100000),    # only level 0 partitioned
                                               (3, 8, 30),        # uneven slabs
                                               (2, 14, 5)])       # every level but the coarsest partitioned (3 of 4):
                                                                  # the shape of the 8-GPU 512^3 run, only the dense
                                                                  # coarse solve stays on rank 0
def test_partitioned_cycle_equals_global(world, grid, thresh):
    r = run(world, grid, thresh, 4, 29500 + world * 10 + grid)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert 'bit-identical x True' in r.stdout
    if (grid, thresh) == (14, 5):
        assert 'partitioned levels 3 of 4' in r.stdout


@pytest.mark.parametrize('world,grid,thresh', [(2, 10, 50), (3, 9, 30)])
def test_partitioned_multicolour_gauss_seidel_equals_global(world, grid, thresh):
    """colour-ordered Gauss-Seidel with GLOBAL colours on the partitioned levels (one halo refresh per
    colour): bit-identical to the single-process sweep"""
    r = run(world, grid, thresh, 3, 29600 + world * 10 + grid, 'mcgs')
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert 'bit-identical x True' in r.stdout


@pytest.mark.parametrize('world,grid,thresh,mode', [(2, 8, 100, 'bsr_jacobi'), (2, 8, 100, 'bsr_mcgs'), (3, 7, 100, 'bsr_mcgs')])
def test_partitioned_block_levels_equal_global(world, grid, thresh, mode):
    """BSR levels (3 x 3 blocks on the finest level, 6 x 6 below, 3 x 6 / 6 x 6 transfer blocks) cut between block rows:
    bsr_matvec / bsr_jacobi / block-row colour-ordered bsr_gauss_seidel on slabs, bit-identical to the single-process cycle"""
    r = run(world, grid, thresh, 3, 29700 + world * 10 + grid + (5 if mode == 'bsr_mcgs' else 0), mode)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert 'bit-identical x True' in r.stdout
    assert 'partitioned levels 2 of 3' in r.stdout
