"""GPU test of the row-partitioned cycle (pyamg_b200/distributed.py) through libb200amg + NCCL:
world = min(2, visible GPUs) processes (a 1-rank world still runs every slab kernel / op handle / the
agglomeration hand-off on the device); gathered x must be bit-identical to the single-GPU engine's."""
import os
import subprocess
import sys

import pytest

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize('grid,thresh,extra', [(24, 200, []), (16, 10 ** 9, []), (24, 200, ['mcgs']),
                                               (14, 100, ['bsr'])])      # block levels (3 x 3, 6 x 6) cut between block rows
def test_partitioned_cycle_on_device(grid, thresh, extra):
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    world = min(2, torch.cuda.device_count())
    port = 29700 + grid + 50 * len(extra) + (7 if extra == ['bsr'] else 0)
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(world),
           '--master-addr', '127.0.0.1', '--master-port', str(port), os.path.join(HERE, 'dist_gpu_worker.py'),
           str(grid), str(thresh), '4'] + extra
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert 'bit-identical x True' in r.stdout
