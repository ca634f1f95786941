"""CPU tests (-m "not gpu") of the bench.py contract pieces that run without a GPU: the reference arm (the
reference's CPU path = oracle/_ref or the C port, hierarchy from the CPU setup restatements) for both benchable
configs, and the byte models the roofline numbers are computed from."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize('extra', [['--grid', '14'], ['--config', '3', '--grid', '60'], ['--config', '4', '--grid', '10']])
def test_reference_arm_line(extra):
    env = dict(os.environ, B200AMG_REF_BUDGET_S='5', CUDA_VISIBLE_DEVICES='')
    r = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--impl', 'reference', '--steps', '2', '--warmup', '1'] + extra,
                       env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert d['impl'] == 'reference' and d['unit'] == 'V-cycles/s' and d['higher_is_better'] is True
    assert d['value'] > 0 and abs(d['value'] - 1000.0 / d['ms_per_step']) <= 1e-6 * d['value']
    assert d['cpu_baseline']['kind'] in ('reference', 'port', 'port+reference-kernels') and d['cpu_baseline']['cores'] == 1
    assert d['cpu_baseline']['value'] == d['value'] and d['e2e']['value'] == d['value']
    assert d['e2e']['h2d_bytes_per_step'] == 0 and d['e2e']['d2h_bytes_per_step'] == 0 and d['gpu_launches'] == 0
    assert d['vs_baseline'] is None and len(d['config']['levels']) >= 2 and 'workload' in d['config']
    # the `config` object is the workload only -- the same keys the GPU arm prints (bench.workload_config)
    assert sorted(d['config']) == ['block_levels', 'colours_per_level', 'grid', 'l2', 'levels', 'workload']
    if d['cpu_baseline']['kind'] == 'reference':
        # the stock reference drove the run: its package came from /root/reference or the installed copy, and (config 4's
        # matrix generator aside) nothing of this repository's product or of torch was imported into the process
        assert 'stock pyamg' in d['cpu_baseline']['sample'] and '__solve' in d['cpu_baseline']['sample']
        if '--config' not in extra or extra[extra.index('--config') + 1] != '4':
            assert d['detail']['modules_loaded'] == []


def test_reference_arm_runs_from_the_installed_copy(tmp_path):
    """baseline/_ref (oracle/install_ref.py) is what the GPU box has instead of /root/reference"""
    if not os.path.isdir(os.path.join(ROOT, 'baseline', '_ref', 'pyamg')):
        pytest.skip('no installed reference copy (build container only)')
    env = dict(os.environ, B200AMG_REF_BUDGET_S='5', CUDA_VISIBLE_DEVICES='', PYAMG_REF_ROOT=os.path.join(ROOT, 'baseline', '_ref'))
    r = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--impl', 'reference', '--steps', '2', '--warmup', '1',
                        '--grid', '14'], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert d['cpu_baseline']['kind'] == 'reference' and 'baseline/_ref' in d['cpu_baseline']['sample']


def test_reference_arm_without_any_reference_package_falls_back_to_the_restatement(tmp_path):
    env = dict(os.environ, B200AMG_REF_BUDGET_S='5', CUDA_VISIBLE_DEVICES='', PYAMG_REF_ROOT=str(tmp_path))
    r = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--impl', 'reference', '--steps', '2', '--warmup', '1',
                        '--grid', '14'], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert d['cpu_baseline']['kind'] in ('port', 'port+reference-kernels') and d['value'] > 0


def test_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK='1', WORLD_SIZE='2', LOCAL_RANK='1', CUDA_VISIBLE_DEVICES='')
    r = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--impl', 'reference', '--gpus', '2'],
                       env=env, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ''


def test_byte_models():
    sys.path.insert(0, ROOT)
    import bench
    n, nnz = 16777216, 117047296
    # CSR SpMV of the 256^3 operator = 1.740 GB; Jacobi adds b (8n)
    assert bench.spmv_bytes(n, n, nnz) == 12 * nnz + 4 * (n + 1) + 16 * n == 1740111876
    assert bench.jacobi_bytes(n, nnz) == bench.spmv_bytes(n, n, nnz) + 8 * n
    assert bench.jacobi_bytes(n, nnz, 5) == 25 * n                       # row-pattern layout: 1 B/row + b, x, x_new
    assert bench.gs_sweep_bytes(n, nnz) == bench.spmv_bytes(n, n, nnz) + 8 * n + 4 * n
    assert bench.grid_dims(256, 8) == (512, 512, 512) and bench.grid_dims(256, 2) == (512, 256, 256)


def test_block_byte_model_reduces_to_the_scalar_one():
    """vcycle_bytes on a hierarchy of 1x1 blocks is SURVEY 8d's scalar formula; 3x3 blocks cost 76 bytes each"""
    sys.path.insert(0, ROOT)
    import bench

    class M:
        def __init__(self, nbr, nbc, nblk, R, C):
            self.n_brow, self.nblocks, self.blocksize = nbr, nblk, (R, C)
            self.shape, self.nnz, self.is_scalar = (nbr * R, nbc * C), nblk * R * C, (R, C) == (1, 1)

    class L:
        pass

    class H:
        pass
    n, nnz, nc, pn = 1000, 6800, 120, 3600
    l0, l1 = L(), L()
    l0.A, l0.P, l0.R = M(n, n, nnz, 1, 1), M(n, nc, pn, 1, 1), M(nc, n, pn, 1, 1)
    l1.A = M(nc, nc, 900, 1, 1)
    ml = H()
    ml.levels = [l0, l1]
    want = 3 * (12 * nnz + 4 * n + 24 * n) + (12 * pn + 4 * nc + 8 * n + 8 * nc) + (12 * pn + 4 * n + 8 * nc + 16 * n) + 8 * nc ** 2
    assert bench.vcycle_bytes(ml) == want
    want_gs = 4 * bench.gs_sweep_bytes(n, nnz) + (12 * nnz + 28 * n) + (12 * pn + 4 * nc + 8 * n + 8 * nc) + (12 * pn + 4 * n + 8 * nc + 16 * n) + 8 * nc ** 2
    assert bench.vcycle_bytes(ml, 2) == want_gs
    assert bench.block_bytes(10, 3, 3) == 760 and bench.block_bytes(10, 1, 1) == 120
    assert bench.bsr_gs_sweep_bytes(100, 1500, 3) == 1500 * 76 + 4 * 101 + 3 * 8 * 300 + 400
