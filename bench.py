#!/usr/bin/env python
"""bench.py -- V-cycles/sec of the solve-phase hot path on BASELINE.json's headline workload:
3-D Poisson 256^3 (7-point), smoothed aggregation, Jacobi(4/3) pre/post smoothing, fp64.

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --steps K --warmup W    # the reference's CPU path (oracle/_ref)

A "step" is ONE V(1,1)-cycle of multilevel_solver.__solve on the resident hierarchy.
`value` = cycles/s with (x, b) resident in HBM; `e2e` = cycles/s through the public
solve(b_host, maxiter=1) call with the host->device copy of b and the device->host copy of x
inside the timed region.  The hierarchy itself is built (untimed) by the device setup
(pyamg_b200.smoothed_aggregation_solver).  Prints ONE JSON line.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

JAC = ('jacobi', {'omega': 4.0 / 3.0})


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--grid', type=int, default=int(os.environ.get('B200AMG_GRID', '256')),
                    help='grid points per dimension (BASELINE config: 256)')
    ap.add_argument('--cpu-cycles', type=int, default=3, help='V-cycles of the CPU baseline sample')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(',')])

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for name, v in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), r[3:7]):
                if v.lower().startswith('active'):
                    reasons.add(name)
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'samples': len(sm), 'reasons': sorted(reasons)}


def measured_peak():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        try:
            return float(json.load(open(p))['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
        except Exception:
            pass
    return 6650.0, 'fallback (B200_PROFILING.md)'


def jacobi_bytes(n, nnz):
    """algorithmic bytes of one CSR Jacobi sweep: A once (12 nnz + 4(n+1)), x gathered
    once (8n), b (8n), x_new written (8n)"""
    return 12 * nnz + 4 * (n + 1) + 24 * n


def spmv_bytes(n_row, n_col, nnz):
    return 12 * nnz + 4 * (n_row + 1) + 8 * n_col + 8 * n_row


def vcycle_bytes(ml):
    """unfused algorithmic bytes of one V(1,1)-cycle (SURVEY 8d formula) over the hierarchy"""
    tot = 0
    for lvl in ml.levels[:-1]:
        n, nnz = lvl.A.shape[0], lvl.A.nnz
        nc = lvl.P.shape[1]
        tot += 3 * (12 * nnz + 4 * n + 24 * n)
        tot += 12 * lvl.R.nnz + 4 * nc + 8 * n + 8 * nc
        tot += 12 * lvl.P.nnz + 4 * n + 8 * nc + 16 * n
    tot += 8 * ml.levels[-1].A.shape[0] ** 2
    return tot


def host_levels(ml):
    """device hierarchy -> scipy matrices + resolved smoother specs for the CPU oracle"""
    from oracle import multilevel as oml
    out = []
    for lvl in ml.levels:
        h = oml.Level()
        h.A = lvl.A.to_scipy()
        if hasattr(lvl, 'P'):
            h.P, h.R = lvl.P.to_scipy(), lvl.R.to_scipy()
            h.pre = ('jacobi', {'omega': lvl.presmoother.omega, 'iterations': lvl.presmoother.iterations})
            h.post = ('jacobi', {'omega': lvl.postsmoother.omega, 'iterations': lvl.postsmoother.iterations})
        out.append(h)
    return out


def time_cpu_cycles(levels, b, cycles, warm=0):
    """the reference's CPU path (scipy SpMV + amg_core smoothers via oracle/_ref when built) as-is,
    i.e. including the fork's second P*coarse_x per level (multilevel.py:548)"""
    import oracle
    from oracle import multilevel as oml
    S = oml.Solver(levels, core=oracle.core, wasted_prolongation=True)
    S.coarse_matrix()
    x = np.zeros_like(b)
    for _ in range(warm):
        S._cycle(0, x, b, 'V')
    t0 = time.perf_counter()
    for _ in range(cycles):
        S._cycle(0, x, b, 'V')
    dt = time.perf_counter() - t0
    return dt / cycles, oracle.kind
