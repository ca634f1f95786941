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
    ap.add_argument('--config', type=int, default=2, choices=[2, 3, 4],
                    help='BASELINE.json configs[] entry (1-based as in SURVEY 8d): 2 = 3D Poisson 256^3 SA + Jacobi (the headline, '
                         'default); 3 = 2D rotated anisotropic diffusion 4000^2, classical RS + symmetric multicolour Gauss-Seidel; '
                         '4 = 3D P1 elasticity on 128^3 vertices (BSR 3x3), SA with the 6 rigid-body modes + symmetric '
                         'block-multicolour bsr_gauss_seidel (3 and 4: extra measurements, single GPU)')
    ap.add_argument('--grid', type=int, default=int(os.environ.get('B200AMG_GRID', '0')),
                    help='grid points per dimension (default: 256 for config 2, 4000 for config 3, 128 for config 4)')
    ap.add_argument('--cpu-cycles', type=int, default=3, help='V-cycles of the CPU baseline sample')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-extras', action='store_true', help='config 2: skip the extra measurements (plain CSR layouts, RAP, config 3)')
    ap.add_argument('--smoother', default='jacobi', choices=['jacobi', 'mcgs'],
                    help='jacobi: the BASELINE headline config; mcgs: same hierarchy with symmetric multicolour Gauss-Seidel '
                         '(gauss_seidel_indexed over MIS colours) pre/post -- an extra, non-headline measurement')
    ap.add_argument('--no-graph', action='store_true', help='N > 1: drive the distributed cycle from Python instead of one CUDA graph per rank')
    ap.add_argument('--agglomerate-below', type=int, default=0,
                    help='N > 1: levels with fewer global rows than this run on rank 0 only (default: 16384 rows per rank)')
    args = ap.parse_args()
    if args.grid <= 0:
        args.grid = {2: 256, 3: 4000, 4: 128}[args.config]
    if args.config in (3, 4):
        args.smoother = 'mcgs'
    return args


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


ENTRY_BYTES = {1: 12, 2: 12, 3: 5, 4: 1, 5: 0, 6: 5, 7: 10}   # bytes per stored entry by engine layout (b200amg_vcycle_layout)
LAYOUT_NAME = {1: 'CSR (int32 column + fp64 value)', 2: 'sliced-ELL copy', 3: 'CSR + 8-bit value codes',
               4: 'CSR rows + 8-bit (offset, value) pattern codes', 5: 'one row-pattern byte per row', 6: 'sliced-ELL copy with 8-bit value codes',
               7: 'sliced-ELL copy with 16-bit column steps'}


def matrix_bytes(n, nnz, layout=1):
    """bytes of operator data one sweep has to stream in the given engine layout"""
    if layout == 5:
        return n
    return ENTRY_BYTES.get(layout, 12) * nnz + 4 * (n + 1)


def jacobi_bytes(n, nnz, layout=1):
    """algorithmic bytes of one Jacobi sweep: A once (CSR: 12 nnz + 4(n+1)), x gathered
    once (8n), b (8n), x_new written (8n).  The engine's coded layouts stream less operator data, and
    the roofline is quoted on what the kernel actually has to move in the layout it runs on"""
    return matrix_bytes(n, nnz, layout) + 24 * n


def gs_sweep_bytes(n, nnz):
    """algorithmic bytes of ONE colour-ordered Gauss-Seidel sweep over all colours: B_spmv (A once, x gathered
    once, x written) + b (8n) + the index list Id (4n)"""
    return 12 * nnz + 4 * (n + 1) + 8 * n + 8 * n + 8 * n + 4 * n


def spmv_bytes(n_row, n_col, nnz):
    return 12 * nnz + 4 * (n_row + 1) + 8 * n_col + 8 * n_row


def block_bytes(nblocks, R, C):
    """stored bytes of the blocks of a BSR operator: fp64 R x C block + int32 block column"""
    return nblocks * (8 * R * C + 4)


def bsr_gs_sweep_bytes(n_brow, nblocks, bs):
    """one colour-ordered block Gauss-Seidel sweep: blocks once, block row pointer, x gathered once, x written, b, and the
    block-row index list Id"""
    n = n_brow * bs
    return block_bytes(nblocks, bs, bs) + 4 * (n_brow + 1) + 8 * n + 8 * n + 8 * n + 4 * n_brow


def vcycle_bytes(ml, gs_sweeps=0):
    """unfused algorithmic bytes of one V(1,1)-cycle (SURVEY 8d formula) over the hierarchy; gs_sweeps > 0: the
    smoothers are colour-ordered Gauss-Seidel with that many sweeps per smoothing step (2 = symmetric).  Block (BSR)
    levels count 8 R C + 4 bytes per block and 4 bytes per BLOCK row of row pointer (the scalar formula at R = C = 1)"""
    tot = 0
    for lvl in ml.levels[:-1]:
        A, P, R = lvl.A, lvl.P, lvl.R
        n, nb = A.shape[0], A.n_brow
        mA = block_bytes(A.nblocks, *A.blocksize)
        nc = P.shape[1]
        if gs_sweeps:
            sweep = gs_sweep_bytes(n, A.nnz) if A.is_scalar else bsr_gs_sweep_bytes(nb, A.nblocks, A.blocksize[0])
            tot += 2 * gs_sweeps * sweep + (mA + 4 * nb + 24 * n)
        else:
            tot += 3 * (mA + 4 * nb + 24 * n)
        tot += block_bytes(R.nblocks, *R.blocksize) + 4 * R.n_brow + 8 * n + 8 * nc
        tot += block_bytes(P.nblocks, *P.blocksize) + 4 * P.n_brow + 8 * nc + 16 * n
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
            blk = not lvl.A.is_scalar
            h.pre, h.post = host_spec(lvl.presmoother, blk), host_spec(lvl.postsmoother, blk)
        out.append(h)
    return out


def host_spec(sm, block=False):
    """resolved device smoother -> the oracle's spec for the same sweep (block levels: the index list holds block rows)"""
    if sm.kind == 1:
        return ('jacobi', {'omega': sm.omega, 'iterations': sm.iterations})
    if sm.kind == 2 and block:
        return ('bsr_gauss_seidel_indexed', {'indices': sm.indices.cpu().numpy(), 'sweep': sm.sweep, 'iterations': sm.iterations})
    if sm.kind == 2:
        return ('gauss_seidel_indexed', {'indices': sm.indices.cpu().numpy(), 'sweep': sm.sweep, 'iterations': sm.iterations})
    if sm.kind in (5, 6, 7):
        Dinv = sm.Dinv.cpu().numpy()
        if sm.kind == 5:
            return ('block_jacobi', {'Dinv': Dinv, 'blocksize': int(Dinv.shape[1]), 'omega': sm.omega, 'iterations': sm.iterations})
        return ('block_gauss_seidel', {'Dinv': Dinv, 'blocksize': int(Dinv.shape[1]), 'sweep': sm.sweep, 'iterations': sm.iterations,
                                       'indices': sm.indices.cpu().numpy() if sm.kind == 6 else None})
    raise SystemExit('bench: smoother kind %r has no CPU-baseline spec' % (sm.kind,))


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


# ---------------------------------------------------------------------------------------------
def dist_env():
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    return rank, world, local


def grid_dims(grid, slabs):
    """g^3 unknowns per GPU: the factors of two of `slabs` go round-robin over the three axes, so
    2 GPUs -> (2g, g, g), 4 -> (2g, 2g, g), 8 -> (2g, 2g, 2g) = BASELINE config 5 (512^3) at g = 256.
    Row slabs are cut along the first (slowest) axis."""
    dims, ax, s = [grid, grid, grid], 0, int(slabs)
    while s % 2 == 0 and s > 1:
        dims[ax % 3] *= 2
        s //= 2
        ax += 1
    dims[0] *= s
    return tuple(dims)


def build_problem(grid, slabs=1):
    """config 2 of BASELINE.json: poisson((g,g,g)) + SA(Jacobi 4/3 pre/post, max_coarse=500), seed 0,
    built on the device by pyamg_b200 (bit-identical to the reference-built hierarchy, tests/).
    slabs > 1: poisson(grid_dims(g, slabs)) -- `slabs` row slabs of g^3 unknowns each (8 x 256^3 = the 512^3 of config 5)"""
    import pyamg_b200 as amg
    A = amg.gallery.poisson(grid_dims(grid, slabs))
    np.random.seed(0)
    ml = amg.smoothed_aggregation_solver(A, presmoother=JAC, postsmoother=JAC, max_coarse=500)
    b = np.random.rand(A.shape[0])
    return ml, b


def build_problem_rs(grid):
    """config 3 of BASELINE.json: stencil_grid(diffusion_stencil_2d(epsilon=0.001, theta=pi/8, 'FE'), (g,g)) +
    ruge_stuben_solver(max_coarse=500) with symmetric multicolour Gauss-Seidel (gauss_seidel_indexed over the MIS colour
    ordering) on every level, all built on the device / its host helpers by pyamg_b200"""
    import pyamg_b200 as amg
    st = amg.gallery.diffusion_stencil_2d(epsilon=0.001, theta=np.pi / 8, type='FE')
    A = amg.gallery.stencil_grid(st, (grid, grid))
    np.random.seed(0)
    mc = ('multicolor_gauss_seidel', {'sweep': 'symmetric'})
    ml = amg.ruge_stuben_solver(A, max_coarse=500, presmoother=mc, postsmoother=mc)
    b = np.random.rand(A.shape[0])
    return ml, b


def build_problem_elasticity(nv):
    """config 4 of BASELINE.json: P1 tetrahedral linear elasticity on nv^3 vertices (E = 1e5, nu = 0.3, face x = 0
    clamped; pyamg_b200.gallery.linear_elasticity_tets = the reference's linear_elasticity_p1 on the structured mesh),
    BSR 3x3, smoothed_aggregation_solver with the 6 rigid-body modes (max_coarse=500 block rows, candidates as given),
    symmetric block-multicolour Gauss-Seidel = bsr_gauss_seidel over the MIS colouring of the block graph"""
    import pyamg_b200 as amg
    A, B = amg.gallery.linear_elasticity_tets(nv)
    np.random.seed(0)
    mc = ('multicolor_gauss_seidel', {'sweep': 'symmetric'})
    ml = amg.smoothed_aggregation_solver(A, B, presmoother=mc, postsmoother=mc, max_coarse=500)
    b = np.random.rand(A.shape[0])
    return ml, b


def build_for(args):
    if args.config == 3:
        return build_problem_rs(args.grid)
    if args.config == 4:
        return build_problem_elasticity(args.grid)
    return build_problem(args.grid)


def workload_name(args):
    if args.config == 4:
        return ('configs[3]: 3D linear elasticity, P1 tetrahedra on %d^3 vertices (E=1e5, nu=0.3, one face clamped), BSR 3x3, '
                'smoothed_aggregation_solver with the 6 rigid-body modes (max_coarse=500), symmetric block-multicolour '
                'bsr_gauss_seidel pre/post (MIS colours of the block graph), fp64, one V(1,1)-cycle per step' % args.grid)
    if args.config == 3:
        return ('configs[2]: 2D rotated anisotropic diffusion %dx%d (eps=0.001, theta=pi/8, FE 9-point), ruge_stuben_solver '
                '(classical strength 0.25, RS splitting, direct interpolation, max_coarse=500), symmetric multicolour '
                'Gauss-Seidel pre/post (gauss_seidel_indexed over MIS colours), fp64, one V(1,1)-cycle per step'
                % (args.grid, args.grid))
    return ('configs[1]: 3D Poisson %d^3 7-point, smoothed_aggregation_solver, %s pre/post, max_coarse=500, fp64, one '
            'V(1,1)-cycle per step' % (args.grid, 'Jacobi(4/3)' if args.smoother == 'jacobi' else
                                       'symmetric multicolour Gauss-Seidel'))


def metric_name(args):
    if args.config == 4:
        return 'V-cycles/sec, 3D elasticity %d^3 BSR 3x3 SA (symmetric block-multicolour bsr_gauss_seidel), V(1,1)' % args.grid
    if args.config == 3:
        return 'V-cycles/sec, 2D rotated anisotropic diffusion %d^2 RS (symmetric multicolour Gauss-Seidel), V(1,1)' % args.grid
    return 'V-cycles/sec, 3D Poisson %d^3 SA (%s), V(1,1)' % (args.grid, 'Jacobi 4/3' if args.smoother == 'jacobi' else
                                                               'symmetric multicolour Gauss-Seidel')


def level_table(ml):
    return [[int(l.A.shape[0]), int(l.A.nnz)] for l in ml.levels]


def block_table(ml):
    """[block rows, blocks, block size] per level (block hierarchies only)"""
    if all(l.A.is_scalar for l in ml.levels):
        return None
    return [[int(l.A.n_brow), int(l.A.nblocks), int(l.A.blocksize[0])] for l in ml.levels]


def cpu_cycles_timed(levels, b, cycles, warm, core=None):
    """-> (seconds per V-cycle, kind, x after `warm + cycles` cycles from x0 = 0)"""
    import oracle
    from oracle import multilevel as oml
    S = oml.Solver(levels, core=core or oracle.core, wasted_prolongation=True)
    S.coarse_matrix()
    x = np.zeros_like(b)
    for _ in range(warm):
        S._cycle(0, x, b, 'V')
    t0 = time.perf_counter()
    for _ in range(cycles):
        S._cycle(0, x, b, 'V')
    dt = time.perf_counter() - t0
    return dt / max(cycles, 1), oracle.kind, x


def run_b200(args):
    import torch
    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device (the product path has no CPU fallback)')
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    import pyamg_b200 as amg

    t0 = time.perf_counter()
    if args.config in (3, 4):
        ml, b = build_for(args)
    else:
        ml, b = build_problem(args.grid)
        if args.smoother == 'mcgs':
            mc = ('multicolor_gauss_seidel', {'sweep': 'symmetric'})
            amg.change_smoothers(ml, mc, mc)
    torch.cuda.synchronize()
    setup_s = time.perf_counter() - t0
    n = b.size
    st = ml._stream if ml._stream is not None else None

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- resident-data timing: K graph-replayed V-cycles on (x, b) living in HBM -------------
    ml.solve(b, maxiter=0, _norms=False)          # builds the engine, loads b, x = 0
    st = ml._stream
    for _ in range(args.warmup):
        ml.cycle_device('V')
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = amg._lib.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.profiler.start()        # no-op unless run under `ncu --profile-from-start off`
    with torch.cuda.stream(st):
        e0.record(st)
        for _ in range(args.steps):
            ml.cycle_device('V')
        e1.record(st)
    barrier()
    torch.cuda.profiler.stop()
    ms = e0.elapsed_time(e1)
    launches = amg._lib.launch_count() - l0
    clocks = sampler.stop() if rank == 0 else None
    if dist is not None:
        t = torch.tensor([ms], dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_step = ms / args.steps
    value = world * 1000.0 / ms_step            # replicas: every rank ran its own K cycles
    return dict(ml=ml, b=b, n=n, ms_step=ms_step, value=value, launches=launches, clocks=clocks,
                setup_s=setup_s, rank=rank, world=world, dist=dist, barrier=barrier)


def run_distributed(args):
    """N > 1: one process per GPU; the level-0/1 operators are row-partitioned (one g^3 slab per GPU),
    halos refreshed by NCCL all-to-all over NVLink, coarse levels agglomerated on rank 0
    (pyamg_b200/distributed.py).  The hierarchy is built on rank 0 and its slabs shipped to the owners."""
    import torch
    import torch.distributed as dist
    rank, world, local = dist_env()
    torch.cuda.set_device(local)
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    import pyamg_b200 as amg
    from pyamg_b200.distributed import DistributedSolver, OP_JACOBI
    t0 = time.perf_counter()
    ml, b = (None, None)
    if rank == 0:
        ml, b = build_problem(args.grid, world)
    S = DistributedSolver.from_ml(ml, agglomerate_below=args.agglomerate_below or None)
    torch.cuda.synchronize()
    dist.barrier()
    setup_s = time.perf_counter() - t0

    def barrier():
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()

    def reduce_max(ms):
        t = torch.tensor([ms], dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    bh = torch.from_numpy(b).pin_memory() if rank == 0 else None
    xh = torch.empty(b.size, dtype=torch.float64).pin_memory() if rank == 0 else None
    S.load(bh)
    ex0, lc0 = S.exchanges, amg._lib.launch_count()
    S.cycle()
    ex_per_cycle, launches_per_cycle = S.exchanges - ex0, amg._lib.launch_count() - lc0
    S.profile_cycle()
    prof = {}
    for _ in range(3):
        for lab, ms in S.profile_cycle():
            prof[lab] = prof.get(lab, 0.0) + ms / 3
    if not args.no_graph:
        S.capture()
    for _ in range(args.warmup):
        S.cycle()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = amg._lib.launch_count()
    cur = torch.cuda.current_stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.profiler.start()
    e0.record(cur)
    for _ in range(args.steps):
        S.cycle()
    e1.record(cur)
    barrier()
    torch.cuda.profiler.stop()
    ms_step = reduce_max(e0.elapsed_time(e1)) / args.steps
    launches = launches_per_cycle * args.steps
    clocks = sampler.stop() if rank == 0 else None

    # end to end: every rank owns its slab of b in pinned host memory and receives its slab of x there -- H2D and D2H
    # over each GPU's own PCIe link, no gather through rank 0 (the way a distributed application holds its vectors)
    S.load(bh)                                              # (untimed) hand every rank its slab once
    Lf = S.levels[0]
    bh_loc = torch.empty(Lf.n_own, dtype=torch.float64).pin_memory()
    xh_loc = torch.empty(Lf.n_own, dtype=torch.float64).pin_memory()
    bh_loc.copy_(Lf.b[:Lf.n_own])
    for _ in range(2):
        S.load_local(bh_loc); S.cycle(); S.store_local(xh_loc)
    barrier()
    t1 = time.perf_counter()
    e0.record(cur)
    for _ in range(args.steps):
        S.load_local(bh_loc)
        S.cycle()
        S.store_local(xh_loc)
        torch.cuda.current_stream().synchronize()
    e1.record(cur)
    barrier()
    e2e_ms = reduce_max(max(e0.elapsed_time(e1), (time.perf_counter() - t1) * 1e3)) / args.steps
    S.check_halo()

    # level-0 Jacobi sweep on this rank's slab, timed alone (roofline)
    L = S.levels[0]
    for _ in range(3):
        L.A.apply(OP_JACOBI, L.x, L.b, L.xs, L.info['pre'][1])
    e0.record(cur)
    for _ in range(10):
        L.A.apply(OP_JACOBI, L.x, L.b, L.xs, L.info['pre'][1])
    e1.record(cur)
    barrier()
    t_j = reduce_max(e0.elapsed_time(e1)) / 10

    # parity inside the run: 2 distributed cycles vs 2 single-GPU engine cycles of the same global problem
    parity = None
    S.load(bh)
    S.cycle(); S.cycle()
    xd = S.gather_x()
    if rank == 0:
        xs = ml.solve(torch.from_numpy(b).cuda(), maxiter=2, _norms=False)
        parity = {'cycles': 2, 'bit_identical_to_single_gpu_engine': bool(torch.equal(xd, xs)),
                  'max_abs_diff': float((xd - xs).abs().max().item())}
    barrier()
    out = None
    if rank == 0:
        n_loc, nnz_loc = L.n_own, int(L.A.dx.numel())
        lay = L.A.layout() if hasattr(L.A, 'layout') else 1
        nB = jacobi_bytes(n_loc, nnz_loc, lay)
        peak, how = measured_peak()
        ach = nB / (t_j * 1e-3) / 1e9
        val = world * 1000.0 / ms_step
        gd = grid_dims(args.grid, world)
        out = {'metric': 'V-cycles/sec, 3D Poisson %d^3 SA (Jacobi 4/3), V(1,1) [x %d row slabs of %d^3: global problem %dx%dx%d, '
                         'value = n_gpus x distributed V-cycles/s]' % (args.grid, world, args.grid, gd[0], gd[1], gd[2]),
               'value': val, 'unit': 'V-cycles/s', 'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
               'ms_per_step': ms_step, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
               'dtype': 'f64', 'data': 'synthetic',
               'config': {'workload': 'configs[1] per GPU (configs[4] = 512^3 in total at 8 GPUs): 3D Poisson %dx%dx%d 7-point (%d row slabs of %d^3 '
                                      'unknowns, one per GPU), smoothed_aggregation_solver, Jacobi(4/3) pre/post, max_coarse=500, fp64, one '
                                      'distributed V(1,1)-cycle per step' % (grid_dims(args.grid, world) + (world, args.grid)),
                          'value_definition': 'n_gpus x distributed V-cycles/s (every distributed cycle sweeps n_gpus slabs of %d^3 '
                                              'unknowns; the same unit as the 1-GPU line)' % args.grid,
                          'distributed_cycles_per_s': 1000.0 / ms_step,
                          'levels': level_table(ml), 'partitioned_levels': len(S.levels),
                          'parallelism': 'row slabs over %d GPUs; halo refresh by %s (%d per cycle); levels >= %d agglomerated on rank 0'
                                         % (world, 'peer-memory push/wait kernels over NVLink' if S._p2p is not None else 'NCCL all-to-all',
                                            ex_per_cycle, S.ka),
                          'halo_wait_timeouts': S.halo_errors(),
                          'l2': 'inputs larger than L2', 'setup_s_rank0_build_and_distribute': round(setup_s, 2),
                          'cuda_graph': not args.no_graph,
                          'eager_cycle_breakdown_ms_rank0': {k: round(v, 4) for k, v in prof.items()}},
               'clocks': clocks,
               'e2e': {'value': world * 1000.0 / e2e_ms, 'unit': 'V-cycles/s', 'h2d_bytes_per_step': 8 * b.size,
                       'd2h_bytes_per_step': 8 * b.size,
                       'call': 'per rank: DistributedSolver.load_local(b_slab_host_pinned) + cycle() + store_local(x_slab_host_pinned): '
                               'every GPU moves its own slab over its own PCIe link, one synchronous step at a time'},
               'gpu_launches': int(launches) * world,
               'roofline': {'bound': 'hbm', 'kernel': 'level-0 Jacobi kernel on one rank\'s slab, timed alone',
                            'layout': LAYOUT_NAME.get(lay, str(lay)),
                            'csr_equivalent_GBs': round(jacobi_bytes(n_loc, nnz_loc) / (t_j * 1e-3) / 1e9, 1),
                            'achieved': round(ach, 1), 'peak': peak, 'peak_source': how, 'unit': 'GB/s',
                            'frac': round(ach / peak, 4), 'algorithmic_bytes_per_launch': nB, 'ms_per_launch': round(t_j, 4),
                            'traffic': None},
               'parity_in_run': parity}
    if out is not None:
        print(json.dumps(out), flush=True)
    # a CUDA graph that holds captured NCCL kernels must be gone before the communicator is torn
    # down (destroy_process_group otherwise blocks forever: seen on 2 x B200, NCCL 2.28.9)
    S.release_graph()
    del S
    torch.cuda.synchronize()
    dist.barrier()
    finish_process_group(dist)


def finish_process_group(dist):
    """destroy_process_group with a 20 s escape hatch (the JSON line is already out)"""
    import threading
    t = threading.Timer(20.0, lambda: os._exit(0))
    t.daemon = True
    t.start()
    dist.destroy_process_group()
    t.cancel()


def measure_e2e(R, args):
    """end to end through the public API with HOST buffers, H2D of every step's b and D2H of every step's x inside the
    timed region.  `value`: multilevel_solver.solve_many(b_hosts, maxiter=1) -- the multiple-right-hand-side form of the
    preconditioner application, three streams: the cycle of b_k overlaps H2D of b_{k+1} and D2H of x_{k-1} (PCIe is
    full duplex; floor = one 134 MB copy per step).  `single_call`: the same step as one synchronous
    x_host = ml.solve(b_host, maxiter=1) call at a time (H2D, cycle, D2H back to back: what a host Krylov loop pays)."""
    import torch
    ml = R['ml']
    n = R['n']
    bh = torch.from_numpy(R['b']).pin_memory()
    ring = [torch.empty(n, dtype=torch.float64).pin_memory() for _ in range(3)]

    def timed(fn):
        R['barrier']()
        cur = torch.cuda.current_stream()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(cur)
        t0 = time.perf_counter()
        fn()
        e1.record(cur)
        R['barrier']()
        ms = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3)   # events and host clock bracket the same calls
        if R['dist'] is not None:
            t = torch.tensor([ms], dtype=torch.float64, device='cuda')
            R['dist'].all_reduce(t, op=R['dist'].ReduceOp.MAX)
            ms = float(t.item())
        return ms

    for _ in range(max(1, min(args.warmup, 3))):
        ml.solve(bh, maxiter=1, _norms=False, _alias_out=True)
    ms_single = timed(lambda: [ml.solve(bh, maxiter=1, _norms=False, _alias_out=True) for _ in range(args.steps)])
    ml.solve_many([bh] * 3, maxiter=1, out=ring)
    ms_pipe = timed(lambda: ml.solve_many([bh] * args.steps, maxiter=1, out=[ring[k % 3] for k in range(args.steps)]))
    return {'value': R['world'] * 1000.0 * args.steps / ms_pipe, 'unit': 'V-cycles/s',
            'h2d_bytes_per_step': 8 * n, 'd2h_bytes_per_step': 8 * n,
            'call': 'multilevel_solver.solve_many([b_host_pinned] * steps, maxiter=1) -> x_hosts: every step copies its b '
                    'host->device and its x device->host; consecutive steps overlap on three streams',
            'single_call': {'value': R['world'] * 1000.0 * args.steps / ms_single, 'unit': 'V-cycles/s',
                            'call': 'x_host = multilevel_solver.solve(b_host_pinned, maxiter=1, _norms=False) per step '
                                    '(= aspreconditioner().matvec: no residual norms), one synchronous call at a time'}}


def measure_solve(R, tol=1e-10, maxiter=400, config=2):
    """solve time (BASELINE metric: "V-cycles/sec & solve time"; SURVEY 8d "time-to-tol=1e-10"): b resident in HBM,
    x0 = 0, ml.solve(b, tol=1e-10) stand-alone (one residual norm per iteration returns to the host, like the reference)
    and with the device PCG around the same cycle (accel='cg').  Wall clock around the synchronous public call."""
    import torch
    ml = R['ml']
    bt = torch.from_numpy(R['b']).cuda()
    out = {'tol': tol, 'call': "multilevel_solver.solve(b_device, tol=1e-10[, accel='cg'])"}
    for key, kw in (('standalone', {}), ('pcg', {'accel': 'cg'})):
        try:
            ml.solve(bt, tol=1e-2, maxiter=2, **kw)                       # warm the code path
            res = []
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            ml.solve(bt, tol=tol, maxiter=maxiter, residuals=res, **kw)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            its = len(res) - 1
            xk = ml.solve(bt, tol=tol, maxiter=maxiter, **kw)                  # untimed repeat for the true residual
            rel = float(torch.linalg.vector_norm(ml.levels[0].A.residual(xk, bt)).item() / np.linalg.norm(R['b']))
            out[key] = {'iterations': its, 'seconds': round(dt, 5), 'true_relative_residual': rel,
                        'stopping_rule': 'sqrt(<r, M r>) < tol * initial (pyamg/krylov/_cg.py)' if kw else '||r|| <= tol * ||b||',
                        'ms_per_iteration': round(1e3 * dt / max(its, 1), 4), 'converged': bool(its < maxiter)}
            if config == 3 and its >= maxiter:
                out[key]['note'] = ('this configuration (classical RS, direct interpolation, multicolour GS on the rotated '
                                    'anisotropic FE stencil) is not a convergent stand-alone iteration: the real reference '
                                    'stagnates the same way (600^2: residual ratio 1.24 after 60 cycles; DESIGN.md section 4); '
                                    'the metric of this line is V-cycles/s')
        except Exception as exc:                                          # noqa: BLE001  (a measurement, not the contract)
            out[key] = {'error': '%s: %s' % (type(exc).__name__, exc)}
    return out


def measure_roofline(R, reps=5):
    """dominant kernel = csr_pipe_kernel<OP_JACOBI> on level 0 (pre- and post-smoothing sweep: 2 of the
    5 level-0 passes, ~36 % of a cycle).
    Its duration is measured live with CUDA events around each operation of un-graphed cycles
    (b200amg_vcycle_profile), on the engine's stream."""
    ml = R['ml']
    ml.profile_cycle('V')                                    # warm
    acc = {}
    tot = 0.0
    for _ in range(reps):
        for op, lev, ms in ml.profile_cycle('V'):
            acc.setdefault((op, lev), []).append(ms)
            tot += ms
    tot /= reps
    share = {('%s@L%d' % k): round(float(np.mean(v)), 4) for k, v in acc.items()}
    A0 = ml.levels[0].A
    lay = ml.layouts()[0]['A']
    peak, how = measured_peak()
    t_r = float(np.mean(acc[('residual', 0)]))
    res_bytes = jacobi_bytes(A0.shape[0], A0.nnz, lay)
    sm0 = ml.levels[0].presmoother
    if sm0.kind == 2 and not A0.is_scalar:
        # block rows by MIS colour of the block graph: one sweep = K launches of bsr_gs_color_kernel<bs> on the BSR arrays
        bs = A0.blocksize[0]
        sweeps = 2 if sm0.sweep == 'symmetric' else 1
        K = int(len(sm0.color_ptr) - 1)
        t_s = float(np.mean(acc[('presmooth', 0)] + acc[('postsmooth', 0)])) / (sweeps * sm0.iterations)
        nB = bsr_gs_sweep_bytes(A0.n_brow, A0.nblocks, bs)
        ach = nB / (t_s * 1e-3) / 1e9
        rB = block_bytes(A0.nblocks, bs, bs) + 4 * (A0.n_brow + 1) + 24 * A0.shape[0]
        return {'bound': 'hbm', 'kernel': 'bsr_gs_color_kernel<%d> level 0: one colour-ordered block Gauss-Seidel sweep = %d '
                                          'launches (one per MIS colour of the block graph)' % (bs, K),
                'achieved': round(ach, 1), 'peak': peak, 'peak_source': how, 'unit': 'GB/s', 'frac': round(ach / peak, 4),
                'algorithmic_bytes_per_launch': nB // K, 'algorithmic_bytes_per_sweep': nB, 'launches_per_sweep': K,
                'ms_per_launch': round(t_s / K, 4), 'ms_per_sweep': round(t_s, 4), 'traffic': None,
                'residual_L0_layout': 'BSR %dx%d (fp64 blocks + int32 block column)' % (bs, bs),
                'residual_L0_GBs': round(rB / (t_r * 1e-3) / 1e9, 1),
                'ungraphed_cycle_ms': round(tot, 4), 'op_ms': share}
    if sm0.kind == 2:
        # colour-ordered Gauss-Seidel: the unit is ONE sweep over all K colours (K launches of sell_gs_kernel on the
        # colour-grouped sliced-ELL copy of A); a symmetric smoothing step is two sweeps
        sweeps = 2 if sm0.sweep == 'symmetric' else 1
        K = int(len(sm0.color_ptr) - 1)
        t_s = float(np.mean(acc[('presmooth', 0)] + acc[('postsmooth', 0)])) / (sweeps * sm0.iterations)
        nB = gs_sweep_bytes(A0.shape[0], A0.nnz)
        ach = nB / (t_s * 1e-3) / 1e9
        return {'bound': 'hbm', 'kernel': 'sell_gs_kernel level 0: one colour-ordered Gauss-Seidel sweep = %d launches '
                                          '(one per MIS colour) over the colour-grouped sliced-ELL copy of A' % K,
                'achieved': round(ach, 1), 'peak': peak, 'peak_source': how, 'unit': 'GB/s', 'frac': round(ach / peak, 4),
                'algorithmic_bytes_per_launch': nB // K, 'algorithmic_bytes_per_sweep': nB, 'launches_per_sweep': K,
                'ms_per_launch': round(t_s / K, 4), 'ms_per_sweep': round(t_s, 4), 'traffic': None,
                'residual_L0_layout': LAYOUT_NAME.get(lay, str(lay)),
                'residual_L0_GBs': round(res_bytes / (t_r * 1e-3) / 1e9, 1),
                'ungraphed_cycle_ms': round(tot, 4), 'op_ms': share}
    nB = jacobi_bytes(A0.shape[0], A0.nnz, lay)
    nB_csr = jacobi_bytes(A0.shape[0], A0.nnz)
    t_j = float(np.mean(acc[('presmooth', 0)] + acc[('postsmooth', 0)]))     # ms per launch
    ach = nB / (t_j * 1e-3) / 1e9
    traffic = None
    try:        # dram__bytes_read.sum + dram__bytes_write.sum of this kernel from the committed ncu --set full capture
        tj = json.load(open(os.path.join(ROOT, 'profiles', 'ncu_traffic.json')))
        if int(tj.get('rows', -1)) == int(A0.shape[0]) and int(tj.get('layout', -1)) == lay:
            traffic = tj['dram_bytes_per_launch']
    except Exception:
        traffic = None
    kname = {5: 'rowpat_tile_kernel<OP_JACOBI> (TMA-staged row-pattern sweep)', 2: 'sell_kernel<OP_JACOBI>'}.get(lay, 'csr_pipe_kernel<OP_JACOBI>')
    return {'bound': 'hbm', 'kernel': kname + ' level 0 (pre- and post-smoothing sweep)',
            'layout': LAYOUT_NAME.get(lay, str(lay)),
            'achieved': round(ach, 1),
            'peak': peak, 'peak_source': how, 'unit': 'GB/s', 'frac': round(ach / peak, 4),
            'algorithmic_bytes_per_launch': nB, 'ms_per_launch': round(t_j, 4), 'traffic': traffic,
            'csr_bytes_per_launch': nB_csr, 'csr_equivalent_GBs': round(nB_csr / (t_j * 1e-3) / 1e9, 1),
            'residual_L0_GBs': round(res_bytes / (t_r * 1e-3) / 1e9, 1),
            'ungraphed_cycle_ms': round(tot, 4), 'op_ms': share}


def measure_extras(R, args):
    """driver-visible breadth inside the ONE bench line (config 2 only; the headline keys are untouched):
    plain_csr -- the same cycle with the engine's coded layouts switched off (B200AMG_VALUE_DICT=0): what an operator
                 that is NOT a constant-coefficient stencil gets (CSR streamed as stored);
    rap       -- the Galerkin product R*A*P of level 0 on the device (time, ratio to 2 x products and to the
                 bytes of R, A, RA, P, out) with scipy's (R*A)*P on one host core beside it;
    config3   -- BASELINE configs[2] (RS + symmetric multicolour Gauss-Seidel, 4000^2) as value / roofline / parity."""
    import torch
    import pyamg_b200 as amg
    from pyamg_b200 import sparse
    ml = R['ml']
    out = {}
    peak, _ = measured_peak()

    def timed_cycles(m, k):
        st = m._stream
        for _ in range(3):
            m.cycle_device('V')
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(k):
            m.cycle_device('V')
        e1.record(st)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / k

    def op_ms(m, reps=3):
        m.profile_cycle('V')
        acc = {}
        for _ in range(reps):
            for op, lev, ms in m.profile_cycle('V'):
                acc.setdefault('%s@L%d' % (op, lev), []).append(ms)
        return {k: float(np.mean(v)) for k, v in acc.items()}

    # ---- plain CSR -----------------------------------------------------------------------------------------------
    try:
        old = os.environ.get('B200AMG_VALUE_DICT')
        os.environ['B200AMG_VALUE_DICT'] = '0'
        ml._destroy_engine()
        ml.solve(R['b'], maxiter=0, _norms=False)
        ms = timed_cycles(ml, args.steps)
        ops = op_ms(ml)
        A0 = ml.levels[0].A
        tj = 0.5 * (ops['presmooth@L0'] + ops['postsmooth@L0'])
        nB = jacobi_bytes(A0.shape[0], A0.nnz)
        out['plain_csr'] = {'value': 1000.0 / ms, 'unit': 'V-cycles/s', 'ms_per_step': ms,
                            'layouts_level0': {k: LAYOUT_NAME.get(v, str(v)) for k, v in ml.layouts()[0].items()},
                            'jacobi_L0': {'kernel': 'csr_pipe_kernel<OP_JACOBI> (persistent TMA pipeline over CSR)',
                                          'ms_per_launch': round(tj, 4), 'algorithmic_bytes_per_launch': nB,
                                          'achieved_GBs': round(nB / (tj * 1e-3) / 1e9, 1), 'frac': round(nB / (tj * 1e-3) / 1e9 / peak, 4)},
                            'residual_L0_frac': round(jacobi_bytes(A0.shape[0], A0.nnz) / (ops['residual@L0'] * 1e-3) / 1e9 / peak, 4),
                            'how': 'same hierarchy, engine rebuilt with B200AMG_VALUE_DICT=0 (no row-pattern / value-code layouts)'}
    except Exception as exc:                                              # noqa: BLE001
        out['plain_csr'] = {'error': '%s: %s' % (type(exc).__name__, exc)}
    finally:
        if old is None:
            os.environ.pop('B200AMG_VALUE_DICT', None)
        else:
            os.environ['B200AMG_VALUE_DICT'] = old
        ml._destroy_engine()
        ml.solve(R['b'], maxiter=0, _norms=False)

    # ---- Galerkin product of level 0 -------------------------------------------------------------------------------
    try:
        L0 = ml.levels[0]
        Rm, Am, Pm = L0.R, L0.A, L0.P
        rowlen = lambda M: (M.indptr[1:] - M.indptr[:-1]).to(torch.int64)   # noqa: E731
        best = None
        for _ in range(2):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            RA = sparse.matmat(Rm, Am, True)
            torch.cuda.synchronize()
            t1 = time.perf_counter()
            Ac = sparse.matmat(RA, Pm, True)
            torch.cuda.synchronize()
            t2 = time.perf_counter()
            if best is None or t2 - t0 < best[0]:
                best = (t2 - t0, t1 - t0, t2 - t1)
            prod1 = int(rowlen(Am)[Rm.indices.long()].sum().item())
            prod2 = int(rowlen(Pm)[RA.indices.long()].sum().item())
            nnz_RA, nnz_out = int(RA.nnz), int(Ac.nnz)
            del RA, Ac
        byt = 12 * (Rm.nnz + Am.nnz + 2 * nnz_RA + Pm.nnz + nnz_out) + 4 * (Rm.shape[0] * 3 + Am.shape[0] + Pm.shape[0])
        rap = {'level': 0, 'device_s': round(best[0], 4), 'device_s_RA': round(best[1], 4), 'device_s_RAP': round(best[2], 4),
               'products': [prod1, prod2], 'flops_2x_products': 2 * (prod1 + prod2),
               'GFLOPs': round(2 * (prod1 + prod2) / best[0] / 1e9, 1),
               'bytes_R_A_RA_P_out': byt, 'GBs_on_those_bytes': round(byt / best[0] / 1e9, 1),
               'nnz': {'R': int(Rm.nnz), 'A': int(Am.nnz), 'RA': nnz_RA, 'P': int(Pm.nnz), 'out': nnz_out},
               'call': 'pyamg_b200.sparse.matmat(matmat(R, A), P) in scipy output order (wall clock, synchronised, best of 2)'}
        if not args.no_cpu_baseline:
            Rs, As, Ps = Rm.to_scipy().tocsr(), Am.to_scipy().tocsr(), Pm.to_scipy().tocsr()
            t0 = time.perf_counter()
            _ = (Rs * As) * Ps
            rap['scipy_host_s'] = round(time.perf_counter() - t0, 2)
            rap['scipy_call'] = '(R * A) * P with scipy.sparse on 1 host core (pyamg/aggregation/aggregation.py:429), once'
            del Rs, As, Ps, _
        out['rap'] = rap
    except Exception as exc:                                              # noqa: BLE001
        out['rap'] = {'error': '%s: %s' % (type(exc).__name__, exc)}

    # ---- config 3 --------------------------------------------------------------------------------------------------
    try:
        import argparse
        a3 = argparse.Namespace(**vars(args))
        a3.config, a3.grid, a3.smoother = 3, 4000, 'mcgs'
        t0 = time.perf_counter()
        m3, b3 = build_problem_rs(a3.grid)
        torch.cuda.synchronize()
        setup3 = time.perf_counter() - t0
        m3.solve(b3, maxiter=0, _norms=False)
        ms3 = timed_cycles(m3, args.steps)
        ops3 = op_ms(m3)
        A0 = m3.levels[0].A
        sm0 = m3.levels[0].presmoother
        K = int(len(sm0.color_ptr) - 1)
        t_s = 0.25 * (ops3['presmooth@L0'] + ops3['postsmooth@L0'])        # one of the two sweeps of a symmetric step
        nB = gs_sweep_bytes(A0.shape[0], A0.nnz)
        c3 = {'workload': workload_name(a3), 'value': 1000.0 / ms3, 'unit': 'V-cycles/s', 'ms_per_step': ms3,
              'levels': level_table(m3), 'colours_per_level': [int(len(l.presmoother.color_ptr) - 1) for l in m3.levels[:-1]],
              'setup_s_device': round(setup3, 2),
              'roofline': {'kernel': 'sell_gs_kernel level 0: one colour-ordered sweep = %d launches' % K,
                           'ms_per_sweep': round(t_s, 4), 'algorithmic_bytes_per_sweep': nB,
                           'achieved_GBs': round(nB / (t_s * 1e-3) / 1e9, 1), 'frac': round(nB / (t_s * 1e-3) / 1e9 / peak, 4)}}
        if not args.no_cpu_baseline:
            lv = host_levels(m3)
            sec, kind, x_cpu = cpu_cycles_timed(lv, b3, 1, warm=1)
            x_gpu = m3.solve(b3, maxiter=2, _norms=False)
            c3['cpu_baseline'] = {'value': 1.0 / sec, 'unit': 'V-cycles/s', 'cores': 1, 'kind': kind,
                                  'sample': '1 V(1,1)-cycle after 1 warm-up cycle on the same hierarchy'}
            c3['bit_identical_x_after_2_cycles'] = bool(np.array_equal(x_gpu, x_cpu))
        out['config3'] = c3
        del m3
    except Exception as exc:                                              # noqa: BLE001
        out['config3'] = {'error': '%s: %s' % (type(exc).__name__, exc)}
    return out


def golden_parity(args):
    """config 2 pinned to the REAL reference at the benchmark's own size: tests/golden/ref_config2_<grid>.json holds what
    the reference (imported in the build container, tests/golden/make_ref_config2.py) produced for this exact protocol --
    level table, both rho(D^-1 A) estimates per level, the residual history of 10 V-cycles, entries of x.  The device
    setup is re-run with the reference's rho fed in (the one quantity the device estimates to 1e-7 only),
    consuming the seeded random stream like the reference, and the 10-cycle history is compared."""
    path = os.path.join(ROOT, 'tests', 'golden', 'ref_config2_%d.json' % args.grid)
    if args.config != 2 or args.smoother != 'jacobi' or not os.path.exists(path):
        return None
    import pyamg_b200 as amg
    g = json.load(open(path))
    fh = float.fromhex
    rho = list(zip([fh(h) for h in g['rho_smooth_hex']], [fh(h) for h in g['rho_relax_hex']]))
    A = amg.gallery.poisson((args.grid,) * 3)
    np.random.seed(0)
    ml = amg.smoothed_aggregation_solver(A, presmoother=JAC, postsmoother=JAC, max_coarse=500, rho_override=rho)
    b = np.random.rand(A.shape[0])
    res = []
    x = ml.solve(b, tol=1e-30, maxiter=len(g['residuals_hex']) - 1, residuals=res)
    x = np.asarray(x)
    ref = np.array([fh(h) for h in g['residuals_hex']])
    res = np.array(res, dtype=np.float64)
    xp = np.array([fh(h) for h in g['x_pick_hex']])
    out = {'fixture': 'tests/golden/ref_config2_%d.json (real reference, build container)' % args.grid,
           'levels_equal': level_table(ml) == g['levels'],
           'b_identical': float(b.sum()).hex() == g['b_sum_hex'],
           'cycles': int(len(ref) - 1),
           'history_vs_reference_max_rel_diff': float(np.max(np.abs(res - ref) / ref)) if len(res) == len(ref) else None,
           'x_entries_max_rel_diff': float(np.max(np.abs(x[g['x_pick_index']] - xp) / np.abs(xp))),
           'x_norm_rel_diff': float(abs(np.linalg.norm(x) - fh(g['x_norm_hex'])) / fh(g['x_norm_hex'])),
           'reference_history_first_last': [float(ref[0]), float(ref[-1])]}
    del ml
    return out


def cpu_baseline(R, args):
    """reference CPU path (scipy SpMV + the reference's compiled amg_core smoothers when
    oracle/_ref is built, else the C port) on this box's host cores, same hierarchy, same b."""
    levels = host_levels(R['ml'])
    if args.config == 4:
        return cpu_baseline_block(R, args, levels)
    sec, kind, x_cpu = cpu_cycles_timed(levels, R['b'], args.cpu_cycles, warm=1)
    # parity of the very same run: GPU x after (1 + cpu_cycles) cycles from zero vs the CPU x
    x_gpu = R['ml'].solve(R['b'], maxiter=1 + args.cpu_cycles, _norms=False)
    err = float(np.max(np.abs(x_gpu - x_cpu)) / max(np.max(np.abs(x_cpu)), 1e-300))
    return ({'value': 1.0 / sec, 'unit': 'V-cycles/s', 'cores': 1, 'host_cores': os.cpu_count(), 'kind': kind,
             'sample': '%d full V(1,1)-cycles of the same %d-level hierarchy after 1 warm-up cycle; serial '
                       'reference code (scipy csr/bsr_matvec + amg_core %s), incl. the fork\'s second '
                       'P*coarse_x per level (multilevel.py:548)'
                       % (args.cpu_cycles, len(levels), 'jacobi' if args.smoother == 'jacobi' else 'gauss_seidel_indexed'),
             's_per_cycle': sec},
            {'cycles': 1 + args.cpu_cycles, 'max_rel_diff_x_vs_cpu_reference': err, 'bit_identical': bool(np.array_equal(x_gpu, x_cpu))})


def lexicographic(levels):
    """the reference's own smoother for config 4: symmetric lexicographic bsr_gauss_seidel on every level"""
    from oracle import multilevel as oml
    out = []
    for src in levels:
        h = oml.Level()
        h.A = src.A
        if hasattr(src, 'P'):
            h.P, h.R = src.P, src.R
            h.pre = h.post = ('gauss_seidel', {'sweep': 'symmetric', 'iterations': 1})
        out.append(h)
    return out


def cpu_baseline_block(R, args, levels):
    """config 4.  TIMED: the reference's CPU cycle with ITS smoother for this config -- symmetric lexicographic
    amg_core.bsr_gauss_seidel (the block-multicolour order only exists on the device; same operators, same bytes per
    sweep).  PARITY: the device's block-multicolour cycles against the same block rows relaxed one by one in the same
    order by the reference row update (oracle.relaxation.bsr_gauss_seidel_indexed), untimed."""
    import oracle
    sec, kind, _ = cpu_cycles_timed(lexicographic(levels), R['b'], args.cpu_cycles, warm=1)
    _, _, x_cpu = cpu_cycles_timed(levels, R['b'], 2, warm=0, core=oracle.port)
    x_gpu = R['ml'].solve(R['b'], maxiter=2, _norms=False)
    err = float(np.max(np.abs(x_gpu - x_cpu)) / max(np.max(np.abs(x_cpu)), 1e-300))
    return ({'value': 1.0 / sec, 'unit': 'V-cycles/s', 'cores': 1, 'host_cores': os.cpu_count(), 'kind': kind,
             'sample': '%d full V(1,1)-cycles of the same %d-level block hierarchy after 1 warm-up cycle; serial reference '
                       'code (scipy bsr_matvec + amg_core bsr_gauss_seidel, symmetric LEXICOGRAPHIC sweeps: the reference '
                       'has no block-multicolour order), incl. the fork\'s second P*coarse_x per level (multilevel.py:548)'
                       % (args.cpu_cycles, len(levels)),
             's_per_cycle': sec},
            {'cycles': 2, 'against': 'block rows relaxed one at a time in the same colour order by the reference row update '
                                     '(oracle C port)',
             'max_rel_diff_x_vs_cpu_reference': err, 'bit_identical': bool(np.array_equal(x_gpu, x_cpu))})


def cpu_problem_elasticity(nv):
    """config 4 without a GPU: the generator's index arithmetic on the CPU device, oracle/sa.py block hierarchy, the
    reference's lexicographic symmetric bsr_gauss_seidel"""
    import torch
    import oracle
    from oracle import sa
    from pyamg_b200 import gallery
    A, B = gallery.linear_elasticity_tets(nv, device=torch.device('cpu'))
    A, B = A.to_scipy(), B.numpy()
    np.random.seed(0)
    gs = ('gauss_seidel', {'sweep': 'symmetric'})
    levels = sa.smoothed_aggregation_hierarchy(A, B, presmoother=gs, postsmoother=gs, max_coarse=500, core=oracle.core)
    return levels, np.random.rand(A.shape[0])


def cpu_problem_rs(grid):
    """config 3 without a GPU: the anisotropic FE stencil matrix, oracle/rs.py hierarchy, MIS colour orderings"""
    import oracle
    from oracle import rs as ors, relaxation as orx
    import scipy.sparse as sp
    eps, th = 0.001, np.pi / 8                                # pyamg/gallery/diffusion.py FE stencil
    C, S = np.cos(th), np.sin(th)
    CS, CC, SS = C * S, C * C, S * S
    a = (-1 * eps - 1) * CC + (-1 * eps - 1) * SS + (3 * eps - 3) * CS
    b_ = (2 * eps - 4) * CC + (-4 * eps + 2) * SS
    c = (-1 * eps - 1) * CC + (-1 * eps - 1) * SS + (-3 * eps + 3) * CS
    d = (-4 * eps + 2) * CC + (2 * eps - 4) * SS
    e = (8 * eps + 8) * CC + (8 * eps + 8) * SS
    st = np.array([[a, b_, c], [d, e, d], [c, b_, a]]) / 6.0
    g = grid
    diags, offs = [], []
    for di in (-1, 0, 1):
        for dj in (-1, 0, 1):
            v = np.full(g * g, st[di + 1, dj + 1])
            jj = np.arange(g * g) % g
            v[(jj + dj < 0) | (jj + dj >= g)] = 0.0
            off = di * g + dj
            diags.append(v[max(0, -off):g * g - max(0, off)])
            offs.append(off)
    A = sp.diags(diags, offs, shape=(g * g, g * g), format='csr')
    A.eliminate_zeros(); A.sort_indices()
    np.random.seed(0)
    levels = ors.ruge_stuben_hierarchy(A, max_coarse=500, core=oracle.core)
    for lvl in levels[:-1]:
        idx, _ = orx.color_ordering(orx.vertex_coloring_mis(lvl.A, oracle.core))
        lvl.pre = lvl.post = ('gauss_seidel_indexed', {'indices': idx, 'sweep': 'symmetric', 'iterations': 1})
    return levels, np.random.rand(A.shape[0])


def load_reference_pyamg():
    """the REAL reference package (pyamg 3.0.2 fork): /root/reference in the build container, else the installed copy
    under baseline/_ref (oracle/install_ref.py; travels to the GPU box).  oracle/refimport.py is loaded by file path so
    that nothing else of this repository -- neither pyamg_b200 nor the oracle's C port -- enters the reference process.
    -> the imported `pyamg` module, or None when neither source exists."""
    import importlib.util
    spec = importlib.util.spec_from_file_location('b200amg_refimport', os.path.join(ROOT, 'oracle', 'refimport.py'))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    if not mod.available():
        return None, 'no reference package (looked at %s)' % mod.REF_ROOT
    import warnings
    warnings.simplefilter('ignore')
    return mod.import_pyamg(), mod.REF_ROOT


def reference_mcgs(pyamg, ml):
    """symmetric multicolour Gauss-Seidel on every level with the reference's own pieces:
    pyamg.graph.vertex_coloring(A, 'MIS') -> rows grouped by colour -> relaxation.gauss_seidel_indexed"""
    from pyamg.graph import vertex_coloring
    from pyamg.relaxation.relaxation import gauss_seidel_indexed
    colours = []
    for lvl in ml.levels[:-1]:
        c = vertex_coloring(lvl.A.tocsr(), 'MIS')
        K = int(c.max()) + 1
        idx = np.concatenate([np.where(c == k)[0] for k in range(K)]).astype('intc')

        def sm(A, x, b, idx=idx):
            gauss_seidel_indexed(A, x, b, idx, iterations=1, sweep='symmetric')
        lvl.presmoother = lvl.postsmoother = sm
        colours.append(K)
    return colours


def reference_problem(pyamg, args):
    """the workload of `args.config` built by the reference's own public API -> (ml, b, colours per level or None)"""
    g = args.grid
    colours = None
    if args.config == 2:
        A = pyamg.gallery.poisson((g, g, g), format='csr')
        np.random.seed(0)
        ml = pyamg.smoothed_aggregation_solver(A, presmoother=JAC, postsmoother=JAC, max_coarse=500)
        if args.smoother == 'mcgs':
            colours = reference_mcgs(pyamg, ml)
    elif args.config == 3:
        from pyamg.gallery.diffusion import diffusion_stencil_2d
        st = diffusion_stencil_2d(epsilon=0.001, theta=np.pi / 8, type='FE')
        A = pyamg.gallery.stencil_grid(st, (g, g), format='csr')
        np.random.seed(0)
        ml = pyamg.ruge_stuben_solver(A, max_coarse=500)
        colours = reference_mcgs(pyamg, ml)
    else:
        # the matrix generator is not on the measured path: the stencil-wise assembly of pyamg_b200.gallery on the CPU
        # device (the reference's element loop in Python would take hours for 12.3 M tetrahedra); solver = the reference's
        import torch
        from pyamg_b200 import gallery
        A, B = gallery.linear_elasticity_tets(g, device=torch.device('cpu'))
        A, B = A.to_scipy(), B.numpy()
        np.random.seed(0)
        gs = ('gauss_seidel', {'sweep': 'symmetric'})
        ml = pyamg.smoothed_aggregation_solver(A, B, presmoother=gs, postsmoother=gs, max_coarse=500)
    b = np.random.rand(ml.levels[0].A.shape[0])
    return ml, b, colours


def scipy_level_tables(levels):
    lv = [[int(l.A.shape[0]), int(l.A.nnz)] for l in levels]
    blocks = None
    if any(getattr(l.A, 'blocksize', (1, 1)) != (1, 1) for l in levels):
        blocks = [[int(l.A.shape[0] // l.A.blocksize[0]), int(l.A.nnz // (l.A.blocksize[0] * l.A.blocksize[1])), int(l.A.blocksize[0])]
                  for l in levels]
    return lv, blocks


def workload_config(args, levels, blocks, colours):
    """the `config` object both arms print: the workload only, so that the two lines carry the SAME object
    (everything arm-specific -- parallelism, setup time, byte counts -- goes to `detail`)"""
    nb0, bs0 = (blocks[0][1], blocks[0][2]) if blocks else (levels[0][1], 1)
    return {'workload': workload_name(args), 'grid': args.grid, 'levels': levels, 'block_levels': blocks,
            'colours_per_level': colours,
            'l2': 'inputs larger than L2 (the level-0 operator alone is %.2f GB)' % (1e-9 * block_bytes(nb0, bs0, bs0))}


def run_reference(args):
    """--impl reference: the reference's OWN stock code path, nothing of this repository's product on it.
    pyamg (imported through oracle/refimport.py from /root/reference or from the installed copy baseline/_ref) builds the
    hierarchy with its public constructor and every step is ONE call of its cycle recursion
    multilevel_solver.__solve(0, x, b, 'V') (pyamg/multilevel.py:473-548) -- scipy SpMV + the reference's compiled
    amg_core smoothers, serial, on the box's host cores.  No torch.cuda, no pyamg_b200, no libb200amg.so in this process
    (config 4 alone borrows pyamg_b200.gallery's matrix generator on the CPU device).
    When no reference package exists the arm falls back to the CPU restatement (oracle/) and says so in `kind`."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    budget_s = float(os.environ.get('B200AMG_REF_BUDGET_S', '200'))
    t0 = time.perf_counter()
    pyamg, where = load_reference_pyamg()
    if pyamg is not None:
        ml, b, colours = reference_problem(pyamg, args)
        setup_s = time.perf_counter() - t0
        levels, blocks = scipy_level_tables(ml.levels)
        cycle = ml._multilevel_solver__solve              # the reference's private recursion, called as solve() calls it
        step = lambda x: cycle(0, x, b, 'V')              # noqa: E731
        kind = 'reference'
        how = ('stock pyamg %s from %s: %s + multilevel_solver.__solve(0, x, b, "V") per step'
               % (getattr(pyamg, '__version__', '3.0.2'), where,
                  {2: 'smoothed_aggregation_solver', 3: 'ruge_stuben_solver + gauss_seidel_indexed over MIS colours',
                   4: 'smoothed_aggregation_solver(B = rigid-body modes), symmetric lexicographic bsr_gauss_seidel'}[args.config]))
        n_levels = len(ml.levels)
        rho = [float(l.A.rho_D_inv) if hasattr(l.A, 'rho_D_inv') else None for l in ml.levels[:-1]]
    else:
        sys.stderr.write('reference arm: %s -- CPU restatement (oracle/) instead\n' % where)
        import oracle
        from oracle import multilevel as oml
        if args.config == 3:
            lv, b = cpu_problem_rs(args.grid)
        elif args.config == 4:
            lv, b = cpu_problem_elasticity(args.grid)
        else:
            from oracle import sa
            import scipy.sparse as sp
            g = args.grid
            I = sp.identity(g, format='csr')
            T = sp.diags([-1.0, 2.0, -1.0], [-1, 0, 1], shape=(g, g), format='csr')
            A = (sp.kron(sp.kron(T, I), I) + sp.kron(sp.kron(I, T), I) + sp.kron(sp.kron(I, I), T)).tocsr()
            A.sort_indices()
            np.random.seed(0)
            lv = sa.smoothed_aggregation_hierarchy(A, presmoother=JAC, postsmoother=JAC, max_coarse=500, core=oracle.core)
            b = np.random.rand(A.shape[0])
        setup_s = time.perf_counter() - t0
        S = oml.Solver(lv, core=oracle.core, wasted_prolongation=True)
        S.coarse_matrix()
        step = lambda x: S._cycle(0, x, b, 'V')           # noqa: E731
        levels, blocks = scipy_level_tables(lv)
        colours = None
        if args.config == 3:
            # rows ascend inside a colour: colours = descents of the index list + 1
            colours = [int(np.flatnonzero(np.diff(l.pre[1]['indices']) < 0).size + 1) for l in lv[:-1]]
        kind = oracle.kind if oracle.kind == 'port' else 'port+reference-kernels'
        how = 'CPU restatement oracle/ (setup oracle/sa.py|rs.py, cycle oracle/multilevel.py over %s + scipy)' % (
            'oracle/_ref/_amg_core.so' if oracle.kind == 'reference' else 'the C port')
        n_levels = len(lv)
        rho = None
    x = np.zeros_like(b)
    hist = [float(np.linalg.norm(b))]                      # x0 = 0: ||b - A x0||
    t0 = time.perf_counter()
    step(x)
    t_first = time.perf_counter() - t0
    warm_done = 1
    while warm_done < args.warmup and (warm_done + 1) * t_first < 0.25 * budget_s:
        step(x)
        warm_done += 1
    timed = max(1, min(args.steps, int(0.75 * budget_s / max(t_first, 1e-9))))
    t0 = time.perf_counter()
    for _ in range(timed):
        step(x)
    sec = (time.perf_counter() - t0) / timed
    val = 1.0 / sec
    sample = ('%d of the %d requested V(1,1)-cycles timed (%d warm-up), each a full cycle over the %d-level hierarchy of '
              'the workload (grid %d), x carried from cycle to cycle; %s; serial code, 1 core'
              % (timed, args.steps, warm_done, n_levels, args.grid, how))
    out = {'impl': 'reference', 'metric': metric_name(args),
           'value': val, 'unit': 'V-cycles/s', 'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup,
           'ms_per_step': sec * 1e3, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
           'dtype': 'f64', 'data': 'synthetic',
           'config': workload_config(args, levels, blocks, colours),
           'cpu_baseline': {'value': val, 'unit': 'V-cycles/s', 'cores': 1, 'host_cores': os.cpu_count(),
                            'kind': kind, 'sample': sample},
           'e2e': {'value': val, 'unit': 'V-cycles/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
           'gpu_launches': 0,
           'detail': {'parallelism': '1 host core (the reference is serial)', 'setup_s_reference': round(setup_s, 1),
                      'timed_steps': timed, 'warmup_done': warm_done, 'rho_D_inv_A': rho,
                      'modules_loaded': sorted(m for m in sys.modules if m.split('.')[0] in ('torch', 'pyamg_b200', 'oracle'))[:8]}}
    print(json.dumps(out))


def main():
    args = parse()
    wd = float(os.environ.get('B200AMG_BENCH_WATCHDOG', '0'))
    if wd > 0:                       # dump every thread's stack if the run is still alive after `wd` seconds
        import faulthandler
        faulthandler.dump_traceback_later(wd, exit=True)
    if args.impl == 'reference':
        run_reference(args)
        return
    if dist_env()[1] > 1 and args.config == 2:
        run_distributed(args)
        return
    R = run_b200(args)
    if args.smoother != 'jacobi' and args.config == 2:
        args.no_cpu_baseline = True
    e2e = measure_e2e(R, args)
    roof = measure_roofline(R) if R['rank'] == 0 else None
    solve = measure_solve(R, config=args.config) if R['rank'] == 0 and R['world'] == 1 else None
    cpu, parity = (None, None)
    if R['rank'] == 0 and R['world'] == 1 and not args.no_cpu_baseline:
        cpu, parity = cpu_baseline(R, args)
    if R['dist'] is not None:
        R['barrier']()
        R['dist'].destroy_process_group()
    if R['rank'] != 0:
        return
    ml = R['ml']
    vb = vcycle_bytes(ml, 2 if args.smoother == 'mcgs' else 0)
    golden = None
    extra = None
    if R['world'] == 1 and args.config == 2 and args.smoother == 'jacobi' and not args.no_extras:
        extra = measure_extras(R, args)
    if R['world'] == 1:
        try:
            golden = golden_parity(args)
        except Exception as exc:                                          # noqa: BLE001  (reported, never hidden)
            golden = {'error': '%s: %s' % (type(exc).__name__, exc)}
    out = {'metric': metric_name(args),
           'value': R['value'], 'unit': 'V-cycles/s', 'n_gpus': R['world'], 'steps': args.steps,
           'warmup': args.warmup, 'ms_per_step': R['ms_step'], 'higher_is_better': True, 'scaling': 'weak',
           'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
           'config': workload_config(args, level_table(ml), block_table(ml),
                                     ([int(len(l.presmoother.color_ptr) - 1) for l in ml.levels[:-1]]
                                      if args.smoother == 'mcgs' else None)),
           'detail': {'parallelism': 'single GPU' if R['world'] == 1 else
                      'replicas: %d independent solves, one per GPU' % R['world'],
                      'setup_s_device': round(R['setup_s'], 2),
                      'engine_layouts': [{k: LAYOUT_NAME.get(v, str(v)) for k, v in L.items()} for L in ml.layouts()],
                      'unfused_algorithmic_bytes_per_cycle': vb,
                      'cycle_GBs_on_unfused_bytes': round(vb / (R['ms_step'] * 1e-3) / 1e9, 1)},
           'clocks': R['clocks'], 'e2e': e2e, 'gpu_launches': int(R['launches']), 'roofline': roof, 'solve': solve}
    if cpu is not None:
        out['cpu_baseline'] = cpu
        out['parity_in_run'] = parity
    if golden is not None:
        out['parity_vs_reference_golden'] = golden
    if extra is not None:
        out['extra'] = extra
    print(json.dumps(out))


if __name__ == '__main__':
    main()
