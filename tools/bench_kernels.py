#!/usr/bin/env python
"""Per-kernel roofline table for every row of SURVEY 8(a): each C-ABI entry point of the hot path timed ALONE on
operators larger than L2, CUDA events on the launching stream, against the measured HBM peak.

    python tools/bench_kernels.py [--out profiles/rNN/kernels.json]

Workloads (synthetic, fp64 / int32):
  csr     3D Poisson 256^3 7-point (16.8 M rows, 117 M entries) -- BASELINE config 2, level 0; the entry points see
          plain CSR (the engine-owned coded layouts of the V-cycle are NOT used here: this is the amg_core-shaped call)
  csr31   level-1 operator of the SA hierarchy of 128^3 (264 k rows, 30 entries / row) x 8 stacked copies = long-row CSR
  bsr3    kron(poisson 128^3, B3): 2.1 M block rows of 3x3 blocks (14.6 M blocks) -- the shape of config 4's level 0
  bsr6    kron(poisson 80^3, B6): 512 k block rows of 6x6 blocks -- the shape of config 4's coarse levels
Algorithmic bytes: SpMV 12 nnz + 4(n+1) + 8 n_col + 8 n_row; BSR 8 bs^2 + 4 per block; Jacobi / residual
= SpMV + 8n (b); colour-ordered GS sweep = SpMV + 8n (b) + 4 n_rows (Id).
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pyamg_b200 as amg                                     # noqa: E402
from pyamg_b200 import relaxation as rx                      # noqa: E402
from pyamg_b200.sparse import DevMat                         # noqa: E402


def peak():
    try:
        return float(json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['hbm_gbs']), 'MEASURED_PEAKS.json'
    except Exception:
        return 6650.0, 'fallback (B200_PROFILING.md)'


def timed(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    st = torch.cuda.current_stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(st)
    for _ in range(reps):
        fn()
    e1.record(st)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def spmv_bytes(A):
    R, C = A.blocksize
    return A.nblocks * (8 * R * C + 4) + 4 * (A.n_brow + 1) + 8 * A.shape[1] + 8 * A.shape[0]


def kron_bsr(A, bs, seed):
    """kron(A, B) as BSR with B a fixed SPD bs x bs block with distinct entries"""
    rng = np.random.RandomState(seed)
    M = rng.rand(bs, bs)
    B = M @ M.T + bs * np.eye(bs)
    Bt = torch.from_numpy(B).to(A.data.device)
    data = (A.data[:, None, None] * Bt[None]).reshape(-1).contiguous()
    return DevMat(A.indptr, A.indices, data, (A.shape[0] * bs, A.shape[1] * bs), (bs, bs), 'bsr', True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--out', default=None)
    ap.add_argument('--grid', type=int, default=256)
    args = ap.parse_args()
    torch.cuda.set_device(0)
    pk, how = peak()
    rows = []

    def add(row_id, what, kernel, nbytes, ms, note=''):
        gbs = nbytes / (ms * 1e-3) / 1e9
        rows.append({'survey_row': row_id, 'entry_point': what, 'kernel': kernel, 'algorithmic_bytes': int(nbytes),
                     'ms': round(ms, 4), 'GBs': round(gbs, 1), 'frac_of_peak': round(gbs / pk, 4), 'note': note})
        print('%-34s %-44s %8.4f ms %8.1f GB/s  %.3f' % (what, kernel, ms, gbs, gbs / pk), flush=True)

    # ---- CSR, short rows: poisson grid^3 ---------------------------------------------------------------------
    g = args.grid
    A = amg.gallery.poisson((g, g, g))
    n = A.shape[0]
    x = torch.rand(n, dtype=torch.float64, device='cuda')
    b = torch.rand(n, dtype=torch.float64, device='cuda')
    y = torch.empty_like(x)
    tmp = torch.empty_like(x)
    tag = 'poisson %d^3: %d rows, %d entries' % (g, n, A.nnz)
    add(1, 'b200amg_csr_matvec (A*x)', 'csr_pipe_kernel<MATVEC>', spmv_bytes(A), timed(lambda: A.matvec(x, out=y)), tag)
    add(1, 'b200amg_csr_residual (b - A*x)', 'csr_pipe_kernel<RESIDUAL>', spmv_bytes(A) + 8 * n,
        timed(lambda: A.residual(x, b, out=y)), tag)
    add(3, 'b200amg_jacobi (amg_core.jacobi)', 'csr_pipe_kernel<JACOBI> + temp copy', spmv_bytes(A) + 8 * n,
        timed(lambda: amg.amg_core.jacobi(A.indptr, A.indices, A.data, x, b, tmp, 0, n, 1, [0.8])),
        tag + '; in-place x like the reference (x -> temp copy inside the call)')
    idx, ptr = rx.color_ordering(rx.vertex_coloring(A, 'MIS'))
    idx_d = torch.from_numpy(idx).cuda()
    K = len(ptr) - 1
    ms = timed(lambda: amg.amg_core.gauss_seidel_indexed(A.indptr, A.indices, A.data, x, b, idx_d, 0, n, 1, ptr))
    add(6, 'b200amg_gauss_seidel_indexed', 'gs_rows_kernel x %d colours (one forward sweep)' % K,
        spmv_bytes(A) + 8 * n + 4 * n, ms, tag + '; CSR as given (the engine uses a colour-grouped SELL copy instead)')
    del idx_d

    # ---- BSR 3x3 and 6x6 ------------------------------------------------------------------------------------
    for bs, gg, sid in ((3, 128, 1), (6, 80, 2)):
        Ab = kron_bsr(amg.gallery.poisson((gg, gg, gg)), bs, sid)
        nb = Ab.shape[0]
        xb = torch.rand(nb, dtype=torch.float64, device='cuda')
        bb = torch.rand(nb, dtype=torch.float64, device='cuda')
        yb = torch.empty_like(xb)
        tb = torch.empty_like(xb)
        tag = 'kron(poisson %d^3, SPD %dx%d): %d block rows, %d blocks' % (gg, bs, bs, Ab.n_brow, Ab.nblocks)
        add(2, 'b200amg_bsr_matvec bs=%d' % bs, 'bsr_matvec_kernel', spmv_bytes(Ab), timed(lambda: Ab.matvec(xb, out=yb)), tag)
        add(4, 'b200amg_bsr_jacobi bs=%d' % bs, 'bsr_jacobi_kernel<%d> + temp copy' % bs, spmv_bytes(Ab) + 8 * nb,
            timed(lambda: amg.amg_core.bsr_jacobi(Ab.indptr, Ab.indices, Ab.data, xb, bb, tb, 0, Ab.n_brow, 1, bs, [0.8])), tag)
        cidx, cptr = rx.color_ordering(rx.vertex_coloring(Ab, 'MIS'))
        cidx_d = torch.from_numpy(cidx).cuda()
        ms = timed(lambda: amg.amg_core.bsr_gauss_seidel(Ab.indptr, Ab.indices, Ab.data, xb, bb, 0, Ab.n_brow, 1, bs, cidx_d, cptr))
        add(8, 'b200amg_bsr_gauss_seidel bs=%d' % bs, 'bsr_gs_color_kernel<%d> x %d colours (one forward sweep)' % (bs, len(cptr) - 1),
            spmv_bytes(Ab) + 8 * nb + 4 * Ab.n_brow, ms, tag + '; block rows by MIS colour of the block graph')
        # block smoothers with stored diagonal-block inverses (the reference's SA default): + 8 bs^2 bytes of Dinv per block row
        Dinv = amg.relaxation.get_block_diag_inverse(Ab).view(-1)
        dbytes = 8 * bs * bs * Ab.n_brow
        add(10, 'b200amg_block_jacobi bs=%d' % bs, 'block_jacobi_kernel<%d> + temp copy' % bs, spmv_bytes(Ab) + 8 * nb + dbytes,
            timed(lambda: amg.amg_core.block_jacobi(Ab.indptr, Ab.indices, Ab.data, xb, bb, Dinv, tb, 0, Ab.n_brow, 1, [0.8], bs)), tag)
        ms = timed(lambda: amg.amg_core.block_gauss_seidel(Ab.indptr, Ab.indices, Ab.data, xb, bb, Dinv, 0, Ab.n_brow, 1, bs, cidx_d, cptr))
        add(10, 'b200amg_block_gauss_seidel bs=%d' % bs, 'block_gs_color_kernel<%d> x %d colours (one forward sweep)' % (bs, len(cptr) - 1),
            spmv_bytes(Ab) + 8 * nb + 4 * Ab.n_brow + dbytes, ms, tag + '; block rows by MIS colour of the block graph')
        del Ab, xb, bb, yb, tb, cidx_d, Dinv

    # ---- norms (row 9) ----------------------------------------------------------------------------------------
    from pyamg_b200 import linalg
    add(9, 'b200amg_dot (norm)', 'dot kernels (two-stage, deterministic)', 8 * n, timed(lambda: linalg.norm(x)), '%d doubles' % n)

    out = {'peak_GBs': pk, 'peak_source': how, 'device': torch.cuda.get_device_name(0), 'rows': rows}
    if args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        json.dump(out, open(args.out, 'w'), indent=1)
    print(json.dumps(out))


if __name__ == '__main__':
    main()
