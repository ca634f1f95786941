// BSR kernels (R x C row-major blocks, 1..8): scipy bsr_matvec and the reference's
// amg_core.bsr_jacobi / bsr_gauss_seidel (pyamg/amg_core/relaxation.h:89-173, :267-360,
// dense block product = linalg.h:359-398 'gemm' started from zero).
// Arithmetic order is the reference's: per off-diagonal block loc = ((0+a0 x0)+a1 x1)+...,
// acc -= loc; then the point sweep inside the diagonal block.
#include <stdlib.h>
#include "internal.cuh"

namespace b200amg {

// one thread per scalar output row (i*R + r)
__global__ void __launch_bounds__(256) bsr_matvec_kernel(int n_brow, int R, int C, const int *__restrict__ Ap,
                                                         const int *__restrict__ Aj, const double *__restrict__ Ax,
                                                         const double *__restrict__ x, const double *__restrict__ b,
                                                         double *y, int mode)
{
    pdl_enter();
    const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= (long long)n_brow * R) return;
    const int i = (int)(g / R), r = (int)(g - (long long)i * R);
    double acc = 0.0;
    const int hi = Ap[i + 1];
    for (int p = Ap[i]; p < hi; ++p) {
        const double *blk = Ax + ((size_t)p * R + r) * C;
        const double *xb = x + (size_t)Aj[p] * C;
        for (int c = 0; c < C; ++c) acc = add_rn(acc, mul_rn(blk[c], xb[c]));
    }
    if (mode == 0) y[g] = acc;
    else if (mode == 1) y[g] = sub_rn(b[g], acc);
    else y[g] = add_rn(b ? b[g] : y[g], acc);        // mode 2: y = (b or y) + A x
}

// amg_core.bsr_jacobi, forward full range: thread per scalar row (i, k); reads xin, writes xout
template <int BS>
__global__ void __launch_bounds__(256) bsr_jacobi_kernel(int nb0, int nb1, const int *__restrict__ Ap,
                                                         const int *__restrict__ Aj, const double *__restrict__ Ax,
                                                         const double *__restrict__ xin, double *__restrict__ xout,
                                                         const double *__restrict__ b, double omega)
{
    pdl_enter();
    const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= (long long)(nb1 - nb0) * BS) return;
    const int i = nb0 + (int)(g / BS), k = (int)(g % BS);
    const size_t row = (size_t)i * BS + k;
    double acc = b[row];
    long long dptr = -1;
    const int hi = Ap[i + 1];
    for (int p = Ap[i]; p < hi; ++p) {
        const int j = Aj[p];
        if (j == i) { dptr = p; continue; }
        const double *blk = Ax + ((size_t)p * BS + k) * BS;
        const double *xb = xin + (size_t)j * BS;
        double loc = 0.0;
#pragma unroll
        for (int c = 0; c < BS; ++c) loc = add_rn(loc, mul_rn(blk[c], xb[c]));
        acc = sub_rn(acc, loc);
    }
    const double xi = xin[row];
    if (dptr < 0) { xout[row] = xi; return; }
    const double *D = Ax + ((size_t)dptr * BS + k) * BS;
    double d = 1.0;
#pragma unroll
    for (int kk = 0; kk < BS; ++kk) {
        if (kk == k) d = D[kk];
        else acc = sub_rn(acc, mul_rn(D[kk], xin[(size_t)i * BS + kk]));
    }
    xout[row] = (d != 0.0) ? add_rn(mul_rn(sub_rn(1.0, omega), xi), div_rn(mul_rn(omega, acc), d)) : xi;
}

// amg_core.bsr_gauss_seidel for one block row (all of it by one thread: the point sweep inside
// the diagonal block is sequential in k).
template <int BS>
__device__ __forceinline__ void bsr_gs_block_row(int i, const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                 const double *__restrict__ Ax, double *x,
                                                 const double *__restrict__ b, bool backward)
{
    double acc[BS];
#pragma unroll
    for (int k = 0; k < BS; ++k) acc[k] = b[(size_t)i * BS + k];
    long long dptr = -1;
    const int hi = Ap[i + 1];
    for (int p = Ap[i]; p < hi; ++p) {
        const int j = Aj[p];
        if (j == i) { dptr = p; continue; }
        const double *blk = Ax + (size_t)p * BS * BS;
        double xv[BS];
#pragma unroll
        for (int c = 0; c < BS; ++c) xv[c] = x[(size_t)j * BS + c];
#pragma unroll
        for (int k = 0; k < BS; ++k) {
            double loc = 0.0;
#pragma unroll
            for (int c = 0; c < BS; ++c) loc = add_rn(loc, mul_rn(blk[k * BS + c], xv[c]));
            acc[k] = sub_rn(acc[k], loc);
        }
    }
    if (dptr < 0) return;
    const double *D = Ax + (size_t)dptr * BS * BS;
    double xl[BS];
#pragma unroll
    for (int k = 0; k < BS; ++k) xl[k] = x[(size_t)i * BS + k];
#pragma unroll
    for (int kq = 0; kq < BS; ++kq) {
        const int k = backward ? BS - 1 - kq : kq;
        double d = 1.0, a = acc[k];
#pragma unroll
        for (int cq = 0; cq < BS; ++cq) {
            const int kk = backward ? BS - 1 - cq : cq;
            if (kk == k) d = D[k * BS + kk];
            else a = sub_rn(a, mul_rn(D[k * BS + kk], xl[kk]));
        }
        if (d != 0.0) { xl[k] = div_rn(a, d); x[(size_t)i * BS + k] = xl[k]; }
    }
}

// One colour of block rows, BS lanes per block row (lane k owns scalar row k of the block row; 32 / BS block rows
// per warp).  Off-diagonal blocks: lane k folds row k of every block exactly like the one-thread form above (the
// BS lanes of a group read the BS*BS doubles of a block as one contiguous piece).  Diagonal block: the point
// sweep is sequential in k, so step kq is done by the lane that owns row k and its result is handed to the
// group by a shuffle before the next step -- the order and the bits of bsr_gs_block_row, with BS times the
// threads in flight (thread-per-block-row reached 0.41-0.53 of HBM peak, profiles/r01b/kernels_roofline.json).
template <int BS, bool PERM>
__global__ void __launch_bounds__(128) bsr_gs_color_kernel(const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                           const double *__restrict__ Ax, double *x,
                                                           const double *__restrict__ b, const int *__restrict__ Id,
                                                           int p0, int p1, bool backward)
{
    pdl_enter();
    constexpr int RPW = 32 / BS;                        // block rows per warp
    const int lane = threadIdx.x & 31, warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int g = lane / BS, k = lane - g * BS;
    const int p = p0 + warp * RPW + g;
    const bool active = g < RPW && p < p1;
    const int i = active ? Id[p] : 0;
    double acc = 0.0;
    long long dptr = -1;
    if (active) {
        acc = b[(size_t)i * BS + k];
        const int rp = PERM ? p : i;                    // PERM: (Ap, Aj, Ax) is the colour-grouped copy, row = position
        const int hi = Ap[rp + 1];
        // U blocks in flight: their column numbers first, then every block row / x_j operand, then the folds in
        // storage order -- one round of memory latency per U blocks instead of two per block
        constexpr int U = BS <= 4 ? 4 : 2;
        for (int q = Ap[rp]; q < hi; q += U) {
            int j[U];
            double bl[U][BS], xv[U][BS];
#pragma unroll
            for (int u = 0; u < U; ++u) j[u] = (q + u < hi) ? Aj[q + u] : -1;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (j[u] >= 0 && j[u] != i) {
                    const double *blk = Ax + ((size_t)(q + u) * BS + k) * BS;
                    const double *xb = x + (size_t)j[u] * BS;
#pragma unroll
                    for (int c = 0; c < BS; ++c) { bl[u][c] = blk[c]; xv[u][c] = xb[c]; }
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (j[u] < 0) continue;
                if (j[u] == i) { dptr = q + u; continue; }
                double loc = 0.0;
#pragma unroll
                for (int c = 0; c < BS; ++c) loc = add_rn(loc, mul_rn(bl[u][c], xv[u][c]));
                acc = sub_rn(acc, loc);
            }
        }
    }
    const bool has_diag = active && dptr >= 0;
    double xl[BS], Dk[BS];
#pragma unroll
    for (int c = 0; c < BS; ++c) {
        xl[c] = has_diag ? x[(size_t)i * BS + c] : 0.0;
        Dk[c] = has_diag ? Ax[((size_t)dptr * BS + k) * BS + c] : 0.0;
    }
#pragma unroll
    for (int kq = 0; kq < BS; ++kq) {
        const int kk0 = backward ? BS - 1 - kq : kq;     // the scalar row relaxed in this step
        double nv = 0.0;
        bool wrote = false;
        if (has_diag && k == kk0) {
            double d = 1.0, a = acc;
#pragma unroll
            for (int cq = 0; cq < BS; ++cq) {
                const int kk = backward ? BS - 1 - cq : cq;
                if (kk == kk0) d = Dk[kk];
                else a = sub_rn(a, mul_rn(Dk[kk], xl[kk]));
            }
            if (d != 0.0) {
                nv = div_rn(a, d);
                wrote = true;
                x[(size_t)i * BS + kk0] = nv;
            }
        }
        // hand the new value (or the unchanged one) to the other lanes of the group
        const int src = g * BS + kk0;
        const double cur = wrote ? nv : xl[kk0];
        const double got = __shfl_sync(0xffffffffu, cur, src < 32 ? src : 0);
        if (has_diag) xl[kk0] = got;
    }
}

// One colour of block rows, ONE WARP per block row: the form for rows with many blocks (coarse elasticity levels:
// 44-55 blocks of 6 x 6 per row).  The lanes-per-row kernel above walks a row's blocks one after the other, two
// dependent global loads per block (Aj, then the block and x_j): 0.8 us per block, 40 us per colour launch whatever the
// colour's size (ncu, profiles/r02).  Here the warp takes 32 / BS blocks of its row per pass -- lane (g, k) forms
// loc = row k of block g times x_j from zero, exactly the reference's gemm (linalg.h:359-398) -- so one pass costs ONE
// round of memory latency; then the pass's locs are subtracted from the row sums in STORAGE order (shuffled one by one
// to the lanes 0..BS-1 that hold acc), which is the order of relaxation.h:139-144: the bits do not change.
// NB passes form a BATCH: the column indices of the whole batch are requested together, then all its block rows and
// x_j, and only then the ordered subtraction runs -- two rounds of memory latency per NB * (32 / BS) blocks instead of
// per 32 / BS (a 55-block row of 6 x 6 blocks: 6 rounds instead of 22; these launches are one wave of warps, so rounds
// of latency are all their run time).  `zero` is 0 at run time: the `tie` words keep the compiler from sinking each load
// next to its use (sell.cu has the story).
template <int BS, bool backward, bool PERM, int NB>
__global__ void __launch_bounds__(128) bsr_gs_color_warp_kernel(const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                                const double *__restrict__ Ax, double *x,
                                                                const double *__restrict__ b, const int *__restrict__ Id,
                                                                int p0, int p1, int zero)
{
    // the operator (index list, row bounds, block columns, blocks) is read BEFORE the dependency wait: the engine passes
    // hierarchy arrays that no kernel of the cycle writes, and launches early only then (bsr_gs_color: static_op)
    pdl_trigger();
    constexpr int G = 32 / BS;                          // blocks per pass
    constexpr unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int p = p0 + (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (p >= p1) { pdl_wait(); return; }                // the whole warp leaves together
    const int i = Id[p];
    const int g = lane / BS, k = lane - g * BS;
    const bool lane_ok = g < G;
    const int lo = Ap[PERM ? p : i], hi = Ap[(PERM ? p : i) + 1];
    double acc = 0.0;
    bool waited = false;
    int dptr = -1;
    for (int q0 = lo; q0 < hi; q0 += G * NB) {
        int j[NB];
#pragma unroll
        for (int t = 0; t < NB; ++t) {
            const int q = q0 + t * G + g;
            j[t] = (lane_ok && q < hi) ? Aj[q] : -1;
        }
        int tie = 0;
        if (NB > 1) {
#pragma unroll
            for (int t = 0; t < NB; ++t) tie |= j[t];
            tie &= zero;
        }
        double bl[NB][BS], xv[NB][BS];
#pragma unroll
        for (int t = 0; t < NB; ++t) {
            const int q = q0 + t * G + g;
            const bool use = j[t] >= 0 && j[t] != i;
            const double *blk = Ax + ((size_t)q * BS + k) * BS + (t == 0 ? tie : 0);
#pragma unroll
            for (int c = 0; c < BS; ++c) bl[t][c] = use ? blk[c] : 0.0;
        }
        if (!waited) {                                  // warp-uniform: x and b are final from here on
            pdl_wait();
            acc = (lane < BS) ? b[(size_t)i * BS + lane] : 0.0;
            waited = true;
        }
#pragma unroll
        for (int t = 0; t < NB; ++t) {
            const bool use = j[t] >= 0 && j[t] != i;
            const double *xb = x + (size_t)(use ? j[t] : 0) * BS;
#pragma unroll
            for (int c = 0; c < BS; ++c) xv[t][c] = use ? xb[c] : 0.0;
        }
        if (NB > 1) {
#pragma unroll
            for (int t = 0; t < NB; ++t)
#pragma unroll
                for (int c = 0; c < BS; ++c) tie |= __double2loint(bl[t][c]) | __double2loint(xv[t][c]);
            tie &= zero;
        }
#pragma unroll
        for (int t = 0; t < NB; ++t) {
            if (q0 + t * G >= hi) break;                // warp-uniform
            const bool have = (t == 0 ? j[t] + tie : j[t]) >= 0;
            const bool isdiag = have && j[t] == i;
            double loc = 0.0;
#pragma unroll
            for (int c = 0; c < BS; ++c) loc = add_rn(loc, mul_rn(bl[t][c], xv[t][c]));
            const unsigned dg = __ballot_sync(FULL, isdiag);
            if (dg) dptr = q0 + t * G + (31 - __clz(dg)) / BS;      // the last stored diagonal block wins (relaxation.h:121)
            // blocks that do not exist / the diagonal block hand over +0.0: acc - (+0.0) == acc bit for bit, also for -0.0
            loc = (have && !isdiag) ? loc : 0.0;
#pragma unroll
            for (int gg = 0; gg < G; ++gg) acc = sub_rn(acc, __shfl_sync(FULL, loc, gg * BS + (lane < BS ? lane : 0)));
        }
    }
    if (!waited) pdl_wait();                            // a row without blocks
    if (dptr < 0) return;
    // point sweep inside the diagonal block (relaxation.h:150-163): lane k < BS owns row k of D
    double xl[BS], Dk[BS];
#pragma unroll
    for (int c = 0; c < BS; ++c) {
        xl[c] = x[(size_t)i * BS + c];
        Dk[c] = (lane < BS) ? Ax[((size_t)dptr * BS + lane) * BS + c] : 0.0;
    }
#pragma unroll
    for (int kq = 0; kq < BS; ++kq) {
        const int kk0 = backward ? BS - 1 - kq : kq;
        double nv = 0.0;
        bool wrote = false;
        if (lane == kk0) {
            double d = 1.0, a = acc;
#pragma unroll
            for (int cq = 0; cq < BS; ++cq) {
                const int kk = backward ? BS - 1 - cq : cq;
                if (kk == kk0) d = Dk[kk];
                else a = sub_rn(a, mul_rn(Dk[kk], xl[kk]));
            }
            if (d != 0.0) {
                nv = div_rn(a, d);
                wrote = true;
                x[(size_t)i * BS + kk0] = nv;
            }
        }
        const double cur = wrote ? nv : xl[kk0];
        xl[kk0] = __shfl_sync(FULL, cur, kk0);
    }
}

// scipy bsr_matvec with ONE WARP per block row (rows with many blocks: the restrictions of an elasticity hierarchy,
// 6 x 3 / 6 x 6 blocks, 30-150 blocks per row; the thread-per-scalar-row kernel above walks them one block at a
// time: 0.5-0.6 ms for 1331 / 27 block rows).  scipy accumulates y[i*R + r] over ALL entries of the scalar row in
// storage order into one running sum, so only the products are free: lane (g, r) multiplies row r of block g of the
// pass by x_j (C separately rounded products), then the running sum of scalar row r travels through the groups in
// block order -- group g adds its C products, hands the sum to group g + 1 by a shuffle -- G shuffles per pass.
// NB passes per batch, their loads requested together (see bsr_gs_color_warp_kernel).
template <int R, int NB>
__global__ void __launch_bounds__(128) bsr_matvec_warp_kernel(int n_brow, int C, const int *__restrict__ Ap,
                                                              const int *__restrict__ Aj, const double *__restrict__ Ax,
                                                              const double *__restrict__ x, const double *__restrict__ b,
                                                              double *y, int mode, int zero)
{
    pdl_enter();
    constexpr int G = 32 / R;
    constexpr unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int i = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (i >= n_brow) return;
    const int g = lane / R, r = lane - g * R;
    const bool lane_ok = g < G;
    const int lo = Ap[i], hi = Ap[i + 1];
    double carry = 0.0;                                   // running sum of scalar row r after the blocks before this pass
    for (int q0 = lo; q0 < hi; q0 += G * NB) {
        int j[NB];
#pragma unroll
        for (int t = 0; t < NB; ++t) {
            const int q = q0 + t * G + g;
            j[t] = (lane_ok && q < hi) ? Aj[q] : -1;
        }
        int tie = 0;
        if (NB > 1) {
#pragma unroll
            for (int t = 0; t < NB; ++t) tie |= j[t];
            tie &= zero;
        }
        double pr[NB][8];
#pragma unroll
        for (int t = 0; t < NB; ++t) {
            const int q = q0 + t * G + g;
            const bool have = j[t] >= 0;
            const double *blk = Ax + ((size_t)q * R + r) * C + (t == 0 ? tie : 0);
            const double *xb = x + (size_t)(have ? j[t] : 0) * C;
            double bl[8], xv[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                bl[c] = (have && c < C) ? blk[c] : 0.0;
                xv[c] = (have && c < C) ? xb[c] : 0.0;
            }
#pragma unroll
            for (int c = 0; c < 8; ++c) pr[t][c] = mul_rn(bl[c], xv[c]);
        }
        if (NB > 1) {
#pragma unroll
            for (int t = 0; t < NB; ++t)
#pragma unroll
                for (int c = 0; c < 8; ++c) tie |= __double2loint(pr[t][c]);
            tie &= zero;
        }
#pragma unroll
        for (int t = 0; t < NB; ++t) {
            if (q0 + t * G >= hi) break;                  // warp-uniform
            const bool have = (t == 0 ? j[t] + tie : j[t]) >= 0;
            double a = carry;
#pragma unroll
            for (int gg = 0; gg < G; ++gg) {
                if (gg > 0) a = __shfl_sync(FULL, a, (gg - 1) * R + (lane_ok ? r : 0));   // sum after group gg - 1
                if (g == gg && have) {
#pragma unroll
                    for (int c = 0; c < 8; ++c)
                        if (c < C) a = add_rn(a, pr[t][c]);
                }
            }
            carry = __shfl_sync(FULL, a, (G - 1) * R + (lane_ok ? r : 0));
        }
    }
    if (lane < R) {
        const size_t row = (size_t)i * R + lane;
        if (mode == 0) y[row] = carry;
        else if (mode == 1) y[row] = sub_rn(b[row], carry);
        else y[row] = add_rn(b ? b[row] : y[row], carry);
    }
}

template <int BS>
__global__ void bsr_gs_sequential_kernel(const int *__restrict__ Ap, const int *__restrict__ Aj,
                                         const double *__restrict__ Ax, double *x, const double *__restrict__ b,
                                         const int *Id, int start, int stop, int step)
{
    pdl_enter();
    for (int p = start; p != stop; p += step) {
        bsr_gs_block_row<BS>(Id ? Id[p] : p, Ap, Aj, Ax, x, b, step < 0);
        __threadfence_block();
    }
}

#define BS_SWITCH(bs, CALL)                                                   \
    switch (bs) {                                                             \
    case 1: { constexpr int BS = 1; CALL; } break;                            \
    case 2: { constexpr int BS = 2; CALL; } break;                            \
    case 3: { constexpr int BS = 3; CALL; } break;                            \
    case 4: { constexpr int BS = 4; CALL; } break;                            \
    case 5: { constexpr int BS = 5; CALL; } break;                            \
    case 6: { constexpr int BS = 6; CALL; } break;                            \
    case 7: { constexpr int BS = 7; CALL; } break;                            \
    case 8: { constexpr int BS = 8; CALL; } break;                            \
    default: return set_err(B200AMG_EINVAL, "blocksize must be in 1..8");     \
    }

#define R_SWITCH(r, CALL)                                                     \
    switch (r) {                                                              \
    case 2: { constexpr int RR = 2; CALL; } break;                            \
    case 3: { constexpr int RR = 3; CALL; } break;                            \
    case 4: { constexpr int RR = 4; CALL; } break;                            \
    case 5: { constexpr int RR = 5; CALL; } break;                            \
    case 6: { constexpr int RR = 6; CALL; } break;                            \
    case 7: { constexpr int RR = 7; CALL; } break;                            \
    default: { constexpr int RR = 8; CALL; } break;                           \
    }

// B200AMG_BSR_WARP=0 keeps the lanes-per-row / thread-per-row kernels everywhere (A/B switch)
static bool bsr_warp_kernels()
{
    static const bool on = [] { const char *e = getenv("B200AMG_BSR_WARP"); return !(e && e[0] == '0'); }();
    return on;
}

// B200AMG_BSR_WARP_BATCH (A/B switch): 0 one pass per round of loads, as in the first version; 2 (default) batches only
// for sub-wave colours; 1 batches of two passes for the larger colours as well -- measured SLOWER on config 4 level 1
// (6 600 rows per colour: 0.674 vs 0.610 ms per smoothing step, profiles/r02/run35_summary.txt): 70 registers instead
// of 50 cost more resident warps than the halved rounds of latency give back
static int bsr_warp_batch()
{
    static const int mode = [] { const char *e = getenv("B200AMG_BSR_WARP_BATCH"); return (e && e[0] >= '0' && e[0] <= '2') ? e[0] - '0' : 2; }();
    return mode;
}

// nblocks: stored blocks of the operator (< 0: unknown) -- rows with at least half a pass of blocks on average go to
// the warp-per-block-row kernel
int bsr_matvec(int n_brow, int R, int C, const int *Ap, const int *Aj, const double *Ax, const double *x,
               const double *b, double *y, int mode, cudaStream_t st, long long nblocks)
{
    if (R < 1 || C < 1 || R > 8 || C > 8) return set_err(B200AMG_EINVAL, "bsr_matvec: block shape must be in 1..8");
    const long long n = (long long)n_brow * R;
    if (n <= 0) return 0;
    // warp per block row only where thread-per-scalar-row cannot fill the machine (few rows) and the rows are long
    // (measured on BASELINE config 4, profiles/r02: 1 331 x 6 rows 0.54 -> 0.13 ms; 2.08 M x 3 rows 0.53 -> 2.3 ms)
    if (R >= 2 && nblocks >= 0 && bsr_warp_kernels() && n < kSMs * 1024 && nblocks >= 2ll * (32 / R) * n_brow) {
        const unsigned grid = (unsigned)((n_brow + 3) / 4);
        if (bsr_warp_batch()) { R_SWITCH(R, B200_LAUNCH_PDL(pdl_class(PDL_OTHER), (bsr_matvec_warp_kernel<RR, 4>), grid, 128, 0, st, n_brow, C, Ap, Aj, Ax, x, b, y, mode, 0)); }
        else { R_SWITCH(R, B200_LAUNCH_PDL(pdl_class(PDL_OTHER), (bsr_matvec_warp_kernel<RR, 1>), grid, 128, 0, st, n_brow, C, Ap, Aj, Ax, x, b, y, mode, 0)); }
        return check_launch("bsr_matvec_warp_kernel");
    }
    B200_LAUNCH_PDL(pdl_class(PDL_OTHER) && n < PDL_SMALL_ROWS, bsr_matvec_kernel, (unsigned)((n + 255) / 256), 256, 0, st, n_brow, R, C, Ap, Aj, Ax, x, b, y, mode);
    return check_launch("bsr_matvec_kernel");
}

int bsr_jacobi(int bs, int nb0, int nb1, const int *Ap, const int *Aj, const double *Ax, const double *xin,
               double *xout, const double *b, double omega, cudaStream_t st)
{
    const long long n = (long long)(nb1 - nb0) * bs;
    if (n <= 0) return 0;
    const unsigned grid = (unsigned)((n + 255) / 256);
    BS_SWITCH(bs, B200_LAUNCH_PDL(pdl_class(PDL_OTHER) && n < PDL_SMALL_ROWS, bsr_jacobi_kernel<BS>, grid, 256, 0, st, nb0, nb1, Ap, Aj, Ax, xin, xout, b, omega));
    return check_launch("bsr_jacobi_kernel");
}

// perm != 0: (Ap, Aj, Ax) is a colour-grouped copy of the operator -- its row p is block row Id[p] (bsr_color_build)
int bsr_gs_color(int bs, const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
                 const int *Id, int p0, int p1, bool backward, cudaStream_t st, double blocks_per_row, bool perm, bool static_op)
{
    if (p1 <= p0) return 0;
    // warp per block row where the colour is too small for lanes-per-row to fill the machine and the rows are long
    // (config 4: level 1, 6 600 rows of 44 blocks per colour, 43 -> 27 us per launch; level 0, 260 k rows of 15: slower)
    if (bs >= 2 && bsr_warp_kernels() && (long long)(p1 - p0) * bs < kSMs * 1024 && blocks_per_row >= 2.0 * (32 / bs)) {
        const unsigned grid = (unsigned)((p1 - p0 + 3) / 4);       // 4 warps = 4 block rows per CTA
#define GS_WARP(BW, PM, NBV) R_SWITCH(bs, B200_LAUNCH_PDL(pdl_class(PDL_SWEEP) && static_op, (bsr_gs_color_warp_kernel<RR, BW, PM, (NBV > 1 && RR <= 3) ? 2 * NBV : NBV>), grid, 128, 0, st, Ap, Aj, Ax, x, b, Id, p0, p1, 0))
        // batches pay while the colour is a fraction of a wave (config 4, run 15: level 2, 48 rows per colour, 0.76 -> 0.56 ms
        // per step; level 1, 6 600 rows per colour, 0.65 -> 0.79 ms: the 128 registers cost resident warps there)
        if (bsr_warp_batch() && p1 - p0 <= kSMs * 8) {
            if (backward) { if (perm) { GS_WARP(true, true, 4); } else { GS_WARP(true, false, 4); } }
            else { if (perm) { GS_WARP(false, true, 4); } else { GS_WARP(false, false, 4); } }
        } else if (bsr_warp_batch() == 1) {
            // a full wave of warps: batches of two passes (about 80 registers) halve the rounds of latency per row
            // without giving up as many resident warps as batches of four
            if (backward) { if (perm) { GS_WARP(true, true, 2); } else { GS_WARP(true, false, 2); } }
            else { if (perm) { GS_WARP(false, true, 2); } else { GS_WARP(false, false, 2); } }
        } else {
            if (backward) { if (perm) { GS_WARP(true, true, 1); } else { GS_WARP(true, false, 1); } }
            else { if (perm) { GS_WARP(false, true, 1); } else { GS_WARP(false, false, 1); } }
        }
#undef GS_WARP
        return check_launch("bsr_gs_color_warp_kernel");
    }
    const int rows_per_cta = (32 / bs) * 4;             // 4 warps, 32 / bs block rows each
    const unsigned grid = (unsigned)((p1 - p0 + rows_per_cta - 1) / rows_per_cta);
    if (perm) { BS_SWITCH(bs, B200_LAUNCH_PDL(pdl_class(PDL_SWEEP) && (long long)(p1 - p0) * bs < PDL_SMALL_ROWS, (bsr_gs_color_kernel<BS, true>), grid, 128, 0, st, Ap, Aj, Ax, x, b, Id, p0, p1, backward)); }
    else { BS_SWITCH(bs, B200_LAUNCH_PDL(pdl_class(PDL_SWEEP) && (long long)(p1 - p0) * bs < PDL_SMALL_ROWS, (bsr_gs_color_kernel<BS, false>), grid, 128, 0, st, Ap, Aj, Ax, x, b, Id, p0, p1, backward)); }
    return check_launch("bsr_gs_color_kernel");
}

int bsr_gs_sequential(int bs, const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
                      const int *Id, int start, int stop, int step, cudaStream_t st)
{
    if (start == stop) return 0;
    BS_SWITCH(bs, B200_LAUNCH_PDL(false, bsr_gs_sequential_kernel<BS>, 1, 1, 0, st, Ap, Aj, Ax, x, b, Id, start, stop, step));
    return check_launch("bsr_gs_sequential_kernel");
}

}  // namespace b200amg
