// Whole smoothing steps of SMALL levels in one launch.
//
// Colour-ordered Gauss-Seidel costs one kernel launch per colour and sweep; on the coarse levels of a hierarchy a
// colour is a few hundred rows and the launch (plus one round of dependent memory latency) is all there is: config 3
// spends 0.15-0.19 ms per smoothing step on each of its levels 5-8 (5 777 ... 255 rows, 12-13 colours, 48 launches),
// config 4 0.8 ms on its 1 331-block-row level 2 (28 colours).  When the level's iterate fits in shared memory
// (<= 28 000 doubles) ONE thread block keeps it there for the whole step -- all sweeps, all colours, a block barrier
// between colours -- and only the operator streams in (from L2 after the first touch).  Same row arithmetic and the
// same colour order as gs_rows_kernel / bsr_gs_color_warp_kernel: bit-identical iterates.
#include "internal.cuh"

namespace b200amg {

constexpr int SL_TPB = 1024;

// dirs: bit q = direction of pass q (0 forward: colours 0..K-1, 1 backward), npass passes in all
template <bool BSRF>
__global__ void __launch_bounds__(SL_TPB, 1) gs_level_cta_kernel(int n, const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                                 const double *__restrict__ Ax, const int *__restrict__ Id,
                                                                 const int *__restrict__ cptr, int K, double *x,
                                                                 const double *__restrict__ b, int npass, unsigned dirs,
                                                                 int x_is_zero)
{
    extern __shared__ double sl_x[];
    for (int i = threadIdx.x; i < n; i += SL_TPB) sl_x[i] = x_is_zero ? 0.0 : x[i];
    __syncthreads();
    for (int q = 0; q < npass; ++q) {
        const bool bwd = (dirs >> q) & 1u;
        for (int cq = 0; cq < K; ++cq) {
            const int c = bwd ? K - 1 - cq : cq;
            const int p1 = cptr[c + 1];
            for (int p = cptr[c] + (int)threadIdx.x; p < p1; p += SL_TPB) {
                const int i = Id[p];
                double acc = BSRF ? b[i] : 0.0, diag = 0.0;
                const int hi = Ap[i + 1];
#pragma unroll 4
                for (int k = Ap[i]; k < hi; ++k) {
                    const int j = Aj[k];
                    const double v = Ax[k];
                    if (j == i) diag = v;
                    else acc = BSRF ? sub_rn(acc, mul_rn(v, sl_x[j])) : add_rn(acc, mul_rn(v, sl_x[j]));
                }
                if (diag != 0.0) sl_x[i] = BSRF ? div_rn(acc, diag) : div_rn(sub_rn(b[i], acc), diag);
            }
            __syncthreads();
        }
    }
    for (int i = threadIdx.x; i < n; i += SL_TPB) x[i] = sl_x[i];
}

// block form: one warp per block row of the colour (the arithmetic of bsr_gs_color_warp_kernel), 32 rows in flight
template <int BS>
__global__ void __launch_bounds__(SL_TPB, 1) bsr_gs_level_cta_kernel(int n, const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                                     const double *__restrict__ Ax, const int *__restrict__ Id,
                                                                     const int *__restrict__ cptr, int K, double *x,
                                                                     const double *__restrict__ b, int npass, unsigned dirs,
                                                                     int x_is_zero)
{
    constexpr int G = 32 / BS;
    constexpr unsigned FULL = 0xffffffffu;
    extern __shared__ double sl_x[];
    for (int i = threadIdx.x; i < n; i += SL_TPB) sl_x[i] = x_is_zero ? 0.0 : x[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane / BS, k = lane - g * BS;
    const bool lane_ok = g < G;
    for (int q = 0; q < npass; ++q) {
        const bool backward = (dirs >> q) & 1u;
        for (int cq = 0; cq < K; ++cq) {
            const int c = backward ? K - 1 - cq : cq;
            const int p1 = cptr[c + 1];
            for (int p = cptr[c] + warp; p < p1; p += SL_TPB / 32) {
                const int i = Id[p];
                const int lo = Ap[i], hi = Ap[i + 1];
                double acc = (lane < BS) ? b[(size_t)i * BS + lane] : 0.0;
                int dptr = -1;
                for (int q0 = lo; q0 < hi; q0 += G) {
                    const int qq = q0 + g;
                    const bool have = lane_ok && qq < hi;
                    const int j = have ? Aj[qq] : -1;
                    const bool isdiag = have && j == i;
                    double loc = 0.0;
                    if (have && !isdiag) {
                        const double *blk = Ax + ((size_t)qq * BS + k) * BS;
                        const double *xb = sl_x + (size_t)j * BS;
#pragma unroll
                        for (int cc = 0; cc < BS; ++cc) loc = add_rn(loc, mul_rn(blk[cc], xb[cc]));
                    }
                    const unsigned dg = __ballot_sync(FULL, isdiag);
                    if (dg) dptr = q0 + (31 - __clz(dg)) / BS;
                    loc = (have && !isdiag) ? loc : 0.0;            // acc - (+0.0) == acc bit for bit (bsr.cu)
#pragma unroll
                    for (int gg = 0; gg < G; ++gg) acc = sub_rn(acc, __shfl_sync(FULL, loc, gg * BS + (lane < BS ? lane : 0)));
                }
                if (dptr < 0) continue;                             // warp-uniform
                double xl[BS], Dk[BS];
#pragma unroll
                for (int cc = 0; cc < BS; ++cc) {
                    xl[cc] = sl_x[(size_t)i * BS + cc];
                    Dk[cc] = (lane < BS) ? Ax[((size_t)dptr * BS + lane) * BS + cc] : 0.0;
                }
#pragma unroll
                for (int kq = 0; kq < BS; ++kq) {
                    double nv[2] = {0.0, 0.0};
                    bool wrote[2] = {false, false};
                    // both candidate rows of this step (forward: kq, backward: BS-1-kq) are written with compile-time
                    // indices; the pass direction picks one
#pragma unroll
                    for (int d = 0; d < 2; ++d) {
                        const int kk0 = d ? BS - 1 - kq : kq;
                        if ((d == 1) != backward) continue;
                        if (lane == kk0) {
                            double dd = 1.0, a = acc;
#pragma unroll
                            for (int cq2 = 0; cq2 < BS; ++cq2) {
                                const int kk = d ? BS - 1 - cq2 : cq2;
                                if (kk == kk0) dd = Dk[kk];
                                else a = sub_rn(a, mul_rn(Dk[kk], xl[kk]));
                            }
                            if (dd != 0.0) {
                                nv[d] = div_rn(a, dd);
                                wrote[d] = true;
                                sl_x[(size_t)i * BS + kk0] = nv[d];
                            }
                        }
                        const double cur = wrote[d] ? nv[d] : xl[kk0];
                        xl[kk0] = __shfl_sync(FULL, cur, kk0);
                    }
                }
            }
            __syncthreads();
        }
    }
    for (int i = threadIdx.x; i < n; i += SL_TPB) x[i] = sl_x[i];
}

// rows (scalar unknowns) whose iterate fits one CTA's shared memory
constexpr int SL_MAX_N = 28000;

static unsigned pass_dirs(int iters, int sweep, int *npass)
{
    unsigned dirs = 0;
    int q = 0;
    for (int it = 0; it < iters && q < 30; ++it) {
        if (sweep == 0 || sweep == 2) ++q;                           // forward pass: bit stays 0
        if (sweep == 1 || sweep == 2) dirs |= 1u << q++;
    }
    *npass = q;
    return dirs;
}

// -1: not eligible (the per-colour launches run), else the launch status.  cptr_dev: K + 1 colour offsets ON THE DEVICE
// One SM streams the whole operator once per pass, so this pays only while the operator is tiny: measured on config 3
// (profiles/r02/run12_summary.txt) 1 778 rows / 29 k entries 0.174 -> 0.096 ms per step, 662 rows 0.141 -> 0.065 ms, but
// 5 777 rows / 101 k entries no gain and 21 373 rows / 400 k entries 0.18 -> 0.96 ms; config 4's 1 331 block rows of
// 55 6x6 blocks (2.6 M values) 0.84 -> 1.3 ms.  Hence the limit on stored values.
constexpr long long SL_MAX_VALUES = 40000;

int gs_level_cta(int bs, bool bsrf, int n, long long stored_values, const int *Ap, const int *Aj, const double *Ax, const int *Id,
                 const int *cptr_dev, int K, double *x, const double *b, int iters, int sweep, bool x_is_zero, cudaStream_t st)
{
    static const bool on = [] { const char *e = getenv("B200AMG_SMALL_LEVEL"); return !(e && e[0] == '0'); }();
    if (!on || !cptr_dev || n > SL_MAX_N || n <= 0 || K <= 0 || iters <= 0 || iters > 15) return -1;
    if (stored_values < 0 || stored_values > SL_MAX_VALUES) return -1;
    int npass = 0;
    const unsigned dirs = pass_dirs(iters, sweep, &npass);
    const int smem = (int)sizeof(double) * n;
    if (bs == 1) {
        static bool set1 = false;
        if (!set1) {
            B200_CUDA(cudaFuncSetAttribute(gs_level_cta_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(double) * SL_MAX_N));
            B200_CUDA(cudaFuncSetAttribute(gs_level_cta_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(double) * SL_MAX_N));
            set1 = true;
        }
        if (bsrf) gs_level_cta_kernel<true><<<1, SL_TPB, smem, st>>>(n, Ap, Aj, Ax, Id, cptr_dev, K, x, b, npass, dirs, x_is_zero);
        else gs_level_cta_kernel<false><<<1, SL_TPB, smem, st>>>(n, Ap, Aj, Ax, Id, cptr_dev, K, x, b, npass, dirs, x_is_zero);
        return check_launch("gs_level_cta_kernel");
    }
#define SL_BSR(BSV)                                                                                                        \
    case BSV: {                                                                                                            \
        static bool setb = false;                                                                                          \
        if (!setb) {                                                                                                       \
            B200_CUDA(cudaFuncSetAttribute(bsr_gs_level_cta_kernel<BSV>, cudaFuncAttributeMaxDynamicSharedMemorySize,      \
                                           (int)sizeof(double) * SL_MAX_N));                                               \
            setb = true;                                                                                                   \
        }                                                                                                                  \
        bsr_gs_level_cta_kernel<BSV><<<1, SL_TPB, smem, st>>>(n, Ap, Aj, Ax, Id, cptr_dev, K, x, b, npass, dirs, x_is_zero); \
    } break;
    switch (bs) {
        SL_BSR(2) SL_BSR(3) SL_BSR(4) SL_BSR(5) SL_BSR(6) SL_BSR(7) SL_BSR(8)
    default: return -1;
    }
#undef SL_BSR
    return check_launch("bsr_gs_level_cta_kernel");
}

}  // namespace b200amg
