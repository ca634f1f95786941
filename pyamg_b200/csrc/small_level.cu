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

constexpr int SL_TPB = 1024;        // block form
constexpr int SL_TPB1 = 256;        // scalar form: few threads, many registers -- a colour of such a level is <= ~300 rows
constexpr int SL_U = 16;            // operator entries of the NEXT colour's row held in registers while this colour folds

// dirs: bit q = direction of pass q (0 forward: colours 0..K-1, 1 backward), npass passes in all.
// A phase (one colour of one pass) used to cost Id -> Ap -> (Aj, Ax) -> fold = 4-6 dependent L2 latencies (~3 us; 26
// phases per symmetric step).  The operator does not change, only x does: each thread requests the row it will relax in
// the NEXT phase -- index, bounds, b, the first SL_U (column, value) pairs -- BEFORE it folds the current one, so after
// the block barrier only shared-memory reads of x and the ordered fold remain on the critical path.
template <bool BSRF>
__global__ void __launch_bounds__(SL_TPB1, 1) gs_level_cta_kernel(int n, const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                                  const double *__restrict__ Ax, const int *__restrict__ Id,
                                                                  const int *__restrict__ cptr, int K, double *x,
                                                                  const double *__restrict__ b, int npass, unsigned dirs,
                                                                  int x_is_zero)
{
    pdl_enter();
    extern __shared__ double sl_x[];
    const int tid = threadIdx.x;
    const int nphase = npass * K;
    auto colour_of = [&](int ph) {
        const int q = ph / K, cq = ph - q * K;
        return ((dirs >> q) & 1u) ? K - 1 - cq : cq;
    };
    // one row walk from global memory (rows beyond the first of a colour, and the tail of rows longer than SL_U)
    auto walk = [&](int i, int k, int hi, double &acc, double &diag) {
#pragma unroll 4
        for (; k < hi; ++k) {
            const int j = Aj[k];
            const double v = Ax[k];
            if (j == i) diag = v;
            else acc = BSRF ? sub_rn(acc, mul_rn(v, sl_x[j])) : add_rn(acc, mul_rn(v, sl_x[j]));
        }
    };
    int ci = -1, clo = 0, chi = 0, cc[SL_U];        // current phase's first row of this thread
    double cb = 0.0, cv[SL_U];
    auto request = [&](int ph, int &ri, int &rlo, int &rhi, double &rb, int (&rc)[SL_U], double (&rv)[SL_U]) {
        ri = -1;
        rlo = rhi = 0;
        rb = 0.0;
        if (ph < nphase) {
            const int c = colour_of(ph);
            const int p = cptr[c] + tid;
            if (p < cptr[c + 1]) {
                ri = Id[p];
                rlo = Ap[ri];
                rhi = Ap[ri + 1];
                rb = b[ri];
            }
        }
#pragma unroll
        for (int u = 0; u < SL_U; ++u) {
            const bool in = rlo + u < rhi;
            rc[u] = in ? Aj[rlo + u] : -1;
            rv[u] = in ? Ax[rlo + u] : 0.0;
        }
    };
    request(0, ci, clo, chi, cb, cc, cv);
    for (int i = tid; i < n; i += SL_TPB1) sl_x[i] = x_is_zero ? 0.0 : x[i];
    __syncthreads();
    for (int ph = 0; ph < nphase; ++ph) {
        int ni, nlo, nhi, nc[SL_U];
        double nb, nv[SL_U];
        request(ph + 1, ni, nlo, nhi, nb, nc, nv);
        __syncwarp();                                   // the requests are issued before the fold below touches x
        if (ci >= 0) {
            double acc = BSRF ? cb : 0.0, diag = 0.0;
#pragma unroll
            for (int u = 0; u < SL_U; ++u) {
                if (cc[u] >= 0) {
                    if (cc[u] == ci) diag = cv[u];
                    else acc = BSRF ? sub_rn(acc, mul_rn(cv[u], sl_x[cc[u]])) : add_rn(acc, mul_rn(cv[u], sl_x[cc[u]]));
                }
            }
            walk(ci, clo + SL_U, chi, acc, diag);
            if (diag != 0.0) sl_x[ci] = BSRF ? div_rn(acc, diag) : div_rn(sub_rn(cb, acc), diag);
            // colours wider than the block: the remaining rows of this thread, straight from memory
            const int c = colour_of(ph);
            const int p1 = cptr[c + 1];
            for (int p = cptr[c] + tid + SL_TPB1; p < p1; p += SL_TPB1) {
                const int i = Id[p];
                double a2 = BSRF ? b[i] : 0.0, d2 = 0.0;
                walk(i, Ap[i], Ap[i + 1], a2, d2);
                if (d2 != 0.0) sl_x[i] = BSRF ? div_rn(a2, d2) : div_rn(sub_rn(b[i], a2), d2);
            }
        }
        __syncthreads();
        ci = ni; clo = nlo; chi = nhi; cb = nb;
#pragma unroll
        for (int u = 0; u < SL_U; ++u) { cc[u] = nc[u]; cv[u] = nv[u]; }
    }
    for (int i = tid; i < n; i += SL_TPB1) x[i] = sl_x[i];
}

// block form: one warp per block row of the colour (the arithmetic of bsr_gs_color_warp_kernel), 32 rows in flight
template <int BS>
__global__ void __launch_bounds__(SL_TPB, 1) bsr_gs_level_cta_kernel(int n, const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                                     const double *__restrict__ Ax, const int *__restrict__ Id,
                                                                     const int *__restrict__ cptr, int K, double *x,
                                                                     const double *__restrict__ b, int npass, unsigned dirs,
                                                                     int x_is_zero)
{
    pdl_enter();
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

// host emulation (tests/emu) has no function attributes: the shared-memory opt-in is a device matter
#ifdef B200_HOST_EMULATION
#define SL_OPT_IN(kernel) ((void)0)
#else
#define SL_OPT_IN(kernel)                                                                                                  \
    do {                                                                                                                   \
        static bool set = false;                                                                                           \
        if (!set) {                                                                                                        \
            B200_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(double) * SL_MAX_N)); \
            set = true;                                                                                                    \
        }                                                                                                                  \
    } while (0)
#endif

int gs_level_cta(int bs, bool bsrf, int n, long long stored_values, const int *Ap, const int *Aj, const double *Ax, const int *Id,
                 const int *cptr_dev, int K, double *x, const double *b, int iters, int sweep, bool x_is_zero, cudaStream_t st)
{
    // B200AMG_SMALL_LEVEL: 0 never, 1 block levels only (default), 2 scalar levels too.  Since the per-colour launches
    // of small colours start early (programmatic dependent launch, operator loads before the wait: ~3 us per colour) they
    // beat the one-block scalar form (~4 us per colour phase): config 3 level 7, 0.104 -> 0.084 ms per step
    // (profiles/r02/run19_summary.txt).
    const char *env = getenv("B200AMG_SMALL_LEVEL");         // (read per call: tests switch it)
    const int mode = (env && env[0] >= '0' && env[0] <= '2') ? env[0] - '0' : 1;
    if (mode == 0 || (mode == 1 && bs == 1)) return -1;
    if (!cptr_dev || n > SL_MAX_N || n <= 0 || K <= 0 || iters <= 0 || iters > 15) return -1;
    if (stored_values < 0 || stored_values > SL_MAX_VALUES) return -1;
    int npass = 0;
    const unsigned dirs = pass_dirs(iters, sweep, &npass);
    const int smem = (int)sizeof(double) * n;
    const int zi = x_is_zero ? 1 : 0;
    if (bs == 1) {
        if (bsrf) {
            SL_OPT_IN(gs_level_cta_kernel<true>);
            B200_LAUNCH_PDL(pdl_class(PDL_SWEEP), gs_level_cta_kernel<true>, 1, SL_TPB1, smem, st, n, Ap, Aj, Ax, Id, cptr_dev, K, x, b, npass, dirs, zi);
        } else {
            SL_OPT_IN(gs_level_cta_kernel<false>);
            B200_LAUNCH_PDL(pdl_class(PDL_SWEEP), gs_level_cta_kernel<false>, 1, SL_TPB1, smem, st, n, Ap, Aj, Ax, Id, cptr_dev, K, x, b, npass, dirs, zi);
        }
        return check_launch("gs_level_cta_kernel");
    }
#define SL_BSR(BSV)                                                                                                          \
    case BSV: {                                                                                                              \
        SL_OPT_IN(bsr_gs_level_cta_kernel<BSV>);                                                                             \
        B200_LAUNCH_PDL(pdl_class(PDL_SWEEP), bsr_gs_level_cta_kernel<BSV>, 1, SL_TPB, smem, st, n, Ap, Aj, Ax, Id, cptr_dev, K, x, b, npass, dirs, zi); \
    } break;
    switch (bs) {
        SL_BSR(2) SL_BSR(3) SL_BSR(4) SL_BSR(5) SL_BSR(6) SL_BSR(7) SL_BSR(8)
    default: return -1;
    }
#undef SL_BSR
    return check_launch("bsr_gs_level_cta_kernel");
}

}  // namespace b200amg
