// Galerkin product rows WITHOUT a sort (the contract is spgemm.cu's: scipy csr_matmat, pyamg/aggregation/aggregation.py:429):
// one warp per row of C = A B, for rows in scipy's output order.
//   stage   every product a_ik * b_kj of the row, in scipy's visiting order (k in A's row order, then B's row order),
//           as (column, value) in shared memory;
//   walk    the staged products 32 at a time IN THAT ORDER: lanes holding the same column form a group
//           (__match_any_sync), the group's lowest lane finds or claims the column's slot of an open-addressing table in
//           shared memory and adds the group's values to the slot's sum in lane order -- so every column receives exactly
//           scipy's `sums[j] += v * Bx[kk]` sequence, starting from 0.0, separately rounded.  A newly claimed slot is
//           appended to a list: the list is scipy's first-touch order;
//   emit    the list backwards (scipy: reverse first touch), sums that are exactly 0.0 dropped as scipy drops them.
// The sorting kernels of spgemm.cu make O(m log^2 m) compare-exchanges per row of m products; this form makes m / 32
// steps.  Rows of up to 512 products stage in shared memory (spgemm_hash_kernel), rows of up to 8192 in their own global
// scratch (spgemm_hash_long_kernel); still longer rows, and column-sorted output, stay with spgemm.cu.
// Measured at 256^3 (profiles/r02/run39, run40): (R A) P of level 0 0.0975 -> 0.0512 s, of level 1 0.1122 -> 0.0704 s, every
// array bit-identical to the sorting kernels'.
#include <cstdlib>

#include "internal.cuh"
#include "setup.cuh"

namespace b200amg {

constexpr int SPH_EMPTY = -1;
// bytes of shared memory per warp: vals[CAP] f64 | sums[2 CAP] f64 | cols[CAP] i32 | keys[2 CAP] i32 | list[CAP] i32
constexpr int SPH_BYTES_PER_PRODUCT = 40;

#ifdef B200_HOST_EMULATION
extern unsigned char sph_smem[];
#endif

template <int CAP, int WPB>
__global__ void __launch_bounds__(32 * WPB) spgemm_hash_kernel(const int *__restrict__ rows, int nrows,
                                                               const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                               const double *__restrict__ Ax, const int *__restrict__ Bp,
                                                               const int *__restrict__ Bj, const double *__restrict__ Bx,
                                                               const long long *__restrict__ ubp, int *Tj, double *Tx, int *cnt)
{
#ifndef B200_HOST_EMULATION
    extern __shared__ __align__(16) unsigned char sph_smem[];
#endif
    const unsigned FULL = 0xffffffffu;
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int r = blockIdx.x * WPB + w;
    if (r >= nrows) return;                                   // (whole warps leave)
    unsigned char *mine = sph_smem + (size_t)w * (CAP * SPH_BYTES_PER_PRODUCT);
    double *vals = reinterpret_cast<double *>(mine);
    double *sums = vals + CAP;
    int *cols = reinterpret_cast<int *>(sums + 2 * CAP);
    int *keys = cols + CAP;
    int *list = keys + 2 * CAP;
    const int i = rows[r];
    const long long o = ubp[i];
    const int m = (int)(ubp[i + 1] - o);                      // products of this row, <= CAP
    int T = 32;                                               // table slots: a power of two >= 2 m
    while (T < 2 * m) T <<= 1;
    const unsigned tmask = (unsigned)T - 1u;
    for (int t = lane; t < T; t += 32) keys[t] = SPH_EMPTY;

    // ---- stage ----
    const int a_lo = Ap[i], a_hi = Ap[i + 1];
    int base = 0;
    for (int k0 = a_lo; k0 < a_hi; k0 += 32) {
        const int k = k0 + lane;
        int len = 0, q0 = 0;
        if (k < a_hi) { const int j = Aj[k]; q0 = Bp[j]; len = Bp[j + 1] - q0; }
        int inc = len;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int up = __shfl_up_sync(FULL, inc, d);
            if (lane >= d) inc += up;
        }
        const int tot = __shfl_sync(FULL, inc, 31);
        if (len > 0) {
            const double a = Ax[k];
            int p = base + inc - len;
            for (int t = 0; t < len; ++t, ++p) {
                cols[p] = Bj[q0 + t];
                vals[p] = mul_rn(a, Bx[q0 + t]);
            }
        }
        base += tot;
    }
    __syncwarp();

    // ---- walk ----
    int ndist = 0;
    for (int p0 = 0; p0 < m; p0 += 32) {
        const int p = p0 + lane;
        const bool act = p < m;
        const int col = act ? cols[p] : -2 - lane;            // idle lanes: groups of their own
        const double v = act ? vals[p] : 0.0;
        const unsigned grp = __match_any_sync(FULL, col);
        const bool lead = act && (__ffs((int)grp) - 1 == lane);
        int slot = 0;
        bool fresh = false;
        if (lead) {
            unsigned h = (((unsigned)col * 0x9E3779B1u) >> 9) & tmask;
            for (;;) {
                const int cur = *reinterpret_cast<volatile int *>(keys + h);
                if (cur == col) break;
                if (cur == SPH_EMPTY) {
                    const int old = atomicCAS(keys + h, SPH_EMPTY, col);
                    if (old == SPH_EMPTY) { fresh = true; break; }
                    if (old == col) break;
                }
                h = (h + 1u) & tmask;
            }
            slot = (int)h;
        }
        const unsigned fb = __ballot_sync(FULL, fresh);
        if (fresh) list[ndist + __popc(fb & ((1u << lane) - 1u))] = slot;
        ndist += __popc(fb);
        double s = (lead && !fresh) ? sums[slot] : 0.0;
        unsigned rest = lead ? grp : 0u;
        while (__any_sync(FULL, rest != 0u)) {                // as many rounds as the largest group has lanes
            const int src = rest ? __ffs((int)rest) - 1 : lane;
            const double pv = __shfl_sync(FULL, v, src);
            if (rest) { s = add_rn(s, pv); rest &= rest - 1u; }
        }
        if (lead) sums[slot] = s;
        __syncwarp();
    }

    // ---- emit ----
    int kept = 0;
    for (int q0 = 0; q0 < ndist; q0 += 32) {
        const int q = q0 + lane;
        bool keep = false;
        int slot = 0;
        double s = 0.0;
        if (q < ndist) { slot = list[ndist - 1 - q]; s = sums[slot]; keep = (s != 0.0); }
        const unsigned kb = __ballot_sync(FULL, keep);
        if (keep) {
            const int pos = kept + __popc(kb & ((1u << lane) - 1u));
            Tj[o + pos] = keys[slot];
            Tx[o + pos] = s;
        }
        kept += __popc(kb);
    }
    if (lane == 0) cnt[i] = kept;
}

// Rows of 513 .. 8192 products: the same walk, one warp per CTA, with the products staged in the row's OWN SCRATCH in global
// memory (Tj / Tx at the row's offset, 12 bytes per product: dead once the walk is over, then overwritten by the row's
// output) and only the table in shared memory: sums[TCAP] f64 | keys[TCAP] i32 | list[TCAP] u16.  The table has
// min(TCAP, >= 2 m) slots; a row of TCAP distinct columns fills it completely, which the probe loop survives (a column is
// either found or an empty slot exists).
template <int TCAP>
__global__ void __launch_bounds__(32) spgemm_hash_long_kernel(const int *__restrict__ rows, int nrows,
                                                              const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                              const double *__restrict__ Ax, const int *__restrict__ Bp,
                                                              const int *__restrict__ Bj, const double *__restrict__ Bx,
                                                              const long long *__restrict__ ubp, int *Tj, double *Tx, int *cnt)
{
#ifndef B200_HOST_EMULATION
    extern __shared__ __align__(16) unsigned char sph_smem[];
#endif
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    double *sums = reinterpret_cast<double *>(sph_smem);
    int *keys = reinterpret_cast<int *>(sums + TCAP);
    unsigned short *list = reinterpret_cast<unsigned short *>(keys + TCAP);
    const int i = rows[blockIdx.x];
    const long long o = ubp[i];
    const int m = (int)(ubp[i + 1] - o);                      // products of this row, <= TCAP
    int *gc = Tj + o;
    double *gv = Tx + o;
    int T = 32;
    while (T < 2 * m && T < TCAP) T <<= 1;
    const unsigned tmask = (unsigned)T - 1u;
    for (int t = lane; t < T; t += 32) keys[t] = SPH_EMPTY;

    // ---- stage (global) ----
    const int a_lo = Ap[i], a_hi = Ap[i + 1];
    int base = 0;
    for (int k0 = a_lo; k0 < a_hi; k0 += 32) {
        const int k = k0 + lane;
        int len = 0, q0 = 0;
        if (k < a_hi) { const int j = Aj[k]; q0 = Bp[j]; len = Bp[j + 1] - q0; }
        int inc = len;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int up = __shfl_up_sync(FULL, inc, d);
            if (lane >= d) inc += up;
        }
        const int tot = __shfl_sync(FULL, inc, 31);
        if (len > 0) {
            const double a = Ax[k];
            int p = base + inc - len;
            for (int t = 0; t < len; ++t, ++p) {
                gc[p] = Bj[q0 + t];
                gv[p] = mul_rn(a, Bx[q0 + t]);
            }
        }
        base += tot;
    }
    __threadfence_block();
    __syncwarp();

    // ---- walk (the next step's products are requested before this step's are folded) ----
    int ndist = 0;
    int ncol = lane < m ? gc[lane] : 0;
    double nval = lane < m ? gv[lane] : 0.0;
    for (int p0 = 0; p0 < m; p0 += 32) {
        const int p = p0 + lane;
        const bool act = p < m;
        const int col = act ? ncol : -2 - lane;
        const double v = act ? nval : 0.0;
        if (p + 32 < m) { ncol = gc[p + 32]; nval = gv[p + 32]; }
        const unsigned grp = __match_any_sync(FULL, col);
        const bool lead = act && (__ffs((int)grp) - 1 == lane);
        int slot = 0;
        bool fresh = false;
        if (lead) {
            unsigned h = (((unsigned)col * 0x9E3779B1u) >> 9) & tmask;
            for (;;) {
                const int cur = *reinterpret_cast<volatile int *>(keys + h);
                if (cur == col) break;
                if (cur == SPH_EMPTY) {
                    const int old = atomicCAS(keys + h, SPH_EMPTY, col);
                    if (old == SPH_EMPTY) { fresh = true; break; }
                    if (old == col) break;
                }
                h = (h + 1u) & tmask;
            }
            slot = (int)h;
        }
        const unsigned fb = __ballot_sync(FULL, fresh);
        if (fresh) list[ndist + __popc(fb & ((1u << lane) - 1u))] = (unsigned short)slot;
        ndist += __popc(fb);
        double s = (lead && !fresh) ? sums[slot] : 0.0;
        unsigned rest = lead ? grp : 0u;
        while (__any_sync(FULL, rest != 0u)) {
            const int src = rest ? __ffs((int)rest) - 1 : lane;
            const double pv = __shfl_sync(FULL, v, src);
            if (rest) { s = add_rn(s, pv); rest &= rest - 1u; }
        }
        if (lead) sums[slot] = s;
        __syncwarp();
    }

    // ---- emit (over the staged products, which are dead now) ----
    int kept = 0;
    for (int q0 = 0; q0 < ndist; q0 += 32) {
        const int q = q0 + lane;
        bool keep = false;
        int slot = 0;
        double s = 0.0;
        if (q < ndist) { slot = list[ndist - 1 - q]; s = sums[slot]; keep = (s != 0.0); }
        const unsigned kb = __ballot_sync(FULL, keep);
        if (keep) {
            const int pos = kept + __popc(kb & ((1u << lane) - 1u));
            Tj[o + pos] = keys[slot];
            Tx[o + pos] = s;
        }
        kept += __popc(kb);
    }
    if (lane == 0) cnt[i] = kept;
}

template <int TCAP>
static int sph_launch_long(const int *rows, int nrows, const int *Ap, const int *Aj, const double *Ax, const int *Bp, const int *Bj,
                           const double *Bx, const long long *ubp, int *Tj, double *Tx, int *cnt, cudaStream_t st)
{
    constexpr int smem = TCAP * 14;
    auto kern = spgemm_hash_long_kernel<TCAP>;
#ifndef B200_HOST_EMULATION
    static bool attr_set = false;
    if (!attr_set) {
        B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_set = true;
    }
#endif
    B200_LAUNCH(kern, nrows, 32, smem, st, rows, nrows, Ap, Aj, Ax, Bp, Bj, Bx, ubp, Tj, Tx, cnt);
    return check_launch("spgemm_hash_long_kernel");
}

// B200AMG_SPGEMM_HASH: 0 = the sorting kernels for every row, 1 = rows of up to 512 products here, 2 = up to 2048,
// 3 = up to 8192 (rows of more than 512 products staged in global memory)
int spgemm_hash_mode()
{
    const char *e = getenv("B200AMG_SPGEMM_HASH");
    return e ? atoi(e) : SPGEMM_HASH_DEFAULT;
}

template <int CAP, int WPB>
static int sph_launch(const int *rows, int nrows, const int *Ap, const int *Aj, const double *Ax, const int *Bp, const int *Bj,
                      const double *Bx, const long long *ubp, int *Tj, double *Tx, int *cnt, cudaStream_t st)
{
    constexpr int smem = CAP * SPH_BYTES_PER_PRODUCT * WPB;
    auto kern = spgemm_hash_kernel<CAP, WPB>;
#ifndef B200_HOST_EMULATION
    static bool attr_set = false;
    if (smem > 48 * 1024 && !attr_set) {
        B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_set = true;
    }
#endif
    B200_LAUNCH(kern, (nrows + WPB - 1) / WPB, 32 * WPB, smem, st, rows, nrows, Ap, Aj, Ax, Bp, Bj, Bx, ubp, Tj, Tx, cnt);
    return check_launch("spgemm_hash_kernel");
}

// the rows of bin `bin` of spgemm_numeric's binning (limits 32 / 128 / 512 / 2048 / 8192 products): 1 = done here,
// 0 = not taken (the caller runs its sorting kernel), < 0 = error
int spgemm_hash_bin(int bin, int mode, const int *rows, int nrows, const int *Ap, const int *Aj, const double *Ax, const int *Bp,
                    const int *Bj, const double *Bx, const long long *ubp, int *Tj, double *Tx, int *cnt, cudaStream_t st)
{
    if (mode <= 0 || bin > (mode >= 3 ? 4 : mode >= 2 ? 3 : 2)) return 0;
    int rc;
    if (mode >= 3 && bin >= 3) {
        rc = bin == 3 ? sph_launch_long<4096>(rows, nrows, Ap, Aj, Ax, Bp, Bj, Bx, ubp, Tj, Tx, cnt, st)
                      : sph_launch_long<8192>(rows, nrows, Ap, Aj, Ax, Bp, Bj, Bx, ubp, Tj, Tx, cnt, st);
        return rc ? rc : 1;
    }
    switch (bin) {
    case 0: rc = sph_launch<32, 8>(rows, nrows, Ap, Aj, Ax, Bp, Bj, Bx, ubp, Tj, Tx, cnt, st); break;
    case 1: rc = sph_launch<128, 8>(rows, nrows, Ap, Aj, Ax, Bp, Bj, Bx, ubp, Tj, Tx, cnt, st); break;
    case 2: rc = sph_launch<512, 4>(rows, nrows, Ap, Aj, Ax, Bp, Bj, Bx, ubp, Tj, Tx, cnt, st); break;
    default: rc = sph_launch<2048, 1>(rows, nrows, Ap, Aj, Ax, Bp, Bj, Bx, ubp, Tj, Tx, cnt, st); break;
    }
    return rc ? rc : 1;
}

}  // namespace b200amg
