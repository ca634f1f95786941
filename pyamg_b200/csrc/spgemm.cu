// Galerkin product building block: row-wise C = A B (scipy csr_matmat semantics;
// pyamg/aggregation/aggregation.py:429, pyamg/classical/classical.py:187 evaluate R*A*P as two
// such products) by expand - sort - compress, one row per warp / CTA:
//   expand   every product a_ik * b_kj of the row, tagged key = (j << 32) | sequence number,
//            the sequence being scipy's visiting order (k in A's row order, then B's row order);
//   sort     the (key, product) pairs with a bitonic network in shared memory (global memory
//            for rows with more than 8192 products);
//   compress each run of equal j is summed IN SEQUENCE ORDER starting from 0.0 with separately
//            rounded adds -- exactly scipy's `sums[j] += v * Bx[kk]` -- and runs whose sum is
//            exactly 0.0 are dropped, as scipy drops them.
// Output rows come out SORTED by column (scipy: reverse first-touch order); after
// sort_indices() pattern and values are bit-identical to scipy's.
#include "internal.cuh"
#include "setup.cuh"

namespace b200amg {

typedef unsigned long long u64;

// ---- group primitives -------------------------------------------------------------------
struct WarpGroup {
    static constexpr bool kWarp = true;
    __device__ static __forceinline__ int tid() { return threadIdx.x & 31; }
    __device__ static __forceinline__ int size() { return 32; }
    __device__ static __forceinline__ void sync() { __syncwarp(); }
    __device__ static __forceinline__ int exscan(int v, int *total, int *) {
        const int lane = threadIdx.x & 31;
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int up = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += up;
        }
        *total = __shfl_sync(0xffffffffu, inc, 31);
        return inc - v;
    }
};

struct CtaGroup {
    static constexpr bool kWarp = false;
    __device__ static __forceinline__ int tid() { return threadIdx.x; }
    __device__ static __forceinline__ int size() { return blockDim.x; }
    __device__ static __forceinline__ void sync() { __syncthreads(); }
    __device__ static __forceinline__ int exscan(int v, int *total, int *s_w /* 33 ints */) {
        const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int up = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += up;
        }
        if (lane == 31) s_w[w] = inc;
        __syncthreads();
        if (w == 0) {
            int ws = lane < nw ? s_w[lane] : 0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int up = __shfl_up_sync(0xffffffffu, ws, o);
                if (lane >= o) ws += up;
            }
            s_w[lane] = ws;
        }
        __syncthreads();
        const int base = w ? s_w[w - 1] : 0;
        *total = s_w[nw - 1];
        __syncthreads();
        return base + inc - v;
    }
};

// normalised bitonic network over m elements (virtual +inf padding up to the next power of 2)
template <class G, bool HASVAL>
__device__ __forceinline__ void bitonic_sort(u64 *keys, double *vals, int m)
{
    int npad = 1;
    while (npad < m) npad <<= 1;
    const int half = npad >> 1;
    for (int size = 2; size <= npad; size <<= 1) {
        for (int t = G::tid(); t < half; t += G::size()) {          // flip step
            const int hs = size >> 1;
            const int blk = t / hs, off = t - blk * hs;
            const int i = blk * size + off, j = blk * size + size - 1 - off;
            if (j < m) {
                const u64 a = keys[i], b = keys[j];
                if (a > b) {
                    keys[i] = b; keys[j] = a;
                    if (HASVAL) { const double v = vals[i]; vals[i] = vals[j]; vals[j] = v; }
                }
            }
        }
        G::sync();
        for (int stride = size >> 2; stride >= 1; stride >>= 1) {
            for (int t = G::tid(); t < half; t += G::size()) {
                const int i = 2 * stride * (t / stride) + (t % stride), j = i + stride;
                if (j < m) {
                    const u64 a = keys[i], b = keys[j];
                    if (a > b) {
                        keys[i] = b; keys[j] = a;
                        if (HASVAL) { const double v = vals[i]; vals[i] = vals[j]; vals[j] = v; }
                    }
                }
            }
            G::sync();
        }
    }
}

// expand + sort + compress one row.  keys/vals: m slots (shared or global); output into
// Tj/Tx (global, at the row's scratch offset).  Returns the number of kept entries.
template <class G>
__device__ __forceinline__ int spgemm_row(int a_lo, int a_hi, const int *__restrict__ Aj, const double *__restrict__ Ax,
                                          const int *__restrict__ Bp, const int *__restrict__ Bj,
                                          const double *__restrict__ Bx, u64 *keys, double *vals, int m,
                                          int *Tj, double *Tx, int *s_scan, bool scipy_order)
{
    const int tid = G::tid(), gs = G::size();
    int base = 0;
    for (int k0 = a_lo; k0 < a_hi; k0 += gs) {
        const int k = k0 + tid;
        int len = 0, j = 0, q0 = 0;
        if (k < a_hi) { j = Aj[k]; q0 = Bp[j]; len = Bp[j + 1] - q0; }
        int tot;
        const int ex = G::exscan(len, &tot, s_scan);
        if (len > 0) {
            const double a = Ax[k];
            for (int t = 0; t < len; ++t) {
                const int p = base + ex + t;
                keys[p] = ((u64)(unsigned)Bj[q0 + t] << 32) | (unsigned)p;
                vals[p] = mul_rn(a, Bx[q0 + t]);
            }
        }
        base += tot;
    }
    G::sync();
    bitonic_sort<G, true>(keys, vals, m);
    // heads fold their run in sequence order; the low word of a head's key keeps its sequence
    // number (= scipy's first touch of that column), non-heads get 0xFFFFFFFF and heads whose
    // sum is exactly 0.0 get 0xFFFFFFFE (dropped).  Only the high word of a neighbour's key
    // is ever read by another thread, and it never changes.
    for (int e = tid; e < m; e += gs) {
        const u64 key = keys[e];
        const unsigned col = (unsigned)(key >> 32);
        unsigned tag = 0xFFFFFFFFu;
        if (e == 0 || (unsigned)(keys[e - 1] >> 32) != col) {
            double s = 0.0;
            int f = e;
            do { s = add_rn(s, vals[f]); ++f; } while (f < m && (unsigned)(keys[f] >> 32) == col);
            vals[e] = s;
            tag = (s != 0.0) ? (unsigned)key : 0xFFFFFFFEu;
        }
        keys[e] = ((u64)col << 32) | tag;
    }
    G::sync();
    // compaction of the kept heads to the front of keys/vals (in place: every element is read
    // into registers before the scan's barrier and written after it, and kept+ex <= e)
    int kept = 0;
    for (int e0 = 0; e0 < m; e0 += gs) {
        const int e = e0 + tid;
        int flag = 0;
        u64 key = 0;
        double v = 0.0;
        if (e < m) {
            key = keys[e];
            flag = ((unsigned)key < 0xFFFFFFFEu);
            v = vals[e];
        }
        int tot;
        const int ex = G::exscan(flag, &tot, s_scan);
        G::sync();
        if (flag) {
            const unsigned col = (unsigned)(key >> 32), seq = (unsigned)key;
            // scipy order = reverse first touch: ascending (0xFFFFFFFF - seq)
            keys[kept + ex] = scipy_order ? (((u64)(0xFFFFFFFFu - seq) << 32) | col) : (u64)col;
            vals[kept + ex] = v;
        }
        kept += tot;
    }
    G::sync();
    if (scipy_order) bitonic_sort<G, true>(keys, vals, kept);
    for (int t = tid; t < kept; t += gs) { Tj[t] = (int)(unsigned)keys[t]; Tx[t] = vals[t]; }
    return kept;
}

// rows with <= CAP products: one warp per row, WPB rows per CTA
template <int CAP, int WPB>
__global__ void __launch_bounds__(32 * WPB) spgemm_warp_kernel(const int *__restrict__ rows, int nrows,
                                                               const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                               const double *__restrict__ Ax, const int *__restrict__ Bp,
                                                               const int *__restrict__ Bj, const double *__restrict__ Bx,
                                                               const long long *__restrict__ ubp, int *Tj, double *Tx,
                                                               int *cnt, bool so)
{
    __shared__ u64 s_k[WPB][CAP];
    __shared__ double s_v[WPB][CAP];
    const int w = threadIdx.x >> 5;
    const int r = blockIdx.x * WPB + w;
    if (r >= nrows) return;
    const int i = rows[r];
    const long long o = ubp[i];
    const int m = (int)(ubp[i + 1] - o);
    const int kept = spgemm_row<WarpGroup>(Ap[i], Ap[i + 1], Aj, Ax, Bp, Bj, Bx, s_k[w], s_v[w], m, Tj + o, Tx + o, nullptr, so);
    if ((threadIdx.x & 31) == 0) cnt[i] = kept;
}

// rows with <= CAP products: one CTA per row, (key,val) in dynamic shared memory
template <int CAP>
__global__ void spgemm_cta_kernel(const int *__restrict__ rows, int nrows, const int *__restrict__ Ap,
                                  const int *__restrict__ Aj, const double *__restrict__ Ax,
                                  const int *__restrict__ Bp, const int *__restrict__ Bj, const double *__restrict__ Bx,
                                  const long long *__restrict__ ubp, int *Tj, double *Tx, int *cnt, bool so)
{
    extern __shared__ __align__(16) unsigned char s_raw[];
    __shared__ int s_scan[33];
    u64 *s_k = reinterpret_cast<u64 *>(s_raw);
    double *s_v = reinterpret_cast<double *>(s_raw + sizeof(u64) * CAP);
    const int i = rows[blockIdx.x];
    const long long o = ubp[i];
    const int m = (int)(ubp[i + 1] - o);
    const int kept = spgemm_row<CtaGroup>(Ap[i], Ap[i + 1], Aj, Ax, Bp, Bj, Bx, s_k, s_v, m, Tj + o, Tx + o, s_scan, so);
    if (threadIdx.x == 0) cnt[i] = kept;
}

// rows with more products than fit in shared memory: sort in the global scratch itself
__global__ void __launch_bounds__(1024) spgemm_big_kernel(const int *__restrict__ rows, int nrows,
                                                          const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                          const double *__restrict__ Ax, const int *__restrict__ Bp,
                                                          const int *__restrict__ Bj, const double *__restrict__ Bx,
                                                          const long long *__restrict__ ubp, u64 *Tk, int *Tj, double *Tx,
                                                          int *cnt, bool so)
{
    __shared__ int s_scan[33];
    const int i = rows[blockIdx.x];
    const long long o = ubp[i];
    const int m = (int)(ubp[i + 1] - o);
    const int kept = spgemm_row<CtaGroup>(Ap[i], Ap[i + 1], Aj, Ax, Bp, Bj, Bx, Tk + o, Tx + o, m, Tj + o, Tx + o, s_scan, so);
    if (threadIdx.x == 0) cnt[i] = kept;
}

__global__ void __launch_bounds__(256) spgemm_count_kernel(int n_row, const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                           const int *__restrict__ Bp, long long *__restrict__ ub)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n_row) return;
    long long s = 0;
    const int hi = Ap[i + 1];
    for (int k = Ap[i]; k < hi; ++k) { const int j = Aj[k]; s += Bp[j + 1] - Bp[j]; }
    ub[i] = s;
}

// one warp per row: copy the kept entries to their final place
__global__ void __launch_bounds__(256) spgemm_compact_kernel(int n_row, const long long *__restrict__ ubp,
                                                             const int *__restrict__ Tj, const double *__restrict__ Tx,
                                                             const int *__restrict__ Cp, int *__restrict__ Cj,
                                                             double *__restrict__ Cx)
{
    const int i = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= n_row) return;
    const long long src = ubp[i];
    const int dst = Cp[i], n = Cp[i + 1] - dst;
    for (int t = lane; t < n; t += 32) { Cj[dst + t] = Tj[src + t]; Cx[dst + t] = Tx[src + t]; }
}

}  // namespace b200amg

using namespace b200amg;
static inline cudaStream_t S(void *s) { return reinterpret_cast<cudaStream_t>(s); }

extern "C" int b200amg_spgemm_count(int n_row, const int *Ap, const int *Aj, const int *Bp, long long *ub, void *stream)
{
    if (n_row < 0) return set_err(B200AMG_EINVAL, "spgemm_count: n_row < 0");
    if (n_row == 0) return 0;
    spgemm_count_kernel<<<(n_row + 255) / 256, 256, 0, S(stream)>>>(n_row, Ap, Aj, Bp, ub);
    return check_launch("spgemm_count_kernel");
}

extern "C" int b200amg_exclusive_scan_i64(long long n, const long long *in, long long *out, long long *scratch, void *stream)
{
    return exclusive_scan<long long, long long>(n, in, out, scratch, S(stream));
}
extern "C" int b200amg_exclusive_scan_i32(long long n, const int *in, int *out, int *scratch, void *stream)
{
    return exclusive_scan<int, int>(n, in, out, scratch, S(stream));
}
extern "C" long long b200amg_scan_scratch_elems(long long n) { return scan_scratch_elems(n); }

extern "C" int b200amg_spgemm_numeric(int n_row, const int *Ap, const int *Aj, const double *Ax, const int *Bp,
                                      const int *Bj, const double *Bx, const long long *ub, const long long *ubp,
                                      unsigned long long *Tk, int *Tj, double *Tx, int *cnt, int *work, int scipy_order,
                                      void *stream)
{
    if (n_row < 0) return set_err(B200AMG_EINVAL, "spgemm_numeric: n_row < 0");
    if (n_row == 0) return 0;
    cudaStream_t st = S(stream);
    B200_CUDA(cudaMemsetAsync(cnt, 0, sizeof(int) * n_row, st));
    int *perm = work + 64;
    const bool so = scipy_order != 0;
    int off[NBINS + 1];
    const BinLimits lim{{32, 128, 512, 2048, 8192}};
    B200_TRY(bin_rows(n_row, ub, lim, perm, work, off, st));
#define ARGS(b) perm + off[b], off[b + 1] - off[b], Ap, Aj, Ax, Bp, Bj, Bx, ubp
    int n;
    // rows of up to 512 (2048, 8192) products in scipy's order can go without a sort (spgemm_hash.cu, B200AMG_SPGEMM_HASH)
    const int hash_mode = so ? spgemm_hash_mode() : 0;
    bool hashed[5] = {false, false, false, false, false};
    for (int b = 0; b < 5 && hash_mode > 0; ++b) {
        if ((n = off[b + 1] - off[b]) <= 0) continue;
        const int took = spgemm_hash_bin(b, hash_mode, ARGS(b), Tj, Tx, cnt, st);
        if (took < 0) return took;
        hashed[b] = took > 0;
    }
    if (!hashed[0] && (n = off[1] - off[0]) > 0) {
        spgemm_warp_kernel<32, 8><<<(n + 7) / 8, 256, 0, st>>>(ARGS(0), Tj, Tx, cnt, so);
        B200_TRY(check_launch("spgemm_warp_kernel<32>"));
    }
    if (!hashed[1] && (n = off[2] - off[1]) > 0) {
        spgemm_warp_kernel<128, 8><<<(n + 7) / 8, 256, 0, st>>>(ARGS(1), Tj, Tx, cnt, so);
        B200_TRY(check_launch("spgemm_warp_kernel<128>"));
    }
    if (!hashed[2] && (n = off[3] - off[2]) > 0) {
        spgemm_cta_kernel<512><<<n, 128, 512 * 16, st>>>(ARGS(2), Tj, Tx, cnt, so);
        B200_TRY(check_launch("spgemm_cta_kernel<512>"));
    }
    if (!hashed[3] && (n = off[4] - off[3]) > 0) {
        spgemm_cta_kernel<2048><<<n, 256, 2048 * 16, st>>>(ARGS(3), Tj, Tx, cnt, so);
        B200_TRY(check_launch("spgemm_cta_kernel<2048>"));
    }
    if (!hashed[4] && (n = off[5] - off[4]) > 0) {
        static bool attr_set = false;
        if (!attr_set) {
            B200_CUDA(cudaFuncSetAttribute(spgemm_cta_kernel<8192>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192 * 16));
            attr_set = true;
        }
        spgemm_cta_kernel<8192><<<n, 512, 8192 * 16, st>>>(ARGS(4), Tj, Tx, cnt, so);
        B200_TRY(check_launch("spgemm_cta_kernel<8192>"));
    }
    if ((n = off[6] - off[5]) > 0) {
        if (!Tk) return set_err(B200AMG_ENOMEM, "spgemm_numeric: rows with > 8192 products need the Tk scratch");
        spgemm_big_kernel<<<n, 1024, 0, st>>>(ARGS(5), Tk, Tj, Tx, cnt, so);
        B200_TRY(check_launch("spgemm_big_kernel"));
    }
#undef ARGS
    return 0;
}

extern "C" int b200amg_spgemm_compact(int n_row, const long long *ubp, const int *Tj, const double *Tx, const int *Cp,
                                      int *Cj, double *Cx, void *stream)
{
    if (n_row <= 0) return 0;
    spgemm_compact_kernel<<<(n_row + 7) / 8, 256, 0, S(stream)>>>(n_row, ubp, Tj, Tx, Cp, Cj, Cx);
    return check_launch("spgemm_compact_kernel");
}
