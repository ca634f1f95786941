// Row-pattern layout for constant-coefficient stencil operators.
//
// On a structured grid every row of a stencil matrix is one of a handful of (column offset, value) sequences:
// the interior stencil plus one variant per boundary face / edge / corner (27 for the 7-point 3D Poisson
// matrix, 9 for a 2D 5- or 9-point stencil).  When an operator has <= 255 distinct row patterns of <= 8
// entries, the engine keeps ONE BYTE PER ROW -- the pattern number -- and a table of the patterns
// (<= 32 KB, in shared memory).  A sweep then streams 1 byte per row of matrix data instead of
// 12 bytes per entry + 4 per row: what is left is the vector traffic (b, x, the result).
//
// The table stores each pattern's entries in the row's STORAGE order with the values' exact bits, so the
// fold below is the same ordered, separately rounded sum as scipy's csr_matvec and the amg_core
// jacobi loops (pyamg/amg_core/relaxation.h:201-239, :267-360): results are bit-identical to the CSR kernels.
// Detection is exact: rows are grouped by a 64-bit hash first, then EVERY row is compared entry by entry with
// its pattern's table entry; any mismatch (hash collision) declines the layout.
#include <string.h>
#include <algorithm>
#include <vector>
#include <cstdlib>
#include "internal.cuh"

namespace b200amg {

constexpr int RP_MAX_PAT = 255;
constexpr int RP_LEN = 8;
constexpr int RP_REPORT = 512;

__device__ __forceinline__ unsigned long long rp_mix(unsigned long long h)
{
    h ^= h >> 30; h *= 0xBF58476D1CE4E5B9ull;
    h ^= h >> 27; h *= 0x94D049BB133111EBull;
    h ^= h >> 31;
    return h;
}

__device__ __forceinline__ unsigned long long rp_row_hash(const int *__restrict__ Aj, const double *__restrict__ Ax,
                                                          int row, int s, int e)
{
    unsigned long long h = rp_mix(0x51ED270B1ull + (unsigned long long)(e - s));
    for (int k = s; k < e; ++k) {
        h = rp_mix(h ^ (unsigned long long)(unsigned)(Aj[k] - row));
        h = rp_mix(h ^ (unsigned long long)__double_as_longlong(Ax[k]));
    }
    return h == ~0ull ? 0ull : h;                       // ~0 is the "empty slot" mark of the report
}

__device__ __forceinline__ int rp_find(const unsigned long long *s_h, int n, unsigned long long h)
{
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (s_h[mid] < h) lo = mid + 1; else hi = mid;
    }
    return (lo < n && s_h[lo] == h) ? lo : -1;
}

// flags[0] = rows whose hash is not in `known`, flags[1] = some row is longer than RP_LEN,
// report[512] = (hash, row) of unknown rows, slot chosen by the hash
__global__ void __launch_bounds__(256) rowpat_scan_kernel(int n_row, const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                          const double *__restrict__ Ax,
                                                          const unsigned long long *__restrict__ known, int nknown,
                                                          int *flags, ulonglong2 *report)
{
    __shared__ unsigned long long s_h[RP_MAX_PAT + 1];
    for (int i = threadIdx.x; i <= RP_MAX_PAT; i += 256) s_h[i] = i < nknown ? known[i] : ~0ull;
    __syncthreads();
    const int stride = gridDim.x * 256;
    for (int row = blockIdx.x * 256 + threadIdx.x; row < n_row; row += stride) {
        const int s = Ap[row], e = Ap[row + 1];
        if (e - s > RP_LEN) { flags[1] = 1; continue; }
        const unsigned long long h = rp_row_hash(Aj, Ax, row, s, e);
        if (rp_find(s_h, nknown, h) < 0) {
            atomicAdd(flags, 1);
            report[h >> 55] = make_ulonglong2(h, (unsigned long long)row);
        }
    }
}

// table[p * 8 + q] = entry q of the representative row of pattern p (len | diagonal mask << 8 in every entry)
__global__ void rowpat_fill_kernel(int npat, const int *__restrict__ rep_rows, const int *__restrict__ Ap,
                                   const int *__restrict__ Aj, const double *__restrict__ Ax, RowPatEntry *table)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npat) return;
    const int row = rep_rows[p], s = Ap[row], len = Ap[row + 1] - s;
    int dm = 0;                                          // bit q set: entry q sits on the diagonal
    for (int q = 0; q < len; ++q) dm |= (Aj[s + q] == row) << q;
    for (int q = 0; q < RP_LEN; ++q) {
        RowPatEntry en;
        en.v = q < len ? Ax[s + q] : 0.0;
        en.off = q < len ? Aj[s + q] - row : 0;
        en.len = len | (dm << 8);
        table[p * RP_LEN + q] = en;
    }
}

// codes[row] = pattern number; flags[2] is raised when a row does not match its pattern entry for entry
__global__ void __launch_bounds__(256) rowpat_encode_kernel(int n_row, const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                            const double *__restrict__ Ax,
                                                            const unsigned long long *__restrict__ known, int nknown,
                                                            const RowPatEntry *__restrict__ table,
                                                            unsigned char *__restrict__ codes, int *flags)
{
    __shared__ unsigned long long s_h[RP_MAX_PAT + 1];
    for (int i = threadIdx.x; i <= RP_MAX_PAT; i += 256) s_h[i] = i < nknown ? known[i] : ~0ull;
    __syncthreads();
    const int stride = gridDim.x * 256;
    for (int row = blockIdx.x * 256 + threadIdx.x; row < n_row; row += stride) {
        const int s = Ap[row], e = Ap[row + 1];
        bool good = (e - s) <= RP_LEN;
        int p = -1;
        if (good) {
            p = rp_find(s_h, nknown, rp_row_hash(Aj, Ax, row, s, e));
            good = p >= 0;
        }
        if (good) {
            good = (table[p * RP_LEN].len & 0xff) == e - s;
            for (int k = s; good && k < e; ++k) {
                const RowPatEntry en = table[p * RP_LEN + (k - s)];
                good = en.off == Aj[k] - row && __double_as_longlong(en.v) == __double_as_longlong(Ax[k]);
            }
        }
        if (good) codes[row] = (unsigned char)p;
        else flags[2] = 1;
        int mc = -1;                                        // flags[3] = largest column referenced (x has more entries
        for (int k = s; k < e; ++k) mc = max(mc, Aj[k]);    // than the operator has rows on a row slab with halo columns)
        if (mc > flags[3]) atomicMax(&flags[3], mc);
    }
}

void rowpat_tile_plan(RowPat *r, const std::vector<RowPatEntry> &ht, int common);   // rowpat_tile.cu

// hist[p] = rows with pattern p (the most frequent pattern is the operator's interior stencil)
__global__ void __launch_bounds__(256) rowpat_hist_kernel(int n, const unsigned char *__restrict__ codes, unsigned *hist)
{
    __shared__ unsigned s_h[256];
    s_h[threadIdx.x] = 0;
    __syncthreads();
    for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) atomicAdd(&s_h[codes[i]], 1u);
    __syncthreads();
    if (s_h[threadIdx.x]) atomicAdd(&hist[threadIdx.x], s_h[threadIdx.x]);
}

void rowpat_free(RowPat *r)
{
    if (r->codes) cudaFree(r->codes);
    if (r->table) cudaFree(r->table);
    if (r->tmask) cudaFree(r->tmask);
    *r = RowPat();
}

int rowpat_build(int n_row, long long nnz, const int *Ap, const int *Aj, const double *Ax, RowPat *out, cudaStream_t st)
{
    *out = RowPat();
    if (nnz < (1 << 16) || n_row < 4096) return 0;
    typedef std::pair<unsigned long long, int> Key;                 // (row hash, representative row)
    std::vector<Key> keys;
    unsigned long long *d_known = nullptr;
    ulonglong2 *d_report = nullptr;
    int *d_flags = nullptr, *d_rep = nullptr;
    unsigned char *codes = nullptr;
    RowPatEntry *table = nullptr;
    cudaError_t e = cudaMalloc(&d_known, sizeof(unsigned long long) * (RP_MAX_PAT + 1));
    if (e == cudaSuccess) e = cudaMalloc(&d_report, sizeof(ulonglong2) * RP_REPORT);
    if (e == cudaSuccess) e = cudaMalloc(&d_flags, sizeof(int) * 4);
    if (e == cudaSuccess) e = cudaMalloc(&d_rep, sizeof(int) * (RP_MAX_PAT + 1));
    int rc = 0;
    bool found = false;
    const int grid = std::min((n_row + 255) / 256, 148 * 16);
    std::vector<unsigned long long> hk(RP_MAX_PAT + 1);
    std::vector<int> hr(RP_MAX_PAT + 1);
    auto upload_keys = [&]() {
        for (size_t i = 0; i < keys.size(); ++i) { hk[i] = keys[i].first; hr[i] = keys[i].second; }
        cudaError_t ee = cudaMemcpyAsync(d_known, hk.data(), sizeof(unsigned long long) * (RP_MAX_PAT + 1), cudaMemcpyHostToDevice, st);
        if (ee == cudaSuccess) ee = cudaMemcpyAsync(d_rep, hr.data(), sizeof(int) * (RP_MAX_PAT + 1), cudaMemcpyHostToDevice, st);
        return ee;
    };
    for (int round = 0; e == cudaSuccess && !rc && !found && round < 10; ++round) {
        e = upload_keys();
        if (e == cudaSuccess) e = cudaMemsetAsync(d_flags, 0, sizeof(int) * 4, st);
        if (e == cudaSuccess) e = cudaMemsetAsync(d_report, 0xFF, sizeof(ulonglong2) * RP_REPORT, st);
        if (e != cudaSuccess) break;
        rowpat_scan_kernel<<<grid, 256, 0, st>>>(n_row, Ap, Aj, Ax, d_known, (int)keys.size(), d_flags, d_report);
        rc = check_launch("rowpat_scan_kernel");
        int flags[4] = {0, 0, 0, 0};
        if (!rc) e = cudaMemcpyAsync(flags, d_flags, sizeof(flags), cudaMemcpyDeviceToHost, st);
        if (!rc && e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (rc || e != cudaSuccess || flags[1]) break;             // a row longer than 8 entries: not this layout
        if (flags[0] == 0) { found = true; break; }
        std::vector<ulonglong2> rep(RP_REPORT);
        e = cudaMemcpy(rep.data(), d_report, sizeof(ulonglong2) * RP_REPORT, cudaMemcpyDeviceToHost);
        for (const ulonglong2 &m : rep)
            if (!(m.x == ~0ull && m.y == ~0ull)) keys.push_back(Key(m.x, (int)m.y));
        std::sort(keys.begin(), keys.end());
        keys.erase(std::unique(keys.begin(), keys.end(), [](const Key &a, const Key &b) { return a.first == b.first; }), keys.end());
        if ((int)keys.size() > RP_MAX_PAT) break;                  // a general matrix fills the report at once
    }
    bool good = false;
    if (found && e == cudaSuccess && !rc && !keys.empty()) {
        const int npat = (int)keys.size();
        e = cudaMalloc(&codes, (size_t)n_row + 64);
        if (e == cudaSuccess) e = cudaMalloc(&table, sizeof(RowPatEntry) * RP_LEN * npat);
        if (e == cudaSuccess) e = upload_keys();
        if (e == cudaSuccess) e = cudaMemsetAsync(d_flags, 0, sizeof(int) * 4, st);
        if (e == cudaSuccess) {
            rowpat_fill_kernel<<<(npat + 63) / 64, 64, 0, st>>>(npat, d_rep, Ap, Aj, Ax, table);
            rc = check_launch("rowpat_fill_kernel");
        }
        if (e == cudaSuccess && !rc) {
            rowpat_encode_kernel<<<grid, 256, 0, st>>>(n_row, Ap, Aj, Ax, d_known, npat, table, codes, d_flags);
            rc = check_launch("rowpat_encode_kernel");
        }
        int flags[4] = {0, 0, 1, 0};
        if (e == cudaSuccess && !rc) e = cudaMemcpyAsync(flags, d_flags, sizeof(flags), cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess && !rc) e = cudaStreamSynchronize(st);
        good = e == cudaSuccess && !rc && flags[2] == 0;
        if (good) {
            std::vector<RowPatEntry> ht((size_t)RP_LEN * npat);
            e = cudaMemcpy(ht.data(), table, sizeof(RowPatEntry) * ht.size(), cudaMemcpyDeviceToHost);
            good = e == cudaSuccess;
            // prefetch distance = farthest neighbour of a TYPICAL row (the row in the middle of the matrix), not of
            // any row: on a row slab of a partitioned operator the boundary planes reach into the halo region
            // (offset ~ n), and prefetching "row + n" from every row hammers one clamped address
            unsigned char mid = 0;
            if (good) good = cudaMemcpy(&mid, codes + n_row / 2, 1, cudaMemcpyDeviceToHost) == cudaSuccess;
            if (good && mid < npat)
            {
                for (int q = 0; q < RP_LEN; ++q) out->far = std::max(out->far, ht[(size_t)mid * RP_LEN + q].off);
                out->common_len = ht[(size_t)mid * RP_LEN].len & 0xff;     // row length of that typical row
            }
            if (good) {
                out->codes = codes;
                out->table = table;
                out->npat = npat;
                out->n = n_row;
                out->xlen = std::max(n_row, flags[3] + 1);
                // segments / masks of the TMA-staged sweep, made from the MOST FREQUENT pattern (the middle row of a
                // structured grid sits on two faces)
                unsigned *d_hist = nullptr, h_hist[256];
                if (cudaMalloc(&d_hist, sizeof(h_hist)) == cudaSuccess) {
                    cudaMemsetAsync(d_hist, 0, sizeof(h_hist), st);
                    rowpat_hist_kernel<<<grid, 256, 0, st>>>(n_row, codes, d_hist);
                    ++g_launches;
                    if (cudaMemcpyAsync(h_hist, d_hist, sizeof(h_hist), cudaMemcpyDeviceToHost, st) == cudaSuccess &&
                        cudaStreamSynchronize(st) == cudaSuccess) {
                        int common = 0;
                        for (int p2 = 1; p2 < npat; ++p2)
                            if (h_hist[p2] > h_hist[common]) common = p2;
                        rowpat_tile_plan(out, ht, common);
                    }
                    cudaGetLastError();
                    cudaFree(d_hist);
                }
            }
        }
    }
    if (d_known) cudaFree(d_known);
    if (d_report) cudaFree(d_report);
    if (d_flags) cudaFree(d_flags);
    if (d_rep) cudaFree(d_rep);
    if (!good) {
        if (codes) cudaFree(codes);
        if (table) cudaFree(table);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return set_err((int)e, "rowpat_build: CUDA error");
        }
        return rc;                                                  // declined: the caller keeps another layout
    }
    return 0;
}

// ---- apply -------------------------------------------------------------------------------------------
// thread = row.  A CTA walks chunks of RP_CHUNK consecutive rows (round-robin over the grid): inside a chunk
// the +-1 and +-nx neighbours are re-read from L1, and because all CTAs advance through the matrix together
// the +-nx*ny neighbours (a few MB away) are still in L2 -- x crosses HBM once.
#ifndef RP_CHUNK_ROWS
#define RP_CHUNK_ROWS 1024
#endif
#ifndef RP_AHEAD
#define RP_AHEAD 1
#endif
#ifndef RP_RESIDENT
#define RP_RESIDENT 4
#endif
constexpr int RP_TPB = 256;
constexpr int RP_CHUNK = RP_CHUNK_ROWS;
constexpr int RP_CTAS_PER_SM = RP_RESIDENT;

struct RowPatArgs {
    int n, npat, far;                // far = largest column offset of any pattern (prefetch distance in x)
    const unsigned char *codes;
    const RowPatEntry *table;
    const double *x, *b;
    double *y;
    double alpha, beta, omega;
};

// The table sits in shared memory as structure of arrays -- 8 column offsets (32 B), 8 values (64 B) and one
// meta word per pattern -- so a row's offsets come with two 128-bit loads and its values with four, instead of
// two loads per entry; gather addresses are formed from the 32-bit column number (one add + one wide
// multiply-add per entry).  Padding slots hold offset 0 / value 0.
struct RowPatSmem {
    const int *off;        // [npat * 8]
    const double *val;     // [npat * 8]
    const int *meta;       // [npat]: len | diagonal mask << 8
};

template <int LEN>
__device__ __forceinline__ void rp_load_offsets(const int *po, int (&off)[RP_LEN])
{
    const int4 a = *reinterpret_cast<const int4 *>(po);
    off[0] = a.x; off[1] = a.y; off[2] = a.z; off[3] = a.w;
    if (LEN > 4) {
        const int4 c = *reinterpret_cast<const int4 *>(po + 4);
        off[4] = c.x; off[5] = c.y; off[6] = c.z; off[7] = c.w;
    }
}

template <int LEN>
__device__ __forceinline__ void rp_load_values(const double *pv, double (&vv)[RP_LEN])
{
#pragma unroll
    for (int q = 0; q < LEN; q += 2) {
        const double2 t = *reinterpret_cast<const double2 *>(pv + q);
        vv[q] = t.x;
        vv[q + 1] = t.y;
    }
}

// ordered fold of a row of exactly LEN entries: all gathers in flight first, then the storage-order sum
template <int OP, int LEN>
__device__ __forceinline__ double rp_fold_fixed(const int *po, const double *pv, const double *__restrict__ xg, int row,
                                                unsigned dm, double acc)
{
    constexpr bool SMOOTH = (OP >= OP_JACOBI);
    constexpr bool BSRF = (OP == OP_JACOBI_BSR1);
    int off[RP_LEN];
    double vv[RP_LEN], xv[RP_LEN];
    rp_load_offsets<LEN>(po, off);
#pragma unroll
    for (int q = 0; q < LEN; ++q) xv[q] = __ldg(xg + (row + off[q]));
    rp_load_values<LEN>(pv, vv);
#pragma unroll
    for (int q = 0; q < LEN; ++q) {
        // a diagonal slot contributes +0.0 instead of being skipped: acc + 0.0 == acc bit for bit (acc is never
        // -0.0 here: it starts at +0.0 / b_i and "-0.0 - (+0.0)" stays -0.0), and the select sits off the
        // dependent chain of the sum
        double pr = mul_rn(vv[q], xv[q]);
        if (SMOOTH && ((dm >> q) & 1u)) pr = 0.0;
        acc = BSRF ? sub_rn(acc, pr) : add_rn(acc, pr);
    }
    return acc;
}

// rows of different lengths in one warp (grid boundaries): every slot predicated on q < len
template <int OP>
__device__ __forceinline__ double rp_fold_mixed(const int *po, const double *pv, const double *__restrict__ xg, int row,
                                                unsigned dm, int len, double acc)
{
    constexpr bool SMOOTH = (OP >= OP_JACOBI);
    constexpr bool BSRF = (OP == OP_JACOBI_BSR1);
    int off[RP_LEN];
    double vv[RP_LEN], xv[RP_LEN];
    rp_load_offsets<RP_LEN>(po, off);
#pragma unroll
    for (int q = 0; q < RP_LEN; ++q)
        if (q < len) xv[q] = __ldg(xg + (row + off[q]));
    rp_load_values<RP_LEN>(pv, vv);
#pragma unroll
    for (int q = 0; q < RP_LEN; ++q)
        if (q < len) {
            const double pr = mul_rn(vv[q], xv[q]);
            if (!SMOOTH || !((dm >> q) & 1u)) acc = BSRF ? sub_rn(acc, pr) : add_rn(acc, pr);
        }
    return acc;
}

template <int OP>
__global__ void __launch_bounds__(RP_TPB, RP_CTAS_PER_SM) rowpat_kernel(const RowPatArgs a)
{
    constexpr bool SMOOTH = (OP >= OP_JACOBI);
    constexpr bool BSRF = (OP == OP_JACOBI_BSR1);
    constexpr bool NEED_B = (OP == OP_RESIDUAL || SMOOTH);
    constexpr bool NEED_Y = (OP == OP_ADD);
    extern __shared__ __align__(16) unsigned char rp_smem[];
    int *s_off = reinterpret_cast<int *>(rp_smem);
    double *s_val = reinterpret_cast<double *>(rp_smem + (size_t)a.npat * RP_LEN * sizeof(int));
    int *s_meta = reinterpret_cast<int *>(rp_smem + (size_t)a.npat * RP_LEN * (sizeof(int) + sizeof(double)));
    pdl_trigger();
    for (int i = threadIdx.x; i < a.npat * RP_LEN; i += RP_TPB) {     // (engine-owned table: complete before any launch)
        const RowPatEntry en = a.table[i];
        s_off[i] = en.off;
        s_val[i] = en.v;
        if ((i & (RP_LEN - 1)) == 0) s_meta[i / RP_LEN] = en.len;
    }
    __syncthreads();
    pdl_wait();
    const double *__restrict__ xg = a.x;
    const int n_chunks = (a.n + RP_CHUNK - 1) / RP_CHUNK;
    // A CTA's steps: step k covers rows [r0, r0 + 256) of chunk blockIdx.x + (k / SPC) * gridDim.x.
    constexpr int SPC = RP_CHUNK / RP_TPB;                       // steps per chunk
    auto step_row = [&](int k) -> int {                          // this thread's row at step k, -1 past the end
        const int ch = (int)blockIdx.x + (k / SPC) * (int)gridDim.x;
        const int row = ch * RP_CHUNK + (k % SPC) * RP_TPB + (int)threadIdx.x;
        return (ch < n_chunks && row < a.n) ? row : -1;
    };
    // Latency hiding.  (1) The pattern number of a row is needed before its gathers can start: it is loaded
    // one step ahead.  (2) What a step pulls from DRAM -- its slice of b and the x entries of its farthest
    // neighbours (first touch of x; every other gather finds x in L1/L2) -- is requested RP_AHEAD steps ahead
    // with prefetch.global.L2, one lane per 128-byte line: the 32 resident warps of an SM cannot keep enough
    // demand loads in flight to cover HBM latency, the prefetches cost no registers.
    const int far = a.far;
    const int my_steps = ((n_chunks - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x) * SPC;
    int code_next = step_row(0) >= 0 ? (int)a.codes[step_row(0)] : -1;
#pragma unroll 1
    for (int k = 0; k < my_steps; ++k) {
        {
            const int row = step_row(k);
            const int code = code_next;
            const bool live = row >= 0;
            {
                const int rn = step_row(k + 1);
                code_next = rn >= 0 ? (int)a.codes[rn] : -1;
            }
#if RP_AHEAD > 0
            if ((threadIdx.x & 15) == 0) {
                const int rp = step_row(k + RP_AHEAD);
                if (rp >= 0) {
                    if (NEED_B) asm volatile("prefetch.global.L2 [%0];" ::"l"(a.b + rp));
                    if (NEED_Y) asm volatile("prefetch.global.L2 [%0];" ::"l"(a.b + rp));
                    // unconditional (clamped) form: a predicate here costs 1 % of the sweep (A/B on one box)
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(xg + min(rp + far, a.n - 1)));
                }
            }
#endif
            const int *po = s_off;
            const double *pv = s_val;
            double bi = 0.0, xi = 0.0, yi = 0.0;
            int len = 0;
            unsigned dm = 0;
            if (live) {
                po = s_off + code * RP_LEN;
                pv = s_val + code * RP_LEN;
                if (NEED_B) bi = __ldg(a.b + row);
                if (SMOOTH) xi = __ldg(xg + row);
                if (NEED_Y) yi = a.b[row];
                const int meta = s_meta[code];
                len = meta & 0xff;
                dm = (unsigned)meta >> 8;
            }
            double acc = BSRF ? bi : 0.0;
            // interior warps: every lane has the same row length -> straight-line code without predicates
            // (a table copy in the constant bank or in registers for warps with ONE pattern was tried: both slower)
            const int l0 = __shfl_sync(0xffffffffu, len, 0);
            if (__all_sync(0xffffffffu, len == l0)) {
                switch (l0) {
                case 1: acc = rp_fold_fixed<OP, 1>(po, pv, xg, row, dm, acc); break;
                case 2: acc = rp_fold_fixed<OP, 2>(po, pv, xg, row, dm, acc); break;
                case 3: acc = rp_fold_fixed<OP, 3>(po, pv, xg, row, dm, acc); break;
                case 4: acc = rp_fold_fixed<OP, 4>(po, pv, xg, row, dm, acc); break;
                case 5: acc = rp_fold_fixed<OP, 5>(po, pv, xg, row, dm, acc); break;
                case 6: acc = rp_fold_fixed<OP, 6>(po, pv, xg, row, dm, acc); break;
                case 7: acc = rp_fold_fixed<OP, 7>(po, pv, xg, row, dm, acc); break;
                case 8: acc = rp_fold_fixed<OP, 8>(po, pv, xg, row, dm, acc); break;
                default: break;
                }
            } else {
                acc = rp_fold_mixed<OP>(po, pv, xg, row, dm, len, acc);
            }
            if (!live) continue;
            double diag = 0.0;
            if (SMOOTH && dm) diag = pv[31 - __clz(dm)];            // last stored diagonal wins
            double *y = a.y;
            if (OP == OP_MATVEC) {
                y[row] = acc;
            } else if (OP == OP_SPMV_AB) {
                y[row] = (a.beta == 0.0) ? a.alpha * acc : a.alpha * acc + a.beta * y[row];
            } else if (OP == OP_RESIDUAL) {
                y[row] = sub_rn(bi, acc);
            } else if (OP == OP_ADD) {
                y[row] = add_rn(yi, acc);
            } else if (OP == OP_JACOBI) {
                y[row] = (diag != 0.0)
                             ? add_rn(mul_rn(sub_rn(1.0, a.omega), xi), mul_rn(a.omega, div_rn(sub_rn(bi, acc), diag)))
                             : xi;
            } else if (OP == OP_JACOBI_BSR1) {
                y[row] = (diag != 0.0) ? add_rn(mul_rn(sub_rn(1.0, a.omega), xi), div_rn(mul_rn(a.omega, acc), diag))
                                       : xi;
            }
        }
    }
}

template <int OP>
static int launch_rowpat(const RowPat &r, const StreamArgs &s, cudaStream_t st)
{
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        B200_CUDA(cudaGetDevice(&dev));
        B200_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    }
    const int n_chunks = (r.n + RP_CHUNK - 1) / RP_CHUNK;
    const int grid = std::min(n_chunks, sms * RP_CTAS_PER_SM);
    RowPatArgs a{r.n, r.npat, r.far, r.codes, r.table, s.x, s.b, s.y, s.alpha, s.beta, s.omega};
    B200_LAUNCH_PDL(pdl_class(PDL_OTHER) && (long long)r.n < PDL_SMALL_ROWS, rowpat_kernel<OP>, grid, RP_TPB, ((sizeof(int) + sizeof(double)) * RP_LEN + sizeof(int)) * r.npat + 16, st, a);
    return check_launch("rowpat_kernel");
}

int rowpat_op(int op, const RowPat &r, const StreamArgs &a0, cudaStream_t st)
{
    StreamArgs a = a0;
    if (op == OP_ADD && !a.b) a.b = a.y;              // y += A x: the kernels read the addend through b
    if (a.row_begin != 0 || a.row_end != r.n) return set_err(B200AMG_EINVAL, "rowpat_op: full row range only");
    // the TMA-staged sweep takes the launch when the operator and the vectors qualify (B200AMG_ROWPAT_TILE=0: A/B switch)
    static const bool tile = [] { const char *e = getenv("B200AMG_ROWPAT_TILE"); return !(e && e[0] == '0'); }();
    if (tile) {
        const int rc = rowpat_tile_op(op, r, a, st);
        if (rc >= 0) return rc;
    }
    switch (op) {
    case OP_MATVEC: return launch_rowpat<OP_MATVEC>(r, a, st);
    case OP_SPMV_AB: return launch_rowpat<OP_SPMV_AB>(r, a, st);
    case OP_RESIDUAL: return launch_rowpat<OP_RESIDUAL>(r, a, st);
    case OP_ADD: return launch_rowpat<OP_ADD>(r, a, st);
    case OP_JACOBI: return launch_rowpat<OP_JACOBI>(r, a, st);
    case OP_JACOBI_BSR1: return launch_rowpat<OP_JACOBI_BSR1>(r, a, st);
    }
    return set_err(B200AMG_EINVAL, "rowpat_op: bad op");
}

}  // namespace b200amg
