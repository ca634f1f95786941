// Sliced-ELL (SELL-32) path for CSR operators with 12..400 stored entries per row: the coarse
// Galerkin operators (31-68 per row) and the level-0 restriction (30 per row).
//
// Why a second layout: for these rows the x gather is what limits a CSR kernel, not HBM.  With
// entries dealt to lanes in CSR order (csr_stream.cu) a warp load of 32 consecutive entries of
// the level-1 operator touches 0.70 cache lines per entry and L1TEX runs at 88 % while DRAM
// sits at 40 % (ncu, profiles/r01).  Dealing ROWS to lanes cuts that to 0.28 lines per entry,
// because neighbouring coarse rows have neighbouring columns -- but then the matrix itself
// must be readable row-per-lane, which is what a slice of 32 rows stored column-major gives:
// lane l of a warp reads slot (k, l); 32 lanes read 128 contiguous bytes of indices and 256 of
// values.  Rows keep their storage order, so every row sum is folded exactly as the reference
// does it (scipy csr_matvec / amg_core jacobi, pyamg/amg_core/relaxation.h:201-239): the bits
// are those of the CSR kernels.  Cost: the copy (engine-owned, built once per level) and
// 7-10 % padding traffic (ncu, profiles/r02); no shared memory, no barriers.
//
// Round 2 added, all as exact re-encodings of the same ordered sums: 16-bit column steps with escape slots for big
// sorted copies (10 instead of 12 bytes per slot), copies WITHOUT the diagonal entries for the smoothers (the sweep never
// adds them; no per-slot diagonal test), few-row forms whose loads of a round are forced in flight (ptxas otherwise
// sinks each load next to its use), and the early launch of the small colour sweeps (common.cuh).
#include <type_traits>
#include "internal.cuh"
#include "setup.cuh"

namespace b200amg {

constexpr int SELL_C = 32;
constexpr int SELL_TPB = 256;
constexpr int SELL_U = 8;
constexpr int SELL_BIG_ROWS = 400000;      // below: too few warps to fill the SMs, the latency-hiding forms run

bool sell_wanted(long long nnz, int nrows)
{
    if (nnz <= 0 || nrows < 4 * SELL_C) return false;
    const double avg = (double)nnz / nrows;
    return avg > 11.0 && avg <= 400.0;
}

// ---- build ----------------------------------------------------------------------------------
// perm (nullable): slot row r holds the matrix row perm[r] (colour-grouped copies for Gauss-Seidel)
// 16-bit column steps: a step of s columns is stored as s / 65535 ESCAPE slots (0xFFFF: "65 535 columns on, no entry")
// followed by the entry's slot with s % 65535 -- a row that reaches across a grid plane of 65 536 points costs a slot
constexpr int SELL_ESC = 0xFFFF;
constexpr int SELL_HAS_ESC = (int)0x80000000;      // flag in row_len: this row holds escape slots
// kept entries of matrix row src -- all of them, or all but the diagonal ones (drop: copies made for a smoother, which
// never adds the diagonal entry to the row sum; the others keep their order) -- and the escape slots that 16-bit steps
// need between consecutive kept entries
__device__ __forceinline__ void sell_row_shape(const int *__restrict__ Ap, const int *__restrict__ Aj, int src, bool drop,
                                               bool steps, int &kept, int &esc)
{
    const int lo = Ap[src], hi = Ap[src + 1];
    kept = hi - lo;
    esc = 0;
    if (!drop && !steps) return;
    kept = 0;
    int prev = 0;
    for (int k = lo; k < hi; ++k) {
        const int j = Aj[k];
        if (drop && j == src) continue;
        if (steps && kept > 0 && j > prev) esc += (j - prev) / SELL_ESC;
        prev = j;
        ++kept;
    }
}

__global__ void __launch_bounds__(256) sell_width_kernel(int n, const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                         const int *__restrict__ perm, bool drop, bool steps,
                                                         long long *__restrict__ width, int *__restrict__ row_len)
{
    const int row = blockIdx.x * 256 + threadIdx.x;
    int len = 0;
    if (row < n) {
        int kept, esc;
        sell_row_shape(Ap, Aj, perm ? perm[row] : row, drop, steps, kept, esc);
        len = kept + esc;
        row_len[row] = esc ? (len | SELL_HAS_ESC) : len;        // the apply kernels take the escape-aware path per slice
    }
    int m = len;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && (row >> 5) < (n + SELL_C - 1) / SELL_C) width[row >> 5] = (long long)m * SELL_C;
}

// dg != nullptr: diagonal entries are left out of the slots, dg[slot row] = the LAST stored one (0.0: none) -- what
// the reference's sweeps use as the diagonal (relaxation.h:201-239, :394-426)
__global__ void __launch_bounds__(256) sell_fill_kernel(int n, const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                        const double *__restrict__ Ax, const int *__restrict__ perm,
                                                        const long long *__restrict__ slice_off, int *__restrict__ cj,
                                                        double *__restrict__ cv, const unsigned char *__restrict__ codes,
                                                        unsigned char *__restrict__ cc, unsigned short *__restrict__ cd,
                                                        int *__restrict__ first, double *__restrict__ dg)
{
    const int row = blockIdx.x * 256 + threadIdx.x;
    const int slice = row >> 5, lane = threadIdx.x & 31;
    if (slice >= (n + SELL_C - 1) / SELL_C) return;
    const long long off = slice_off[slice];
    const int w = (int)((slice_off[slice + 1] - off) / SELL_C);
    int lo = 0, len = 0, src = -1;
    if (row < n) {
        src = perm ? perm[row] : row;
        lo = Ap[src];
        len = Ap[src + 1] - lo;
    }
    int slot = 0, prev = 0;
    double diag = 0.0;
    for (int k = 0; k < len; ++k) {
        const int j = Aj[lo + k];
        if (dg && j == src) {
            diag = Ax[lo + k];
            continue;
        }
        if (cd) {                                       // 16-bit steps (never with value codes)
            const int step = (slot > 0 && j > prev) ? j - prev : 0;
            if (slot == 0) first[row] = j;
            for (int e = step / SELL_ESC; e > 0; --e, ++slot) {
                cd[off + (long long)slot * SELL_C + lane] = (unsigned short)SELL_ESC;
                cv[off + (long long)slot * SELL_C + lane] = 0.0;
            }
            cd[off + (long long)slot * SELL_C + lane] = (unsigned short)(step % SELL_ESC);
        } else {
            cj[off + (long long)slot * SELL_C + lane] = j;
        }
        if (cc) cc[off + (long long)slot * SELL_C + lane] = codes[lo + k];
        else cv[off + (long long)slot * SELL_C + lane] = Ax[lo + k];
        prev = j;
        ++slot;
    }
    if (cd && row < n && slot == 0) first[row] = 0;
    for (; slot < w; ++slot) {
        if (cd) cd[off + (long long)slot * SELL_C + lane] = 0;
        else cj[off + (long long)slot * SELL_C + lane] = 0;
        if (cc) cc[off + (long long)slot * SELL_C + lane] = 0;
        else cv[off + (long long)slot * SELL_C + lane] = 0.0;
    }
    if (dg && row < n) dg[row] = diag;
}

// 16-bit column steps need rows stored by ascending column (flag[0] = 1 otherwise); flag[1] += escape slots needed
__global__ void __launch_bounds__(256) sell_steps_check_kernel(int n, const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                               bool drop, int *__restrict__ flag)
{
    const int row = blockIdx.x * 256 + threadIdx.x;
    bool bad = false;
    int esc = 0;
    if (row < n) {
        const int hi = Ap[row + 1];
        for (int k = Ap[row] + 1; k < hi; ++k) bad |= Aj[k] < Aj[k - 1];
        int kept;
        sell_row_shape(Ap, Aj, row, drop, true, kept, esc);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) esc += __shfl_xor_sync(0xffffffffu, esc, o);
    if (bad) flag[0] = 1;
    if ((threadIdx.x & 31) == 0 && esc > 0) atomicAdd(flag + 1, min(esc, 1 << 20));
}

void sell_free(Sell *s)
{
    if (s->slice_off) cudaFree(s->slice_off);
    if (s->cj) cudaFree(s->cj);
    if (s->cd) cudaFree(s->cd);
    if (s->first) cudaFree(s->first);
    if (s->dg) cudaFree(s->dg);
    if (s->cv) cudaFree(s->cv);
    if (s->cc) cudaFree(s->cc);
    if (s->dict) cudaFree(s->dict);
    if (s->row_len) cudaFree(s->row_len);
    *s = Sell();
}

int sell_build(int n, long long nnz, const int *Ap, const int *Aj, const double *Ax, Sell *out, cudaStream_t st,
               const int *perm, const unsigned char *codes, const double *dict, bool drop_diag)
{
    const bool coded = codes != nullptr && dict != nullptr;     // 8-bit value codes instead of fp64 values
    // 16-bit column steps (10 instead of 12 bytes per slot) where bytes are what the pass costs: big uncoded operators
    // (B200AMG_SELL_COL16: 0 never, 2 for every size -- tests; read per build)
    const char *env = getenv("B200AMG_SELL_COL16");
    bool steps = !(env && env[0] == '0') && !coded && (n >= SELL_BIG_ROWS || (env && env[0] == '2'));
    *out = Sell();
    if (n <= 0) return 0;
    const int ns = (n + SELL_C - 1) / SELL_C;
    long long *width = nullptr, *scratch = nullptr;
    Sell s;
    s.n = n;
    cudaError_t e = cudaMalloc(&width, sizeof(long long) * ns);
    if (e == cudaSuccess) e = cudaMalloc(&scratch, sizeof(long long) * (size_t)scan_scratch_elems(ns));
    if (e == cudaSuccess) e = cudaMalloc(&s.slice_off, sizeof(long long) * (ns + 1));
    if (e == cudaSuccess) e = cudaMalloc(&s.row_len, sizeof(int) * (size_t)n);
    int rc = 0;
    if (e == cudaSuccess && steps) {
        // sorted rows, and escape slots for steps >= 65 535 a small fraction of the entries (a 512-point-wide grid
        // would need four per plane crossing: 32-bit columns are cheaper then)
        int *flag = reinterpret_cast<int *>(width);
        int got[2] = {0, 0};
        e = cudaMemsetAsync(flag, 0, 2 * sizeof(int), st);
        if (e == cudaSuccess) {
            B200_LAUNCH(sell_steps_check_kernel, (n + 255) / 256, 256, 0, st, n, Ap, Aj, drop_diag, flag);
            rc = check_launch("sell_steps_check_kernel");
        }
        if (!rc && e == cudaSuccess) e = cudaMemcpyAsync(got, flag, 2 * sizeof(int), cudaMemcpyDeviceToHost, st);
        if (!rc && e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (got[0] || (long long)got[1] * 32 > nnz) steps = false;
    }
    if (e == cudaSuccess && !rc) {
        B200_LAUNCH(sell_width_kernel, (n + 255) / 256, 256, 0, st, n, Ap, Aj, perm, drop_diag, steps, width, s.row_len);
        rc = check_launch("sell_width_kernel");
        if (!rc) rc = exclusive_scan<long long, long long>(ns, width, s.slice_off, scratch, st);
        if (!rc) {
            e = cudaMemcpyAsync(&s.padded, s.slice_off + ns, sizeof(long long), cudaMemcpyDeviceToHost, st);
            if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        }
    }
    // refuse pathological padding (one long row per slice): the CSR kernels handle those
    if (e == cudaSuccess && !rc && s.padded > 0 && s.padded <= 2 * nnz + 64 * SELL_C) {
        if (steps) {
            e = cudaMalloc(&s.cd, sizeof(unsigned short) * (size_t)s.padded + 64);
            if (e == cudaSuccess) e = cudaMalloc(&s.first, sizeof(int) * (size_t)n);
        } else {
            e = cudaMalloc(&s.cj, sizeof(int) * (size_t)s.padded);
        }
        if (e == cudaSuccess && drop_diag) e = cudaMalloc(&s.dg, sizeof(double) * (size_t)n);
        if (e == cudaSuccess && !coded) e = cudaMalloc(&s.cv, sizeof(double) * (size_t)s.padded);
        if (e == cudaSuccess && coded) e = cudaMalloc(&s.cc, (size_t)s.padded + 64);
        if (e == cudaSuccess && coded) e = cudaMalloc(&s.dict, sizeof(double) * 256);
        if (e == cudaSuccess && coded) e = cudaMemcpyAsync(s.dict, dict, sizeof(double) * 256, cudaMemcpyDeviceToDevice, st);
        if (e == cudaSuccess) {
            B200_LAUNCH(sell_fill_kernel, (n + 255) / 256, 256, 0, st, n, Ap, Aj, Ax, perm, s.slice_off, s.cj, s.cv, codes, s.cc,
                        s.cd, s.first, s.dg);
            rc = check_launch("sell_fill_kernel");
            // the apply kernels request slots of this copy BEFORE they wait for their predecessor in the stream
            // (programmatic dependent launch): the copy must be complete when the build returns
            if (!rc) e = cudaStreamSynchronize(st);
        }
    }
    if (width) cudaFree(width);
    if (scratch) cudaFree(scratch);
    if (e != cudaSuccess || rc || !(s.cj || s.cd) || !(s.cv || s.cc)) {
        sell_free(&s);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return set_err((int)e, "sell_build: device allocation failed");
        }
        return rc;      // rc == 0 with no arrays: layout declined, caller keeps CSR
    }
    *out = s;
    return 0;
}

// ---- apply ----------------------------------------------------------------------------------
// PIPE: prefetch the next round's slots (2 CTAs/SM at U = 8) -- pays when the matrix has too few rows to fill the
// SMs with warps anyway (a warp is then a chain of w / U dependent memory latencies); large matrices run the plain
// form (64 registers, 4 CTAs/SM).  U: slots per lane per round.
// FMT 0: 32-bit columns + fp64 values (12 bytes per slot)
// FMT 1: the value slots hold 8-bit codes into a table of <= 256 doubles (dict.cu) -- 5 bytes per slot
// FMT 2: 16-bit column STEPS + fp64 values (10 bytes per slot): column = first[row] + running sum of the steps
// DF: a copy made for smoothing (Sell::dg): no diagonal entries among the slots, the diagonal comes from dg[row]
template <int OP, int U, bool PIPE, int FMT, bool TIE = false, bool DF = false>
__global__ void __launch_bounds__(SELL_TPB) sell_kernel(const int n, const long long *__restrict__ slice_off,
                                                        const void *__restrict__ cols, const void *__restrict__ cvals,
                                                        const double *__restrict__ dict, const int *__restrict__ first,
                                                        const int *__restrict__ row_len, const double *__restrict__ dg,
                                                        const StreamArgs a)
{
    static_assert(!DF || OP >= OP_JACOBI, "diagonal-free copies serve the smoothing ops only");
    constexpr bool CODED = (FMT == 1), STEPS = (FMT == 2);
    static_assert(!(PIPE && FMT != 0), "the prefetching form keeps 32-bit columns and fp64 values");
    static_assert(PIPE || !TIE, "only the prefetching form ties its loads");
    using V = typename std::conditional<CODED, unsigned char, double>::type;
    using J = typename std::conditional<STEPS, unsigned short, int>::type;
    const V *__restrict__ cv = reinterpret_cast<const V *>(cvals);
    const J *__restrict__ cj = reinterpret_cast<const J *>(cols);
    __shared__ double s_dict[CODED ? 256 : 1];
    pdl_trigger();
    if (CODED) {
        s_dict[threadIdx.x] = dict[threadIdx.x];               // coded launches use SELL_TPB == 256 threads
        __syncthreads();
    }
    auto value = [&](V raw) -> double {
        if constexpr (CODED) return s_dict[raw];
        else return raw;
    };
    constexpr bool SMOOTH = (OP >= OP_JACOBI);
    constexpr bool BSRF = (OP == OP_JACOBI_BSR1);
    constexpr bool NEED_B = (OP == OP_RESIDUAL || SMOOTH);
    const int row = blockIdx.x * blockDim.x + threadIdx.x;
    const int slice = row >> 5, lane = threadIdx.x & 31;
    if (slice >= (n + SELL_C - 1) / SELL_C) return;            // whole warp
    const long long off = slice_off[slice];
    const int w = (int)((slice_off[slice + 1] - off) / SELL_C);
    const bool live = row < n;
    const int len_raw = live ? row_len[row] : 0;
    const int len = len_raw & ~SELL_HAS_ESC;
    // escape slots are rare (rows that cross a grid plane of >= 65 535 points): slices without any take a loop that
    // does not test for them (the test costs the level-1 Jacobi sweep of config 2 a quarter of its speed)
    const bool slice_esc = STEPS && __any_sync(0xffffffffu, len_raw < 0);
    const double *__restrict__ xg = a.x;
    double bi = 0.0, xi = 0.0, yi = 0.0;
    int run = (STEPS && live) ? __ldg(first + row) : 0;        // STEPS: column of the last slot walked
    // x, b, y may be what the previous kernel of the stream wrote: nothing of them is read before pdl_wait()
    auto load_vectors = [&] {
        if (live) {
            if (NEED_B) bi = __ldg(a.b + row);
            if (SMOOTH) xi = __ldg(xg + row);
            if (OP == OP_ADD) yi = a.b[row];
        }
    };
    double acc = 0.0, diag = (DF && live) ? __ldg(dg + row) : 0.0;
    const J *pj = cj + off + lane;
    const V *pv = cv + off + lane;
    if constexpr (PIPE) {
    // software pipeline: the (index, value) slots of round r+1 are requested before round r's x
    // gathers are consumed, so a round costs one memory latency, not two
    int c[U], cn[U];
    double v[U], vn[U], xv[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const bool in = u < len;
        c[u] = in ? (int)__ldcs(pj + (long long)u * SELL_C) : -1;
        v[u] = in ? value(__ldcs(pv + (long long)u * SELL_C)) : 0.0;
    }
    pdl_wait();                                                 // the first round of slots is already on its way
    load_vectors();
    acc = BSRF ? bi : 0.0;
    for (int k0 = 0; k0 < w; k0 += U) {
#pragma unroll
        for (int u = 0; u < U; ++u) xv[u] = (c[u] >= 0) ? __ldg(xg + c[u]) : 0.0;
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int k = k0 + U + u;
            const bool in = k < len;
            cn[u] = in ? (int)__ldcs(pj + (long long)k * SELL_C) : -1;
            vn[u] = in ? value(__ldcs(pv + (long long)k * SELL_C)) : 0.0;
        }
        // ptxas otherwise sinks every load next to its use to save registers (3-7 gathers in flight, the prefetch
        // AFTER the fold: a round then costs four latencies).  Making the first fold step wait for a word that depends
        // on every load of the round forces all of them to be issued first; `tie` is 0 at run time (row_begin == 0:
        // sell_op takes full row ranges only), which the compiler cannot know.
        int tie = 0;
        if (TIE) {
#pragma unroll
            for (int u = 0; u < U; ++u) tie |= __double2loint(xv[u]) | cn[u] | __double2loint(vn[u]);
            tie &= a.row_begin;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if ((u == 0 ? c[u] + tie : c[u]) >= 0) {
                const double pr = mul_rn(v[u], xv[u]);
                if (!DF && SMOOTH && c[u] == row) diag = v[u];          // last stored diagonal wins
                else acc = BSRF ? sub_rn(acc, pr) : add_rn(acc, pr);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            c[u] = cn[u];
            v[u] = vn[u];
        }
    }
    } else {
    pdl_wait();
    load_vectors();
    acc = BSRF ? bi : 0.0;
    auto rounds = [&](auto esc_tag) {
        constexpr bool ESC = decltype(esc_tag)::value;
        for (int k0 = 0; k0 < w; k0 += U) {
            int c[U];
            J raw[U];
            double v[U], xv[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const bool in = k0 + u < len;
                raw[u] = in ? __ldcs(pj + (long long)(k0 + u) * SELL_C) : (J)0;
                v[u] = in ? value(__ldcs(pv + (long long)(k0 + u) * SELL_C)) : 0.0;
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const bool in = k0 + u < len;
                const bool esc = ESC && (int)raw[u] == SELL_ESC;        // "65 535 columns on, no entry"
                if (STEPS) run += (int)raw[u];                          // padding slots hold step 0
                c[u] = (in && !esc) ? (STEPS ? run : (int)raw[u]) : -1;
            }
#pragma unroll
            for (int u = 0; u < U; ++u) xv[u] = (c[u] >= 0) ? __ldg(xg + c[u]) : 0.0;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (c[u] >= 0) {
                    const double pr = mul_rn(v[u], xv[u]);
                    if (!DF && SMOOTH && c[u] == row) diag = v[u];      // last stored diagonal wins
                    else acc = BSRF ? sub_rn(acc, pr) : add_rn(acc, pr);
                }
            }
        }
    };
    if (STEPS && slice_esc) rounds(std::true_type{});
    else rounds(std::false_type{});
    }
    if (!live) return;
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
        y[row] = (diag != 0.0) ? add_rn(mul_rn(sub_rn(1.0, a.omega), xi), div_rn(mul_rn(a.omega, acc), diag)) : xi;
    }
}

// ---- colour-ordered Gauss-Seidel on a colour-grouped SELL copy -----------------------------------
// amg_core.gauss_seidel_indexed (relaxation.h:394-426) / bsr_gauss_seidel with bs = 1 (:89-173) for the
// positions [p0, p1) of the index list `Id` = ONE colour: rows of a colour never read each other, so the
// parallel update equals the sequential sweep bit for bit.  Slot row r of the copy holds matrix row Id[r]
// (columns keep the original numbering); x is updated in place.
// U slots per lane per round; `zero` is 0 at run time and ties every round's loads together (see sell_kernel): a
// colour of a coarse level is a few thousand rows -- one wave of warps whose run time is the number of dependent
// memory latencies in a row walk (2 per round), so short rows want ONE round.
// STEPS: 16-bit column steps instead of 32-bit columns (Sell::cd / first)
// DF: the copy holds no diagonal entries, dg[slot row] is the diagonal (sell_build drop_diag)
template <bool BSRF, int U, bool STEPS, bool DF>
__global__ void __launch_bounds__(SELL_TPB) sell_gs_kernel(const int p0, const int p1, const long long *__restrict__ slice_off,
                                                           const void *__restrict__ cols, const double *__restrict__ cv,
                                                           const int *__restrict__ first,
                                                           const int *__restrict__ row_len, const int *__restrict__ Id,
                                                           const double *__restrict__ dg,
                                                           double *x, const double *__restrict__ b, const int zero)
{
    constexpr bool TIE = U > 8;
    using J = typename std::conditional<STEPS, unsigned short, int>::type;
    const J *__restrict__ cj = reinterpret_cast<const J *>(cols);
    // Programmatic dependent launch (common.cuh): the NEXT colour's kernel may start as soon as every block of this one
    // has, and everything up to pdl_wait() below reads only what no kernel of the cycle writes (the colour-grouped copy
    // of the operator, the index list).  A colour of a coarse level is launch latency plus a chain of dependent loads;
    // this takes the launch and the operator part of the chain off the critical path.
    pdl_trigger();
    const int r = (p0 & ~31) + blockIdx.x * blockDim.x + threadIdx.x;       // slot row (slices may straddle colours)
    const int slice = r >> 5, lane = threadIdx.x & 31;
    if ((slice << 5) >= p1) return;                                         // whole warp
    const bool live = r >= p0 && r < p1;
    const long long off = slice_off[slice];
    const int w = (int)((slice_off[slice + 1] - off) / SELL_C);
    const int len = live ? (row_len[r] & ~SELL_HAS_ESC) : 0;
    const int row = live ? Id[r] : -1;
    const J *pj = cj + off + lane;
    const double *pv = cv + off + lane;
    int run = (STEPS && live) ? first[r] : 0;
    double bi = 0.0, acc = 0.0, diag = (DF && live) ? dg[r] : 0.0;
    for (int k0 = 0; k0 < w; k0 += U) {
        int c[U];
        J raw[U];
        double v[U], xv[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const bool in = k0 + u < len;
            raw[u] = in ? __ldcs(pj + (long long)(k0 + u) * SELL_C) : (J)0;
            v[u] = in ? __ldcs(pv + (long long)(k0 + u) * SELL_C) : 0.0;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            // (tested per slot here: a per-slice switch like sell_kernel's made THIS kernel slower -- config 3 level 0,
            // 1.02 -> 1.21 ms per step, profiles/r02/run22_summary.txt)
            const bool esc = STEPS && (int)raw[u] == SELL_ESC;
            if (STEPS) run += (int)raw[u];
            c[u] = (k0 + u < len && !esc) ? (STEPS ? run : (int)raw[u]) : -1;
        }
        int tie = 0;
        if (TIE) {
#pragma unroll
            for (int u = 0; u < U; ++u) tie |= (int)raw[u] | __double2loint(v[u]);
            tie &= zero;
        }
        if (k0 == 0) {
            pdl_wait();                                                     // x and b are final from here on
            bi = live ? b[row] : 0.0;
            acc = BSRF ? bi : 0.0;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) xv[u] = (c[u] >= 0) ? x[c[u] + (u == 0 ? tie : 0)] : 0.0;
        if (TIE) {
#pragma unroll
            for (int u = 0; u < U; ++u) tie |= __double2loint(xv[u]);
            tie &= zero;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if ((u == 0 ? c[u] + tie : c[u]) >= 0) {
                if (!DF && c[u] == row) diag = v[u];
                else {
                    const double pr = mul_rn(v[u], xv[u]);
                    acc = BSRF ? sub_rn(acc, pr) : add_rn(acc, pr);
                }
            }
        }
    }
    if (w == 0) {                                                           // a slice of empty rows
        pdl_wait();
        bi = live ? b[row] : 0.0;
    }
    if (live && diag != 0.0) x[row] = BSRF ? div_rn(acc, diag) : div_rn(sub_rn(bi, acc), diag);
}

template <bool BSRF>
static int launch_sell_gs(const Sell &s, const int *Id, int p0, int p1, double *x, const double *b, cudaStream_t st)
{
    // B200AMG_SELL_GS_U: 8 / 32 for every colour (A/B); default by the size of the colour
    static const int forced = [] { const char *e = getenv("B200AMG_SELL_GS_U"); return e ? atoi(e) : 0; }();
    const int rows = p1 - (p0 & ~31);
    // config 3 (run 15): colours below 64 k rows 3-7 % faster with one 32-slot round; 121 k rows: 16 slots SLOWER than 8
    const int U = forced ? forced : (rows < 64 * 1024 ? 32 : 8);
    const void *cols = s.cd ? (const void *)s.cd : (const void *)s.cj;
    const int *firstp = s.cd ? s.first : nullptr;
#define SELL_GS_ONE(EARLY, UU, TPB, ST, DFV)                                                                                     \
    B200_LAUNCH_PDL(EARLY, (sell_gs_kernel<BSRF, UU, ST, DFV>), (unsigned)((rows + TPB - 1) / TPB), TPB, 0, st, p0, p1,          \
                    (const long long *)s.slice_off, cols, (const double *)s.cv, firstp, (const int *)s.row_len, Id,             \
                    (const double *)s.dg, x, b, 0)
#define SELL_GS_LAUNCH(EARLY, UU, TPB)                                                                                           \
    do {                                                                                                                         \
        if (s.cd) { if (s.dg) SELL_GS_ONE(EARLY, UU, TPB, true, true); else SELL_GS_ONE(EARLY, UU, TPB, true, false); }          \
        else { if (s.dg) SELL_GS_ONE(EARLY, UU, TPB, false, true); else SELL_GS_ONE(EARLY, UU, TPB, false, false); }             \
    } while (0)
    if (U == 32) SELL_GS_LAUNCH(pdl_class(PDL_SWEEP), 32, 64);
    else SELL_GS_LAUNCH(false, 8, SELL_TPB);
#undef SELL_GS_LAUNCH
#undef SELL_GS_ONE
    return check_launch("sell_gs_kernel");
}

int sell_gs_color(bool bsrf, const Sell &s, const int *Id, int p0, int p1, double *x, const double *b, cudaStream_t st)
{
    if (p1 <= p0) return 0;
    return bsrf ? launch_sell_gs<true>(s, Id, p0, p1, x, b, st) : launch_sell_gs<false>(s, Id, p0, p1, x, b, st);
}

// few-row operators (< SELL_BIG_ROWS).  B200AMG_SELL_SMALL=0: the round-1 form for every op (A/B switch)
static bool sell_small_deep()
{
    static const bool on = [] { const char *e = getenv("B200AMG_SELL_SMALL"); return !(e && e[0] == '0'); }();
    return on;
}

template <int OP>
static int launch_sell(const Sell &s, const StreamArgs &a, cudaStream_t st)
{
    const int grid = (s.n + SELL_TPB - 1) / SELL_TPB;
    const double *nodg = nullptr;
    if (s.dg) {
        // a copy without diagonal entries: smoothing ops only (the engine builds it next to the full copy for levels of
        // >= SELL_BIG_ROWS rows; any size works)
        if constexpr (OP >= OP_JACOBI) {
            if (s.cc) return set_err(B200AMG_ESTATE, "sell_op: diagonal-free copy of a coded operator");
            if (s.cd)
                B200_LAUNCH_PDL(false, (sell_kernel<OP, SELL_U, false, 2, false, true>), grid, SELL_TPB, 0, st, s.n, s.slice_off, s.cd, s.cv, nullptr, s.first, s.row_len, s.dg, a);
            else
                B200_LAUNCH_PDL(false, (sell_kernel<OP, SELL_U, false, 0, false, true>), grid, SELL_TPB, 0, st, s.n, s.slice_off, s.cj, s.cv, nullptr, nullptr, s.row_len, s.dg, a);
            return check_launch("sell_kernel");
        } else {
            return set_err(B200AMG_ESTATE, "sell_op: a diagonal-free copy serves the smoothing ops only");
        }
    }
    if (s.cc) {
        B200_LAUNCH_PDL(pdl_class(PDL_OTHER) && s.n < PDL_SMALL_ROWS, (sell_kernel<OP, SELL_U, false, 1>), grid, SELL_TPB, 0, st, s.n, s.slice_off, s.cj, s.cc, s.dict, nullptr, s.row_len, nodg, a);
    } else if (s.cd) {
        B200_LAUNCH_PDL(pdl_class(PDL_OTHER) && s.n < PDL_SMALL_ROWS, (sell_kernel<OP, SELL_U, false, 2>), grid, SELL_TPB, 0, st, s.n, s.slice_off, s.cd, s.cv, nullptr, s.first, s.row_len, nodg, a);
    } else if (s.n < SELL_BIG_ROWS) {
        // a warp is a chain of w / U dependent latencies and there are too few warps to hide it: 16 slots per round,
        // all forced in flight, and 64-thread blocks so that the few warps spread over all SMs.  Measured on config 2
        // (profiles/r02/run15_summary.txt): restriction of level 1 (44 629 rows x 240) 0.088 -> 0.047 ms, level-2
        // residual 0.021 -> 0.017 ms; 8 slots tied 0.052, 32 slots in two phases 0.064, 256-thread blocks 0.064.  The
        // smoothing ops already kept their loads together (the diagonal test splits the basic block) and are fastest
        // as they were (0.021 vs 0.028 ms).
        if (OP < OP_JACOBI && sell_small_deep())
            B200_LAUNCH_PDL(pdl_class(PDL_SELL), (sell_kernel<OP, 16, true, 0, true>), (s.n + 63) / 64, 64, 0, st, s.n, s.slice_off, s.cj, s.cv, nullptr, nullptr, s.row_len, nodg, a);
        else
            B200_LAUNCH_PDL(pdl_class(PDL_SELL), (sell_kernel<OP, SELL_U, true, 0>), grid, SELL_TPB, 0, st, s.n, s.slice_off, s.cj, s.cv, nullptr, nullptr, s.row_len, nodg, a);
    } else {
        B200_LAUNCH_PDL(pdl_class(PDL_OTHER) && s.n < PDL_SMALL_ROWS, (sell_kernel<OP, SELL_U, false, 0>), grid, SELL_TPB, 0, st, s.n, s.slice_off, s.cj, s.cv, nullptr, nullptr, s.row_len, nodg, a);
    }
    return check_launch("sell_kernel");
}

int sell_op(int op, const Sell &s, const StreamArgs &a0, cudaStream_t st)
{
    StreamArgs a = a0;
    if (op == OP_ADD && !a.b) a.b = a.y;              // y += A x: the kernels read the addend through b
    if (a.row_begin != 0 || a.row_end != s.n) return set_err(B200AMG_EINVAL, "sell_op: full row range only");
    switch (op) {
    case OP_MATVEC: return launch_sell<OP_MATVEC>(s, a, st);
    case OP_SPMV_AB: return launch_sell<OP_SPMV_AB>(s, a, st);
    case OP_RESIDUAL: return launch_sell<OP_RESIDUAL>(s, a, st);
    case OP_ADD: return launch_sell<OP_ADD>(s, a, st);
    case OP_JACOBI: return launch_sell<OP_JACOBI>(s, a, st);
    case OP_JACOBI_BSR1: return launch_sell<OP_JACOBI_BSR1>(s, a, st);
    }
    return set_err(B200AMG_EINVAL, "sell_op: bad op");
}

}  // namespace b200amg
