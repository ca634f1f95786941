// Persistent, TMA-fed CSR pipeline: the fine-level SpMV / residual / prolongation / Jacobi pass.
//
// Same arithmetic contract as csr_stream.cu (storage-order, separately rounded: bit-identical
// to scipy csr_matvec and amg_core jacobi / bsr_jacobi(bs=1), pyamg/amg_core/relaxation.h:201-239,
// :267-360), different schedule: csr_stream.cu is latency-bound (ncu: long_scoreboard + barrier
// stalls, 40-57 % of DRAM peak) because every CTA runs the dependent chain
// row pointers -> indices/values -> x gather with nothing in flight behind it.  Here
//   * the grid is persistent (3 CTAs per SM); CTA c owns tiles c, c+grid, c+2*grid, ...;
//     a tile is `rpt` consecutive rows (256 for short rows, fewer for long ones);
//   * one thread drives two rings of bulk async copies (cp.async.bulk = TMA engine, mbarrier
//     completion): row-pointer slices 3 tiles ahead, and -- once a slice has landed and the
//     tile's nonzero range is known -- its (Aj, Ax) range 1 tile ahead, straight into shared
//     memory.  The <= 3 ragged tail entries travel as 4/8-byte cp.async on the same mbarrier,
//     so no load ever reads past the arrays;
//   * eight consumer warps (thread t = row t of the tile) take tile i from shared memory while
//     tile i+1 streams in; slots are handed back through "empty" mbarriers, one arrival per
//     warp, so warps drift apart freely and no CTA-wide barrier sits on the path:
//       row-parallel (<= 12 nnz/row): thread t walks row t, gathering x[Aj] (neighbouring rows
//         of a banded operator touch neighbouring x: 2-3 cache lines per warp load);
//       entry-parallel (long rows): all threads turn values into rounded products in place,
//         then thread t folds row t in storage order.
// A tile whose range does not fit a stage (rows much longer than the matrix average) is
// summed straight from global memory by its row threads: slower, same bits.
#include <type_traits>
#include "internal.cuh"
#include "pipe.cuh"

namespace b200amg {

constexpr int P_ROWS = 256;        // consumer threads = max rows per tile
constexpr int P_TPB = P_ROWS + 32; // + one producer warp
constexpr int P_CAP = 2816;        // entries per data stage (256 rows x 11)
constexpr int P_PTRS = 4;          // row-pointer ring
constexpr int P_DATA = 2;          // (Aj, Ax) ring
constexpr int P_CTAS_PER_SM = 3;
constexpr int P_ROWPAR_MAX_AVG = 12;
constexpr int P_CWARPS = P_ROWS / 32;

// MODE 0: (Aj, Ax) as stored.
// MODE 1: the operator's values are 8-bit codes into a table of <= 256 distinct doubles -- 5 bytes per entry
//         instead of 12 cross HBM; the decoded value is the same double, so the arithmetic and its bits are
//         unchanged.  The smaller stages pay for a deeper (4-slot) data ring.
// MODE 2: 8-bit codes into a table of (column offset, value) pairs (constant-coefficient stencils): 1 byte per
//         entry; the column is row + offset.
struct __align__(16) PatEntry {
    double v;
    int off, pad;
};

template <int MODE>
struct __align__(16) PipeSmemT {
#ifndef PIPE_DICT_ND
#define PIPE_DICT_ND 4
#endif
    static constexpr bool CODED = MODE != 0;
    static constexpr int ND = CODED ? PIPE_DICT_ND : P_DATA;     // data ring slots
    static constexpr int NP = CODED ? PIPE_DICT_ND + 2 : P_PTRS; // row-pointer ring slots
    static constexpr int CAP = CODED ? 2064 : P_CAP;             // entries per slot (coded: 256 rows x 8 + alignment)
    using V = typename std::conditional<CODED, unsigned char, double>::type;
    V v[ND][CAP];
    int c[MODE == 2 ? 1 : ND][MODE == 2 ? 4 : CAP];
    int ptr[NP][P_ROWS + 8];
    PatEntry pat[MODE == 2 ? 256 : 1];
    double dict[MODE == 1 ? 256 : 2];
    uint64_t ptr_full[NP], ptr_empty[NP];
    uint64_t data_full[ND], data_empty[ND];
};

__device__ __forceinline__ void consumer_sync()
{
    asm volatile("bar.sync 1, %0;" ::"n"(P_ROWS) : "memory");
}

template <int OP, int MODE>
__global__ void __launch_bounds__(P_TPB, P_CTAS_PER_SM) csr_pipe_kernel(const StreamArgs a, const int rpt,
                                                                        const int num_tiles)
{
    using Smem = PipeSmemT<MODE>;
    using V = typename Smem::V;
    constexpr int ND = Smem::ND, NP = Smem::NP, CAP = Smem::CAP;
    constexpr bool DICT = MODE != 0;                     // coded entries
    constexpr bool HAVE_C = MODE != 2;                   // column indices travel with the entries
    constexpr int ALIGN = DICT ? 16 : 4;                 // entries per 16 bytes of the narrowest staged array
    constexpr bool SMOOTH = (OP >= OP_JACOBI);
    constexpr bool BSRF = (OP == OP_JACOBI_BSR1);
    constexpr bool NEED_B = (OP == OP_RESIDUAL || SMOOTH);
    constexpr bool NEED_XI = SMOOTH;
    constexpr bool NEED_Y = (OP == OP_ADD);
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem &S = *reinterpret_cast<Smem *>(smem_raw);
    const int t = threadIdx.x;
    pdl_trigger();
    if (t == 0) {
        for (int i = 0; i < NP; ++i) { mbar_init(&S.ptr_full[i], 2); mbar_init(&S.ptr_empty[i], P_CWARPS); }
        for (int i = 0; i < ND; ++i) { mbar_init(&S.data_full[i], 2); mbar_init(&S.data_empty[i], P_CWARPS); }
        mbar_fence_init();
    }
    if (MODE == 1 && t < 256) S.dict[t] = a.dict[t];           // (engine-owned tables: complete before any launch)
    if (MODE == 2 && t < 256) S.pat[t] = PatEntry{a.dict[t], a.offs[t], 0};
    __syncthreads();
    pdl_wait();                 // set-up overlaps the previous kernel of the stream; the operator and the vectors wait for it

    const int my_tiles = (num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    const int *__restrict__ Ap = a.Ap;
    const int *__restrict__ Aj = a.Aj;
    const double *__restrict__ Ax = a.Ax;
    const V *__restrict__ Av = DICT ? reinterpret_cast<const V *>(a.codes) : reinterpret_cast<const V *>(a.Ax);
    const double *__restrict__ xg = a.x;
    // entries [0, arr_end) of Aj / Ax exist (known when the row range starts at row 0)
    const long long arr_end = (a.row_begin == 0) ? a.nnz : -1;

    auto tile_rows = [&](int i, int &r0, int &nr) {
        const int tile = (int)blockIdx.x + i * (int)gridDim.x;
        r0 = a.row_begin + tile * rpt;
        nr = min(rpt, a.row_end - r0);
    };

    if (t >= P_ROWS) {
        // ---------------- producer warp: one lane drives both copy rings -----------------------
        if (t != P_ROWS) return;
        // row pointers run LA tiles ahead of the data copies; LA = NP - ND keeps the pointer ring from
        // being the binding constraint (slot of tile j+LA is free once tile j+LA-NP = j-ND is consumed,
        // which the data slot of tile j needs anyway)
        constexpr int LA = NP - ND;
        for (int step = 0; step < my_tiles + LA; ++step) {
            const int jd = step - LA;
            if (jd >= 0) {                                 // (Aj, Ax) range of tile jd
                int r0, nr;
                tile_rows(jd, r0, nr);
                mbar_wait(&S.ptr_full[jd % NP], (jd / NP) & 1);
                if (jd >= ND) mbar_wait(&S.data_empty[jd % ND], ((jd / ND) - 1) & 1);
                const int *sp = S.ptr[jd % NP];
                const int s = sp[0], e = sp[nr];
                const int base = s & ~(ALIGN - 1);
                const int st = jd % ND;
                uint64_t *bar = &S.data_full[st];
                if (e - base > CAP) {                    // does not fit: rows read global memory
                    mbar_arrive(bar);
                    mbar_arrive(bar);
                } else {
                    // whole 16-byte groups travel as bulk copies; when the rounded-up range still lies inside
                    // the arrays (every tile but the matrix' last) there is no tail at all
                    const int e_pad = (e + ALIGN - 1) & ~(ALIGN - 1);
                    const bool over = (long long)e_pad <= arr_end;
                    const int nb = over ? e_pad - base : (e - base) & ~(ALIGN - 1);
                    mbar_arrive_expect_tx(bar, (uint32_t)nb * ((HAVE_C ? 4u : 0u) + (uint32_t)sizeof(V)));
                    if (nb) {
                        if (HAVE_C) bulk_g2s(S.c[st], Aj + base, (uint32_t)nb * 4u, bar);
                        bulk_g2s(S.v[st], Av + base, (uint32_t)nb * (uint32_t)sizeof(V), bar);
                    }
                    if (!over) {
                        if (HAVE_C)
                            for (int k = base + nb; k < e; ++k) cp_async_4(&S.c[st][k - base], Aj + k);
                        if (DICT) {                        // codes travel in 4-byte groups (the code array is padded)
                            for (int k = base + nb; k < e; k += 4) cp_async_4(&S.v[st][k - base], Av + k);
                        } else {
                            for (int k = base + nb; k < e; ++k) cp_async_8(&S.v[st][k - base], Av + k);
                        }
                    }
                    mbar_arrive_after_cp_async(bar);
                }
            }
            const int jp = step;
            if (jp < my_tiles) {                           // row-pointer slice of tile jp
                int r0, nr;
                tile_rows(jp, r0, nr);
                const int st = jp % NP;
                if (jp >= NP) mbar_wait(&S.ptr_empty[st], ((jp / NP) - 1) & 1);
                uint64_t *bar = &S.ptr_full[st];
                const int cnt = nr + 1, nb = cnt & ~3;
                mbar_arrive_expect_tx(bar, (uint32_t)nb * 4u);
                if (nb) bulk_g2s(S.ptr[st], Ap + r0, (uint32_t)nb * 4u, bar);
                for (int q = nb; q < cnt; ++q) cp_async_4(&S.ptr[st][q], Ap + r0 + q);
                mbar_arrive_after_cp_async(bar);
            }
        }
        return;
    }

    // -------------------- consumer warps: thread t owns row t of every tile -----------------------
    for (int i = 0; i < my_tiles; ++i) {
        int r0, nr;
        tile_rows(i, r0, nr);
        const int row = r0 + t;
        // vector operands of the epilogue are requested first: their latency hides behind the row sum
        double bi = 0.0, xi = 0.0, yi = 0.0;
        if (t < nr) {
            if (NEED_B) bi = __ldg(a.b + row);
            if (NEED_XI) xi = __ldg(xg + row);
            if (NEED_Y) yi = a.b[row];                         // OP_ADD: y = b + A x (b == y: in place)
        }
        mbar_wait(&S.ptr_full[i % NP], (i / NP) & 1);
        const int *sp = S.ptr[i % NP];
        const int s = sp[0], e = sp[nr];
        const int base = s & ~(ALIGN - 1);
        const int *sc = S.c[HAVE_C ? i % ND : 0];
        V *sv = S.v[i % ND];
        auto fetch = [&](int k, int &c, double &v) {     // entry k of the staged range: column and value
            if constexpr (MODE == 2) {
                const PatEntry pe = S.pat[sv[k]];
                c = row + pe.off;
                v = pe.v;
            } else if constexpr (MODE == 1) {
                c = sc[k];
                v = S.dict[sv[k]];
            } else {
                c = sc[k];
                v = sv[k];
            }
        };
        const bool rowpar = (e - s) <= P_ROWPAR_MAX_AVG * nr;
        int lo = 0, hi = 0;
        if (t < nr) {
            lo = sp[t];
            hi = sp[t + 1];
        }
        double acc = BSRF ? bi : 0.0, diag = 0.0;
        mbar_wait(&S.data_full[i % ND], (i / ND) & 1);
        if (e - base > CAP || (DICT && !rowpar)) {
            for (int k = lo; k < hi; ++k) {
                const int c = Aj[k];
                const double v = Ax[k];
                const double pr = mul_rn(v, xg[c]);
                if (SMOOTH && c == row) diag = v;
                else acc = BSRF ? sub_rn(acc, pr) : add_rn(acc, pr);
            }
        } else if (rowpar) {
            const int k0 = lo - base, len = hi - lo;
            if (len <= 8) {
                // all gathers of the row in flight at once, then the ordered fold.  Jacobi only records WHICH
                // slots hold the diagonal (bit mask): they are skipped by the fold, the last one is the divisor.
                double vv[8], xv[8];
                unsigned dm = 0;
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    if (q < len) {
                        int c;
                        fetch(k0 + q, c, vv[q]);
                        if (SMOOTH && c == row) dm |= 1u << q;
                        xv[q] = __ldg(xg + c);
                    }
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    if (q < len) {
                        const double pr = mul_rn(vv[q], xv[q]);
                        if (!SMOOTH || !((dm >> q) & 1u)) acc = BSRF ? sub_rn(acc, pr) : add_rn(acc, pr);
                    }
                if (SMOOTH && dm) {                                  // last stored diagonal wins
                    int c;
                    fetch(k0 + 31 - __clz(dm), c, diag);
                }
            } else {
#pragma unroll 4
                for (int k = k0; k < k0 + len; ++k) {
                    int c;
                    double v;
                    fetch(k, c, v);
                    const double pr = mul_rn(v, __ldg(xg + c));
                    if (SMOOTH && c == row) diag = v;
                    else acc = BSRF ? sub_rn(acc, pr) : add_rn(acc, pr);
                }
            }
        } else {
            if constexpr (MODE == 0) {
                for (int k = s - base + t; k < e - base; k += P_ROWS) sv[k] = mul_rn(sv[k], __ldg(xg + sc[k]));
                consumer_sync();
                for (int k = lo - base; k < hi - base; ++k) {
                    if (SMOOTH && sc[k] == row) diag = __ldg(Ax + base + k);
                    else acc = BSRF ? sub_rn(acc, sv[k]) : add_rn(acc, sv[k]);
                }
                // the in-place products were generic-proxy writes: order them before the next bulk refill
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            }
        }
        // this warp is done with both slots (reads only on the other paths: the mbarrier orders them)
        __syncwarp();
        if ((t & 31) == 0) {
            mbar_arrive(&S.data_empty[i % ND]);
            mbar_arrive(&S.ptr_empty[i % NP]);
        }
        if (t < nr) {
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
                             ? add_rn(mul_rn(sub_rn(1.0, a.omega), xi),
                                      mul_rn(a.omega, div_rn(sub_rn(bi, acc), diag)))
                             : xi;
            } else if (OP == OP_JACOBI_BSR1) {
                y[row] = (diag != 0.0)
                             ? add_rn(mul_rn(sub_rn(1.0, a.omega), xi), div_rn(mul_rn(a.omega, acc), diag))
                             : xi;
            }
        }
    }
}

// rows per tile from the average row length: a tile should fill most of one stage
int pipe_rows_per_tile(long long nnz, int nrows)
{
    if (nnz < 0 || nrows <= 0) return 0;                 // unknown: caller uses csr_stream
    const double avg = (double)nnz / nrows;
    return avg <= 11.0 ? P_ROWS : 0;                     // long rows: sliced-ELL (engine) or csr_stream
}

static int g_sms = 0;

template <int OP, int MODE>
static int launch_pipe(const StreamArgs &a, int rpt, cudaStream_t st)
{
    using Smem = PipeSmemT<MODE>;
    static bool attr_done = false;
    if (!attr_done) {
        B200_CUDA(cudaFuncSetAttribute(csr_pipe_kernel<OP, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)sizeof(Smem)));
        attr_done = true;
    }
    if (g_sms == 0) {
        int dev = 0;
        B200_CUDA(cudaGetDevice(&dev));
        B200_CUDA(cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, dev));
    }
    const int nrows = a.row_end - a.row_begin;
    const int num_tiles = (nrows + rpt - 1) / rpt;
    const int grid = min(num_tiles, g_sms * P_CTAS_PER_SM);
    B200_LAUNCH_PDL(pdl_class(PDL_OTHER) && (long long)(a.row_end - a.row_begin) < PDL_SMALL_ROWS, (csr_pipe_kernel<OP, MODE>), grid, P_TPB, sizeof(Smem), st, a, rpt, num_tiles);
    return check_launch("csr_pipe_kernel");
}

// -> 1 if launched, 0 if this case is not eligible (caller falls back to csr_stream), <0 / >1 never
int pipe_op(int op, const StreamArgs &a, long long nnz, cudaStream_t st, int *rc)
{
    const int nrows = a.row_end - a.row_begin;
    const int rpt = pipe_rows_per_tile(nnz, nrows);
    if (rpt == 0 || nrows <= 0) return 0;
    if (!aligned16(a.Aj) || !aligned16(a.Ax) || !aligned16(a.Ap + a.row_begin)) return 0;
    const bool coded = a.codes != nullptr && a.dict != nullptr && aligned16(a.codes);
    const int mode = !coded ? 0 : (a.offs != nullptr ? 2 : 1);
#define PIPE_CASE(OPC)                                                              \
    case OPC:                                                                       \
        *rc = mode == 2 ? launch_pipe<OPC, 2>(a, rpt, st)                           \
              : mode == 1 ? launch_pipe<OPC, 1>(a, rpt, st) : launch_pipe<OPC, 0>(a, rpt, st); \
        return 1;
    switch (op) {
        PIPE_CASE(OP_MATVEC)
        PIPE_CASE(OP_SPMV_AB)
        PIPE_CASE(OP_RESIDUAL)
        PIPE_CASE(OP_ADD)
        PIPE_CASE(OP_JACOBI)
        PIPE_CASE(OP_JACOBI_BSR1)
    }
#undef PIPE_CASE
    return 0;
}

}  // namespace b200amg
