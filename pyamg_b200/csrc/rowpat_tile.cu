// Row-pattern sweep with the x tiles, b and the pattern bytes of a chunk staged in shared memory by the TMA engine.
//
// rowpat.cu (thread = row, gathers through L1/L2) is bound by the dependent chain of one step per warp, not by DRAM
// (ncu, profiles/r01b: DRAM 35 %, issue 67 %, ~200 warp instructions per 32 rows).  On a constant-coefficient stencil
// the gathers of a chunk of CH consecutive rows are CONTIGUOUS ranges of x, one per column offset -- x[r0 + off_q,
// r0 + off_q + CH) -- and ranges whose offsets lie within CH of each other merge (7-point 3-D: three segments,
// {-nx*ny}, {-nx .. +nx}, {+nx*ny}).  So:
//   * one producer lane streams, per chunk, those x segments, the chunk's slice of b and its CH pattern bytes into a
//     ring of shared-memory stages with bulk async copies (cp.async.bulk + mbarrier: UBLKCP in SASS), TT_STAGES deep;
//   * 8 consumer warps take a landed stage: thread t owns rows t, t + 256, ... of the chunk, reads its operands with
//     conflict-free 8-byte shared loads (consecutive lanes, consecutive addresses), the stencil's values and segment
//     offsets come from the kernel's parameter bank (no table loads), and the CH / 256 rows of a thread are
//     independent, so their loads and folds interleave;
//   * a row whose pattern is the operator's common pattern with some entries missing (grid boundaries: same offsets,
//     same values, a subsequence) runs the same code with an 8-bit presence mask looked up from its pattern byte --
//     no divergent warps at x-line ends; any other row (e.g. halo columns of a row slab) takes a per-row global path.
// Arithmetic per row is the reference's: storage-order fold, separately rounded multiply and add, the diagonal entry
// skipped by the smoothers, IEEE division (pyamg/amg_core/relaxation.h:201-239, :267-360; scipy csr_matvec) -- the
// results are bit-identical to rowpat.cu / the CSR kernels.
#include <algorithm>
#include <cstdlib>
#include <vector>
#include "internal.cuh"
#include "pipe.cuh"

#ifndef TT_CHUNK
#define TT_CHUNK 1024
#endif
#ifndef TT_STAGES
#define TT_STAGES 4          // most ring stages the plan may use; it takes as many (>= 2) as fit the CTA's shared memory
#endif
#ifndef TT_RESIDENT
#define TT_RESIDENT 2
#endif
#ifndef TT_CONSUMERS
#define TT_CONSUMERS 256
#endif

namespace b200amg {

constexpr int TT_CH = TT_CHUNK;
constexpr int TT_MAX_ST = TT_STAGES;
constexpr int TT_CONS = TT_CONSUMERS;           // consumer threads
constexpr int TT_TPB = TT_CONS + 32;            // + producer warp
constexpr int TT_RPT = TT_CH / TT_CONS;         // rows per consumer thread and chunk
constexpr int TT_MAXSEG = RP_TILE_MAXSEG;
constexpr unsigned TT_NOT_SUBSET = 0x100u;
constexpr int TT_SMEM_LIMIT = 227 * 1024 / TT_RESIDENT - 1024;

struct TileArgs {
    int n, nseg;
    int nst;                                                          // ring stages (2..TT_MAX_ST)
    int xlen;                                                         // entries of x (>= n: halo columns of a row slab)
    int seg_lo[TT_MAXSEG], seg_len[TT_MAXSEG], seg_base[TT_MAXSEG];   // elements; lo relative to the chunk's first row
    int xs_elems;                                                     // doubles of x per stage
    int sbase[RP_TILE_LEN];                                           // x of entry q for chunk row i: xs[sbase[q] + i]
    double val[RP_TILE_LEN];                                          // the common pattern's values, storage order
    const unsigned short *mask;                                       // [256] presence mask per pattern (| TT_NOT_SUBSET)
    const unsigned char *codes;
    const RowPatEntry *table;
    const double *x, *b;
    double *y;
    double omega;
};

// one row outside the common pattern: table and x from global memory (rowpat.cu arithmetic)
template <int OP>
__device__ __noinline__ void tile_generic_row(const TileArgs &a, int row, int code)
{
    constexpr bool SMOOTH = (OP >= OP_JACOBI);
    constexpr bool BSRF = (OP == OP_JACOBI_BSR1);
    constexpr bool NEED_B = (OP == OP_RESIDUAL || SMOOTH);
    const RowPatEntry *te = a.table + (size_t)code * RP_TILE_LEN;
    const int meta = te[0].len, len = meta & 0xff;
    const unsigned dm = (unsigned)meta >> 8;
    const double bi = NEED_B ? a.b[row] : 0.0;
    double acc = BSRF ? bi : 0.0, diag = 0.0;
    for (int q = 0; q < len; ++q) {
        const double v = te[q].v;
        if (SMOOTH && ((dm >> q) & 1u)) { diag = v; continue; }
        const double pr = mul_rn(v, a.x[row + te[q].off]);
        acc = BSRF ? sub_rn(acc, pr) : add_rn(acc, pr);
    }
    if (OP == OP_MATVEC) {
        a.y[row] = acc;
    } else if (OP == OP_RESIDUAL) {
        a.y[row] = sub_rn(bi, acc);
    } else {
        const double xi = a.x[row];
        if (diag == 0.0) a.y[row] = xi;
        else if (OP == OP_JACOBI)
            a.y[row] = add_rn(mul_rn(sub_rn(1.0, a.omega), xi), mul_rn(a.omega, div_rn(sub_rn(bi, acc), diag)));
        else
            a.y[row] = add_rn(mul_rn(sub_rn(1.0, a.omega), xi), div_rn(mul_rn(a.omega, acc), diag));
    }
}

template <int OP, int LEN>
__global__ void __launch_bounds__(TT_TPB, TT_RESIDENT) rowpat_tile_kernel(const __grid_constant__ TileArgs a)
{
    constexpr bool SMOOTH = (OP >= OP_JACOBI);
    constexpr bool BSRF = (OP == OP_JACOBI_BSR1);
    constexpr bool NEED_B = (OP == OP_RESIDUAL || SMOOTH);
    constexpr int QD = LEN / 2;                                  // the diagonal sits in the middle of the common pattern
    extern __shared__ __align__(128) unsigned char tt_smem[];
    // stage s: [x: xs_elems doubles][b: CH doubles][codes: CH bytes]; then mask[256], full[nst], empty[nst]
    const int TT_ST = a.nst;
    const int stage_bytes = (a.xs_elems + (NEED_B ? TT_CH : 0)) * 8 + TT_CH;
    unsigned short *s_mask = reinterpret_cast<unsigned short *>(tt_smem + (size_t)TT_ST * stage_bytes);
    uint64_t *full = reinterpret_cast<uint64_t *>(tt_smem + (size_t)TT_ST * stage_bytes + 512);
    uint64_t *empty = full + TT_ST;
    const int t = threadIdx.x;
    pdl_trigger();
    if (t == 0) {
        for (int s = 0; s < TT_ST; ++s) { mbar_init(&full[s], 2); mbar_init(&empty[s], TT_CONS / 32); }
        mbar_fence_init();
    }
    for (int i = t; i < 256; i += TT_TPB) s_mask[i] = a.mask[i];
    __syncthreads();
    pdl_wait();                 // barriers and masks are set up while the previous kernel drains; x and b wait for it
    const int n_chunks = (a.n + TT_CH - 1) / TT_CH;
    const int my_chunks = (n_chunks - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

    if (t >= TT_CONS) {
        // ---------------- producer: one lane feeds the stage ring ----------------
        if (t != TT_CONS) return;
        int st = 0, round = 0;                              // ring slot and how often it has been filled
        for (int k = 0; k < my_chunks; ++k) {
            if (round > 0) mbar_wait(&empty[st], (round - 1) & 1);
            const long long r0 = (long long)((int)blockIdx.x + k * (int)gridDim.x) * TT_CH;
            unsigned char *sb = tt_smem + (size_t)st * stage_bytes;
            double *xs = reinterpret_cast<double *>(sb);
            uint64_t *bar = &full[st];
            // the byte count must be announced before the copies can complete: clipped ranges twice (cheap, nseg <= 8)
            uint32_t bytes = 0;
            for (int s = 0; s < a.nseg; ++s) {
                const long long g0 = r0 + a.seg_lo[s], g1 = g0 + a.seg_len[s];
                const long long c0 = g0 < 0 ? 0 : g0, c1 = g1 < a.xlen ? g1 : (long long)a.xlen;
                if (c1 > c0) bytes += (uint32_t)((c1 - c0) & ~1ll) * 8u;
            }
            const int nb = (int)(a.n - r0 < TT_CH ? a.n - r0 : TT_CH);          // rows of this chunk
            if (NEED_B) bytes += (uint32_t)(nb & ~1) * 8u;
            const uint32_t code_bytes = (uint32_t)((nb + 15) & ~15);             // the code array is padded by 64 bytes
            bytes += code_bytes;
            mbar_arrive_expect_tx(bar, bytes);
            for (int s = 0; s < a.nseg; ++s) {
                const long long g0 = r0 + a.seg_lo[s], g1 = g0 + a.seg_len[s];
                const long long c0 = g0 < 0 ? 0 : g0, c1 = g1 < a.xlen ? g1 : (long long)a.xlen;
                if (c1 <= c0) continue;
                double *dst = xs + a.seg_base[s] + (c0 - g0);
                const long long cnt = c1 - c0, even = cnt & ~1ll;
                if (even) bulk_g2s(dst, a.x + c0, (uint32_t)even * 8u, bar);
                if (cnt & 1) cp_async_8(dst + even, a.x + c0 + even);            // odd length: the last element on its own
            }
            if (NEED_B) {
                double *bs = xs + a.xs_elems;
                if (nb & ~1) bulk_g2s(bs, a.b + r0, (uint32_t)(nb & ~1) * 8u, bar);
                if (nb & 1) cp_async_8(bs + (nb & ~1), a.b + r0 + (nb & ~1));
            }
            bulk_g2s(sb + (size_t)(a.xs_elems + (NEED_B ? TT_CH : 0)) * 8, a.codes + r0, code_bytes, bar);
            mbar_arrive_after_cp_async(bar);
            if (++st == TT_ST) { st = 0; ++round; }
        }
        return;
    }

    // ---------------- consumers: thread t owns rows t, t + TT_CONS, ... of every chunk ----------------
    const double diag = a.val[QD];
    const double one_m_omega = sub_rn(1.0, a.omega);
    int st = 0, round = 0;
    for (int k = 0; k < my_chunks; ++k, st = (st + 1 == TT_ST ? 0 : st + 1), round += (st == 0)) {
        const int r0 = ((int)blockIdx.x + k * (int)gridDim.x) * TT_CH;
        const unsigned char *sb = tt_smem + (size_t)st * stage_bytes;
        const double *xs = reinterpret_cast<const double *>(sb);
        const double *bs = xs + a.xs_elems;
        const unsigned char *cs = sb + (size_t)(a.xs_elems + (NEED_B ? TT_CH : 0)) * 8;
        mbar_wait(&full[st], round & 1);
        double xv[TT_RPT][LEN], bi[TT_RPT];
        unsigned m[TT_RPT];
        int code[TT_RPT];
#pragma unroll
        for (int j = 0; j < TT_RPT; ++j) {
            const int i = j * TT_CONS + t;
            code[j] = cs[i];
            m[j] = s_mask[code[j]];
#pragma unroll
            for (int q = 0; q < LEN; ++q) xv[j][q] = xs[a.sbase[q] + i];
            bi[j] = NEED_B ? bs[i] : 0.0;
        }
        double out[TT_RPT];
        bool fast[TT_RPT];
#pragma unroll
        for (int j = 0; j < TT_RPT; ++j) {
            const unsigned mm = m[j];
            fast[j] = !(mm & TT_NOT_SUBSET);
            double acc = BSRF ? bi[j] : 0.0;
#pragma unroll
            for (int q = 0; q < LEN; ++q) {
                if (SMOOTH && q == QD) continue;
                if ((mm >> q) & 1u) {
                    const double pr = mul_rn(a.val[q], xv[j][q]);
                    acc = BSRF ? sub_rn(acc, pr) : add_rn(acc, pr);
                }
            }
            if (OP == OP_MATVEC) {
                out[j] = acc;
            } else if (OP == OP_RESIDUAL) {
                out[j] = sub_rn(bi[j], acc);
            } else {
                const double xi = xv[j][QD];
                const bool has_diag = ((mm >> QD) & 1u) && diag != 0.0;
                if (OP == OP_JACOBI)
                    out[j] = has_diag ? add_rn(mul_rn(one_m_omega, xi), mul_rn(a.omega, div_rn(sub_rn(bi[j], acc), diag))) : xi;
                else
                    out[j] = has_diag ? add_rn(mul_rn(one_m_omega, xi), div_rn(mul_rn(a.omega, acc), diag)) : xi;
            }
        }
        // the stage's shared memory is no longer needed: hand it back before the stores
        __syncwarp();
        if ((t & 31) == 0) mbar_arrive(&empty[st]);
#pragma unroll
        for (int j = 0; j < TT_RPT; ++j) {
            const int row = r0 + j * TT_CONS + t;
            if (row < a.n) {
                if (fast[j]) a.y[row] = out[j];
                else tile_generic_row<OP>(a, row, code[j]);
            }
        }
    }
}

// ---- plan: segments and masks from the host copy of the pattern table --------------------------------------------
void rowpat_tile_plan(RowPat *r, const std::vector<RowPatEntry> &ht, int common)
{
    r->tile_ok = false;
    const int L = ht[(size_t)common * RP_TILE_LEN].len & 0xff;
    if (L != 3 && L != 5 && L != 7) return;
    int off[RP_TILE_LEN];
    double val[RP_TILE_LEN];
    for (int q = 0; q < L; ++q) { off[q] = ht[(size_t)common * RP_TILE_LEN + q].off; val[q] = ht[(size_t)common * RP_TILE_LEN + q].v; }
    for (int q = 1; q < L; ++q)
        if (off[q] <= off[q - 1]) return;                             // sorted rows, distinct columns
    if (off[L / 2] != 0) return;                                      // diagonal in the middle
    // segments: ranges [off, off + CH) that touch or overlap merge; starts rounded down / lengths up to 16 bytes
    int nseg = 0, lo[TT_MAXSEG], hi[TT_MAXSEG], seg_of[RP_TILE_LEN];
    for (int q = 0; q < L; ++q) {
        if (nseg && off[q] <= hi[nseg - 1]) {
            hi[nseg - 1] = off[q] + TT_CH;
        } else {
            if (nseg == TT_MAXSEG) return;
            lo[nseg] = off[q] & ~1;
            hi[nseg] = off[q] + TT_CH;
            ++nseg;
        }
        seg_of[q] = nseg - 1;
    }
    int base = 0;
    for (int s = 0; s < nseg; ++s) {
        r->seg_lo[s] = lo[s];
        r->seg_len[s] = (hi[s] - lo[s] + 1) & ~1;
        r->seg_base[s] = base;
        base += r->seg_len[s];
    }
    r->nseg = nseg;
    r->xs_elems = base;
    // as many ring stages as fit (3 for the 7-point stencil on 256-point grid lines, 2 on 512-point lines).  A second
    // launch shape for large stages -- 512 consumers, one CTA per SM, 5 stages -- was measured at 512^3 and is slower
    // (0.863 vs 0.729 ms per Jacobi sweep, profiles/r02/run14_summary.txt): not kept
    int nst = TT_MAX_ST;
    while (nst >= 2 && (long long)nst * ((long long)(base + TT_CH) * 8 + TT_CH) + 512 + 2 * nst * 8 > TT_SMEM_LIMIT) --nst;
    if (nst < 2) return;
    r->tile_stages = nst;
    for (int q = 0; q < L; ++q) {
        r->sbase[q] = r->seg_base[seg_of[q]] + (off[q] - r->seg_lo[seg_of[q]]);
        r->cval[q] = val[q];
    }
    r->tile_len = L;
    // presence masks: pattern p is the common pattern with entries missing (same order, same offsets, same value bits)
    std::vector<unsigned short> mask(256, (unsigned short)TT_NOT_SUBSET);
    for (int p = 0; p < r->npat; ++p) {
        const RowPatEntry *e = &ht[(size_t)p * RP_TILE_LEN];
        const int len = e[0].len & 0xff;
        unsigned mk = 0;
        int c = 0;
        bool ok = true;
        for (int q = 0; q < len && ok; ++q) {
            while (c < L && off[c] != e[q].off) ++c;
            ok = c < L && memcmp(&val[c], &e[q].v, sizeof(double)) == 0;
            if (ok) mk |= 1u << c++;
        }
        if (ok) mask[p] = (unsigned short)mk;
    }
    if (cudaMalloc(&r->tmask, 512) != cudaSuccess) { cudaGetLastError(); r->tmask = nullptr; return; }
    if (cudaMemcpy(r->tmask, mask.data(), 512, cudaMemcpyHostToDevice) != cudaSuccess) {
        cudaGetLastError();
        cudaFree(r->tmask);
        r->tmask = nullptr;
        return;
    }
    r->tile_ok = true;
}

template <int OP, int LEN>
static int launch_tile(const RowPat &r, const StreamArgs &s, cudaStream_t st)
{
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        B200_CUDA(cudaGetDevice(&dev));
        B200_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    }
    constexpr bool NEED_B = (OP == OP_RESIDUAL || OP >= OP_JACOBI);
    const int smem = r.tile_stages * ((r.xs_elems + (NEED_B ? TT_CH : 0)) * 8 + TT_CH) + 512 + 2 * r.tile_stages * 8;
    static bool attr_set = false;
    if (!attr_set) {
        B200_CUDA(cudaFuncSetAttribute(rowpat_tile_kernel<OP, LEN>, cudaFuncAttributeMaxDynamicSharedMemorySize, TT_SMEM_LIMIT));
        attr_set = true;
    }
    TileArgs a;
    a.n = r.n;
    a.nst = r.tile_stages;
    a.xlen = r.xlen;
    a.nseg = r.nseg;
    for (int i = 0; i < TT_MAXSEG; ++i) { a.seg_lo[i] = r.seg_lo[i]; a.seg_len[i] = r.seg_len[i]; a.seg_base[i] = r.seg_base[i]; }
    a.xs_elems = r.xs_elems;
    for (int q = 0; q < RP_TILE_LEN; ++q) { a.sbase[q] = r.sbase[q]; a.val[q] = r.cval[q]; }
    a.mask = r.tmask;
    a.codes = r.codes;
    a.table = r.table;
    a.x = s.x; a.b = s.b; a.y = s.y;
    a.omega = s.omega;
    const int n_chunks = (r.n + TT_CH - 1) / TT_CH;
    const int grid = std::min(n_chunks, sms * TT_RESIDENT);
    B200_LAUNCH_PDL(pdl_class(PDL_OTHER) && (long long)a.n < PDL_SMALL_ROWS, (rowpat_tile_kernel<OP, LEN>), grid, TT_TPB, smem, st, a);
    return check_launch("rowpat_tile_kernel");
}

template <int OP>
static int launch_tile_len(const RowPat &r, const StreamArgs &s, cudaStream_t st)
{
    switch (r.tile_len) {
    case 3: return launch_tile<OP, 3>(r, s, st);
    case 5: return launch_tile<OP, 5>(r, s, st);
    case 7: return launch_tile<OP, 7>(r, s, st);
    }
    return -1;
}

// -1: not eligible (the caller runs rowpat.cu's kernel), else the launch status
int rowpat_tile_op(int op, const RowPat &r, const StreamArgs &a, cudaStream_t st)
{
    if (!r.tile_ok || r.n < 4 * TT_CH) return -1;
    if (!aligned16(a.x) || (a.b && !aligned16(a.b))) return -1;
    if (a.x == a.y) return -1;                                        // out of place only (x is re-read by other CTAs)
    switch (op) {
    case OP_MATVEC: return launch_tile_len<OP_MATVEC>(r, a, st);
    case OP_RESIDUAL: return launch_tile_len<OP_RESIDUAL>(r, a, st);
    case OP_JACOBI: return launch_tile_len<OP_JACOBI>(r, a, st);
    case OP_JACOBI_BSR1: return launch_tile_len<OP_JACOBI_BSR1>(r, a, st);
    }
    return -1;
}

}  // namespace b200amg
