// EXPERIMENT, off by default (B200AMG_ROWPAT_ILP2=1 turns it on; rowpat.cu::rowpat_op): verified bit-identical on the
// host emulation (tests/test_emulated_kernels.py::test_prepared_rowpat_ilp2_kernel_is_bit_identical), compiles to 79-80
// registers without spills, NOT YET RUN ON HARDWARE (written after round 1's GPU minutes were spent).
//
// Row-pattern sweep with TWO rows in flight per thread.  Evidence (profiles/r01b/rejected/README.md): the default
// kernel (rowpat.cu) is bound by the ~2400-cycle dependent chain of one step per warp, not by DRAM.
// Here a thread owns rows r and r + 256 of a chunk at the same time: for warps whose 64 rows all have the operator's
// common row length LEN (the interior stencil) the 2 x LEN gathers are issued back to back before either fold starts;
// other warps (grid boundaries) take the one-row-at-a-time path.  Arithmetic per row is that of rowpat.cu (same order,
// same roundings).
#include <algorithm>
#include "internal.cuh"

#ifndef B200_PREFETCH_L2
#define B200_PREFETCH_L2(p) asm volatile("prefetch.global.L2 [%0];" ::"l"(p))
#endif

namespace b200amg {

constexpr int R2_TPB = 256;
constexpr int R2_CHUNK = 1024;
constexpr int R2_LEN = 8;                                   // table row length (== RP_LEN)
constexpr int R2_CTAS_PER_SM = 3;                           // 85 registers per thread

struct R2Args {
    int n, npat, far;
    const unsigned char *codes;
    const RowPatEntry *table;
    const double *x, *b;
    double *y;
    double omega;
};

template <int LEN>
__device__ __forceinline__ void r2_offsets(const int *po, int (&off)[R2_LEN])
{
    const int4 a = *reinterpret_cast<const int4 *>(po);
    off[0] = a.x; off[1] = a.y; off[2] = a.z; off[3] = a.w;
    if (LEN > 4) {
        const int4 c = *reinterpret_cast<const int4 *>(po + 4);
        off[4] = c.x; off[5] = c.y; off[6] = c.z; off[7] = c.w;
    }
}

template <int LEN>
__device__ __forceinline__ void r2_values(const double *pv, double (&vv)[R2_LEN])
{
#pragma unroll
    for (int q = 0; q < LEN; q += 2) {
        const double2 t = *reinterpret_cast<const double2 *>(pv + q);
        vv[q] = t.x;
        vv[q + 1] = t.y;
    }
}

template <int OP, int LEN>
__device__ __forceinline__ double r2_fold(const double *pv, const double (&xv)[R2_LEN], unsigned dm, double acc)
{
    constexpr bool SMOOTH = (OP >= OP_JACOBI);
    constexpr bool BSRF = (OP == OP_JACOBI_BSR1);
    double vv[R2_LEN];
    r2_values<LEN>(pv, vv);
#pragma unroll
    for (int q = 0; q < LEN; ++q) {
        double pr = mul_rn(vv[q], xv[q]);
        if (SMOOTH && ((dm >> q) & 1u)) pr = 0.0;            // acc + 0.0 == acc bit for bit (rowpat.cu)
        acc = BSRF ? sub_rn(acc, pr) : add_rn(acc, pr);
    }
    return acc;
}

// one row, any length (grid-boundary warps): every slot predicated
template <int OP>
__device__ __forceinline__ double r2_fold_any(const int *po, const double *pv, const double *__restrict__ xg, int row,
                                              unsigned dm, int len, double acc)
{
    constexpr bool SMOOTH = (OP >= OP_JACOBI);
    constexpr bool BSRF = (OP == OP_JACOBI_BSR1);
    int off[R2_LEN];
    double vv[R2_LEN], xv[R2_LEN];
    r2_offsets<R2_LEN>(po, off);
#pragma unroll
    for (int q = 0; q < R2_LEN; ++q)
        if (q < len) xv[q] = __ldg(xg + (row + off[q]));
    r2_values<R2_LEN>(pv, vv);
#pragma unroll
    for (int q = 0; q < R2_LEN; ++q)
        if (q < len) {
            const double pr = mul_rn(vv[q], xv[q]);
            if (!SMOOTH || !((dm >> q) & 1u)) acc = BSRF ? sub_rn(acc, pr) : add_rn(acc, pr);
        }
    return acc;
}

template <int OP>
__device__ __forceinline__ void r2_store(const R2Args &a, int row, double acc, double bi, double xi, unsigned dm,
                                         const double *pv)
{
    constexpr bool SMOOTH = (OP >= OP_JACOBI);
    double diag = 0.0;
    if (SMOOTH && dm) diag = pv[31 - __clz(dm)];                 // last stored diagonal wins
    if (OP == OP_MATVEC) {
        a.y[row] = acc;
    } else if (OP == OP_RESIDUAL) {
        a.y[row] = sub_rn(bi, acc);
    } else if (OP == OP_JACOBI) {
        a.y[row] = (diag != 0.0) ? add_rn(mul_rn(sub_rn(1.0, a.omega), xi), mul_rn(a.omega, div_rn(sub_rn(bi, acc), diag)))
                                 : xi;
    } else if (OP == OP_JACOBI_BSR1) {
        a.y[row] = (diag != 0.0) ? add_rn(mul_rn(sub_rn(1.0, a.omega), xi), div_rn(mul_rn(a.omega, acc), diag)) : xi;
    }
}

template <int OP, int LEN>
__global__ void __launch_bounds__(R2_TPB, R2_CTAS_PER_SM) rowpat_ilp2_kernel(const R2Args a)
{
    constexpr bool SMOOTH = (OP >= OP_JACOBI);
    constexpr bool BSRF = (OP == OP_JACOBI_BSR1);
    constexpr bool NEED_B = (OP == OP_RESIDUAL || SMOOTH);
    extern __shared__ __align__(16) unsigned char r2_smem[];
    int *s_off = reinterpret_cast<int *>(r2_smem);
    double *s_val = reinterpret_cast<double *>(r2_smem + (size_t)a.npat * R2_LEN * sizeof(int));
    int *s_meta = reinterpret_cast<int *>(r2_smem + (size_t)a.npat * R2_LEN * (sizeof(int) + sizeof(double)));
    for (int i = threadIdx.x; i < a.npat * R2_LEN; i += R2_TPB) {
        const RowPatEntry en = a.table[i];
        s_off[i] = en.off;
        s_val[i] = en.v;
        if ((i & (R2_LEN - 1)) == 0) s_meta[i / R2_LEN] = en.len;
    }
    __syncthreads();
    const double *__restrict__ xg = a.x;
    const int n = a.n, far = a.far;
    const int n_chunks = (n + R2_CHUNK - 1) / R2_CHUNK;
    constexpr int PPC = R2_CHUNK / (2 * R2_TPB);                 // row pairs per chunk and thread (2)
    // pair k of this CTA: rows (rA, rA + 256) of chunk blockIdx.x + (k / PPC) * gridDim.x; -1 past the end
    auto pair_row = [&](int k) -> int {
        const int ch = (int)blockIdx.x + (k / PPC) * (int)gridDim.x;
        const int row = ch * R2_CHUNK + (k % PPC) * 2 * R2_TPB + (int)threadIdx.x;
        return (ch < n_chunks && row < n) ? row : -1;
    };
    const int my_pairs = ((n_chunks - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x) * PPC;
    auto code_of = [&](int row) -> int { return (row >= 0 && row < n) ? (int)a.codes[row] : -1; };
    int cA_next = code_of(pair_row(0));
    int cB_next = pair_row(0) >= 0 ? code_of(pair_row(0) + R2_TPB) : -1;
#pragma unroll 1
    for (int k = 0; k < my_pairs; ++k) {
        const int rowA = pair_row(k);
        const int rowB = (rowA >= 0 && rowA + R2_TPB < n) ? rowA + R2_TPB : -1;
        const int codeA = cA_next, codeB = cB_next;
        {
            const int rn = pair_row(k + 1);                       // pattern numbers one pair ahead
            cA_next = code_of(rn);
            cB_next = rn >= 0 ? code_of(rn + R2_TPB) : -1;
            if ((threadIdx.x & 15) == 0 && rn >= 0) {             // next pair's DRAM lines into L2, one lane per line
                if (NEED_B) {
                    B200_PREFETCH_L2(a.b + rn);
                    B200_PREFETCH_L2(a.b + min(rn + R2_TPB, n - 1));
                }
                B200_PREFETCH_L2(xg + min(rn + far, n - 1));
                B200_PREFETCH_L2(xg + min(rn + R2_TPB + far, n - 1));
            }
        }
        const bool liveA = rowA >= 0, liveB = rowB >= 0;
        const int *poA = s_off, *poB = s_off;
        const double *pvA = s_val, *pvB = s_val;
        int lenA = 0, lenB = 0;
        unsigned dmA = 0, dmB = 0;
        double biA = 0.0, biB = 0.0, xiA = 0.0, xiB = 0.0;
        if (liveA) {
            poA = s_off + codeA * R2_LEN; pvA = s_val + codeA * R2_LEN;
            const int meta = s_meta[codeA];
            lenA = meta & 0xff; dmA = (unsigned)meta >> 8;
            if (NEED_B) biA = __ldg(a.b + rowA);
            if (SMOOTH) xiA = __ldg(xg + rowA);
        }
        if (liveB) {
            poB = s_off + codeB * R2_LEN; pvB = s_val + codeB * R2_LEN;
            const int meta = s_meta[codeB];
            lenB = meta & 0xff; dmB = (unsigned)meta >> 8;
            if (NEED_B) biB = __ldg(a.b + rowB);
            if (SMOOTH) xiB = __ldg(xg + rowB);
        }
        double accA = BSRF ? biA : 0.0, accB = BSRF ? biB : 0.0;
        if (__all_sync(0xffffffffu, lenA == LEN && lenB == LEN)) {
            // interior warps: 2 x LEN gathers in flight before the first fold
            int offA[R2_LEN], offB[R2_LEN];
            double xvA[R2_LEN], xvB[R2_LEN];
            r2_offsets<LEN>(poA, offA);
            r2_offsets<LEN>(poB, offB);
#pragma unroll
            for (int q = 0; q < LEN; ++q) xvA[q] = __ldg(xg + (rowA + offA[q]));
#pragma unroll
            for (int q = 0; q < LEN; ++q) xvB[q] = __ldg(xg + (rowB + offB[q]));
            accA = r2_fold<OP, LEN>(pvA, xvA, dmA, accA);
            r2_store<OP>(a, rowA, accA, biA, xiA, dmA, pvA);
            accB = r2_fold<OP, LEN>(pvB, xvB, dmB, accB);
            r2_store<OP>(a, rowB, accB, biB, xiB, dmB, pvB);
        } else {
            accA = r2_fold_any<OP>(poA, pvA, xg, rowA, dmA, lenA, accA);
            if (liveA) r2_store<OP>(a, rowA, accA, biA, xiA, dmA, pvA);
            accB = r2_fold_any<OP>(poB, pvB, xg, rowB, dmB, lenB, accB);
            if (liveB) r2_store<OP>(a, rowB, accB, biB, xiB, dmB, pvB);
        }
    }
}

template <int OP, int LEN>
static int launch_ilp2(const RowPat &r, const StreamArgs &s, cudaStream_t st)
{
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        B200_CUDA(cudaGetDevice(&dev));
        B200_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    }
    const size_t smem = ((sizeof(int) + sizeof(double)) * R2_LEN + sizeof(int)) * (size_t)r.npat + 16;
    const int n_chunks = (r.n + R2_CHUNK - 1) / R2_CHUNK;
    const int grid = std::min(n_chunks, sms * R2_CTAS_PER_SM);
    R2Args a{r.n, r.npat, r.far, r.codes, r.table, s.x, s.b, s.y, s.omega};
    B200_LAUNCH((rowpat_ilp2_kernel<OP, LEN>), grid, R2_TPB, smem, st, a);
    return check_launch("rowpat_ilp2_kernel");
}

// common_len: row length of the operator's typical (middle) row -- 7 for a 3-D 7-point stencil, 5 for a 2-D 5-point one.
// returns -1 when the case is not eligible (the caller runs the default kernel), else the launch status
int rowpat_ilp2_op(int op, const RowPat &r, const StreamArgs &a, int common_len, cudaStream_t st)
{
    if (r.n < 64 * R2_CHUNK) return -1;
#define R2_CASE(OPC)                                                       \
    case OPC:                                                              \
        if (common_len == 7) return launch_ilp2<OPC, 7>(r, a, st);         \
        if (common_len == 5) return launch_ilp2<OPC, 5>(r, a, st);         \
        return -1;
    switch (op) {
        R2_CASE(OP_MATVEC)
        R2_CASE(OP_RESIDUAL)
        R2_CASE(OP_JACOBI)
        R2_CASE(OP_JACOBI_BSR1)
    }
#undef R2_CASE
    return -1;
}

}  // namespace b200amg

#ifdef B200_TEST_HOOKS
// host-emulation entry (tests/test_emulated_kernels.py): the kernel on a caller-built row-pattern table, any size
extern "C" int b200amg_test_rowpat_ilp2(int op, int n, int npat, int far, const unsigned char *codes, const void *table,
                                        int common_len, const double *x, const double *b, double *y, double omega)
{
    using namespace b200amg;
    RowPat r;
    r.codes = const_cast<unsigned char *>(codes);
    r.table = reinterpret_cast<RowPatEntry *>(const_cast<void *>(table));
    r.npat = npat; r.n = n; r.far = far;
    StreamArgs s{};
    s.x = x; s.b = b; s.y = y; s.omega = omega;
#define R2_HOOK(OPC)                                                                 \
    case OPC:                                                                        \
        if (common_len == 7) return launch_ilp2<OPC, 7>(r, s, nullptr);              \
        if (common_len == 5) return launch_ilp2<OPC, 5>(r, s, nullptr);              \
        return -1;
    switch (op) {
        R2_HOOK(OP_MATVEC)
        R2_HOOK(OP_RESIDUAL)
        R2_HOOK(OP_JACOBI)
        R2_HOOK(OP_JACOBI_BSR1)
    }
#undef R2_HOOK
    return -1;
}
#endif
