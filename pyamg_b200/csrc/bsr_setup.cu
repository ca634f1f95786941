// Block (BSR) pieces of the smoothed-aggregation setup (SURVEY 8f-2 for BASELINE config 4: 3x3 elasticity
// blocks with 6 rigid-body candidates).  Each kernel restates one step of the reference with its fp64
// operation order, so that the block hierarchy equals the reference-built one bit for bit:
//   fit_candidates        amg_core.fit_candidates, any K1 x K2     pyamg/amg_core/smoothed_aggregation.h:322-448
//   bsr_scale_rows        scipy bsr_scale_rows [+ scalar factor]   pyamg/util/utils.py:133-200, smooth.py:164
//   bsr_sub               scipy bsr_binop_bsr (canonical / general) for P = T - U, pyamg/aggregation/smooth.py:203-206
//   bsr_matmat            scipy bsr_matmat for D^-1 A T and (R A) P, pyamg/aggregation/aggregation.py:429:
//                         blocks come out in FIRST-TOUCH order, zero blocks are kept, each block accumulates its
//                         products in visiting order with scipy's gemm inner order (dot = C[r,c]; dot += A[r,n] B[n,c])
//   bsr_gather_blocks     block permutation (+ per-block transpose) behind bsr transpose / sort_indices
#include "internal.cuh"

namespace b200amg {

static inline cudaStream_t S(void *s) { return reinterpret_cast<cudaStream_t>(s); }

// ---- bsr_scale_rows: out[blk][r][c] = (Ax[blk][r][c] * v[R*i + r]) [* c] -- one warp per block row ----------
__global__ void __launch_bounds__(256) bsr_scale_rows_kernel(int n_brow, int R, int C, const int *__restrict__ Ap,
                                                             const double *__restrict__ Ax, const double *__restrict__ v,
                                                             double c, int use_c, double *__restrict__ out)
{
    const int i = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= n_brow) return;
    const int RC = R * C;
    const long long lo = (long long)Ap[i] * RC, hi = (long long)Ap[i + 1] * RC;
    for (long long k = lo + lane; k < hi; k += 32) {
        const int r = (int)((k - lo) % RC) / C;
        double a = mul_rn(Ax[k], v[(long long)i * R + r]);
        if (use_c) a = mul_rn(a, c);
        out[k] = a;
    }
}

// ---- block gather: out block q = in block perm[q] (R x C row-major), transposed to C x R when `tr` ----------
// perm arrives as doubles: it is the value array of a csr_transpose run over block numbers 0..nblk-1
__global__ void __launch_bounds__(256) bsr_gather_blocks_kernel(long long total, int R, int C, const double *__restrict__ perm,
                                                                const double *__restrict__ in, double *__restrict__ out, int tr)
{
    const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
    if (idx >= total) return;
    const int RC = R * C;
    const long long q = idx / RC;
    const int e = (int)(idx - q * RC);
    const long long p = (long long)perm[q];
    // e addresses the OUTPUT block: (c, r) of a C x R block when transposing, (r, c) otherwise
    const int src = tr ? (e % R) * C + (e / R) : e;
    out[idx] = in[p * RC + src];
}

// ---- fit_candidates: one thread per aggregate, Gram-Schmidt over the K2 candidate columns restricted to the
// aggregate, in place in Qx (member order = CSC order of AggOp: (Mp, Mi)), R factors into Rx[j][K2][K2] --------
__global__ void __launch_bounds__(128) fit_candidates_kernel(int n_agg, int K1, int K2, const int *__restrict__ Mp,
                                                             const int *__restrict__ Mi, const double *__restrict__ B,
                                                             double tol, double *__restrict__ Qx, double *__restrict__ Rx)
{
    const int j = blockIdx.x * 128 + threadIdx.x;
    if (j >= n_agg) return;
    const int BS = K1 * K2;
    const int lo = Mp[j], hi = Mp[j + 1];
    double *Q = Qx + (long long)lo * BS;
    for (int p = lo; p < hi; ++p) {
        const double *src = B + (long long)Mi[p] * BS;
        double *dst = Qx + (long long)p * BS;
        for (int e = 0; e < BS; ++e) dst[e] = src[e];
    }
    double *Rj = Rx + (long long)j * K2 * K2;
    for (int e = 0; e < K2 * K2; ++e) Rj[e] = 0.0;
    const int rows = (hi - lo) * K1;                       // the chunk is a (rows x K2) row-major matrix
    for (int bj = 0; bj < K2; ++bj) {
        double nrm = 0.0;
        for (int t = 0; t < rows; ++t) { const double x = Q[(long long)t * K2 + bj]; nrm = add_rn(nrm, mul_rn(x, x)); }
        nrm = sqrt(nrm);
        const double thresh = mul_rn(tol, nrm);
        for (int bi = 0; bi < bj; ++bi) {
            double d = 0.0;
            for (int t = 0; t < rows; ++t) d = add_rn(d, mul_rn(Q[(long long)t * K2 + bj], Q[(long long)t * K2 + bi]));
            for (int t = 0; t < rows; ++t)
                Q[(long long)t * K2 + bj] = sub_rn(Q[(long long)t * K2 + bj], mul_rn(d, Q[(long long)t * K2 + bi]));
            Rj[K2 * bi + bj] = d;
        }
        nrm = 0.0;
        for (int t = 0; t < rows; ++t) { const double x = Q[(long long)t * K2 + bj]; nrm = add_rn(nrm, mul_rn(x, x)); }
        nrm = sqrt(nrm);
        double scale = 0.0;
        if (nrm > thresh) { scale = div_rn(1.0, nrm); Rj[K2 * bj + bj] = nrm; }
        else Rj[K2 * bj + bj] = 0.0;
        for (int t = 0; t < rows; ++t) Q[(long long)t * K2 + bj] = mul_rn(Q[(long long)t * K2 + bj], scale);
    }
}

// member p of the CSC order is node Mi[p]; its K1 x K2 block goes to block Tp[node] of T (one block per aggregated row)
__global__ void __launch_bounds__(256) tentative_scatter_kernel(long long total, int BS, const int *__restrict__ Mi,
                                                                const int *__restrict__ Tp, const double *__restrict__ Qx,
                                                                double *__restrict__ Tx)
{
    const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
    if (idx >= total) return;
    const long long p = idx / BS;
    const int e = (int)(idx - p * BS);
    Tx[(long long)Tp[Mi[p]] * BS + e] = Qx[idx];
}

// ---- C = A - B on blocks (scipy bsr_binop_bsr), one thread per block row; rows are short (T: one block, U: a few).
// A block whose RC results are all exactly 0.0 is not stored.  pass 0: cnt[i]; pass 1: fill at Cp[i].
// GENERAL: columns threaded in first-touch order (A's row, then B's row), emitted in REVERSE; values are
// (0 + a) - (0 + b) as the dense row buffers of the reference make them.  Otherwise the canonical merge of two
// sorted duplicate-free rows.
__device__ __forceinline__ bool bsr_sub_block(const double *a, const double *b, int RC, bool general, double *out)
{
    bool nz = false;
    for (int e = 0; e < RC; ++e) {
        double av = a ? a[e] : 0.0, bv = b ? b[e] : 0.0;
        if (general) { if (a) av = add_rn(0.0, av); if (b) bv = add_rn(0.0, bv); }
        const double r = sub_rn(av, bv);
        nz |= (r != 0.0);
        if (out) out[e] = r;
    }
    return nz;
}

template <bool GENERAL>
__global__ void __launch_bounds__(128) bsr_sub_kernel(int n_brow, int RC, const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                      const double *__restrict__ Ax, const int *__restrict__ Bp,
                                                      const int *__restrict__ Bj, const double *__restrict__ Bx, int pass,
                                                      int *__restrict__ cnt, const int *__restrict__ Cp, int *__restrict__ Cj,
                                                      double *__restrict__ Cx)
{
    const int i = blockIdx.x * 128 + threadIdx.x;
    if (i >= n_brow) return;
    const int a0 = Ap[i], a1 = Ap[i + 1], b0 = Bp[i], b1 = Bp[i + 1];
    int o = pass ? Cp[i] : 0;
    auto emit = [&](int col, int pa, int pb) {
        const double *a = pa >= 0 ? Ax + (long long)pa * RC : nullptr;
        const double *b = pb >= 0 ? Bx + (long long)pb * RC : nullptr;
        // test first, store second: a dropped block must not touch slot o (it may belong to the next row)
        if (bsr_sub_block(a, b, RC, GENERAL, nullptr)) {
            if (pass) { bsr_sub_block(a, b, RC, GENERAL, Cx + (long long)o * RC); Cj[o] = col; }
            ++o;
        }
    };
    if (GENERAL) {
        for (int q = b1 - 1; q >= b0; --q) {                 // B-only columns, last touched first
            const int col = Bj[q];
            bool inA = false;
            for (int p = a0; p < a1; ++p) inA |= (Aj[p] == col);
            if (!inA) emit(col, -1, q);
        }
        for (int p = a1 - 1; p >= a0; --p) {                 // A's columns, with B's block when present
            const int col = Aj[p];
            int qb = -1;
            for (int q = b0; q < b1; ++q) if (Bj[q] == col) qb = q;
            emit(col, p, qb);
        }
    } else {
        int pa = a0, pb = b0;
        while (pa < a1 || pb < b1) {
            const int ca = pa < a1 ? Aj[pa] : 0x7fffffff, cb = pb < b1 ? Bj[pb] : 0x7fffffff;
            if (ca == cb) { emit(ca, pa, pb); ++pa; ++pb; }
            else if (ca < cb) { emit(ca, pa, -1); ++pa; }
            else { emit(cb, -1, pb); ++pb; }
        }
    }
    if (!pass) cnt[i] = o;
}

// ---- bsr_matmat ----------------------------------------------------------------------------------------------
// The block pattern of C = A B comes from the scalar SpGEMM run on the two block patterns: once column-sorted
// (Sj) and once in scipy's csr order = REVERSE first touch (Lj), both over the same row pointer Cp.  scipy's
// bsr_matmat stores blocks in first-touch order, i.e. slot s of row i holds column Lj[Cp[i+1]-1-s].
// slot_map: for every sorted entry its slot; Cj: the columns in slot order.  One warp per row.
__device__ __forceinline__ int find_sorted(const int *__restrict__ a, int n, int key)
{
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (a[mid] < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__global__ void __launch_bounds__(256) bsr_slot_map_kernel(int n_brow, const int *__restrict__ Cp, const int *__restrict__ Sj,
                                                           const int *__restrict__ Lj, int *__restrict__ slot,
                                                           int *__restrict__ Cj)
{
    const int i = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= n_brow) return;
    const int lo = Cp[i], n = Cp[i + 1] - lo;
    for (int q = lane; q < n; q += 32) {
        const int col = Lj[lo + q];
        const int s = n - 1 - q;
        slot[lo + find_sorted(Sj + lo, n, col)] = s;
        Cj[lo + s] = col;
    }
}

// numeric phase, one warp per block row of A (blocks R x N) times B (blocks N x C): the blocks of A's row are
// visited one after the other (a later one may add to the same output block); inside one visit the blocks of B's
// row hit DIFFERENT output blocks (rows of B hold no duplicate columns), so lanes share (block of B, r, c).
// Cx must be zero on entry.
__global__ void __launch_bounds__(256) bsr_matmat_numeric_kernel(int n_brow, int R, int N, int C, const int *__restrict__ Ap,
                                                                 const int *__restrict__ Aj, const double *__restrict__ Ax,
                                                                 const int *__restrict__ Bp, const int *__restrict__ Bj,
                                                                 const double *__restrict__ Bx, const int *__restrict__ Cp,
                                                                 const int *__restrict__ Sj, const int *__restrict__ slot,
                                                                 double *Cx)
{
    const int i = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= n_brow) return;
    const int RC = R * C;
    const int c_lo = Cp[i], c_n = Cp[i + 1] - c_lo;
    const int a_hi = Ap[i + 1];
    for (int jj = Ap[i]; jj < a_hi; ++jj) {
        const int j = Aj[jj];
        const int b_lo = Bp[j], items = (Bp[j + 1] - b_lo) * RC;
        const double *Ablk = Ax + (long long)jj * R * N;
        for (int it = lane; it < items; it += 32) {
            const int kk = b_lo + it / RC, e = it % RC;
            const int r = e / C, c = e - r * C;
            const int s = slot[c_lo + find_sorted(Sj + c_lo, c_n, Bj[kk])];
            double *dst = Cx + (long long)(c_lo + s) * RC + e;
            const double *Bblk = Bx + (long long)kk * N * C;
            double dot = *dst;
            for (int n = 0; n < N; ++n) dot = add_rn(dot, mul_rn(Ablk[r * N + n], Bblk[n * C + c]));
            *dst = dot;
        }
        __syncwarp();
    }
}

}  // namespace b200amg

using namespace b200amg;

extern "C" int b200amg_bsr_scale_rows(int n_brow, int R, int C, const int *Ap, const double *Ax, const double *v, double c,
                                      int use_c, double *out, void *stream)
{
    if (R < 1 || C < 1) return set_err(B200AMG_EINVAL, "bsr_scale_rows: blocksize < 1");
    if (n_brow <= 0) return 0;
    B200_LAUNCH(bsr_scale_rows_kernel, (n_brow + 7) / 8, 256, 0, S(stream), n_brow, R, C, Ap, Ax, v, c, use_c, out);
    return check_launch("bsr_scale_rows_kernel");
}

extern "C" int b200amg_bsr_gather_blocks(long long n_blocks, int R, int C, const double *perm, const double *in, double *out,
                                         int transpose_blocks, void *stream)
{
    if (R < 1 || C < 1) return set_err(B200AMG_EINVAL, "bsr_gather_blocks: blocksize < 1");
    const long long total = n_blocks * R * C;
    if (total <= 0) return 0;
    B200_LAUNCH(bsr_gather_blocks_kernel, (unsigned)((total + 255) / 256), 256, 0, S(stream), total, R, C, perm, in, out,
                                                                                     transpose_blocks);
    return check_launch("bsr_gather_blocks_kernel");
}

extern "C" int b200amg_fit_candidates(int n_agg, int K1, int K2, const int *Mp, const int *Mi, const double *B, double tol,
                                      double *Qx, double *Rx, void *stream)
{
    if (K1 < 1 || K2 < 1) return set_err(B200AMG_EINVAL, "fit_candidates: K1, K2 must be positive");
    if (n_agg <= 0) return 0;
    B200_LAUNCH(fit_candidates_kernel, (n_agg + 127) / 128, 128, 0, S(stream), n_agg, K1, K2, Mp, Mi, B, tol, Qx, Rx);
    return check_launch("fit_candidates_kernel");
}

extern "C" int b200amg_tentative_scatter(long long n_members, int K1, int K2, const int *Mi, const int *Tp, const double *Qx,
                                         double *Tx, void *stream)
{
    const long long total = n_members * K1 * K2;
    if (total <= 0) return 0;
    B200_LAUNCH(tentative_scatter_kernel, (unsigned)((total + 255) / 256), 256, 0, S(stream), total, K1 * K2, Mi, Tp, Qx, Tx);
    return check_launch("tentative_scatter_kernel");
}

extern "C" int b200amg_bsr_sub(int n_brow, int R, int C, const int *Ap, const int *Aj, const double *Ax, const int *Bp,
                               const int *Bj, const double *Bx, int general, int pass, int *cnt, const int *Cp, int *Cj,
                               double *Cx, void *stream)
{
    if (R < 1 || C < 1) return set_err(B200AMG_EINVAL, "bsr_sub: blocksize < 1");
    if (n_brow <= 0) return 0;
    const int grid = (n_brow + 127) / 128;
    if (general) B200_LAUNCH(bsr_sub_kernel<true>, grid, 128, 0, S(stream), n_brow, R * C, Ap, Aj, Ax, Bp, Bj, Bx, pass, cnt, Cp, Cj, Cx);
    else B200_LAUNCH(bsr_sub_kernel<false>, grid, 128, 0, S(stream), n_brow, R * C, Ap, Aj, Ax, Bp, Bj, Bx, pass, cnt, Cp, Cj, Cx);
    return check_launch("bsr_sub_kernel");
}

extern "C" int b200amg_bsr_slot_map(int n_brow, const int *Cp, const int *Sj, const int *Lj, int *slot, int *Cj, void *stream)
{
    if (n_brow <= 0) return 0;
    B200_LAUNCH(bsr_slot_map_kernel, (n_brow + 7) / 8, 256, 0, S(stream), n_brow, Cp, Sj, Lj, slot, Cj);
    return check_launch("bsr_slot_map_kernel");
}

extern "C" int b200amg_bsr_matmat_numeric(int n_brow, int R, int N, int C, const int *Ap, const int *Aj, const double *Ax,
                                          const int *Bp, const int *Bj, const double *Bx, const int *Cp, const int *Sj,
                                          const int *slot, long long n_out_blocks, double *Cx, void *stream)
{
    if (R < 1 || N < 1 || C < 1) return set_err(B200AMG_EINVAL, "bsr_matmat_numeric: blocksize < 1");
    if (n_brow <= 0 || n_out_blocks <= 0) return 0;
    B200_CUDA(cudaMemsetAsync(Cx, 0, sizeof(double) * (size_t)n_out_blocks * R * C, S(stream)));
    B200_LAUNCH(bsr_matmat_numeric_kernel, (n_brow + 7) / 8, 256, 0, S(stream), n_brow, R, N, C, Ap, Aj, Ax, Bp, Bj, Bx, Cp, Sj, slot, Cx);
    return check_launch("bsr_matmat_numeric_kernel");
}
