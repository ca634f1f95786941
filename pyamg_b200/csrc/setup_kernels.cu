// Small setup-phase kernels around the Galerkin product (smoothed-aggregation prolongator for
// scalar problems).  Each restates one numpy / scipy / amg_core step with the same fp64
// operation order, so the hierarchy the device builds equals the reference-built one:
//   csr_scale_rows      scipy csr_scale_rows (pyamg/util/utils.py:133-200) [+ scalar factor]
//   tentative_scalar    amg_core.fit_candidates with one candidate, K1 = K2 = 1
//                       (pyamg/amg_core/smoothed_aggregation.h:322-448)
//   csr_sub_sorted      scipy csr_binop_csr (canonical) for P = T - U (pyamg/aggregation/smooth.py:203-206)
#include "internal.cuh"

namespace b200amg {

__global__ void __launch_bounds__(256) csr_scale_rows_kernel(int n, const int *__restrict__ Ap, const double *__restrict__ Ax,
                                                             const double *__restrict__ v, double c, int use_c,
                                                             double *__restrict__ out)
{
    // one warp per row
    const int i = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= n) return;
    const double s = v[i];
    const int hi = Ap[i + 1];
    for (int k = Ap[i] + lane; k < hi; k += 32) {
        double a = mul_rn(Ax[k], s);
        if (use_c) a = mul_rn(a, c);
        out[k] = a;
    }
}

// members of aggregate j = Mp[j]..Mp[j+1] of (Mi, Mb) (ascending node index, value b_i).
// norm_j = sqrt(sum b_i^2) in that order; kept if norm_j > tol*norm_j (i.e. norm_j > 0 for
// tol < 1), scale = 1/norm_j; Bc[j] = norm_j (0 if dropped); Tx[node] = b_i * scale.
__global__ void __launch_bounds__(256) tentative_scalar_kernel(int nagg, const int *__restrict__ Mp,
                                                               const int *__restrict__ Mi, const double *__restrict__ Mb,
                                                               double tol, double *__restrict__ Bc,
                                                               double *__restrict__ node_val)
{
    const int j = blockIdx.x * 256 + threadIdx.x;
    if (j >= nagg) return;
    const int lo = Mp[j], hi = Mp[j + 1];
    double s = 0.0;
    for (int p = lo; p < hi; ++p) s = add_rn(s, mul_rn(Mb[p], Mb[p]));
    const double nrm = sqrt(s);
    const double thresh = mul_rn(tol, nrm);
    double scale = 0.0, r = 0.0;
    if (nrm > thresh) { scale = div_rn(1.0, nrm); r = nrm; }
    Bc[j] = r;
    for (int p = lo; p < hi; ++p) node_val[Mi[p]] = mul_rn(Mb[p], scale);
}

// C = A - B for canonical (sorted, duplicate-free) CSR operands; results equal to 0 are not
// stored (scipy csr_binop_csr_canonical).  pass 0: cnt[i]; pass 1: fill at Cp[i].
__global__ void __launch_bounds__(256) csr_sub_sorted_kernel(int n, const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                             const double *__restrict__ Ax, const int *__restrict__ Bp,
                                                             const int *__restrict__ Bj, const double *__restrict__ Bx,
                                                             int pass, int *__restrict__ cnt, const int *__restrict__ Cp,
                                                             int *__restrict__ Cj, double *__restrict__ Cx)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    int pa = Ap[i], pb = Bp[i];
    const int ea = Ap[i + 1], eb = Bp[i + 1];
    int o = pass ? Cp[i] : 0;
    while (pa < ea || pb < eb) {
        int col;
        double r;
        const int ca = pa < ea ? Aj[pa] : 0x7fffffff, cb = pb < eb ? Bj[pb] : 0x7fffffff;
        if (ca == cb) { col = ca; r = sub_rn(Ax[pa], Bx[pb]); ++pa; ++pb; }
        else if (ca < cb) { col = ca; r = sub_rn(Ax[pa], 0.0); ++pa; }
        else { col = cb; r = sub_rn(0.0, Bx[pb]); ++pb; }
        if (r != 0.0) {
            if (pass) { Cj[o] = col; Cx[o] = r; }
            ++o;
        }
    }
    if (!pass) cnt[i] = o;
}


// C = A - B for rows in ANY order without duplicates, restating scipy's csr_binop_csr_general
// (taken by `T - U` when U comes unsorted out of csr_matmat, pyamg/aggregation/smooth.py:203-206):
// columns are threaded on a list in first-touch order (A's row, then B's row) and emitted in
// REVERSE of that order; zero results are not stored.  Rows are short (A: 1 entry, B: a few).
__global__ void __launch_bounds__(256) csr_sub_general_kernel(int n, const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                              const double *__restrict__ Ax, const int *__restrict__ Bp,
                                                              const int *__restrict__ Bj, const double *__restrict__ Bx,
                                                              int pass, int *__restrict__ cnt, const int *__restrict__ Cp,
                                                              int *__restrict__ Cj, double *__restrict__ Cx)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const int a0 = Ap[i], a1 = Ap[i + 1], b0 = Bp[i], b1 = Bp[i + 1];
    int o = pass ? Cp[i] : 0;
    for (int q = b1 - 1; q >= b0; --q) {                 // B-only columns, last touched first
        const int col = Bj[q];
        bool inA = false;
        for (int p = a0; p < a1; ++p) inA |= (Aj[p] == col);
        if (inA) continue;
        const double r = sub_rn(0.0, add_rn(0.0, Bx[q]));
        if (r != 0.0) { if (pass) { Cj[o] = col; Cx[o] = r; } ++o; }
    }
    for (int p = a1 - 1; p >= a0; --p) {                 // A's columns (with B's value if present)
        const int col = Aj[p];
        double bv = 0.0;
        for (int q = b0; q < b1; ++q) if (Bj[q] == col) bv = add_rn(bv, Bx[q]);
        const double r = sub_rn(add_rn(0.0, Ax[p]), bv);
        if (r != 0.0) { if (pass) { Cj[o] = col; Cx[o] = r; } ++o; }
    }
    if (!pass) cnt[i] = o;
}

// T (n x nagg, at most one entry per row): row i has (agg[i], node_val[i]) unless agg[i] < 0.
// Tp must already hold the exclusive scan of (agg >= 0).
__global__ void __launch_bounds__(256) tentative_fill_kernel(int n, const int *__restrict__ agg, const double *__restrict__ node_val,
                                                             const int *__restrict__ Tp, int *__restrict__ Tj, double *__restrict__ Tx)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n || agg[i] < 0) return;
    Tj[Tp[i]] = agg[i];
    Tx[Tp[i]] = node_val[i];
}

__global__ void __launch_bounds__(256) flag_nonneg_kernel(int n, const int *__restrict__ agg, int *__restrict__ flag)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i < n) flag[i] = agg[i] >= 0 ? 1 : 0;
}

}  // namespace b200amg

using namespace b200amg;
static inline cudaStream_t S(void *s) { return reinterpret_cast<cudaStream_t>(s); }

extern "C" int b200amg_csr_scale_rows(int n_row, const int *Ap, const double *Ax, const double *v, double c, int use_c,
                                      double *out, void *stream)
{
    if (n_row <= 0) return 0;
    B200_LAUNCH(csr_scale_rows_kernel, (n_row + 7) / 8, 256, 0, S(stream), n_row, Ap, Ax, v, c, use_c, out);
    return check_launch("csr_scale_rows_kernel");
}

extern "C" int b200amg_tentative_scalar(int n_agg, const int *Mp, const int *Mi, const double *Mb, double tol, double *Bc,
                                        double *node_val, void *stream)
{
    if (n_agg <= 0) return 0;
    B200_LAUNCH(tentative_scalar_kernel, (n_agg + 255) / 256, 256, 0, S(stream), n_agg, Mp, Mi, Mb, tol, Bc, node_val);
    return check_launch("tentative_scalar_kernel");
}

extern "C" int b200amg_flag_nonneg(int n, const int *agg, int *flag, void *stream)
{
    if (n <= 0) return 0;
    B200_LAUNCH(flag_nonneg_kernel, (n + 255) / 256, 256, 0, S(stream), n, agg, flag);
    return check_launch("flag_nonneg_kernel");
}

extern "C" int b200amg_tentative_fill(int n, const int *agg, const double *node_val, const int *Tp, int *Tj, double *Tx,
                                      void *stream)
{
    if (n <= 0) return 0;
    B200_LAUNCH(tentative_fill_kernel, (n + 255) / 256, 256, 0, S(stream), n, agg, node_val, Tp, Tj, Tx);
    return check_launch("tentative_fill_kernel");
}

extern "C" int b200amg_csr_sub_sorted(int n_row, const int *Ap, const int *Aj, const double *Ax, const int *Bp,
                                      const int *Bj, const double *Bx, int pass, int *cnt, const int *Cp, int *Cj,
                                      double *Cx, void *stream)
{
    if (n_row <= 0) return 0;
    B200_LAUNCH(csr_sub_sorted_kernel, (n_row + 255) / 256, 256, 0, S(stream), n_row, Ap, Aj, Ax, Bp, Bj, Bx, pass, cnt, Cp, Cj, Cx);
    return check_launch("csr_sub_sorted_kernel");
}

extern "C" int b200amg_csr_sub_general(int n_row, const int *Ap, const int *Aj, const double *Ax, const int *Bp,
                                       const int *Bj, const double *Bx, int pass, int *cnt, const int *Cp, int *Cj,
                                       double *Cx, void *stream)
{
    if (n_row <= 0) return 0;
    B200_LAUNCH(csr_sub_general_kernel, (n_row + 255) / 256, 256, 0, S(stream), n_row, Ap, Aj, Ax, Bp, Bj, Bx, pass, cnt, Cp, Cj, Cx);
    return check_launch("csr_sub_general_kernel");
}
