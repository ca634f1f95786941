// Symmetric strength of connection on the device: amg_core.symmetric_strength_of_connection,
// pyamg/amg_core/smoothed_aggregation.h:48-99, with the glue of pyamg/strength.py:213-318 around it:
//   * CSR levels: the stored values;
//   * BSR levels (every coarse level of a smoothed-aggregation hierarchy, 1 x 1 blocks included): one value per block
//     = sum of the squared block entries -- the SQUARED Frobenius norm, no root (strength.py:300-305) -- summed in the
//     order numpy's `.sum(axis=1)` uses on a contiguous row (pairwise_sum: eight running sums, then the tail);
//   * an off-diagonal entry is strong when a_ij^2 >= (theta^2 |a_ii|) |a_jj|, the diagonal always is; |a_ii| sums
//     duplicate diagonal entries.
// Only the PATTERN of S reaches the aggregation (aggregate.py:77-85), so the row scaling of strength.py:311-316 is
// not done.  theta == 0 keeps every stored entry and never comes here (aggregation.py uses A's own pattern).
#include "internal.cuh"

namespace b200amg {

static inline cudaStream_t S(void *s) { return reinterpret_cast<cudaStream_t>(s); }

// numpy pairwise_sum for n <= 128 (numpy/core/src/umath/loops_utils.h.src): the order add.reduce uses on a contiguous axis
__device__ __forceinline__ double numpy_row_sum_of_squares(const double *a, int n)
{
    if (n < 8) {
        double res = 0.0;
        for (int i = 0; i < n; ++i) res = add_rn(res, mul_rn(a[i], a[i]));
        return res;
    }
    double r[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = mul_rn(a[j], a[j]);
    int i = 8;
    for (; i < n - (n % 8); i += 8)
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = add_rn(r[j], mul_rn(a[i + j], a[i + j]));
    double res = add_rn(add_rn(add_rn(r[0], r[1]), add_rn(r[2], r[3])), add_rn(add_rn(r[4], r[5]), add_rn(r[6], r[7])));
    for (; i < n; ++i) res = add_rn(res, mul_rn(a[i], a[i]));
    return res;
}

__global__ void __launch_bounds__(256) block_sqnorm_kernel(long long n_blocks, int RC, const double *__restrict__ Ax,
                                                           double *__restrict__ out)
{
    const long long q = (long long)blockIdx.x * 256 + threadIdx.x;
    if (q >= n_blocks) return;
    out[q] = numpy_row_sum_of_squares(Ax + q * RC, RC);
}

// diag[i] = | sum of the stored entries (i, i) |
__global__ void __launch_bounds__(256) abs_diag_kernel(int n, const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                       const double *__restrict__ Ax, double *__restrict__ diag)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    double d = 0.0;
    for (int p = Ap[i]; p < Ap[i + 1]; ++p) if (Aj[p] == i) d = add_rn(d, Ax[p]);
    diag[i] = fabs(d);
}

// pass 0: cnt[i] = strong entries of row i; pass 1: their columns / values at Sp[i], in storage order
__global__ void __launch_bounds__(256) sym_strength_kernel(int n, double theta, const int *__restrict__ Ap,
                                                           const int *__restrict__ Aj, const double *__restrict__ Ax,
                                                           const double *__restrict__ diag, int pass, int *__restrict__ cnt,
                                                           const int *__restrict__ Sp, int *__restrict__ Sj,
                                                           double *__restrict__ Sx)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const double eps_ii = mul_rn(mul_rn(theta, theta), diag[i]);
    int o = pass ? Sp[i] : 0;
    for (int p = Ap[i]; p < Ap[i + 1]; ++p) {
        const int j = Aj[p];
        const double a = Ax[p];
        if (j == i || mul_rn(a, a) >= mul_rn(eps_ii, diag[j])) {
            if (pass) { Sj[o] = j; Sx[o] = a; }
            ++o;
        }
    }
    if (!pass) cnt[i] = o;
}

}  // namespace b200amg

using namespace b200amg;

extern "C" int b200amg_block_sqnorm(long long n_blocks, int R, int C, const double *Ax, double *out, void *stream)
{
    if (R < 1 || C < 1 || R * C > 128) return set_err(B200AMG_EINVAL, "block_sqnorm: block of 1..128 entries expected");
    if (n_blocks <= 0) return 0;
    B200_LAUNCH(block_sqnorm_kernel, (unsigned)((n_blocks + 255) / 256), 256, 0, S(stream), n_blocks, R * C, Ax, out);
    return check_launch("block_sqnorm_kernel");
}

extern "C" int b200amg_abs_diag(int n_row, const int *Ap, const int *Aj, const double *Ax, double *diag, void *stream)
{
    if (n_row <= 0) return 0;
    B200_LAUNCH(abs_diag_kernel, (n_row + 255) / 256, 256, 0, S(stream), n_row, Ap, Aj, Ax, diag);
    return check_launch("abs_diag_kernel");
}

extern "C" int b200amg_symmetric_strength(int n_row, double theta, const int *Ap, const int *Aj, const double *Ax,
                                          const double *diag, int pass, int *cnt, const int *Sp, int *Sj, double *Sx,
                                          void *stream)
{
    if (theta < 0.0) return set_err(B200AMG_EINVAL, "symmetric_strength: expected a positive theta");
    if (n_row <= 0) return 0;
    B200_LAUNCH(sym_strength_kernel, (n_row + 255) / 256, 256, 0, S(stream), n_row, theta, Ap, Aj, Ax, diag, pass, cnt, Sp, Sj, Sx);
    return check_launch("sym_strength_kernel");
}
