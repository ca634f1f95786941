// Pseudo-inverses of small dense blocks, one thread per block: amg_core.pinv_array (pyamg/amg_core/linalg.h:888-954)
// over svd_jacobi (linalg.h:455-780, one-sided Jacobi SVD, real case).  The reference inverts the diagonal blocks of
// the block smoothers with it whenever the block is smaller than 7 (pyamg/util/utils.py:601-690, get_block_diag with
// inv_flag=True), so restating it operation for operation makes Dinv -- and with it block_jacobi / block_gauss_seidel
// cycles -- equal to the reference's to the bit.
// THIS FILE IS COMPILED WITH -fmad=false (csrc/Makefile): the arithmetic below is written as plain expressions and must
// not be contracted into FMAs; sqrt and division are IEEE-rounded by default.
#include "internal.cuh"

namespace b200amg {

static inline cudaStream_t S(void *s) { return reinterpret_cast<cudaStream_t>(s); }

__device__ __forceinline__ double pv_dot(const double *x, const double *y, int n)
{
    double s = 0.0;
    for (int i = 0; i < n; ++i) s += x[i] * y[i];
    return s;
}
__device__ __forceinline__ double pv_norm(const double *x, int n) { return sqrt(pv_dot(x, x, n)); }

// column-major m x n (m >= n): U (m x n), V (n x n), S (n); returns like the reference (0 ok, 1 sweep limit)
__device__ int svd_jacobi_dev(const double *A, double *U, double *V, double *Sv, int m, int n)
{
    const double eps = 2.220446049250313e-16;
    if (n == 1 && m == 1) {
        const double normA = fabs(A[0]);
        V[0] = 1.0;
        Sv[0] = normA;
        U[0] = (normA == 0.0) ? 1.0 : A[0] / normA;
        return 0;
    }
    const int nsq = n * n;
    int count = 1, sweep = 0;
    const int sweepmax = 15 * n > 30 ? 15 * n : 30;
    const double tolerance = sqrt((double)m) * eps;
    for (int i = 0; i < nsq; ++i) V[i] = 0.0;
    for (int i = 0; i < nsq; i += n + 1) V[i] = 1.0;
    for (int i = 0; i < m * n; ++i) U[i] = A[i];
    for (int j = 0; j < n; ++j) Sv[j] = eps * pv_norm(U + j * m, m);
    while (count > 0 && sweep <= sweepmax) {
        count = n * (n - 1) / 2;
        int jm = 0, jn = 0;
        for (int j = 0; j < n - 1; ++j) {
            int km = (j + 1) * m, kn = (j + 1) * n;
            for (int k = j + 1; k < n; ++k) {
                const double a = pv_norm(U + jm, m), b = pv_norm(U + km, m);
                const double d = pv_dot(U + jm, U + km, m);
                const double norm_d = fabs(d);
                const double abserr_a = Sv[j], abserr_b = Sv[k];
                const bool sorted = (a >= b), orthog = (norm_d <= tolerance * a * b);
                const bool noisya = (a < abserr_a), noisyb = (b < abserr_b);
                if (sorted && (orthog || noisya || noisyb)) {
                    count--;
                } else if (!sorted || (norm_d == 0.0 && a == b)) {
                    Sv[j] = abserr_b;
                    Sv[k] = abserr_a;
                    for (int i = 0; i < m; ++i) { const double uj = U[jm + i], uk = U[km + i]; U[jm + i] = -uk; U[km + i] = uj; }
                    for (int i = 0; i < n; ++i) { const double vj = V[jn + i], vk = V[kn + i]; V[jn + i] = -vk; V[kn + i] = vj; }
                } else {
                    const double tau = (b * b - a * a) / (2.0 * norm_d);
                    const double t = (tau < 0.0 ? -1.0 : 1.0) / (fabs(tau) + sqrt(1.0 + tau * tau));
                    const double c = 1.0 / (sqrt(1.0 + t * t));
                    const double sn = d * (t * c / norm_d);
                    const double nsn = -sn;
                    const double norm_sin = fabs(sn);
                    Sv[j] = fabs(c) * abserr_a + norm_sin * abserr_b;
                    Sv[k] = norm_sin * abserr_a + fabs(c) * abserr_b;
                    for (int i = 0; i < m; ++i) {
                        const double uj = U[jm + i], uk = U[km + i];
                        U[jm + i] = uj * c + nsn * uk;
                        U[km + i] = sn * uj + uk * c;
                    }
                    for (int i = 0; i < n; ++i) {
                        const double vj = V[jn + i], vk = V[kn + i];
                        V[jn + i] = vj * c + nsn * vk;
                        V[kn + i] = sn * vj + vk * c;
                    }
                }
                km += m;
                kn += n;
            }
            jm += m;
            jn += n;
        }
        sweep++;
    }
    double sigma_tol = 0.0;
    int iszero = n;
    for (int j = 0; j < n; ++j) {
        const double cur = pv_norm(U + j * m, m);
        if (j == 0) { const double alpha = 50.0 / sqrt(sqrt(eps)); sigma_tol = alpha * cur * eps; }
        if (cur <= sigma_tol) {
            iszero--;
            Sv[j] = 0.0;
            for (int i = 0; i < m; ++i) U[j * m + i] = 0.0;
        } else {
            Sv[j] = cur;
            for (int i = 0; i < m; ++i) U[j * m + i] = U[j * m + i] / cur;
        }
    }
    if (iszero == 0) {
        for (int i = 0; i < nsq; ++i) V[i] = 0.0;
        for (int i = 0; i < nsq; i += n + 1) V[i] = 1.0;
        for (int i = 0; i < n * m; i += m + 1) U[i] = 1.0;
        return 0;
    }
    return count > 0 ? 1 : 0;
}

// AA[blk] (n x n, row-major when transA, column-major otherwise) <- its pseudo-inverse, row-major (pinv_array)
template <int N>
__global__ void __launch_bounds__(64) pinv_array_kernel(double *AA, int m, int transA)
{
    const int blk = blockIdx.x * 64 + threadIdx.x;
    if (blk >= m) return;
    constexpr int n = N, nsq = N * N;
    double Tran[nsq], U[nsq], V[nsq], SinvUh[nsq], Sv[N], A[nsq];
    double *dst = AA + (size_t)blk * nsq;
    for (int i = 0; i < nsq; ++i) A[i] = dst[i];
    if (transA) {
        for (int r = 0; r < n; ++r) for (int c = 0; c < n; ++c) Tran[c * n + r] = A[r * n + c];
        svd_jacobi_dev(Tran, U, V, Sv, n, n);
    } else {
        svd_jacobi_dev(A, U, V, Sv, n, n);
    }
    for (int j = 0; j < n; ++j) if (Sv[j] != 0.0) Sv[j] = 1.0 / Sv[j];
    int counter = 0;
    for (int j = 0; j < n; ++j) {                   // Sinv(k) * U(j,k) -> SinvUh(k,j), column major
        int uo = j;
        for (int k = 0; k < n; ++k) { SinvUh[counter] = U[uo] * Sv[k]; counter++; uo += n; }
    }
    for (int r = 0; r < n; ++r) for (int c = 0; c < n; ++c) Tran[c * n + r] = V[r * n + c];
    // A^+ = V * SinvUh: gemm(row-major Tran, column-major SinvUh, row-major result, overwrite), linalg.h:399-420
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) {
            double sum = 0.0;
            for (int k = 0; k < n; ++k) sum += Tran[i * n + k] * SinvUh[j * n + k];
            dst[i * n + j] = sum;
        }
}

// out[i] = the stored diagonal block of block row i, zeros when there is none (get_block_diag, utils.py:655-668)
__global__ void __launch_bounds__(256) block_diag_extract_kernel(long long total, int bs, const int *__restrict__ Ap,
                                                                 const int *__restrict__ Aj, const double *__restrict__ Ax,
                                                                 double *__restrict__ out)
{
    const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
    if (idx >= total) return;
    const int bsq = bs * bs;
    const int i = (int)(idx / bsq), e = (int)(idx - (long long)i * bsq);
    long long dptr = -1;
    for (int q = Ap[i]; q < Ap[i + 1]; ++q) if (Aj[q] == i) dptr = q;
    out[idx] = dptr >= 0 ? Ax[dptr * bsq + e] : 0.0;
}

int pinv_array(double *AA, int m, int n, int transA, cudaStream_t st)
{
    if (m <= 0) return 0;
    const unsigned grid = (unsigned)((m + 63) / 64);
    switch (n) {
    case 1: B200_LAUNCH(pinv_array_kernel<1>, grid, 64, 0, st, AA, m, transA); break;
    case 2: B200_LAUNCH(pinv_array_kernel<2>, grid, 64, 0, st, AA, m, transA); break;
    case 3: B200_LAUNCH(pinv_array_kernel<3>, grid, 64, 0, st, AA, m, transA); break;
    case 4: B200_LAUNCH(pinv_array_kernel<4>, grid, 64, 0, st, AA, m, transA); break;
    case 5: B200_LAUNCH(pinv_array_kernel<5>, grid, 64, 0, st, AA, m, transA); break;
    case 6: B200_LAUNCH(pinv_array_kernel<6>, grid, 64, 0, st, AA, m, transA); break;
    default: return set_err(B200AMG_EINVAL, "pinv_array: block size must be in 1..6 (the reference switches to LAPACK above)");
    }
    return check_launch("pinv_array_kernel");
}

}  // namespace b200amg

using namespace b200amg;

extern "C" int b200amg_pinv_array(double *AA, int m, int n, int transA, void *stream)
{
    return pinv_array(AA, m, n, transA, S(stream));
}

extern "C" int b200amg_block_diag_inverse(int n_brow, int blocksize, const int *Ap, const int *Aj, const double *Ax,
                                          double *Dinv, void *stream)
{
    if (n_brow <= 0) return 0;
    if (blocksize < 1 || blocksize > 8) return set_err(B200AMG_EINVAL, "block_diag_inverse: blocksize must be in 1..8");
    if (blocksize >= 7)                                  // the reference calls LAPACK gelss here: no bit contract
        return block_diag_inverse_gj(n_brow, blocksize, Ap, Aj, Ax, Dinv, S(stream));
    const long long total = (long long)n_brow * blocksize * blocksize;
    B200_LAUNCH(block_diag_extract_kernel, (unsigned)((total + 255) / 256), 256, 0, S(stream), total, blocksize, Ap, Aj, Ax, Dinv);
    B200_TRY(check_launch("block_diag_extract_kernel"));
    return pinv_array(Dinv, n_brow, blocksize, 1, S(stream));
}
