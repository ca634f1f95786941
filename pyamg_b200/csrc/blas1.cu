// Vector kernels: deterministic dot / norm, axpy, the zero-guess Jacobi shortcut, diagonal
// extraction, and the dense coarse-grid mat-vec (coarse_grid_solver('pinv2'):
// pyamg/multilevel.py:608-612 applies a cached dense pseudo-inverse with np.dot).
#include "internal.cuh"

namespace b200amg {

constexpr int DOT_TPB = 256;
constexpr int DOT_MAX_GRID = 1024;   // == doubles of `work` the caller must provide

__device__ __forceinline__ double block_sum(double v)
{
    __shared__ double s_w[DOT_TPB / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) s_w[w] = v;
    __syncthreads();
    if (w == 0) {
        v = lane < DOT_TPB / 32 ? s_w[lane] : 0.0;
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    }
    return v;   // valid in thread 0
}

__global__ void __launch_bounds__(DOT_TPB) dot_partial_kernel(long long n, const double *__restrict__ x,
                                                              const double *__restrict__ y, double *__restrict__ part)
{
    pdl_enter();
    double s = 0.0;
    const long long stride = (long long)gridDim.x * DOT_TPB;
    long long i = (long long)blockIdx.x * DOT_TPB + threadIdx.x;
    const long long n2 = n >> 1;
    if (((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) & 15u) == 0) {
        const double2 *x2 = reinterpret_cast<const double2 *>(x), *y2 = reinterpret_cast<const double2 *>(y);
        for (; i < n2; i += stride) {
            const double2 a = x2[i], b = y2[i];
            s += a.x * b.x + a.y * b.y;
        }
        if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) s += x[n - 1] * y[n - 1];
    } else {
        for (; i < n; i += stride) s += x[i] * y[i];
    }
    s = block_sum(s);
    if (threadIdx.x == 0) part[blockIdx.x] = s;
}

__global__ void __launch_bounds__(DOT_TPB) dot_final_kernel(int m, const double *__restrict__ part, double *out)
{
    pdl_enter();
    double s = 0.0;
    for (int i = threadIdx.x; i < m; i += DOT_TPB) s += part[i];
    s = block_sum(s);
    if (threadIdx.x == 0) out[0] = s;
}

int dot(long long n, const double *x, const double *y, double *out, double *work, cudaStream_t st)
{
    if (n < 0) return set_err(B200AMG_EINVAL, "dot: n < 0");
    long long want = (n / 2 + DOT_TPB - 1) / DOT_TPB;
    int grid = (int)(want < 1 ? 1 : (want > DOT_MAX_GRID ? DOT_MAX_GRID : want));
    B200_LAUNCH_PDL(pdl_class(PDL_OTHER) && n < PDL_SMALL_ROWS, dot_partial_kernel, grid, DOT_TPB, 0, st, n, x, y, work);
    B200_TRY(check_launch("dot_partial_kernel"));
    B200_LAUNCH_PDL(pdl_class(PDL_OTHER), dot_final_kernel, 1, DOT_TPB, 0, st, grid, work, out);
    return check_launch("dot_final_kernel");
}

// ---- b200amg_dot_noalloc: the same two kernels with the partial sums in a library-owned array -----------------
// For callers that cannot allocate (CUDA graph capture, serving loops): the DOT_MAX_GRID partial sums live in a
// file-scope device array (8 KB on every GPU that loads the library, held until exit) instead of a caller-provided
// `work`.  In HOST code `g_dot_scratch` is only the symbol's host-side stand-in, never a device address: the address
// the kernels get comes from cudaGetSymbolAddress, looked up once per device with that device current.
// Nothing has to start at zero (dot_partial_kernel overwrites the `grid` cells dot_final_kernel reads); every call
// shares the array, so two calls that overlap on the GPU (different streams, no ordering) race.
constexpr long long DOT_NOALLOC_MAX_N = 1ll << 28;          // 268 435 456 doubles per vector (2 GiB)
constexpr int DOT_NOALLOC_MAX_DEVICES = 64;
static __device__ __align__(256) double g_dot_scratch[DOT_MAX_GRID];
static double *g_dot_scratch_ptr[DOT_NOALLOC_MAX_DEVICES];   // device address of g_dot_scratch, per device ordinal

static int dot_scratch_address(double **out)
{
    int dev = -1;
    B200_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= DOT_NOALLOC_MAX_DEVICES) return set_err(B200AMG_EINVAL, "dot_noalloc: device ordinal out of range");
    if (!g_dot_scratch_ptr[dev]) {
        double *ptr = nullptr;
        B200_CUDA(cudaGetSymbolAddress((void **)&ptr, g_dot_scratch));
        if (!ptr) return set_err(B200AMG_ESTATE, "dot_noalloc: cudaGetSymbolAddress returned a null address");
        g_dot_scratch_ptr[dev] = ptr;
    }
    *out = g_dot_scratch_ptr[dev];
    return 0;
}

int dot_noalloc(long long n, const double *x, const double *y, double *out, cudaStream_t st)
{
    if (n > DOT_NOALLOC_MAX_N)
        return set_err(B200AMG_EINVAL, "b200amg_dot_noalloc: n exceeds the maximum of 268435456 (2^28) elements");
    double *scratch = nullptr;
    B200_TRY(dot_scratch_address(&scratch));
    return dot(n, x, y, out, scratch, st);
}

// ---- preconditioned CG without the host in the loop (pyamg/krylov/_cg.py:109-170) ---------------------------------
// The iteration's scalars live in a small device array `st` (doubles):
//   [0] rz  [1] pAp  [2] rz_new  [3] alpha  [4] beta  [5] tol (absolute, preconditioned norm)  [6] flag  [7] iter
//   [8] maxiter  [9] capacity of hist  [16 + k] hist[k] = sqrt(rz) after iteration k (k = 0: initial)
// flag: 0 running, 1 converged (normr < tol), 2 maxiter reached, -1 indefinite matrix (pAp < 0), -2 indefinite
// preconditioner (rz < 0), -3 singular preconditioner (rz == 0 without convergence).  Once flag != 0 every kernel of
// a later iteration is a no-op, so the host may enqueue iterations ahead and read the flag every few of them.
constexpr int PCG_HIST0 = 16;

// which == 0: after pAp = <Ap, p>; which == 1: after rz_new = <r, z>
__global__ void pcg_scalars_kernel(double *st, int which)
{
    pdl_enter();
    if (st[6] != 0.0) return;
    if (which == 0) {
        if (st[1] < 0.0) { st[6] = -1.0; return; }               // _cg.py:120-122
        st[3] = st[0] / st[1];                                   // alpha = rz / pAp
        return;
    }
    const double rz_new = st[2];
    if (rz_new < 0.0) { st[6] = -2.0; return; }                  // _cg.py:135-137
    st[4] = rz_new / st[0];                                      // beta = rz / rz_old
    st[0] = rz_new;
    const int it = (int)st[7] + 1;
    st[7] = (double)it;
    const double normr = sqrt(rz_new);
    if (it < (int)st[9]) st[PCG_HIST0 + it] = normr;
    if (normr < st[5]) st[6] = 1.0;                              // _cg.py:152-153
    else if (rz_new == 0.0) st[6] = -3.0;                        // :154-159
    else if (it >= (int)st[8]) st[6] = 2.0;                      // :161-162
}

// x += alpha p;  r -= alpha Ap (when update_r != 0: the iterations that do not recompute r = b - A x)
__global__ void __launch_bounds__(256) pcg_update_xr_kernel(long long n, const double *__restrict__ st,
                                                            const double *__restrict__ p, const double *__restrict__ Ap,
                                                            double *x, double *r, int update_r)
{
    pdl_enter();
    if (st[6] != 0.0) return;
    const double alpha = st[3];
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    x[i] = add_rn(x[i], mul_rn(alpha, p[i]));
    if (update_r) r[i] = sub_rn(r[i], mul_rn(alpha, Ap[i]));
}

// p = beta p + z
__global__ void __launch_bounds__(256) pcg_update_p_kernel(long long n, const double *__restrict__ st,
                                                           const double *__restrict__ z, double *p)
{
    pdl_enter();
    if (st[6] != 0.0) return;
    const double beta = st[4];
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i < n) p[i] = add_rn(mul_rn(p[i], beta), z[i]);
}

int pcg_scalars(double *st, int which, cudaStream_t s)
{
    B200_LAUNCH_PDL(pdl_class(PDL_OTHER), pcg_scalars_kernel, 1, 1, 0, s, st, which);
    return check_launch("pcg_scalars_kernel");
}

int pcg_update_xr(long long n, const double *st, const double *p, const double *Ap, double *x, double *r, int update_r,
                  cudaStream_t s)
{
    if (n <= 0) return 0;
    B200_LAUNCH_PDL(pdl_class(PDL_OTHER) && n < PDL_SMALL_ROWS, pcg_update_xr_kernel, (unsigned)((n + 255) / 256), 256, 0, s, n, st, p, Ap, x, r, update_r);
    return check_launch("pcg_update_xr_kernel");
}

int pcg_update_p(long long n, const double *st, const double *z, double *p, cudaStream_t s)
{
    if (n <= 0) return 0;
    B200_LAUNCH_PDL(pdl_class(PDL_OTHER) && n < PDL_SMALL_ROWS, pcg_update_p_kernel, (unsigned)((n + 255) / 256), 256, 0, s, n, st, z, p);
    return check_launch("pcg_update_p_kernel");
}

__global__ void __launch_bounds__(256) axpy_kernel(long long n, double a, const double *__restrict__ x, double *y)
{
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i < n) y[i] = add_rn(mul_rn(a, x[i]), y[i]);
}

int axpy(long long n, double a, const double *x, double *y, cudaStream_t st)
{
    if (n <= 0) return 0;
    axpy_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, a, x, y);
    return check_launch("axpy_kernel");
}

// y = a * x (the polynomial smoother's h = c * residual, pyamg/relaxation/relaxation.py:663-666)
__global__ void __launch_bounds__(256) scal_kernel(long long n, double a, const double *__restrict__ x, double *__restrict__ y)
{
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i < n) y[i] = mul_rn(a, x[i]);
}

int scal(long long n, double a, const double *x, double *y, cudaStream_t st)
{
    if (n <= 0) return 0;
    scal_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, a, x, y);
    return check_launch("scal_kernel");
}

// diag[i] = last stored a_ii of row i (0 when absent): relaxation.h:228-229 semantics
__global__ void __launch_bounds__(256) extract_diag_kernel(int nb, int bs, const int *__restrict__ Ap,
                                                           const int *__restrict__ Aj, const double *__restrict__ Ax,
                                                           double *__restrict__ diag)
{
    const long long g = (long long)blockIdx.x * 256 + threadIdx.x;
    if (g >= (long long)nb * bs) return;
    const int i = (int)(g / bs), k = (int)(g % bs);
    double d = 0.0;
    const int hi = Ap[i + 1];
    for (int p = Ap[i]; p < hi; ++p)
        if (Aj[p] == i) d = Ax[((size_t)p * bs + k) * bs + k];
    diag[g] = d;
}

int extract_diag(int nb, int bs, const int *Ap, const int *Aj, const double *Ax, double *diag, cudaStream_t st)
{
    const long long n = (long long)nb * bs;
    if (n <= 0) return 0;
    extract_diag_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(nb, bs, Ap, Aj, Ax, diag);
    return check_launch("extract_diag_kernel");
}

// One Jacobi sweep from a ZERO initial guess collapses to a diagonal scaling, bit for bit:
// CSR arithmetic  x = (1-w)*0 + w*((b - 0)/d) = w*(b/d) ;  BSR arithmetic  x = (w*b)/d.
// (only valid for scalar blocks; rows with d == 0 keep x = 0)
__global__ void __launch_bounds__(256) scale_diag_kernel(long long n, double omega, const double *__restrict__ b,
                                                         const double *__restrict__ d, double *__restrict__ x, int bsrf)
{
    pdl_enter();
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const double di = d[i];
    double v = 0.0;
    if (di != 0.0) v = bsrf ? div_rn(mul_rn(omega, b[i]), di) : mul_rn(omega, div_rn(b[i], di));
    x[i] = v;
}

int scale_diag(long long n, double omega, const double *b, const double *d, double *x, int bsrf, cudaStream_t st)
{
    if (n <= 0) return 0;
    B200_LAUNCH_PDL(pdl_class(PDL_OTHER) && n < PDL_SMALL_ROWS, scale_diag_kernel, (unsigned)((n + 255) / 256), 256, 0, st, n, omega, b, d, x, bsrf);
    return check_launch("scale_diag_kernel");
}

// y = M x for a small dense row-major M: one warp per row
__global__ void __launch_bounds__(256) dense_matvec_kernel(int n, const double *__restrict__ M,
                                                           const double *__restrict__ x, double *__restrict__ y)
{
    pdl_enter();
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= n) return;
    const double *m = M + (size_t)row * n;
    double s = 0.0;
    for (int c = lane; c < n; c += 32) s += m[c] * x[c];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    if (lane == 0) y[row] = s;
}

int dense_matvec(int n, const double *M, const double *x, double *y, cudaStream_t st)
{
    if (n <= 0) return 0;
    B200_LAUNCH_PDL(pdl_class(PDL_OTHER), dense_matvec_kernel, (n + 7) / 8, 256, 0, st, n, M, x, y);
    return check_launch("dense_matvec_kernel");
}

}  // namespace b200amg
