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
    dot_partial_kernel<<<grid, DOT_TPB, 0, st>>>(n, x, y, work);
    B200_TRY(check_launch("dot_partial_kernel"));
    dot_final_kernel<<<1, DOT_TPB, 0, st>>>(grid, work, out);
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
    scale_diag_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, omega, b, d, x, bsrf);
    return check_launch("scale_diag_kernel");
}

// y = M x for a small dense row-major M: one warp per row
__global__ void __launch_bounds__(256) dense_matvec_kernel(int n, const double *__restrict__ M,
                                                           const double *__restrict__ x, double *__restrict__ y)
{
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
    dense_matvec_kernel<<<(n + 7) / 8, 256, 0, st>>>(n, M, x, y);
    return check_launch("dense_matvec_kernel");
}

}  // namespace b200amg
