// Block relaxation with stored inverses of the diagonal blocks (the reference's SA default smoother):
//   amg_core.block_jacobi         pyamg/amg_core/relaxation.h:661-727
//   amg_core.block_gauss_seidel   pyamg/amg_core/relaxation.h:753-806
// Per block row i:  rsum = sum over OFF-diagonal blocks of (A_ij x_j), each small product formed from zero in column
// order and added to rsum (gemm, linalg.h:359-449);  rsum = b_i - rsum;  v = Dinv_i rsum (from zero, column order);
// Gauss-Seidel: x_i = v in place; Jacobi: x_i = (1 - omega) temp_i + omega v with the products read from temp.
// Every multiply and add separately rounded -- the bits of the serial reference.
// Layout of the work: BS lanes per block row (lane k owns scalar row k of the block row), rsum handed round the
// group by shuffles for the Dinv product; 32 / BS block rows per warp.  Gauss-Seidel runs the rows of one colour in
// parallel (colour-ordered index list, like b200amg_bsr_gauss_seidel) or, without colours, sequentially in one group.
#include "internal.cuh"

namespace b200amg {

static inline cudaStream_t S(void *s) { return reinterpret_cast<cudaStream_t>(s); }

// off-diagonal part of scalar row k of block row i, then b - rsum
template <int BS>
__device__ __forceinline__ double block_row_rhs(const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                const double *__restrict__ Ax, const double *xsrc,
                                                const double *__restrict__ b, int i, int k)
{
    double rs = 0.0;
    const int hi = Ap[i + 1];
    for (int q = Ap[i]; q < hi; ++q) {
        const int j = Aj[q];
        if (j == i) continue;
        const double *blk = Ax + ((size_t)q * BS + k) * BS;
        const double *xb = xsrc + (size_t)j * BS;
        double v = 0.0;
#pragma unroll
        for (int c = 0; c < BS; ++c) v = add_rn(v, mul_rn(blk[c], xb[c]));
        rs = add_rn(rs, v);
    }
    return sub_rn(b[(size_t)i * BS + k], rs);
}

// v_k = sum_c Dinv_i[k][c] * rhs_c, rhs_c living in lane (group base + c); every lane of the warp takes part
template <int BS>
__device__ __forceinline__ double block_apply_dinv(const double *__restrict__ Dinv, bool active, int i, int g, int k, double rhs)
{
    double v = 0.0;
#pragma unroll
    for (int c = 0; c < BS; ++c) {
        const int src = g * BS + c;
        const double rc = __shfl_sync(0xffffffffu, rhs, src < 32 ? src : 0);
        if (active) v = add_rn(v, mul_rn(Dinv[((size_t)i * BS + k) * BS + c], rc));
    }
    return v;
}

// rows Id[p0..p1) (Id == nullptr: rows p0..p1) relaxed independently of each other: one colour of a colour-ordered sweep
template <int BS>
__global__ void __launch_bounds__(128) block_gs_color_kernel(const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                             const double *__restrict__ Ax, double *x,
                                                             const double *__restrict__ b, const double *__restrict__ Dinv,
                                                             const int *__restrict__ Id, int p0, int p1)
{
    constexpr int RPW = 32 / BS;
    const int lane = threadIdx.x & 31, warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int g = lane / BS, k = lane - g * BS;
    const int p = p0 + warp * RPW + g;
    const bool active = g < RPW && p < p1;
    const int i = active ? (Id ? Id[p] : p) : 0;
    const double rhs = active ? block_row_rhs<BS>(Ap, Aj, Ax, x, b, i, k) : 0.0;
    const double v = block_apply_dinv<BS>(Dinv, active, i, g, k, rhs);
    if (active) x[(size_t)i * BS + k] = v;
}

// any order, repeats allowed: positions start, start+step, ... (stop exclusive) of Id (or the rows themselves), one
// group of BS lanes, one row after the other -- the semantic definition (slow)
template <int BS>
__global__ void __launch_bounds__(32) block_gs_sequential_kernel(const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                                 const double *__restrict__ Ax, double *x,
                                                                 const double *__restrict__ b, const double *__restrict__ Dinv,
                                                                 const int *__restrict__ Id, int start, int stop, int step)
{
    const int lane = threadIdx.x & 31;
    const bool active = lane < BS;
    for (int p = start; p != stop; p += step) {
        const int i = Id ? Id[p] : p;
        const double rhs = active ? block_row_rhs<BS>(Ap, Aj, Ax, x, b, i, lane) : 0.0;
        const double v = block_apply_dinv<BS>(Dinv, active, i, 0, active ? lane : 0, rhs);
        __syncwarp();                                   // every lane has read x before the row is rewritten
        if (active) x[(size_t)i * BS + lane] = v;
        __syncwarp();
    }
}

// block rows start + t*step, t = 0..count-1: x <- (1 - omega) temp + omega Dinv (b - offdiag(A) temp)
template <int BS>
__global__ void __launch_bounds__(128) block_jacobi_kernel(const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                           const double *__restrict__ Ax, double *__restrict__ x,
                                                           const double *__restrict__ b, const double *__restrict__ Dinv,
                                                           const double *__restrict__ temp, int start, int step, int count,
                                                           double omega)
{
    constexpr int RPW = 32 / BS;
    const int lane = threadIdx.x & 31, warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int g = lane / BS, k = lane - g * BS;
    const int t = warp * RPW + g;
    const bool active = g < RPW && t < count;
    const int i = active ? start + t * step : 0;
    const double rhs = active ? block_row_rhs<BS>(Ap, Aj, Ax, temp, b, i, k) : 0.0;
    const double v = block_apply_dinv<BS>(Dinv, active, i, g, k, rhs);
    if (active)
        x[(size_t)i * BS + k] = add_rn(mul_rn(sub_rn(1.0, omega), temp[(size_t)i * BS + k]), mul_rn(omega, v));
}

// temp <- x on the block rows of the sweep (relaxation.h:686-688)
__global__ void __launch_bounds__(256) block_copy_rows_kernel(const double *__restrict__ x, double *__restrict__ temp, int bs,
                                                              int start, int step, long long total)
{
    const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
    if (idx >= total) return;
    const long long t = idx / bs;
    const long long at = ((long long)start + t * step) * bs + (idx - t * bs);
    temp[at] = x[at];
}

// Dinv_i = inverse of the stored diagonal block of block row i (zero block when there is none or when it is
// numerically singular): Gauss-Jordan with partial pivoting on [D | I], one thread per block row
template <int BS>
__global__ void __launch_bounds__(128) block_diag_inverse_kernel(int n_brow, const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                                 const double *__restrict__ Ax, double *__restrict__ Dinv)
{
    const int i = blockIdx.x * 128 + threadIdx.x;
    if (i >= n_brow) return;
    double M[BS][2 * BS];
    long long dptr = -1;
    for (int q = Ap[i]; q < Ap[i + 1]; ++q) if (Aj[q] == i) dptr = q;
    double big = 0.0;
#pragma unroll
    for (int r = 0; r < BS; ++r)
#pragma unroll
        for (int c = 0; c < BS; ++c) {
            const double a = dptr >= 0 ? Ax[((size_t)dptr * BS + r) * BS + c] : 0.0;
            M[r][c] = a;
            M[r][BS + c] = r == c ? 1.0 : 0.0;
            big = fmax(big, fabs(a));
        }
    bool singular = dptr < 0 || big == 0.0;
    const double tiny = big * 2.220446049250313e-16 * 1e6;       // the cut-off the reference hands to gelss (eps * 1e6)
#pragma unroll
    for (int col = 0; col < BS; ++col) {
        int piv = col;
        double best = fabs(M[col][col]);
#pragma unroll
        for (int r = col + 1; r < BS; ++r) if (fabs(M[r][col]) > best) { best = fabs(M[r][col]); piv = r; }
        if (!(best > tiny)) { singular = true; break; }
#pragma unroll
        for (int c = 0; c < 2 * BS; ++c) {
            // dynamic row index: pick row `piv` without indexing the register array dynamically
            double pv = 0.0;
#pragma unroll
            for (int r = 0; r < BS; ++r) if (r == piv) pv = M[r][c];
            const double cur = M[col][c];
#pragma unroll
            for (int r = 0; r < BS; ++r) if (r == piv) M[r][c] = cur;
            M[col][c] = pv;
        }
        const double inv = 1.0 / M[col][col];
#pragma unroll
        for (int c = 0; c < 2 * BS; ++c) M[col][c] *= inv;
#pragma unroll
        for (int r = 0; r < BS; ++r) {
            if (r == col) continue;
            const double f = M[r][col];
#pragma unroll
            for (int c = 0; c < 2 * BS; ++c) M[r][c] -= f * M[col][c];
        }
    }
#pragma unroll
    for (int r = 0; r < BS; ++r)
#pragma unroll
        for (int c = 0; c < BS; ++c) Dinv[((size_t)i * BS + r) * BS + c] = singular ? 0.0 : M[r][BS + c];
}

#define BLOCK_BS_SWITCH(bs, CALL)                                             \
    switch (bs) {                                                             \
    case 1: { constexpr int BS = 1; CALL; } break;                            \
    case 2: { constexpr int BS = 2; CALL; } break;                            \
    case 3: { constexpr int BS = 3; CALL; } break;                            \
    case 4: { constexpr int BS = 4; CALL; } break;                            \
    case 5: { constexpr int BS = 5; CALL; } break;                            \
    case 6: { constexpr int BS = 6; CALL; } break;                            \
    case 7: { constexpr int BS = 7; CALL; } break;                            \
    case 8: { constexpr int BS = 8; CALL; } break;                            \
    default: return set_err(B200AMG_EINVAL, "blocksize must be in 1..8");     \
    }

static int sweep_count(int start, int stop, int step)
{
    if (step == 0) return -1;
    const long long span = (long long)stop - start;
    if (span == 0) return 0;
    if ((span > 0) != (step > 0) || span % step != 0) return -1;     // the reference's `i != row_stop` loop would not end
    return (int)(span / step);
}

int block_jacobi_pp(int bs, int nb, const int *Ap, const int *Aj, const double *Ax, const double *xin, double *xout,
                    const double *b, const double *Dinv, double omega, cudaStream_t st)
{
    if (nb <= 0) return 0;
    BLOCK_BS_SWITCH(bs, {
        const int rows_per_cta = (32 / BS) * 4;
        B200_LAUNCH(block_jacobi_kernel<BS>, (unsigned)((nb + rows_per_cta - 1) / rows_per_cta), 128, 0, st, Ap, Aj, Ax, xout, b,
                    Dinv, xin, 0, 1, nb, omega);
    });
    return check_launch("block_jacobi_kernel");
}

int block_gs_color(int bs, const int *Ap, const int *Aj, const double *Ax, double *x, const double *b, const double *Dinv,
                   const int *Id, int p0, int p1, cudaStream_t st)
{
    if (p1 <= p0) return 0;
    BLOCK_BS_SWITCH(bs, {
        const int rows_per_cta = (32 / BS) * 4;
        B200_LAUNCH(block_gs_color_kernel<BS>, (unsigned)((p1 - p0 + rows_per_cta - 1) / rows_per_cta), 128, 0, st, Ap, Aj, Ax, x, b,
                    Dinv, Id, p0, p1);
    });
    return check_launch("block_gs_color_kernel");
}

int block_gs_sequential(int bs, const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
                        const double *Dinv, const int *Id, int start, int stop, int step, cudaStream_t st)
{
    if (start == stop) return 0;
    BLOCK_BS_SWITCH(bs, B200_LAUNCH(block_gs_sequential_kernel<BS>, 1, 32, 0, st, Ap, Aj, Ax, x, b, Dinv, Id, start, stop, step));
    return check_launch("block_gs_sequential_kernel");
}

// Gauss-Jordan inverses: used for blocks of 7 x 7 and 8 x 8, where the reference itself leaves its own SVD for LAPACK
int block_diag_inverse_gj(int n_brow, int bs, const int *Ap, const int *Aj, const double *Ax, double *Dinv, cudaStream_t st)
{
    if (n_brow <= 0) return 0;
    BLOCK_BS_SWITCH(bs, B200_LAUNCH(block_diag_inverse_kernel<BS>, (n_brow + 127) / 128, 128, 0, st, n_brow, Ap, Aj, Ax, Dinv));
    return check_launch("block_diag_inverse_kernel");
}

}  // namespace b200amg

using namespace b200amg;

extern "C" int b200amg_block_jacobi(const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
                                    const double *Dinv, double *temp, int row_start, int row_stop, int row_step,
                                    double omega, int blocksize, void *stream)
{
    const int count = sweep_count(row_start, row_stop, row_step);
    if (count < 0) return set_err(B200AMG_EINVAL, "block_jacobi: (row_start, row_stop, row_step) is not a finite sweep");
    if (count == 0) return 0;
    const long long total = (long long)count * blocksize;
    B200_LAUNCH(block_copy_rows_kernel, (unsigned)((total + 255) / 256), 256, 0, S(stream), x, temp, blocksize, row_start, row_step,
                total);
    B200_TRY(check_launch("block_copy_rows_kernel"));
    BLOCK_BS_SWITCH(blocksize, {
        const int rows_per_cta = (32 / BS) * 4;
        B200_LAUNCH(block_jacobi_kernel<BS>, (unsigned)((count + rows_per_cta - 1) / rows_per_cta), 128, 0, S(stream), Ap, Aj, Ax, x,
                    b, Dinv, temp, row_start, row_step, count, omega);
    });
    return check_launch("block_jacobi_kernel");
}

extern "C" int b200amg_block_gauss_seidel(const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
                                          const double *Dinv, int row_start, int row_stop, int row_step, int blocksize,
                                          const int *Id, const int *color_ptr_host, int K, void *stream)
{
    const int count = sweep_count(row_start, row_stop, row_step);
    if (count < 0) return set_err(B200AMG_EINVAL, "block_gauss_seidel: (row_start, row_stop, row_step) is not a finite sweep");
    if (count == 0) return 0;
    if (!color_ptr_host)
        return block_gs_sequential(blocksize, Ap, Aj, Ax, x, b, Dinv, Id, row_start, row_stop, row_step, S(stream));
    if (!Id) return set_err(B200AMG_EINVAL, "block_gauss_seidel: colour pointer without an index list");
    // the sweep must cover whole colours: positions [cp[0], cp[K]) forward or backward
    int c0, c1, dir;
    if (row_step == 1 && row_start == color_ptr_host[0] && row_stop == color_ptr_host[K]) { c0 = 0; c1 = K; dir = 1; }
    else if (row_step == -1 && row_start == color_ptr_host[K] - 1 && row_stop == color_ptr_host[0] - 1) { c0 = K - 1; c1 = -1; dir = -1; }
    else return set_err(B200AMG_EINVAL, "block_gauss_seidel: sweep range must cover whole colours");
    for (int c = c0; c != c1; c += dir) {
        B200_TRY(block_gs_color(blocksize, Ap, Aj, Ax, x, b, Dinv, Id, color_ptr_host[c], color_ptr_host[c + 1], S(stream)));
    }
    return 0;
}

extern "C" int b200amg_block_diag_inverse_gj(int n_brow, int blocksize, const int *Ap, const int *Aj, const double *Ax,
                                             double *Dinv, void *stream)
{
    return block_diag_inverse_gj(n_brow, blocksize, Ap, Aj, Ax, Dinv, S(stream));
}
