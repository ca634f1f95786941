// CSR "stream" kernels: one pass over (Ap, Aj, Ax) at HBM speed, bit-exact row sums.
//
// One CTA owns 256 consecutive rows and walks their contiguous nonzero range in chunks of
// 2048 entries: (1) all threads stream Aj/Ax with 128-bit coalesced loads, gather x[Aj]
// (L1/L2 hits for banded matrices) and park the rounded products in shared memory;
// (2) thread t then folds row t's products IN STORAGE ORDER with separately rounded adds.
// That is the reference's serial arithmetic (scipy csr_matvec; amg_core jacobi /
// gauss_seidel_indexed, pyamg/amg_core/relaxation.h:201-239, :394-426), so results are
// identical to the bit while the matrix is read once, fully coalesced, regardless of the
// row length (7 nnz/row on level 0, 30-70 on the coarse levels).
#include "internal.cuh"

namespace b200amg {

char g_err[512] = "";
long long g_launches = 0;

constexpr int TPB = 256;
constexpr int CHUNK = 2048;

template <int OP, bool VEC>
__global__ void __launch_bounds__(TPB) csr_stream_kernel(const StreamArgs a)
{
    constexpr bool SMOOTH = (OP >= OP_JACOBI);
    constexpr bool BSRF = (OP == OP_JACOBI_BSR1 || OP == OP_GS_BSR1);
    __shared__ int s_ptr[TPB + 1];
    __shared__ __align__(16) double s_p[CHUNK];
    __shared__ __align__(16) int s_c[SMOOTH ? CHUNK : 4];

    const int t = threadIdx.x;
    const int r0 = a.row_begin + blockIdx.x * TPB;
    const int nr = min(TPB, a.row_end - r0);
    for (int q = t; q <= nr; q += TPB) s_ptr[q] = a.Ap[r0 + q];
    __syncthreads();
    const int s = s_ptr[0], e = s_ptr[nr];
    int lo = 0, hi = 0, row = 0;
    double acc = 0.0, diag = 0.0;
    if (t < nr) {
        lo = s_ptr[t];
        hi = s_ptr[t + 1];
        row = a.map ? a.map[r0 + t] : r0 + t;
        if (BSRF) acc = a.b[row];
    }
    const double *__restrict__ Ax = a.Ax;
    const int *__restrict__ Aj = a.Aj;
    const double *xg = a.x;

    for (int base = s & ~3; base < e; base += CHUNK) {
#pragma unroll
        for (int u = 0; u < CHUNK / (4 * TPB); ++u) {
            const int off = (u * TPB + t) * 4;
            const int k = base + off;
            if (VEC && k >= s && k + 4 <= e) {
                const int4 c = *reinterpret_cast<const int4 *>(Aj + k);
                const double2 v0 = *reinterpret_cast<const double2 *>(Ax + k);
                const double2 v1 = *reinterpret_cast<const double2 *>(Ax + k + 2);
                const double x0 = xg[c.x], x1 = xg[c.y], x2 = xg[c.z], x3 = xg[c.w];
                *reinterpret_cast<double2 *>(&s_p[off]) = make_double2(mul_rn(v0.x, x0), mul_rn(v0.y, x1));
                *reinterpret_cast<double2 *>(&s_p[off + 2]) = make_double2(mul_rn(v1.x, x2), mul_rn(v1.y, x3));
                if (SMOOTH) *reinterpret_cast<int4 *>(&s_c[off]) = c;
            } else {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int kk = k + q;
                    if (kk >= s && kk < e) {
                        const int c = Aj[kk];
                        s_p[off + q] = mul_rn(Ax[kk], xg[c]);
                        if (SMOOTH) s_c[off + q] = c;
                    }
                }
            }
        }
        __syncthreads();
        if (t < nr) {
            const int k0 = max(lo, base) - base;
            const int k1 = min(hi, base + CHUNK) - base;
            for (int k = k0; k < k1; ++k) {
                if (SMOOTH) {
                    if (s_c[k] == row) diag = Ax[base + k];           // last stored diagonal wins
                    else acc = BSRF ? sub_rn(acc, s_p[k]) : add_rn(acc, s_p[k]);
                } else {
                    acc = add_rn(acc, s_p[k]);
                }
            }
        }
        __syncthreads();
    }
    if (t >= nr) return;
    double *y = a.y;
    if (OP == OP_MATVEC) {
        y[row] = acc;
    } else if (OP == OP_SPMV_AB) {
        y[row] = (a.beta == 0.0) ? a.alpha * acc : a.alpha * acc + a.beta * y[row];
    } else if (OP == OP_RESIDUAL) {
        y[row] = sub_rn(a.b[row], acc);
    } else if (OP == OP_ADD) {
        y[row] = add_rn(y[row], acc);
    } else if (OP == OP_JACOBI) {
        const double xi = xg[row];
        y[row] = (diag != 0.0)
                     ? add_rn(mul_rn(sub_rn(1.0, a.omega), xi),
                              mul_rn(a.omega, div_rn(sub_rn(a.b[row], acc), diag)))
                     : xi;
    } else if (OP == OP_JACOBI_BSR1) {
        const double xi = xg[row];
        y[row] = (diag != 0.0)
                     ? add_rn(mul_rn(sub_rn(1.0, a.omega), xi), div_rn(mul_rn(a.omega, acc), diag))
                     : xi;
    } else if (OP == OP_GS) {
        if (diag != 0.0) y[row] = div_rn(sub_rn(a.b[row], acc), diag);
    } else if (OP == OP_GS_BSR1) {
        if (diag != 0.0) y[row] = div_rn(acc, diag);
    }
}

template <int OP>
static int launch_stream(const StreamArgs &a, cudaStream_t st)
{
    const int nrows = a.row_end - a.row_begin;
    if (nrows <= 0) return 0;
    const int grid = (nrows + TPB - 1) / TPB;
    if (aligned16(a.Aj) && aligned16(a.Ax))
        csr_stream_kernel<OP, true><<<grid, TPB, 0, st>>>(a);
    else
        csr_stream_kernel<OP, false><<<grid, TPB, 0, st>>>(a);
    return check_launch("csr_stream_kernel");
}

int stream_op(int op, const StreamArgs &a, cudaStream_t st)
{
    switch (op) {
    case OP_MATVEC: return launch_stream<OP_MATVEC>(a, st);
    case OP_SPMV_AB: return launch_stream<OP_SPMV_AB>(a, st);
    case OP_RESIDUAL: return launch_stream<OP_RESIDUAL>(a, st);
    case OP_ADD: return launch_stream<OP_ADD>(a, st);
    case OP_JACOBI: return launch_stream<OP_JACOBI>(a, st);
    case OP_JACOBI_BSR1: return launch_stream<OP_JACOBI_BSR1>(a, st);
    case OP_GS: return launch_stream<OP_GS>(a, st);
    case OP_GS_BSR1: return launch_stream<OP_GS_BSR1>(a, st);
    }
    return set_err(B200AMG_EINVAL, "stream_op: bad op");
}

// ---- row-gather kernel: rows named by an index list (not contiguous in storage) -----------
// Used by gauss_seidel_indexed on the caller's ORIGINAL matrix (one colour per launch): one
// thread per listed row, storage-order accumulation.  The colour-grouped copy built by the
// cycle engine goes through csr_stream_kernel instead.
template <bool BSRF>
__global__ void __launch_bounds__(128) gs_rows_kernel(const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                      const double *__restrict__ Ax, double *x,
                                                      const double *__restrict__ b,
                                                      const int *__restrict__ Id, int p0, int p1)
{
    const int p = p0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= p1) return;
    const int i = Id[p];
    double acc = BSRF ? b[i] : 0.0, diag = 0.0;
    const int hi = Ap[i + 1];
    for (int k = Ap[i]; k < hi; ++k) {
        const int j = Aj[k];
        const double v = Ax[k];
        if (j == i) diag = v;
        else acc = BSRF ? sub_rn(acc, mul_rn(v, x[j])) : add_rn(acc, mul_rn(v, x[j]));
    }
    if (diag != 0.0) x[i] = BSRF ? div_rn(acc, diag) : div_rn(sub_rn(b[i], acc), diag);
}

// ---- strictly sequential sweep (one warp): reference semantics for ANY index list ----------
// lanes fetch 32 entries of the current row at a time; the sum is folded in storage order.
template <bool BSRF>
__global__ void __launch_bounds__(32) gs_sequential_kernel(const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                           const double *__restrict__ Ax, double *x,
                                                           const double *__restrict__ b, const int *Id,
                                                           int start, int stop, int step)
{
    const int lane = threadIdx.x;
    for (int p = start; p != stop; p += step) {
        const int i = Id ? Id[p] : p;
        const int lo = Ap[i], hi = Ap[i + 1];
        double acc = BSRF ? b[i] : 0.0, diag = 0.0;
        for (int k0 = lo; k0 < hi; k0 += 32) {
            const int k = k0 + lane;
            int j = -1;
            double v = 0.0, pr = 0.0;
            if (k < hi) {
                j = Aj[k];
                v = Ax[k];
                pr = mul_rn(v, __ldcg(x + j));
            }
            const int cnt = min(32, hi - k0);
            for (int q = 0; q < cnt; ++q) {
                const int jq = __shfl_sync(0xffffffffu, j, q);
                const double pq = __shfl_sync(0xffffffffu, pr, q);
                const double vq = __shfl_sync(0xffffffffu, v, q);
                if (jq == i) diag = vq;
                else acc = BSRF ? sub_rn(acc, pq) : add_rn(acc, pq);
            }
        }
        if (lane == 0 && diag != 0.0)
            __stcg(x + i, BSRF ? div_rn(acc, diag) : div_rn(sub_rn(b[i], acc), diag));
        __syncwarp();
    }
}

int gs_rows(bool bsrf, const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
            const int *Id, int p0, int p1, cudaStream_t st)
{
    if (p1 <= p0) return 0;
    const int grid = (p1 - p0 + 127) / 128;
    if (bsrf) gs_rows_kernel<true><<<grid, 128, 0, st>>>(Ap, Aj, Ax, x, b, Id, p0, p1);
    else gs_rows_kernel<false><<<grid, 128, 0, st>>>(Ap, Aj, Ax, x, b, Id, p0, p1);
    return check_launch("gs_rows_kernel");
}

int gs_sequential(bool bsrf, const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
                  const int *Id, int start, int stop, int step, cudaStream_t st)
{
    if (start == stop) return 0;
    if (bsrf) gs_sequential_kernel<true><<<1, 32, 0, st>>>(Ap, Aj, Ax, x, b, Id, start, stop, step);
    else gs_sequential_kernel<false><<<1, 32, 0, st>>>(Ap, Aj, Ax, x, b, Id, start, stop, step);
    return check_launch("gs_sequential_kernel");
}

}  // namespace b200amg
