// CSR "stream" kernels: one pass over (Ap, Aj, Ax) at HBM speed, bit-exact row sums.
//
// One CTA owns `rows_per_cta` (<= 256) consecutive rows and walks their contiguous nonzero range
// in chunks of 2048 entries that all threads stage into shared memory with 128-bit coalesced
// loads.  Every row is then accumulated IN STORAGE ORDER with separately rounded multiply and
// add -- the reference's serial arithmetic (scipy csr_matvec; amg_core jacobi,
// pyamg/amg_core/relaxation.h:201-239) -- so results are identical to the bit, while the
// matrix is read exactly once, fully coalesced, whatever the row length.
// Two schedules, chosen per CTA from its own rows (uniform branch):
//   * row-parallel (short rows, <= 12 nnz/row on average: level-0 stencils, prolongators):
//     raw (Aj, Ax) are staged; thread t then walks row t, gathering x[Aj].  Neighbouring rows of
//     a banded matrix hit neighbouring x entries, so a warp's gather touches 2-3 cache lines
//     instead of ~14 and L1TEX stops being the limiter (ncu, profiles/r01).
//   * entry-parallel (long rows: coarse operators, restriction): all threads gather x and stage
//     the rounded PRODUCTS; thread t folds row t's products from shared memory.
#include "internal.cuh"

namespace b200amg {

char g_err[512] = "";
long long g_launches = 0;

constexpr int TPB = 256;
constexpr int CHUNK = 2048;
constexpr int ROWPAR_MAX_AVG = 12;

template <int OP, bool VEC>
__global__ void __launch_bounds__(TPB) csr_stream_kernel(const StreamArgs a, const int rows_per_cta)
{
    constexpr bool SMOOTH = (OP >= OP_JACOBI);
    constexpr bool BSRF = (OP == OP_JACOBI_BSR1);
    __shared__ int s_ptr[TPB + 1];
    __shared__ __align__(16) double s_p[CHUNK];
    __shared__ __align__(16) int s_c[CHUNK];

    const int t = threadIdx.x;
    pdl_enter();
    const int rpb = rows_per_cta;
    const int r0 = a.row_begin + blockIdx.x * rpb;
    const int nr = min(rpb, a.row_end - r0);
    for (int q = t; q <= nr; q += TPB) s_ptr[q] = a.Ap[r0 + q];
    __syncthreads();
    const int s = s_ptr[0], e = s_ptr[nr];
    const bool rowpar = (e - s) <= ROWPAR_MAX_AVG * nr;
    int lo = 0, hi = 0;
    const int row = r0 + t;
    double acc = 0.0, diag = 0.0;
    if (t < nr) {
        lo = s_ptr[t];
        hi = s_ptr[t + 1];
        if (BSRF) acc = a.b[row];
    }
    const double *__restrict__ Ax = a.Ax;
    const int *__restrict__ Aj = a.Aj;
    const double *__restrict__ xg = a.x;

    for (int base = s & ~3; base < e; base += CHUNK) {
        if (rowpar) {
#pragma unroll
            for (int u = 0; u < CHUNK / (4 * TPB); ++u) {
                const int off = (u * TPB + t) * 4;
                const int k = base + off;
                if (VEC && k + 4 <= e) {
                    *reinterpret_cast<int4 *>(&s_c[off]) = __ldg(reinterpret_cast<const int4 *>(Aj + k));
                    *reinterpret_cast<double2 *>(&s_p[off]) = __ldg(reinterpret_cast<const double2 *>(Ax + k));
                    *reinterpret_cast<double2 *>(&s_p[off + 2]) = __ldg(reinterpret_cast<const double2 *>(Ax + k + 2));
                } else {
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        if (k + q < e) { s_c[off + q] = Aj[k + q]; s_p[off + q] = Ax[k + q]; }
                }
            }
            __syncthreads();
            if (t < nr) {
                const int k0 = max(lo, base) - base;
                const int k1 = min(hi, base + CHUNK) - base;
#pragma unroll 4
                for (int k = k0; k < k1; ++k) {
                    const int c = s_c[k];
                    const double v = s_p[k];
                    const double pr = mul_rn(v, __ldg(xg + c));
                    if (SMOOTH) {
                        if (c == row) diag = v;                            // last stored diagonal wins
                        else acc = BSRF ? sub_rn(acc, pr) : add_rn(acc, pr);
                    } else {
                        acc = add_rn(acc, pr);
                    }
                }
            }
        } else {
#pragma unroll
            for (int u = 0; u < CHUNK / (4 * TPB); ++u) {
                const int off = (u * TPB + t) * 4;
                const int k = base + off;
                if (VEC && k >= s && k + 4 <= e) {
                    const int4 c = __ldg(reinterpret_cast<const int4 *>(Aj + k));
                    const double2 v0 = __ldg(reinterpret_cast<const double2 *>(Ax + k));
                    const double2 v1 = __ldg(reinterpret_cast<const double2 *>(Ax + k + 2));
                    const double x0 = __ldg(xg + c.x), x1 = __ldg(xg + c.y), x2 = __ldg(xg + c.z), x3 = __ldg(xg + c.w);
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
        y[row] = add_rn(a.b[row], acc);                            // b == y: in place
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
    }
}

// rows per CTA: 256 for short rows; long rows get fewer rows per CTA so that a CTA owns about
// one chunk of entries and the serial per-row folds of different rows run in different CTAs
int rows_per_cta_for(long long nnz, int nrows)
{
    if (nnz <= 0 || nrows <= 0) return TPB;
    const double avg = (double)nnz / nrows;
    if (avg <= ROWPAR_MAX_AVG) return TPB;
    int r = TPB;
    if (avg <= 100.0) return TPB;      // measured: splitting 30-nnz rows over more CTAs only adds prologues
    while (r > 4 && r * avg > 1.25 * CHUNK) r >>= 1;
    return r;
}

template <int OP>
static int launch_stream(const StreamArgs &a0, cudaStream_t st)
{
    const StreamArgs &a = a0;
    const int nrows = a.row_end - a.row_begin;
    if (nrows <= 0) return 0;
    const int rpb = rows_per_cta_for(a.nnz, nrows);
    const int grid = (nrows + rpb - 1) / rpb;
    if (aligned16(a.Aj) && aligned16(a.Ax))
        B200_LAUNCH_PDL(pdl_class(PDL_OTHER) && (long long)(a.row_end - a.row_begin) < PDL_SMALL_ROWS, (csr_stream_kernel<OP, true>), grid, TPB, 0, st, a, rpb);
    else
        B200_LAUNCH_PDL(pdl_class(PDL_OTHER) && (long long)(a.row_end - a.row_begin) < PDL_SMALL_ROWS, (csr_stream_kernel<OP, false>), grid, TPB, 0, st, a, rpb);
    return check_launch("csr_stream_kernel");
}

int stream_op(int op, const StreamArgs &a0, cudaStream_t st)
{
    StreamArgs a = a0;
    if (op == OP_ADD && !a.b) a.b = a.y;              // y += A x: the kernels read the addend through b
    int rc = 0;
    if (a.nnz >= 0 && pipe_op(op, a, a.nnz, st, &rc)) return rc;
    switch (op) {
    case OP_MATVEC: return launch_stream<OP_MATVEC>(a, st);
    case OP_SPMV_AB: return launch_stream<OP_SPMV_AB>(a, st);
    case OP_RESIDUAL: return launch_stream<OP_RESIDUAL>(a, st);
    case OP_ADD: return launch_stream<OP_ADD>(a, st);
    case OP_JACOBI: return launch_stream<OP_JACOBI>(a, st);
    case OP_JACOBI_BSR1: return launch_stream<OP_JACOBI_BSR1>(a, st);
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
    pdl_enter();
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

// ---- nnz of a matrix handed to a stand-alone entry point ------------------------------------
// The amg_core / scipy signatures carry no nnz; tile sizing needs the average row length.  It is
// read once per (Ap, n_row) with a 4-byte D2H copy (never while the stream is being captured)
// and cached.  A stale entry (buffer re-used for another matrix) can only mis-size tiles:
// results do not depend on it.
long long nnz_lookup(const int *Ap, int n_row, cudaStream_t st)
{
    struct Slot { const int *Ap; int n_row; long long nnz; };
    static Slot cache[64];
    static int used = 0, next = 0;
    if (!Ap || n_row <= 0) return -1;
    for (int i = 0; i < used; ++i)
        if (cache[i].Ap == Ap && cache[i].n_row == n_row) return cache[i].nnz;
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(st, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) return -1;
    int ends[2] = {0, 0};
    if (cudaMemcpyAsync(&ends[0], Ap, sizeof(int), cudaMemcpyDeviceToHost, st) != cudaSuccess ||
        cudaMemcpyAsync(&ends[1], Ap + n_row, sizeof(int), cudaMemcpyDeviceToHost, st) != cudaSuccess ||
        cudaStreamSynchronize(st) != cudaSuccess) {
        cudaGetLastError();
        return -1;
    }
    Slot sl{Ap, n_row, (long long)ends[1] - ends[0]};
    cache[next] = sl;
    next = (next + 1) % 64;
    if (used < 64) ++used;
    return sl.nnz;
}

int gs_rows(bool bsrf, const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
            const int *Id, int p0, int p1, cudaStream_t st)
{
    if (p1 <= p0) return 0;
    const int grid = (p1 - p0 + 127) / 128;
    if (bsrf) B200_LAUNCH_PDL(pdl_class(PDL_SWEEP) && p1 - p0 < 64 * 1024, gs_rows_kernel<true>, grid, 128, 0, st, Ap, Aj, Ax, x, b, Id, p0, p1);
    else B200_LAUNCH_PDL(pdl_class(PDL_SWEEP) && p1 - p0 < 64 * 1024, gs_rows_kernel<false>, grid, 128, 0, st, Ap, Aj, Ax, x, b, Id, p0, p1);
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
