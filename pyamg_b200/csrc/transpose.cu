// Explicit transpose B = A^T in CSR (scipy csr_tocsc: R = P.T.tocsr(), pyamg/classical/classical.py:174;
// R = P.H for SA, pyamg/aggregation/aggregation.py:400-401).  Entries of every output row end up
// ascending in the original row index, exactly like the serial counting transpose: entries
// are scattered with atomics as keys (row << 32 | position) and each output row's keys are
// then sorted (bitonic, shared memory) before the (row, value) pairs are emitted.
#include "internal.cuh"
#include "setup.cuh"

namespace b200amg {

typedef unsigned long long u64;

__global__ void __launch_bounds__(256) col_count_kernel(long long nnz, const int *__restrict__ Aj, int *cnt)
{
    const long long k = (long long)blockIdx.x * 256 + threadIdx.x;
    if (k < nnz) atomicAdd(&cnt[Aj[k]], 1);
}

__global__ void __launch_bounds__(256) tr_fill_kernel(int n_row, const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                      const int *__restrict__ Bp, int *cursor, u64 *keys)
{
    // one warp per input row
    const int i = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= n_row) return;
    const int hi = Ap[i + 1];
    for (int k = Ap[i] + lane; k < hi; k += 32) {
        const int j = Aj[k];
        const int slot = Bp[j] + atomicAdd(&cursor[j], 1);
        keys[slot] = ((u64)(unsigned)i << 32) | (unsigned)k;
    }
}

__device__ __forceinline__ void sort_small(u64 *k, int m, int tid, int gs, bool warp)
{
    int npad = 1;
    while (npad < m) npad <<= 1;
    const int half = npad >> 1;
    for (int size = 2; size <= npad; size <<= 1) {
        for (int stride = size >> 1, first = 1; stride >= 1; stride >>= 1, first = 0) {
            for (int t = tid; t < half; t += gs) {
                int i, j;
                if (first) { const int blk = t / stride, off = t - blk * stride; i = blk * size + off; j = blk * size + size - 1 - off; }
                else { i = 2 * stride * (t / stride) + (t % stride); j = i + stride; }
                if (j < m) { const u64 a = k[i], b = k[j]; if (a > b) { k[i] = b; k[j] = a; } }
            }
            if (warp) __syncwarp(); else __syncthreads();
        }
    }
}

// output rows with <= CAP entries: warp per row (CAP <= 128) staged in shared memory
template <int CAP, int WPB>
__global__ void __launch_bounds__(32 * WPB) tr_emit_warp_kernel(const int *__restrict__ rows, int nrows,
                                                                const int *__restrict__ Bp, const u64 *__restrict__ keys,
                                                                const double *__restrict__ Ax, int *__restrict__ Bi,
                                                                double *__restrict__ Bx)
{
    __shared__ u64 s_k[WPB][CAP];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int r = blockIdx.x * WPB + w;
    if (r >= nrows) return;
    const int j = rows[r], lo = Bp[j], m = Bp[j + 1] - lo;
    for (int t = lane; t < m; t += 32) s_k[w][t] = keys[lo + t];
    __syncwarp();
    sort_small(s_k[w], m, lane, 32, true);
    for (int t = lane; t < m; t += 32) {
        const u64 key = s_k[w][t];
        Bi[lo + t] = (int)(key >> 32);
        Bx[lo + t] = Ax[(unsigned)key];
    }
}

// longer output rows: one CTA per row, sort in place in global memory
__global__ void __launch_bounds__(512) tr_emit_cta_kernel(const int *__restrict__ rows, int nrows, const int *__restrict__ Bp,
                                                          u64 *keys, const double *__restrict__ Ax, int *__restrict__ Bi,
                                                          double *__restrict__ Bx)
{
    const int j = rows[blockIdx.x], lo = Bp[j], m = Bp[j + 1] - lo;
    sort_small(keys + lo, m, threadIdx.x, blockDim.x, false);
    for (int t = threadIdx.x; t < m; t += blockDim.x) {
        const u64 key = keys[lo + t];
        Bi[lo + t] = (int)(key >> 32);
        Bx[lo + t] = Ax[(unsigned)key];
    }
}

__global__ void __launch_bounds__(256) row_len_kernel(int n, const int *__restrict__ p, long long *__restrict__ len)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i < n) len[i] = p[i + 1] - p[i];
}

}  // namespace b200amg

using namespace b200amg;

// work layout (bytes): keys u64[nnz] | len i64[n_col] | cnt/cursor int[n_col+1] | perm int[n_col+64] | scan scratch
extern "C" long long b200amg_csr_transpose_work_bytes(int n_row, int n_col, long long nnz)
{
    return 8 * (nnz + 2) + 8 * ((long long)n_col + 2) + 4 * ((long long)n_col + 2) * 2 + 4 * 128 +
           4 * scan_scratch_elems(n_col + 1) + 64;
}

extern "C" int b200amg_csr_transpose(int n_row, int n_col, long long nnz, const int *Ap, const int *Aj, const double *Ax,
                                     int *Bp, int *Bi, double *Bx, void *work, void *stream)
{
    if (n_row < 0 || n_col < 0 || nnz < 0) return set_err(B200AMG_EINVAL, "csr_transpose: negative size");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    unsigned char *w = reinterpret_cast<unsigned char *>(work);
    u64 *keys = reinterpret_cast<u64 *>(w); w += 8 * (nnz + 2);
    long long *len = reinterpret_cast<long long *>(w); w += 8 * ((long long)n_col + 2);
    int *cnt = reinterpret_cast<int *>(w); w += 4 * ((long long)n_col + 2);
    int *perm = reinterpret_cast<int *>(w); w += 4 * ((long long)n_col + 2) + 4 * 128;
    int *scr = reinterpret_cast<int *>(w);
    B200_CUDA(cudaMemsetAsync(cnt, 0, sizeof(int) * ((size_t)n_col + 1), st));
    if (nnz > 0) {
        col_count_kernel<<<(unsigned)((nnz + 255) / 256), 256, 0, st>>>(nnz, Aj, cnt);
        B200_TRY(check_launch("col_count_kernel"));
    }
    B200_TRY((exclusive_scan<int, int>(n_col, cnt, Bp, scr, st)));
    if (nnz == 0 || n_col == 0) return 0;
    B200_CUDA(cudaMemsetAsync(cnt, 0, sizeof(int) * ((size_t)n_col + 1), st));
    tr_fill_kernel<<<(n_row + 7) / 8, 256, 0, st>>>(n_row, Ap, Aj, Bp, cnt, keys);
    B200_TRY(check_launch("tr_fill_kernel"));
    row_len_kernel<<<(n_col + 255) / 256, 256, 0, st>>>(n_col, Bp, len);
    B200_TRY(check_launch("row_len_kernel"));
    int off[NBINS + 1];
    const BinLimits lim{{32, 128, 128, 128, 128}};   // bins 0,1 warp kernels; everything longer -> bin 5
    B200_TRY(bin_rows(n_col, len, lim, perm + 64, perm, off, st));
    int n;
    if ((n = off[1] - off[0]) > 0) {
        tr_emit_warp_kernel<32, 8><<<(n + 7) / 8, 256, 0, st>>>(perm + 64 + off[0], n, Bp, keys, Ax, Bi, Bx);
        B200_TRY(check_launch("tr_emit_warp_kernel<32>"));
    }
    if ((n = off[2] - off[1]) > 0) {
        tr_emit_warp_kernel<128, 8><<<(n + 7) / 8, 256, 0, st>>>(perm + 64 + off[1], n, Bp, keys, Ax, Bi, Bx);
        B200_TRY(check_launch("tr_emit_warp_kernel<128>"));
    }
    if ((n = off[NBINS] - off[2]) > 0) {
        tr_emit_cta_kernel<<<n, 512, 0, st>>>(perm + 64 + off[2], n, Bp, keys, Ax, Bi, Bx);
        B200_TRY(check_launch("tr_emit_cta_kernel"));
    }
    return 0;
}
