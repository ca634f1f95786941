// Device-wide exclusive scans and row binning (setup-phase utilities for the Galerkin product,
// the explicit transpose and the colour-grouped matrix copies).
#include "internal.cuh"
#include "setup.cuh"

namespace b200amg {

constexpr int SCAN_TPB = 256;
constexpr int SCAN_ITEMS = 4;
constexpr int SCAN_TILE = SCAN_TPB * SCAN_ITEMS;

template <typename T>
__device__ __forceinline__ T block_exclusive_scan(T v, T *total)
{
    __shared__ T s_w[SCAN_TPB / 32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    T inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        T up = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += up;
    }
    if (lane == 31) s_w[w] = inc;
    __syncthreads();
    if (w == 0) {
        T ws = lane < SCAN_TPB / 32 ? s_w[lane] : 0;
#pragma unroll
        for (int o = 1; o < SCAN_TPB / 32; o <<= 1) {
            T up = __shfl_up_sync(0xffffffffu, ws, o);
            if (lane >= o) ws += up;
        }
        if (lane < SCAN_TPB / 32) s_w[lane] = ws;
    }
    __syncthreads();
    const T base = w ? s_w[w - 1] : 0;
    *total = s_w[SCAN_TPB / 32 - 1];
    __syncthreads();
    return base + inc - v;
}

// phase 1: per-tile local exclusive scan + tile sum
template <typename TI, typename TO>
__global__ void __launch_bounds__(SCAN_TPB) scan_tiles_kernel(long long n, const TI *__restrict__ in, TO *__restrict__ out,
                                                              TO *__restrict__ tile_sums)
{
    const long long base = (long long)blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    TO v[SCAN_ITEMS], s = 0;
#pragma unroll
    for (int q = 0; q < SCAN_ITEMS; ++q) {
        v[q] = (base + q < n) ? (TO)in[base + q] : (TO)0;
        s += v[q];
    }
    TO total;
    TO ex = block_exclusive_scan<TO>(s, &total);
#pragma unroll
    for (int q = 0; q < SCAN_ITEMS; ++q) {
        if (base + q < n) out[base + q] = ex;
        ex += v[q];
    }
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

// phase 3: add scanned tile offsets; element n (the grand total) is written by the last tile
template <typename TO>
__global__ void __launch_bounds__(SCAN_TPB) scan_add_kernel(long long n, TO *__restrict__ out, const TO *__restrict__ tile_off,
                                                            const TO *__restrict__ tile_sums, long long ntiles)
{
    const long long base = (long long)blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    const TO off = tile_off[blockIdx.x];
#pragma unroll
    for (int q = 0; q < SCAN_ITEMS; ++q)
        if (base + q < n) out[base + q] += off;
    if (blockIdx.x == ntiles - 1 && threadIdx.x == 0) out[n] = off + tile_sums[blockIdx.x];
}

template <typename TO>
__global__ void scan_single_total_kernel(TO *out) { out[0] = 0; }

// out[0..n] = exclusive scan of in[0..n) (out[n] = total).  scratch: >= 2*(n/1024+2)*1.01 TO's.
template <typename TI, typename TO>
int exclusive_scan(long long n, const TI *in, TO *out, TO *scratch, cudaStream_t st)
{
    if (n <= 0) {
        scan_single_total_kernel<TO><<<1, 1, 0, st>>>(out);
        return check_launch("scan_single_total_kernel");
    }
    const long long ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    TO *sums = scratch, *offs = scratch + ntiles, *next = scratch + 2 * ntiles + 1;
    scan_tiles_kernel<TI, TO><<<(unsigned)ntiles, SCAN_TPB, 0, st>>>(n, in, out, sums);
    B200_TRY(check_launch("scan_tiles_kernel"));
    B200_TRY((exclusive_scan<TO, TO>(ntiles == 1 ? 0 : ntiles, sums, offs, next, st)));
    if (ntiles == 1) { /* offs[0] = 0 written by the n<=0 branch */ }
    scan_add_kernel<TO><<<(unsigned)ntiles, SCAN_TPB, 0, st>>>(n, out, offs, sums, ntiles);
    return check_launch("scan_add_kernel");
}

template int exclusive_scan<int, int>(long long, const int *, int *, int *, cudaStream_t);
template int exclusive_scan<int, long long>(long long, const int *, long long *, long long *, cudaStream_t);
template int exclusive_scan<long long, long long>(long long, const long long *, long long *, long long *, cudaStream_t);

long long scan_scratch_elems(long long n) { return 2 * ((n + SCAN_TILE - 1) / SCAN_TILE) + 4096; }

// ---- binning of rows by a size measure ------------------------------------------------------
// bin b holds rows with size in (limit[b-1], limit[b]]; bin NBINS-1 is "larger than all limits".
__global__ void __launch_bounds__(256) bin_count_kernel(int n, const long long *__restrict__ size, BinLimits lim, int *counts)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const long long s = size[i];
    if (s <= 0) return;                 // empty rows need no work
    int b = 0;
    while (b < NBINS - 1 && s > lim.v[b]) ++b;
    atomicAdd(&counts[b], 1);
}

__global__ void __launch_bounds__(256) bin_fill_kernel(int n, const long long *__restrict__ size, BinLimits lim,
                                                       const int *__restrict__ offsets, int *cursor, int *perm)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const long long s = size[i];
    if (s <= 0) return;
    int b = 0;
    while (b < NBINS - 1 && s > lim.v[b]) ++b;
    perm[offsets[b] + atomicAdd(&cursor[b], 1)] = i;
}

// Host-synchronous: returns bin offsets (NBINS+1) on the host, row lists in perm (device).
// work: 2*NBINS+2 ints of device scratch.
int bin_rows(int n, const long long *size, const BinLimits &lim, int *perm, int *work, int *offsets_host, cudaStream_t st)
{
    int *counts = work, *cursor = work + NBINS, *d_off = work + 2 * NBINS;
    B200_CUDA(cudaMemsetAsync(work, 0, sizeof(int) * (3 * NBINS + 1), st));
    for (int b = 0; b <= NBINS; ++b) offsets_host[b] = 0;
    if (n <= 0) return 0;
    bin_count_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, size, lim, counts);
    B200_TRY(check_launch("bin_count_kernel"));
    int h[NBINS];
    B200_CUDA(cudaMemcpyAsync(h, counts, sizeof(int) * NBINS, cudaMemcpyDeviceToHost, st));
    B200_CUDA(cudaStreamSynchronize(st));
    for (int b = 0; b < NBINS; ++b) offsets_host[b + 1] = offsets_host[b] + h[b];
    B200_CUDA(cudaMemcpyAsync(d_off, offsets_host, sizeof(int) * (NBINS + 1), cudaMemcpyHostToDevice, st));
    bin_fill_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, size, lim, d_off, cursor, perm);
    return check_launch("bin_fill_kernel");
}

}  // namespace b200amg
