// TEST INFRASTRUCTURE ONLY -- a stand-in for <cuda_runtime.h> that lets g++ compile a libb200amg kernel source
// (pyamg_b200/csrc/*.cu, unchanged) for the HOST, so that thread-per-row / warp-per-row kernels can be checked
// against scipy in this GPU-less container before they ever see hardware (tests/test_emulated_kernels.py).
// Every CUDA thread of a block is a real OS thread; __syncthreads / __syncwarp are barriers; blocks run one after
// the other.  Supports what the emulated kernels use: barriers, full-warp __shfl_sync / __shfl_xor_sync / __all_sync / __any_sync /
// __ballot_sync, __ldg / __ldcs / __ldcg, int atomicAdd, cudaMalloc / cudaFree / cudaMemcpyAsync as host calls, int4 / double2 loads;
// dynamic shared memory only as a global array the test provides; no static shared memory, no atomics.  The product never
// loads anything built from this header (pyamg_b200 needs CUDA tensors and raises without a device).
#pragma once
#define B200_HOST_EMULATION 1
#include <cstdlib>
#include <algorithm>
#include <atomic>
#include <barrier>
#include <cmath>
#include <cstddef>
#include <cstring>
#include <memory>
#include <thread>
#include <vector>

struct dim3 { unsigned x = 1, y = 1, z = 1; dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {} };
struct uint3emu { unsigned x, y, z; };
inline thread_local uint3emu threadIdx, blockIdx;
inline thread_local dim3 blockDim, gridDim;

struct alignas(16) int4 { int x, y, z, w; };
struct alignas(16) double2 { double x, y; };
using std::max;
using std::min;

typedef int cudaError_t;
typedef void *cudaStream_t;
enum { cudaSuccess = 0 };
inline cudaError_t cudaGetLastError() { return cudaSuccess; }
inline const char *cudaGetErrorString(cudaError_t) { return "emulated"; }
enum cudaDeviceAttr { cudaDevAttrMultiProcessorCount = 16 };
inline cudaError_t cudaGetDevice(int *d) { *d = 0; return cudaSuccess; }
inline cudaError_t cudaDeviceGetAttribute(int *v, cudaDeviceAttr, int) { *v = 2; return cudaSuccess; }   // a 2-SM "device"
inline cudaError_t cudaMemsetAsync(void *p, int v, size_t n, cudaStream_t) { memset(p, v, n); return cudaSuccess; }
// "device" memory is host memory here (layout builders such as sell_build allocate their copies)
template <class T> inline cudaError_t cudaMalloc(T **p, size_t n) { *p = static_cast<T *>(malloc(n ? n : 1)); return *p ? cudaSuccess : 2; }
inline cudaError_t cudaFree(void *p) { free(p); return cudaSuccess; }
enum cudaMemcpyKind { cudaMemcpyHostToHost, cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice };
inline cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, cudaMemcpyKind, cudaStream_t) { memcpy(d, s, n); return cudaSuccess; }
inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __shared__
#define __align__(n) __attribute__((aligned(n)))

// separately rounded arithmetic: the volatile keeps the compiler from contracting or reassociating
inline double __dmul_rn(double a, double b) { volatile double r = a * b; return r; }
inline double __dadd_rn(double a, double b) { volatile double r = a + b; return r; }
inline double __dsub_rn(double a, double b) { volatile double r = a - b; return r; }
inline double __ddiv_rn(double a, double b) { volatile double r = a / b; return r; }

namespace cuda_emu {
struct BlockSync {
    std::barrier<> block;
    std::vector<std::unique_ptr<std::barrier<>>> warp;
    std::vector<unsigned long long> slots;           // one 8-byte exchange slot per thread (warp shuffles)
    explicit BlockSync(int nthreads) : block(nthreads), slots(((nthreads + 31) / 32) * 32) {
        for (int w = 0; w < (nthreads + 31) / 32; ++w)
            warp.emplace_back(new std::barrier<>(std::min(32, nthreads - 32 * w)));
    }
};
inline thread_local BlockSync *cur = nullptr;

template <class F>
void launch(dim3 grid, dim3 block, F body)
{
    const int nt = (int)block.x;
    for (unsigned b = 0; b < grid.x; ++b) {
        BlockSync sync(nt);
        std::vector<std::thread> ts;
        ts.reserve(nt);
        for (int t = 0; t < nt; ++t)
            ts.emplace_back([&, t] {
                threadIdx = {(unsigned)t, 0, 0};
                blockIdx = {b, 0, 0};
                blockDim = block;
                gridDim = grid;
                cur = &sync;
                body();
                // a thread that has left the kernel no longer takes part in barriers
                sync.warp[t / 32]->arrive_and_drop();
                sync.block.arrive_and_drop();
            });
        for (auto &th : ts) th.join();
    }
}
}  // namespace cuda_emu

inline void __syncthreads() { cuda_emu::cur->block.arrive_and_wait(); }
inline void __syncwarp(unsigned = 0xffffffffu) { cuda_emu::cur->warp[threadIdx.x / 32]->arrive_and_wait(); }

// full-warp shuffle: every lane of the warp that is still in the kernel must call it (true of the kernels emulated
// here: lanes leave only at the end)
template <class T>
inline T __shfl_sync(unsigned, T v, int src, int width = 32)
{
    static_assert(sizeof(T) <= 8, "shuffle payload");
    cuda_emu::BlockSync *s = cuda_emu::cur;
    const unsigned tid = threadIdx.x, w = tid / 32, lane = tid % 32;
    unsigned long long raw = 0;
    memcpy(&raw, &v, sizeof(T));
    s->slots[tid] = raw;
    s->warp[w]->arrive_and_wait();
    const unsigned from = (lane / width) * width + ((unsigned)src % width);
    raw = s->slots[w * 32 + from];
    s->warp[w]->arrive_and_wait();
    T out;
    memcpy(&out, &raw, sizeof(T));
    return out;
}

// warp vote over the lanes of the calling thread's warp (all of them still in the kernel)
inline int __all_sync(unsigned, int pred)
{
    cuda_emu::BlockSync *s = cuda_emu::cur;
    const unsigned tid = threadIdx.x, w = tid / 32;
    s->slots[tid] = pred ? 1ull : 0ull;
    s->warp[w]->arrive_and_wait();
    const unsigned lanes = std::min(32u, blockDim.x - 32 * w);
    int all = 1;
    for (unsigned l = 0; l < lanes; ++l) all &= (int)s->slots[w * 32 + l];
    s->warp[w]->arrive_and_wait();
    return all;
}
inline unsigned __ballot_sync(unsigned, int pred)
{
    cuda_emu::BlockSync *s = cuda_emu::cur;
    const unsigned tid = threadIdx.x, w = tid / 32;
    s->slots[tid] = pred ? 1ull : 0ull;
    s->warp[w]->arrive_and_wait();
    const unsigned lanes = std::min(32u, blockDim.x - 32 * w);
    unsigned m = 0;
    for (unsigned l = 0; l < lanes; ++l) m |= (unsigned)s->slots[w * 32 + l] << l;
    s->warp[w]->arrive_and_wait();
    return m;
}
template <class T> inline T __ldg(const T *p) { return *p; }
template <class T> inline T __ldcs(const T *p) { return *p; }
template <class T> inline T __ldcg(const T *p) { return *p; }
template <class T>
inline T __shfl_xor_sync(unsigned m, T v, int lane_mask, int width = 32)
{
    return __shfl_sync(m, v, (int)((threadIdx.x % 32) ^ (unsigned)lane_mask), width);
}
inline int __any_sync(unsigned m, int pred) { return __ballot_sync(m, pred) != 0u; }
template <class T>
inline T __shfl_up_sync(unsigned m, T v, unsigned delta, int width = 32)
{
    const unsigned lane = threadIdx.x % 32;
    const T got = __shfl_sync(m, v, (int)(lane >= delta ? lane - delta : lane), width);
    return lane % (unsigned)width >= delta ? got : v;
}
// lanes of the calling thread's warp that hold the same 32-bit value
inline unsigned __match_any_sync(unsigned, int v)
{
    cuda_emu::BlockSync *s = cuda_emu::cur;
    const unsigned tid = threadIdx.x, w = tid / 32;
    s->slots[tid] = (unsigned long long)(unsigned)v;
    s->warp[w]->arrive_and_wait();
    const unsigned lanes = std::min(32u, blockDim.x - 32 * w);
    unsigned m = 0;
    for (unsigned l = 0; l < lanes; ++l) m |= (unsigned)(s->slots[w * 32 + l] == (unsigned long long)(unsigned)v) << l;
    s->warp[w]->arrive_and_wait();
    return m;
}
inline int __popc(unsigned v) { return __builtin_popcount(v); }
inline int __ffs(int v) { return __builtin_ffs(v); }
inline int atomicAdd(int *p, int v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
inline int atomicCAS(int *p, int expected, int desired)
{
    __atomic_compare_exchange_n(p, &expected, desired, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST);
    return expected;                                         // (the value found, like the device's)
}
inline void __threadfence_block() { std::atomic_thread_fence(std::memory_order_seq_cst); }
inline void __threadfence() { std::atomic_thread_fence(std::memory_order_seq_cst); }
inline int __clz(unsigned v) { return v ? __builtin_clz(v) : 32; }
inline int __double2loint(double d) { unsigned long long r; memcpy(&r, &d, 8); return (int)(unsigned)(r & 0xffffffffull); }
#define B200_PREFETCH_L2(p) ((void)(p))

#define B200_LAUNCH(kernel, grid, block, smem, stream, ...) \
    cuda_emu::launch(dim3(grid), dim3(block), [&] { kernel(__VA_ARGS__); })
#define B200_LAUNCH_PDL(early, kernel, grid, block, smem, stream, ...) \
    cuda_emu::launch(dim3(grid), dim3(block), [&] { kernel(__VA_ARGS__); })
