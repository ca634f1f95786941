// Halo refresh over NVLink peer memory (multi-GPU row-partitioned cycle, pyamg_b200/distributed.py).
//
// Every rank's level vectors live in a SYMMETRIC allocation (torch symmetric memory: the same
// buffer mapped into every peer's address space).  Refreshing a halo is then two tiny kernels
// instead of an NCCL all-to-all (41 us each, 10 per cycle, measured on 2 x B200):
//   push : this rank scatters the owned entries its neighbours need STRAIGHT INTO THEIR halo
//          regions (st.global over NVLink), fences system-wide, and the last CTA to finish
//          raises this rank's epoch in a flag cell on every neighbour (st.release.sys);
//   wait : one warp spins (ld.acquire.sys, bounded) until every neighbour's flag has reached
//          the expected epoch; the kernels that follow in the stream then read the halo.
// Epoch counters live in device memory, so both kernels are replay-safe inside a CUDA graph.
// A neighbour may only overwrite a halo region after this rank has consumed it: guaranteed by
// the cycle's order of exchanges (two pushes into the same buffer are always separated by an
// exchange on another buffer, whose wait orders them behind the consumer; every push signals and
// every wait covers the whole neighbour set).
#include "internal.cuh"

namespace b200amg {

__device__ __forceinline__ void st_release_sys_u32(unsigned *p, unsigned v)
{
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ unsigned ld_acquire_sys_u32(const unsigned *p)
{
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// send_idx[j]: owned entry to send; dst_rank[j]: row of `bases` (peer buffer base); dst_off[j]: element
// offset inside the peer's buffer.  sig_ptrs[i]: flag cell of this rank on neighbour i.
// state[0] = epoch of the last completed push, state[1] = CTA arrival counter.
__global__ void __launch_bounds__(256) halo_push_kernel(const double *__restrict__ v, int n_send,
                                                        const int *__restrict__ send_idx,
                                                        const int *__restrict__ dst_rank,
                                                        const long long *__restrict__ dst_off,
                                                        const unsigned long long *__restrict__ bases,
                                                        const unsigned long long *__restrict__ sig_ptrs, int n_nbr,
                                                        unsigned *state)
{
    const int j = blockIdx.x * 256 + threadIdx.x;
    if (j < n_send) {
        double *dst = reinterpret_cast<double *>(bases[dst_rank[j]]) + dst_off[j];
        *dst = v[send_idx[j]];
    }
    // one system-scope fence per CTA, by the thread that goes on to signal: the barrier orders the CTA's peer stores
    // before it and the fence is cumulative (the grid-sync pattern) -- a fence per thread made every thread wait for
    // its own NVLink round trip
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        const unsigned prev = atomicAdd(&state[1], 1u);
        if (prev == gridDim.x - 1) {                    // last CTA: every CTA's stores are fenced
            state[1] = 0;
            const unsigned e = state[0] + 1;
            state[0] = e;
            __threadfence_system();
            for (int i = 0; i < n_nbr; ++i) st_release_sys_u32(reinterpret_cast<unsigned *>(sig_ptrs[i]), e);
        }
    }
}

// flags[i]: cell written by neighbour i; state[0] = epoch of the last completed wait; state[1] = error flag
__global__ void __launch_bounds__(32) halo_wait_kernel(const unsigned long long *__restrict__ flag_ptrs, int n_nbr,
                                                       unsigned *state, long long max_spins)
{
    const unsigned e = state[0] + 1;
    const int i = threadIdx.x;
    if (i < n_nbr) {
        const unsigned *f = reinterpret_cast<const unsigned *>(flag_ptrs[i]);
        long long spins = 0;
        while ((int)(ld_acquire_sys_u32(f) - e) < 0) {
            if (++spins > max_spins) { atomicExch(&state[1], 1u); break; }     // never hang the GPU
            __nanosleep(20);
        }
    }
    __syncwarp();
    if (i == 0) state[0] = e;
}

// push and wait in ONE kernel node: the CTAs scatter and fence as in halo_push_kernel; the last CTA to arrive raises
// this rank's epoch on the neighbours and then ALSO waits for theirs (lane i spins on neighbour i's flag cell), so the
// kernel retires when the halo of this rank is complete.  The push never depends on a peer, so two ranks running this
// kernel cannot block each other.  One launch per refresh instead of two (13 refreshes per distributed cycle).
constexpr int HALO_EPT = 4;      // plan entries per thread
__global__ void __launch_bounds__(256) halo_exchange_kernel(const double *__restrict__ v, int n_send,
                                                            const int *__restrict__ send_idx,
                                                            const int *__restrict__ dst_rank,
                                                            const long long *__restrict__ dst_off,
                                                            const unsigned long long *__restrict__ bases,
                                                            const unsigned long long *__restrict__ sig_ptrs,
                                                            const unsigned long long *__restrict__ flag_ptrs, int n_nbr,
                                                            unsigned *push_state, unsigned *wait_state, long long max_spins)
{
    __shared__ int s_last;
    // The refresh sits between two slab kernels 13 times per cycle: it is launched early (programmatic dependent launch,
    // common.cuh) and resolves its part of the PLAN -- source index, peer base, destination offset: set-up data no
    // kernel writes -- while the kernel that produces v still runs; only the payload waits for it.
    pdl_trigger();
    double *dst[HALO_EPT];
    int src[HALO_EPT];
#pragma unroll
    for (int q = 0; q < HALO_EPT; ++q) {
        const int j = (blockIdx.x * HALO_EPT + q) * 256 + threadIdx.x;
        src[q] = -1;
        dst[q] = nullptr;
        if (j < n_send) {
            src[q] = send_idx[j];
            dst[q] = reinterpret_cast<double *>(bases[dst_rank[j]]) + dst_off[j];
        }
    }
    pdl_wait();
#pragma unroll
    for (int q = 0; q < HALO_EPT; ++q)
        if (src[q] >= 0) *dst[q] = v[src[q]];
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        const unsigned prev = atomicAdd(&push_state[1], 1u);
        s_last = (prev == gridDim.x - 1);
        if (s_last) {
            push_state[1] = 0;
            const unsigned e = push_state[0] + 1;
            push_state[0] = e;
            __threadfence_system();
            for (int i = 0; i < n_nbr; ++i) st_release_sys_u32(reinterpret_cast<unsigned *>(sig_ptrs[i]), e);
        }
    }
    __syncthreads();
    if (!s_last || threadIdx.x >= 32) return;
    const unsigned e = wait_state[0] + 1;
    const int i = threadIdx.x;
    if (i < n_nbr) {
        const unsigned *f = reinterpret_cast<const unsigned *>(flag_ptrs[i]);
        long long spins = 0;
        while ((int)(ld_acquire_sys_u32(f) - e) < 0) {
            if (++spins > max_spins) { atomicExch(&wait_state[1], 1u); break; }
            __nanosleep(20);
        }
    }
    __syncwarp();
    if (i == 0) wait_state[0] = e;
}

}  // namespace b200amg

using namespace b200amg;

extern "C" int b200amg_halo_exchange(const double *v, int n_send, const int *send_idx, const int *dst_rank,
                                     const long long *dst_off, const unsigned long long *bases,
                                     const unsigned long long *sig_ptrs, const unsigned long long *flag_ptrs, int n_nbr,
                                     unsigned *push_state, unsigned *wait_state, long long max_spins, void *stream)
{
    if (n_send < 0 || n_nbr < 0 || n_nbr > 32) return set_err(B200AMG_EINVAL, "halo_exchange: bad sizes (0..32 neighbours)");
    const int per_cta = 256 * HALO_EPT;
    const int grid = n_send > 0 ? (n_send + per_cta - 1) / per_cta : 1;
    // B200AMG_HALO_PDL=1 launches the refresh early.  Off by default: measured on 2 B200 it changes nothing inside a cycle
    // (1.325 vs 1.322 ms) and little back to back (8-11 vs 11-13 us per refresh, tools/halo_microbench.py)
    static const bool early = [] { const char *e = getenv("B200AMG_HALO_PDL"); return e && e[0] == '1'; }();
    const cudaError_t ce = launch_pdl(early, halo_exchange_kernel, dim3(grid), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream),
                                      v, n_send, send_idx, dst_rank, dst_off, bases, sig_ptrs, flag_ptrs, n_nbr, push_state,
                                      wait_state, max_spins);
    if (ce != cudaSuccess) return set_err((int)ce, "halo_exchange_kernel: launch failed");
    return check_launch("halo_exchange_kernel");
}

extern "C" int b200amg_halo_push(const double *v, int n_send, const int *send_idx, const int *dst_rank,
                                 const long long *dst_off, const unsigned long long *bases,
                                 const unsigned long long *sig_ptrs, int n_nbr, unsigned *state, void *stream)
{
    if (n_send < 0 || n_nbr < 0) return set_err(B200AMG_EINVAL, "halo_push: negative size");
    const int grid = n_send > 0 ? (n_send + 255) / 256 : 1;
    halo_push_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(v, n_send, send_idx, dst_rank, dst_off,
                                                                               bases, sig_ptrs, n_nbr, state);
    return check_launch("halo_push_kernel");
}

extern "C" int b200amg_halo_wait(const unsigned long long *flag_ptrs, int n_nbr, unsigned *state, long long max_spins,
                                 void *stream)
{
    if (n_nbr < 0 || n_nbr > 32) return set_err(B200AMG_EINVAL, "halo_wait: 0..32 neighbours");
    halo_wait_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(flag_ptrs, n_nbr, state, max_spins);
    return check_launch("halo_wait_kernel");
}
