// Shared helpers for libb200amg (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include "../../include/b200amg.h"

namespace b200amg {

extern char g_err[512];
extern long long g_launches;

inline int set_err(int code, const char *what) {
    snprintf(g_err, sizeof(g_err), "%s", what);
    return code;
}

inline int check_launch(const char *what) {
    ++g_launches;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
        return (int)e;
    }
    return 0;
}

#define B200_CUDA(call)                                                         \
    do {                                                                        \
        cudaError_t e__ = (call);                                               \
        if (e__ != cudaSuccess) {                                               \
            snprintf(b200amg::g_err, sizeof(b200amg::g_err), "%s: %s", #call,   \
                     cudaGetErrorString(e__));                                  \
            return (int)e__;                                                    \
        }                                                                       \
    } while (0)

#define B200_TRY(expr)                \
    do {                              \
        int rc__ = (expr);            \
        if (rc__ != 0) return rc__;   \
    } while (0)

// Separately rounded fp64 ops: the reference's serial loops are compiled for baseline x86-64
// (no FMA), so parity to the bit needs mul and add rounded one after the other.
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double sub_rn(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double div_rn(double a, double b) { return __ddiv_rn(a, b); }

// ---- programmatic dependent launch (PDL) -----------------------------------------------------------------------
// A cycle is a chain of 30-400 dependent kernels; on the coarse levels a kernel is a few microseconds of work behind
// 2-3 us of launch latency and a chain of dependent first loads.  A kernel launched through launch_pdl(early = true)
// may be scheduled while its predecessor in the stream still runs; it calls pdl_wait() BEFORE its first access to
// anything an earlier kernel of the stream may have written (griddepcontrol.wait: returns when the predecessor has
// finished and flushed its writes) -- whatever it does before that (barrier set-up, loads from engine-owned operator
// copies, which are complete before any launch) overlaps the predecessor.  pdl_trigger() at the top of a kernel lets
// ITS successor be scheduled (griddepcontrol.launch_dependents).  Launched the ordinary way both do nothing.
//
// One thing an early start takes away: a kernel normally begins with an L1 that holds nothing of its predecessor's
// era; blocks of two kernels sharing an SM means lines the predecessor pulled in (x, say, before another SM updated
// it) are still there after the wait -- measured: wrong iterates from the first cycle on a 1 728-row problem.  So the
// wait is followed by a gpu-scope fence, which the compiler turns into MEMBAR + CCTL.IVALL (L1 invalidate).  Only
// small grids are launched early (that is where launch latency matters) and only small grids pay the fence:
// both sides use the same bound on gridDim.x.  B200AMG_PDL=0 turns the attribute off (A/B switch); =2 also launches
// the big persistent kernels early.  Works under stream capture (programmatic graph edges).
constexpr unsigned PDL_MAX_GRID = 2048;
constexpr long long PDL_SMALL_ROWS = 400000;      // launches over fewer rows are taken to be latency-bound
#ifdef B200_HOST_EMULATION
inline void pdl_enter() {}
inline void pdl_trigger() {}
inline void pdl_wait() {}
#else
__device__ __forceinline__ void pdl_trigger() { cudaTriggerProgrammaticLaunchCompletion(); }
__device__ __forceinline__ void pdl_wait()
{
    cudaGridDependencySynchronize();
    if (gridDim.x <= PDL_MAX_GRID) {                  // (call sites are warp-uniform; the invalidate is SM-wide)
        if ((threadIdx.x & 31) == 0) asm volatile("fence.acq_rel.gpu;" ::: "memory");
        __syncwarp();
    }
}
__device__ __forceinline__ void pdl_enter() { pdl_trigger(); pdl_wait(); }
inline int pdl_level()
{
    static const int lv = [] { const char *e = getenv("B200AMG_PDL"); return (e && e[0] >= '0' && e[0] <= '2') ? e[0] - '0' : 1; }();
    return lv;
}
// Kernel classes (B200AMG_PDL_CLASSES bit mask): 1 colour sweeps, 2 few-row sliced-ELL forms, 4 everything else.
// Only the colour sweeps launch early by default -- measured under graph replay (profiles/r02/run20_pdl_classes_summary.txt):
// config 3 (300 colour launches per cycle, operator loads before the wait) 6.10 -> 5.40 ms with class 1 alone; config 2
// (35 kernels of 5-190 us) 0.979 ms without, 0.994 with class 2, 1.001 with class 4, 1.019 with both: where the wait sits
// at the top of a kernel there is nothing to overlap and the fence is pure cost.
constexpr int PDL_SWEEP = 1, PDL_SELL = 2, PDL_OTHER = 4;
inline bool pdl_class(int bit)
{
    static const int mask = [] { const char *e = getenv("B200AMG_PDL_CLASSES"); return e ? atoi(e) : PDL_SWEEP; }();
    return (mask & bit) != 0;
}
// early: the caller's judgement that the launch is latency-bound (few rows); level 2 ignores it
template <class... P, class... A>
inline cudaError_t launch_pdl(bool early, void (*kernel)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, A &&...args)
{
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cfg.attrs = attr;
    const int lv = pdl_level();
    cfg.numAttrs = (grid.x <= PDL_MAX_GRID && grid.y == 1 && grid.z == 1 && (lv == 2 || (lv == 1 && early))) ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<A &&>(args)...);
}
#endif

// kernel launch through one macro, so that tests/emu can compile a kernel source for the host and run it there
#ifndef B200_LAUNCH
#define B200_LAUNCH(kernel, grid, block, smem, stream, ...) kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#endif
// ... and the same for kernels that call pdl_enter() / pdl_wait() (see above)
#ifndef B200_LAUNCH_PDL
#define B200_LAUNCH_PDL(early, kernel, grid, block, smem, stream, ...) \
    (void)b200amg::launch_pdl((early), kernel, dim3(grid), dim3(block), (size_t)(smem), (stream), __VA_ARGS__)
#endif

constexpr int kSMs = 148;  // B200

inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace b200amg
