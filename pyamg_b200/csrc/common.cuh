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
// A cycle is a chain of 30-400 dependent kernels; between two of them the GPU otherwise idles for the launch
// (2-3 us) plus the first dependent loads of the next kernel.  A kernel launched through launch_pdl() may be
// scheduled while its predecessor in the stream still runs; it calls pdl_enter() BEFORE its first access to anything
// an earlier kernel of the stream may have written (griddepcontrol.wait: returns when the predecessor has finished
// and its writes are visible) -- whatever it does before that (barrier setup, loads of operator data that no kernel
// of the cycle writes) overlaps the predecessor.  pdl_enter() also lets the NEXT kernel be scheduled early
// (griddepcontrol.launch_dependents).  Launched the ordinary way both instructions do nothing.  B200AMG_PDL=0 turns
// the launch attribute off (A/B switch).  Works under stream capture (programmatic graph edges).
#ifdef B200_HOST_EMULATION
inline void pdl_enter() {}
inline void pdl_trigger() {}
inline void pdl_wait() {}
#else
__device__ __forceinline__ void pdl_trigger() { cudaTriggerProgrammaticLaunchCompletion(); }
__device__ __forceinline__ void pdl_wait() { cudaGridDependencySynchronize(); }
__device__ __forceinline__ void pdl_enter() { pdl_trigger(); pdl_wait(); }
inline bool pdl_enabled()
{
    static const bool on = [] { const char *e = getenv("B200AMG_PDL"); return !(e && e[0] == '0'); }();
    return on;
}
template <class... P, class... A>
inline cudaError_t launch_pdl(void (*kernel)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, A &&...args)
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
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<A &&>(args)...);
}
#endif

// kernel launch through one macro, so that tests/emu can compile a kernel source for the host and run it there
#ifndef B200_LAUNCH
#define B200_LAUNCH(kernel, grid, block, smem, stream, ...) kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#endif
// ... and the same for kernels that call pdl_enter() / pdl_wait() (see above)
#ifndef B200_LAUNCH_PDL
#define B200_LAUNCH_PDL(kernel, grid, block, smem, stream, ...) \
    (void)b200amg::launch_pdl(kernel, dim3(grid), dim3(block), (size_t)(smem), (stream), __VA_ARGS__)
#endif

constexpr int kSMs = 148;  // B200

inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace b200amg
