// Shared helpers for libb200amg (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
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

// kernel launch through one macro, so that tests/emu can compile a kernel source for the host and run it there
#ifndef B200_LAUNCH
#define B200_LAUNCH(kernel, grid, block, smem, stream, ...) kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#endif

constexpr int kSMs = 148;  // B200

inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace b200amg
