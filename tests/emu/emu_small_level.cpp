// TEST INFRASTRUCTURE ONLY: C entry point over the one-launch smoothing step of pyamg_b200/csrc/small_level.cu when that
// file is compiled for the host (tests/test_emulated_kernels.py), plus the backing store of its `extern __shared__` array.
#include "internal.cuh"

namespace b200amg { alignas(16) double sl_x[28000]; }

using namespace b200amg;

extern "C" int emu_gs_level_cta(int bs, int bsrf, int n, long long stored_values, const int *Ap, const int *Aj, const double *Ax,
                                const int *Id, const int *cptr, int K, double *x, const double *b, int iters, int sweep,
                                int x_is_zero)
{
    return gs_level_cta(bs, bsrf != 0, n, stored_values, Ap, Aj, Ax, Id, cptr, K, x, b, iters, sweep, x_is_zero != 0, nullptr);
}
