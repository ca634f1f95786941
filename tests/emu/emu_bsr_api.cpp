// TEST INFRASTRUCTURE ONLY: C entry points over the block kernels of pyamg_b200/csrc/bsr.cu when that file is compiled
// for the host (tests/test_emulated_kernels.py).  They restate, for block sizes > 1, the dispatch api.cu does around the
// same internal functions (the 1 x 1 case goes to the CSR kernels there, which the emulation does not cover).
#include "internal.cuh"
#include <cstring>

using namespace b200amg;

extern "C" int emu_bsr_matvec(int n_brow, int R, int C, const int *Ap, const int *Aj, const double *Ax, const double *x,
                              const double *b, double *y, int mode)
{
    return bsr_matvec(n_brow, R, C, Ap, Aj, Ax, x, b, y, mode, nullptr);
}

extern "C" int emu_bsr_jacobi(const int *Ap, const int *Aj, const double *Ax, double *x, const double *b, double *temp,
                              int row_start, int row_stop, int bs, double omega)
{
    memcpy(temp, x, sizeof(double) * (size_t)(row_stop - row_start) * bs);       // api.cu: the leading entries, like the reference
    return bsr_jacobi(bs, row_start, row_stop, Ap, Aj, Ax, temp, x, b, omega, nullptr);
}

extern "C" int emu_bsr_gauss_seidel(const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
                                    int row_start, int row_stop, int row_step, int bs, const int *Id, const int *cp, int K)
{
    if (!cp) return bsr_gs_sequential(bs, Ap, Aj, Ax, x, b, Id, row_start, row_stop, row_step, nullptr);
    int c0, c1, dir;
    if (row_step == 1 && row_start == cp[0] && row_stop == cp[K]) { c0 = 0; c1 = K; dir = 1; }
    else if (row_step == -1 && row_start == cp[K] - 1 && row_stop == cp[0] - 1) { c0 = K - 1; c1 = -1; dir = -1; }
    else return -1;
    for (int c = c0; c != c1; c += dir) {
        const int rc = bsr_gs_color(bs, Ap, Aj, Ax, x, b, Id, cp[c], cp[c + 1], dir < 0, nullptr);
        if (rc) return rc;
    }
    return 0;
}

// the warp-per-block-row schedules (picked by the block-count hint in the product: long rows)
extern "C" int emu_bsr_matvec_warp(int n_brow, int R, int C, const int *Ap, const int *Aj, const double *Ax, const double *x,
                                   const double *b, double *y, int mode)
{
    return bsr_matvec(n_brow, R, C, Ap, Aj, Ax, x, b, y, mode, nullptr, 1000000000ll);
}

extern "C" int emu_bsr_gauss_seidel_warp(const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
                                         int row_start, int row_stop, int row_step, int bs, const int *Id, const int *cp, int K)
{
    int c0, c1, dir;
    if (row_step == 1 && row_start == cp[0] && row_stop == cp[K]) { c0 = 0; c1 = K; dir = 1; }
    else if (row_step == -1 && row_start == cp[K] - 1 && row_stop == cp[0] - 1) { c0 = K - 1; c1 = -1; dir = -1; }
    else return -1;
    for (int c = c0; c != c1; c += dir) {
        const int rc = bsr_gs_color(bs, Ap, Aj, Ax, x, b, Id, cp[c], cp[c + 1], dir < 0, nullptr, 1.0e9);
        if (rc) return rc;
    }
    return 0;
}
