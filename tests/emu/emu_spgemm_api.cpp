// TEST INFRASTRUCTURE ONLY: C entry point over the sort-free Galerkin-product rows of pyamg_b200/csrc/spgemm_hash.cu when
// that file is compiled for the host (tests/test_emulated_kernels.py), plus the backing store of its shared-memory array.
// The caller bins the rows by their number of products like b200amg_spgemm_numeric does (32 / 128 / 512 / 2048).
#include "internal.cuh"
#include "setup.cuh"

namespace b200amg { alignas(16) unsigned char sph_smem[1 << 20]; }

using namespace b200amg;

extern "C" int emu_spgemm_hash_bin(int bin, int mode, const int *rows, int nrows, const int *Ap, const int *Aj, const double *Ax,
                                   const int *Bp, const int *Bj, const double *Bx, const long long *ubp, int *Tj, double *Tx,
                                   int *cnt)
{
    return spgemm_hash_bin(bin, mode, rows, nrows, Ap, Aj, Ax, Bp, Bj, Bx, ubp, Tj, Tx, cnt, nullptr);
}
