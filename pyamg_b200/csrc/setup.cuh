// Setup-phase utilities shared by scan.cu / spgemm.cu / permute.cu.
#pragma once
#include "common.cuh"

namespace b200amg {

constexpr int NBINS = 6;
struct BinLimits { long long v[NBINS - 1]; };

template <typename TI, typename TO>
int exclusive_scan(long long n, const TI *in, TO *out, TO *scratch, cudaStream_t st);
long long scan_scratch_elems(long long n);
int bin_rows(int n, const long long *size, const BinLimits &lim, int *perm, int *work, int *offsets_host,
             cudaStream_t st);

// spgemm_hash.cu: rows of C = A B in scipy's output order without a sort (one warp per row); mode = B200AMG_SPGEMM_HASH
constexpr int SPGEMM_HASH_DEFAULT = 3;
int spgemm_hash_mode();
int spgemm_hash_bin(int bin, int mode, const int *rows, int nrows, const int *Ap, const int *Aj, const double *Ax, const int *Bp,
                    const int *Bj, const double *Bx, const long long *ubp, int *Tj, double *Tx, int *cnt, cudaStream_t st);

}  // namespace b200amg
