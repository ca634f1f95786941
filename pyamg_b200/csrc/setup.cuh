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

}  // namespace b200amg
