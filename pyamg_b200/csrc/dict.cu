// Value dictionary for operators with few distinct values (constant-coefficient stencils: 2 values for
// Poisson, <= 9 for the rotated anisotropic diffusion stencil).  codes[k] indexes a table of <= 256 doubles
// with dict[codes[k]] bit-identical to Ax[k] (keys are the 64-bit patterns, so -0.0 / NaN payloads survive).
// The level-0 operator then costs 5 bytes per entry of HBM traffic instead of 12 (csr_pipe.cu, DICT path).
// Built once per level by the cycle engine; declined (codes == nullptr) when the operator has more values.
#include <string.h>
#include <algorithm>
#include <vector>
#include "internal.cuh"

namespace b200amg {

constexpr int DICT_MAX = 256;
constexpr int MISS_CAP = 512;

// miss[0] = number of entries whose value is not in the table; miss_vals[512]: hashed report of such values
__global__ void __launch_bounds__(256) dict_encode_kernel(long long nnz, const double *__restrict__ Ax,
                                                          const unsigned long long *__restrict__ keys, int nkeys,
                                                          unsigned char *__restrict__ codes, int *miss,
                                                          unsigned long long *miss_vals)
{
    __shared__ unsigned long long s_k[DICT_MAX];
    for (int i = threadIdx.x; i < DICT_MAX; i += 256) s_k[i] = i < nkeys ? keys[i] : ~0ull;
    __syncthreads();
    const long long stride = (long long)gridDim.x * 256;
    for (long long k = (long long)blockIdx.x * 256 + threadIdx.x; k < nnz; k += stride) {
        const unsigned long long key = (unsigned long long)__double_as_longlong(Ax[k]);
        int lo = 0, hi = nkeys;                      // first index with s_k[idx] >= key
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (s_k[mid] < key) lo = mid + 1; else hi = mid;
        }
        if (lo < nkeys && s_k[lo] == key) {
            codes[k] = (unsigned char)lo;
        } else {
            // report the value in a hashed slot: the same missing value always lands in the same slot, so a
            // value that occurs millions of times cannot crowd out the others
            atomicAdd(miss, 1);
            miss_vals[(key * 0x9E3779B97F4A7C15ull) >> 55] = key;
        }
    }
}

void dict_free(ValueDict *d)
{
    if (d->codes) cudaFree(d->codes);
    if (d->dict) cudaFree(d->dict);
    *d = ValueDict();
}

int dict_build(long long nnz, const double *Ax, ValueDict *out, cudaStream_t st)
{
    *out = ValueDict();
    if (nnz < (1 << 16)) return 0;                   // not worth a second layout
    std::vector<unsigned long long> keys;
    {   // seed the table from three samples of the value stream
        const long long seg = 1 << 14;
        std::vector<double> h(3 * seg);
        const long long starts[3] = {0, (nnz - seg) / 2, nnz - seg};
        for (int q = 0; q < 3; ++q)
            B200_CUDA(cudaMemcpyAsync(h.data() + q * seg, Ax + starts[q], sizeof(double) * seg, cudaMemcpyDeviceToHost, st));
        B200_CUDA(cudaStreamSynchronize(st));
        for (double v : h) {
            unsigned long long k;
            memcpy(&k, &v, 8);
            keys.push_back(k);
        }
        std::sort(keys.begin(), keys.end());
        keys.erase(std::unique(keys.begin(), keys.end()), keys.end());
        if ((int)keys.size() > DICT_MAX) return 0;
    }
    unsigned char *codes = nullptr;
    unsigned long long *d_keys = nullptr, *d_missv = nullptr;
    int *d_miss = nullptr;
    cudaError_t e = cudaMalloc(&codes, (size_t)nnz + 64);
    if (e == cudaSuccess) e = cudaMalloc(&d_keys, sizeof(unsigned long long) * DICT_MAX);
    if (e == cudaSuccess) e = cudaMalloc(&d_missv, sizeof(unsigned long long) * MISS_CAP);
    if (e == cudaSuccess) e = cudaMalloc(&d_miss, sizeof(int));
    int rc = 0;
    bool done = false;
    for (int round = 0; e == cudaSuccess && !rc && !done && round < 8; ++round) {
        e = cudaMemcpyAsync(d_keys, keys.data(), sizeof(unsigned long long) * keys.size(), cudaMemcpyHostToDevice, st);
        if (e == cudaSuccess) e = cudaMemsetAsync(d_miss, 0, sizeof(int), st);
        if (e == cudaSuccess) e = cudaMemsetAsync(d_missv, 0xFF, sizeof(unsigned long long) * MISS_CAP, st);
        if (e != cudaSuccess) break;
        const int grid = (int)std::min<long long>((nnz + 255) / 256, 148 * 16);
        dict_encode_kernel<<<grid, 256, 0, st>>>(nnz, Ax, d_keys, (int)keys.size(), codes, d_miss, d_missv);
        rc = check_launch("dict_encode_kernel");
        int miss = 0;
        if (!rc) e = cudaMemcpyAsync(&miss, d_miss, sizeof(int), cudaMemcpyDeviceToHost, st);
        if (!rc && e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (rc || e != cudaSuccess) break;
        if (miss == 0) { done = true; break; }
        std::vector<unsigned long long> mv(MISS_CAP);
        e = cudaMemcpy(mv.data(), d_missv, sizeof(unsigned long long) * MISS_CAP, cudaMemcpyDeviceToHost);
        for (unsigned long long k : mv)
            if (k != ~0ull) keys.push_back(k);
        std::sort(keys.begin(), keys.end());
        keys.erase(std::unique(keys.begin(), keys.end()), keys.end());
        if ((int)keys.size() > DICT_MAX) break;
    }
    double *dict = nullptr;
    if (done && e == cudaSuccess) {
        e = cudaMalloc(&dict, sizeof(double) * DICT_MAX);
        if (e == cudaSuccess) {
            std::vector<unsigned long long> full(DICT_MAX, keys.back());
            std::copy(keys.begin(), keys.end(), full.begin());
            e = cudaMemcpy(dict, full.data(), sizeof(double) * DICT_MAX, cudaMemcpyHostToDevice);
        }
    }
    if (d_keys) cudaFree(d_keys);
    if (d_missv) cudaFree(d_missv);
    if (d_miss) cudaFree(d_miss);
    if (!done || e != cudaSuccess || rc) {
        if (codes) cudaFree(codes);
        if (dict) cudaFree(dict);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return set_err((int)e, "dict_build: CUDA error");
        }
        return rc;                                                       // declined: caller keeps fp64 values
    }
    out->codes = codes;
    out->dict = dict;
    return 0;
}

}  // namespace b200amg
