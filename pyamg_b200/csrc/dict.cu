// Value dictionary for operators with few distinct values (constant-coefficient stencils: 2 values for
// Poisson, <= 9 for the rotated anisotropic diffusion stencil).  codes[k] indexes a table of <= 256 doubles
// with dict[codes[k]] bit-identical to Ax[k] (keys are the 64-bit patterns, so -0.0 / NaN payloads survive).
// The level-0 operator then costs 5 bytes per entry of HBM traffic instead of 12 (csr_pipe.cu, DICT path).
// Built once per level by the cycle engine; declined (codes == nullptr) when the operator has more values.
//
// Pattern dictionary (pattern_build): the same idea over (Aj[k] - row, Ax[k]) PAIRS of a square operator.  A
// constant-coefficient stencil on a structured grid has one pair per stencil point and boundary variant
// (7 for the 7-point Poisson matrix), so the whole (Aj, Ax) stream shrinks to 1 byte per entry; storage order,
// and with it the summation order, is untouched.
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

// Pattern dictionary: one thread per row looks up (Aj[k] - row, Ax[k]) in a table sorted by (offset, value bits).
// miss_pairs[512] (value bits, offset) is the hashed report of pairs that are not in the table yet.
__global__ void __launch_bounds__(256) pattern_encode_kernel(int n_row, const int *__restrict__ Ap,
                                                             const int *__restrict__ Aj, const double *__restrict__ Ax,
                                                             const unsigned long long *__restrict__ key_val,
                                                             const int *__restrict__ key_off, int nkeys,
                                                             unsigned char *__restrict__ codes, int *miss,
                                                             ulonglong2 *miss_pairs)
{
    __shared__ unsigned long long s_v[DICT_MAX];
    __shared__ int s_o[DICT_MAX];
    for (int i = threadIdx.x; i < DICT_MAX; i += 256) {
        s_v[i] = i < nkeys ? key_val[i] : ~0ull;
        s_o[i] = i < nkeys ? key_off[i] : 0x7fffffff;
    }
    __syncthreads();
    const int stride = gridDim.x * 256;
    for (int row = blockIdx.x * 256 + threadIdx.x; row < n_row; row += stride) {
        for (int k = Ap[row]; k < Ap[row + 1]; ++k) {
            const int off = Aj[k] - row;
            const unsigned long long val = (unsigned long long)__double_as_longlong(Ax[k]);
            int lo = 0, hi = nkeys;                  // first index with key >= (off, val)
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (s_o[mid] < off || (s_o[mid] == off && s_v[mid] < val)) lo = mid + 1; else hi = mid;
            }
            if (lo < nkeys && s_o[lo] == off && s_v[lo] == val) {
                codes[k] = (unsigned char)lo;
            } else {
                atomicAdd(miss, 1);
                const unsigned long long h = (val ^ ((unsigned long long)(unsigned)off * 0xD6E8FEB86659FD93ull)) * 0x9E3779B97F4A7C15ull;
                miss_pairs[h >> 55] = make_ulonglong2(val, (unsigned long long)(unsigned)off);   // one 16-byte store
            }
        }
    }
}

void dict_free(ValueDict *d)
{
    if (d->codes) cudaFree(d->codes);
    if (d->dict) cudaFree(d->dict);
    if (d->offs) cudaFree(d->offs);
    *d = ValueDict();
}

int pattern_build(int n_row, long long nnz, const int *Ap, const int *Aj, const double *Ax, ValueDict *out,
                  cudaStream_t st)
{
    *out = ValueDict();
    if (nnz < (1 << 16) || n_row <= 0) return 0;
    typedef std::pair<int, unsigned long long> Key;    // (offset, value bits): the table order
    std::vector<Key> keys;                             // starts empty: the refinement rounds discover the pairs
    unsigned char *codes = nullptr;
    unsigned long long *d_kv = nullptr;
    int *d_ko = nullptr, *d_miss = nullptr;
    ulonglong2 *d_pairs = nullptr;
    cudaError_t e = cudaMalloc(&codes, (size_t)nnz + 64);
    if (e == cudaSuccess) e = cudaMalloc(&d_kv, sizeof(unsigned long long) * DICT_MAX);
    if (e == cudaSuccess) e = cudaMalloc(&d_ko, sizeof(int) * DICT_MAX);
    if (e == cudaSuccess) e = cudaMalloc(&d_pairs, sizeof(ulonglong2) * MISS_CAP);
    if (e == cudaSuccess) e = cudaMalloc(&d_miss, sizeof(int));
    int rc = 0;
    bool done = false;
    std::vector<unsigned long long> kv(DICT_MAX);
    std::vector<int> ko(DICT_MAX);
    for (int round = 0; e == cudaSuccess && !rc && !done && round < 10; ++round) {
        for (size_t i = 0; i < keys.size(); ++i) { ko[i] = keys[i].first; kv[i] = keys[i].second; }
        e = cudaMemcpyAsync(d_kv, kv.data(), sizeof(unsigned long long) * DICT_MAX, cudaMemcpyHostToDevice, st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(d_ko, ko.data(), sizeof(int) * DICT_MAX, cudaMemcpyHostToDevice, st);
        if (e == cudaSuccess) e = cudaMemsetAsync(d_miss, 0, sizeof(int), st);
        if (e == cudaSuccess) e = cudaMemsetAsync(d_pairs, 0xFF, sizeof(ulonglong2) * MISS_CAP, st);
        if (e != cudaSuccess) break;
        const int grid = std::min((n_row + 255) / 256, 148 * 16);
        pattern_encode_kernel<<<grid, 256, 0, st>>>(n_row, Ap, Aj, Ax, d_kv, d_ko, (int)keys.size(), codes, d_miss, d_pairs);
        rc = check_launch("pattern_encode_kernel");
        int miss = 0;
        if (!rc) e = cudaMemcpyAsync(&miss, d_miss, sizeof(int), cudaMemcpyDeviceToHost, st);
        if (!rc && e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (rc || e != cudaSuccess) break;
        if (miss == 0) { done = true; break; }
        std::vector<ulonglong2> mp(MISS_CAP);
        e = cudaMemcpy(mp.data(), d_pairs, sizeof(ulonglong2) * MISS_CAP, cudaMemcpyDeviceToHost);
        for (const ulonglong2 &m : mp)
            if (!(m.x == ~0ull && m.y == ~0ull)) keys.push_back(Key((int)(unsigned)m.y, m.x));
        std::sort(keys.begin(), keys.end());
        keys.erase(std::unique(keys.begin(), keys.end()), keys.end());
        if ((int)keys.size() > DICT_MAX) break;          // a general matrix fills the report in the first round
    }
    double *dict = nullptr;
    int *offs = nullptr;
    if (done && e == cudaSuccess) {
        e = cudaMalloc(&dict, sizeof(double) * DICT_MAX);
        if (e == cudaSuccess) e = cudaMalloc(&offs, sizeof(int) * DICT_MAX);
        if (e == cudaSuccess) {
            for (size_t i = 0; i < (size_t)DICT_MAX; ++i) {
                const Key &k = keys[std::min(i, keys.size() - 1)];
                ko[i] = k.first; kv[i] = k.second;
            }
            e = cudaMemcpy(dict, kv.data(), sizeof(double) * DICT_MAX, cudaMemcpyHostToDevice);
            if (e == cudaSuccess) e = cudaMemcpy(offs, ko.data(), sizeof(int) * DICT_MAX, cudaMemcpyHostToDevice);
        }
    }
    if (d_kv) cudaFree(d_kv);
    if (d_ko) cudaFree(d_ko);
    if (d_pairs) cudaFree(d_pairs);
    if (d_miss) cudaFree(d_miss);
    if (!done || e != cudaSuccess || rc) {
        if (codes) cudaFree(codes);
        if (dict) cudaFree(dict);
        if (offs) cudaFree(offs);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return set_err((int)e, "pattern_build: CUDA error");
        }
        return rc;                                       // declined
    }
    out->codes = codes;
    out->dict = dict;
    out->offs = offs;
    return 0;
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
