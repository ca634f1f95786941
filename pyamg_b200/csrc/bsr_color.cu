// Engine-owned colour-grouped copy of a BSR operator for colour-ordered block Gauss-Seidel (kernels: bsr.cu).
#include "internal.cuh"
#include "setup.cuh"

namespace b200amg {

// ---- colour-grouped copy of a BSR operator: row p of the copy = block row Id[p], blocks and block columns unchanged.
// A colour's rows are then one contiguous run of blocks (the lanes of a warp stream neighbouring rows) instead of
// 1 KB pieces scattered over the whole operator.  Exact re-encoding: same blocks, same order inside a row.
__global__ void __launch_bounds__(256) bsr_perm_len_kernel(int n, const int *__restrict__ Ap, const int *__restrict__ Id, int *len)
{
    const int p = blockIdx.x * 256 + threadIdx.x;
    if (p < n) len[p] = Ap[Id[p] + 1] - Ap[Id[p]];
}

// one warp per row: block columns, then the row's blocks as one contiguous piece of doubles
__global__ void __launch_bounds__(256) bsr_perm_copy_kernel(int n, int bs2, const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                            const double *__restrict__ Ax, const int *__restrict__ Id,
                                                            const int *__restrict__ rp, int *__restrict__ cj, double *__restrict__ cx)
{
    const int p = (int)((blockIdx.x * 256 + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (p >= n) return;
    const int src = Ap[Id[p]], dst = rp[p], len = rp[p + 1] - dst;
    for (int q = lane; q < len; q += 32) cj[dst + q] = Aj[src + q];
    const double *s = Ax + (size_t)src * bs2;
    double *d = cx + (size_t)dst * bs2;
    for (long long e = lane; e < (long long)len * bs2; e += 32) d[e] = s[e];
}

void bsr_color_free(BsrColor *c)
{
    if (c->rp) cudaFree(c->rp);
    if (c->cj) cudaFree(c->cj);
    if (c->cx) cudaFree(c->cx);
    *c = BsrColor();
}

int bsr_color_build(int nb, int bs, long long nblocks, const int *Ap, const int *Aj, const double *Ax, const int *Id,
                    BsrColor *out, cudaStream_t st)
{
    *out = BsrColor();
    if (nb <= 0 || nblocks <= 0) return 0;
    BsrColor c;
    int *len = nullptr, *scratch = nullptr;
    cudaError_t e = cudaMalloc(&len, sizeof(int) * (size_t)nb);
    if (e == cudaSuccess) e = cudaMalloc(&scratch, sizeof(int) * (size_t)scan_scratch_elems(nb));
    if (e == cudaSuccess) e = cudaMalloc(&c.rp, sizeof(int) * ((size_t)nb + 1));
    if (e == cudaSuccess) e = cudaMalloc(&c.cj, sizeof(int) * (size_t)nblocks);
    if (e == cudaSuccess) e = cudaMalloc(&c.cx, sizeof(double) * (size_t)nblocks * bs * bs);
    int rc = 0;
    if (e == cudaSuccess) {
        bsr_perm_len_kernel<<<(nb + 255) / 256, 256, 0, st>>>(nb, Ap, Id, len);
        rc = check_launch("bsr_perm_len_kernel");
        if (!rc) rc = exclusive_scan<int, int>(nb, len, c.rp, scratch, st);
        if (!rc) {
            bsr_perm_copy_kernel<<<(unsigned)(((long long)nb * 32 + 255) / 256), 256, 0, st>>>(nb, bs * bs, Ap, Aj, Ax, Id, c.rp, c.cj, c.cx);
            rc = check_launch("bsr_perm_copy_kernel");
        }
        if (!rc) e = cudaStreamSynchronize(st);
    }
    if (len) cudaFree(len);
    if (scratch) cudaFree(scratch);
    if (e != cudaSuccess || rc) {
        bsr_color_free(&c);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return 0;                                   // no memory for the copy: the caller keeps the operator as stored
        }
        return rc;
    }
    *out = c;
    return 0;
}

}  // namespace b200amg
