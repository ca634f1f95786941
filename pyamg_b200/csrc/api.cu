// extern "C" entry points of libb200amg.so for the stand-alone kernels (include/b200amg.h).
#include <stdlib.h>
#include "internal.cuh"

using namespace b200amg;

static inline cudaStream_t S(void *s) { return reinterpret_cast<cudaStream_t>(s); }

extern "C" const char *b200amg_last_error(void) { return g_err; }
extern "C" int b200amg_version(void) { return 100; }
extern "C" long long b200amg_launch_count(void) { return g_launches; }

extern "C" int b200amg_csr_matvec(int n_row, int n_col, const int *Ap, const int *Aj, const double *Ax,
                                  const double *x, double *y, void *stream)
{
    if (n_row < 0 || n_col < 0) return set_err(B200AMG_EINVAL, "csr_matvec: negative size");
    StreamArgs a{Ap, Aj, Ax, x, nullptr, y, nnz_lookup(Ap, n_row, S(stream)), 0, n_row, 1.0, 0.0, 0.0};
    return stream_op(OP_MATVEC, a, S(stream));
}

extern "C" int b200amg_csr_spmv(int n_row, int n_col, const int *Ap, const int *Aj, const double *Ax,
                                const double *x, double *y, double alpha, double beta, void *stream)
{
    if (n_row < 0 || n_col < 0) return set_err(B200AMG_EINVAL, "csr_spmv: negative size");
    StreamArgs a{Ap, Aj, Ax, x, nullptr, y, nnz_lookup(Ap, n_row, S(stream)), 0, n_row, alpha, beta, 0.0};
    return stream_op(OP_SPMV_AB, a, S(stream));
}

extern "C" int b200amg_csr_residual(int n_row, const int *Ap, const int *Aj, const double *Ax, const double *x,
                                    const double *b, double *r, void *stream)
{
    if (n_row < 0) return set_err(B200AMG_EINVAL, "csr_residual: negative size");
    StreamArgs a{Ap, Aj, Ax, x, b, r, nnz_lookup(Ap, n_row, S(stream)), 0, n_row, 1.0, 0.0, 0.0};
    return stream_op(OP_RESIDUAL, a, S(stream));
}

extern "C" int b200amg_csr_matvec_add(int n_row, const int *Ap, const int *Aj, const double *Ax, const double *xc,
                                      double *x, void *stream)
{
    if (n_row < 0) return set_err(B200AMG_EINVAL, "csr_matvec_add: negative size");
    StreamArgs a{Ap, Aj, Ax, xc, nullptr, x, nnz_lookup(Ap, n_row, S(stream)), 0, n_row, 1.0, 0.0, 0.0};
    return stream_op(OP_ADD, a, S(stream));
}

extern "C" int b200amg_bsr_matvec(int n_brow, int n_bcol, int R, int C, const int *Ap, const int *Aj,
                                  const double *Ax, const double *x, const double *b, double *y, int mode,
                                  void *stream)
{
    if (n_brow < 0 || n_bcol < 0 || mode < 0 || mode > 2) return set_err(B200AMG_EINVAL, "bsr_matvec: bad arguments");
    if (R == 1 && C == 1) {   // scipy routes 1x1 blocks through csr_matvec
        StreamArgs a{Ap, Aj, Ax, x, b, y, nnz_lookup(Ap, n_brow, S(stream)), 0, n_brow, 1.0, 0.0, 0.0};
        return stream_op(mode == 0 ? OP_MATVEC : (mode == 1 ? OP_RESIDUAL : OP_ADD), a, S(stream));
    }
    return bsr_matvec(n_brow, R, C, Ap, Aj, Ax, x, b, y, mode, S(stream), nnz_lookup(Ap, n_brow, S(stream)));
}

extern "C" int b200amg_jacobi(const int *Ap, const int *Aj, const double *Ax, double *x, const double *b, double *temp,
                              int row_start, int row_stop, int row_step, double omega, void *stream)
{
    if (row_step != 1 || row_start < 0 || row_stop < row_start)
        return set_err(B200AMG_EINVAL, "jacobi: only forward contiguous ranges (row_step == 1)");
    if (row_stop == row_start) return 0;
    // temp <- x on the swept range (relaxation.h:213-215), then x <- f(temp)
    B200_CUDA(cudaMemcpyAsync(temp + row_start, x + row_start, sizeof(double) * (row_stop - row_start),
                              cudaMemcpyDeviceToDevice, S(stream)));
    StreamArgs a{Ap, Aj, Ax, temp, b, x, row_start == 0 ? nnz_lookup(Ap, row_stop, S(stream)) : -1, row_start, row_stop, 1.0, 0.0, omega};
    return stream_op(OP_JACOBI, a, S(stream));
}

extern "C" int b200amg_bsr_jacobi(const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
                                  double *temp, int row_start, int row_stop, int row_step, int blocksize,
                                  double omega, void *stream)
{
    if (row_step != 1 || row_start < 0 || row_stop < row_start)
        return set_err(B200AMG_EINVAL, "bsr_jacobi: only forward contiguous ranges (row_step == 1)");
    if (blocksize < 1 || blocksize > 8) return set_err(B200AMG_EINVAL, "bsr_jacobi: blocksize must be in 1..8");
    if (row_stop == row_start) return 0;
    // the reference copies the LEADING |row_stop-row_start|*bs entries (relaxation.h:301-303)
    const size_t ncopy = (size_t)(row_stop - row_start) * blocksize;
    B200_CUDA(cudaMemcpyAsync(temp, x, sizeof(double) * ncopy, cudaMemcpyDeviceToDevice, S(stream)));
    if (blocksize == 1) {
        StreamArgs a{Ap, Aj, Ax, temp, b, x, row_start == 0 ? nnz_lookup(Ap, row_stop, S(stream)) : -1, row_start, row_stop, 1.0, 0.0, omega};
        return stream_op(OP_JACOBI_BSR1, a, S(stream));
    }
    return bsr_jacobi(blocksize, row_start, row_stop, Ap, Aj, Ax, temp, x, b, omega, S(stream));
}

static int colour_loop(int row_start, int row_stop, int row_step, const int *cp, int K, int *c_first, int *c_last)
{
    // the sweep must cover whole colours: positions [cp[0], cp[K]) forward or backward
    if (row_step == 1 && row_start == cp[0] && row_stop == cp[K]) { *c_first = 0; *c_last = K; return 1; }
    if (row_step == -1 && row_start == cp[K] - 1 && row_stop == cp[0] - 1) { *c_first = K - 1; *c_last = -1; return -1; }
    return 0;
}

extern "C" int b200amg_gauss_seidel_indexed(const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
                                            const int *Id, int row_start, int row_stop, int row_step,
                                            const int *color_ptr_host, int K, void *stream)
{
    if (row_step == 0) return set_err(B200AMG_EINVAL, "gauss_seidel_indexed: row_step == 0");
    if (!color_ptr_host) return gs_sequential(false, Ap, Aj, Ax, x, b, Id, row_start, row_stop, row_step, S(stream));
    int c0, c1;
    const int dir = colour_loop(row_start, row_stop, row_step, color_ptr_host, K, &c0, &c1);
    if (!dir) return set_err(B200AMG_EINVAL, "gauss_seidel_indexed: sweep range must cover whole colours");
    for (int c = c0; c != c1; c += dir)
        B200_TRY(gs_rows(false, Ap, Aj, Ax, x, b, Id, color_ptr_host[c], color_ptr_host[c + 1], S(stream)));
    return 0;
}

extern "C" int b200amg_gauss_seidel(const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
                                    int row_start, int row_stop, int row_step, void *stream)
{
    if (row_step == 0) return set_err(B200AMG_EINVAL, "gauss_seidel: row_step == 0");
    return gs_sequential(false, Ap, Aj, Ax, x, b, 0, row_start, row_stop, row_step, S(stream));
}

extern "C" int b200amg_bsr_gauss_seidel(const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
                                        int row_start, int row_stop, int row_step, int blocksize, const int *Id,
                                        const int *color_ptr_host, int K, void *stream)
{
    if (row_step == 0) return set_err(B200AMG_EINVAL, "bsr_gauss_seidel: row_step == 0");
    if (blocksize < 1 || blocksize > 8) return set_err(B200AMG_EINVAL, "bsr_gauss_seidel: blocksize must be in 1..8");
    if (!color_ptr_host) {
        if (blocksize == 1) return gs_sequential(true, Ap, Aj, Ax, x, b, Id, row_start, row_stop, row_step, S(stream));
        return bsr_gs_sequential(blocksize, Ap, Aj, Ax, x, b, Id, row_start, row_stop, row_step, S(stream));
    }
    if (!Id) return set_err(B200AMG_EINVAL, "bsr_gauss_seidel: colour pointer without an index list");
    int c0, c1;
    const int dir = colour_loop(row_start, row_stop, row_step, color_ptr_host, K, &c0, &c1);
    if (!dir) return set_err(B200AMG_EINVAL, "bsr_gauss_seidel: sweep range must cover whole colours");
    // schedule hint: blocks per block row over the first cp[K] rows (all rows when Id is a permutation; cached read)
    const int nrows_hint = color_ptr_host[K];
    const double bpr = (blocksize > 1 && nrows_hint > 0) ? (double)nnz_lookup(Ap, nrows_hint, S(stream)) / nrows_hint : -1.0;
    for (int c = c0; c != c1; c += dir) {
        const int p0 = color_ptr_host[c], p1 = color_ptr_host[c + 1];
        if (blocksize == 1) B200_TRY(gs_rows(true, Ap, Aj, Ax, x, b, Id, p0, p1, S(stream)));
        else B200_TRY(bsr_gs_color(blocksize, Ap, Aj, Ax, x, b, Id, p0, p1, dir < 0, S(stream), bpr));
    }
    return 0;
}

extern "C" int b200amg_dot(long long n, const double *x, const double *y, double *out_dev, double *work, void *stream)
{
    return dot(n, x, y, out_dev, work, S(stream));
}

extern "C" int b200amg_dot_noalloc(long long n, const double *x, const double *y, double *out_dev, void *stream)
{
    return dot_noalloc(n, x, y, out_dev, S(stream));
}

extern "C" int b200amg_pcg_scalars(double *state, int which, void *stream)
{
    if (!state || which < 0 || which > 1) return set_err(B200AMG_EINVAL, "pcg_scalars: bad arguments");
    return pcg_scalars(state, which, S(stream));
}

extern "C" int b200amg_pcg_update_xr(long long n, const double *state, const double *p, const double *Ap, double *x,
                                     double *r, int update_r, void *stream)
{
    return pcg_update_xr(n, state, p, Ap, x, r, update_r, S(stream));
}

extern "C" int b200amg_pcg_update_p(long long n, const double *state, const double *z, double *p, void *stream)
{
    return pcg_update_p(n, state, z, p, S(stream));
}

extern "C" int b200amg_axpy(long long n, double a, const double *x, double *y, void *stream)
{
    return axpy(n, a, x, y, S(stream));
}

extern "C" int b200amg_scal(long long n, double a, const double *x, double *y, void *stream)
{
    return scal(n, a, x, y, S(stream));
}

extern "C" int b200amg_scale_diag(long long n, double omega, const double *b, const double *diag, double *x,
                                  int bsr_flavor, void *stream)
{
    return scale_diag(n, omega, b, diag, x, bsr_flavor, S(stream));
}

extern "C" int b200amg_extract_diag(int n_brow, int blocksize, const int *Ap, const int *Aj, const double *Ax,
                                    double *diag, void *stream)
{
    if (blocksize < 1) return set_err(B200AMG_EINVAL, "extract_diag: blocksize < 1");
    return extract_diag(n_brow, blocksize, Ap, Aj, Ax, diag, S(stream));
}

extern "C" int b200amg_dense_matvec(int n, const double *M, const double *x, double *y, void *stream)
{
    return dense_matvec(n, M, x, y, S(stream));
}

// ---- operator handles: one level operator bound to its best layout ---------------------------
struct OpHandle {
    int n_row;
    long long nnz;
    const int *Ap, *Aj;
    const double *Ax;
    Sell sell;
    ValueDict vd;
    RowPat rp;
    int R = 1, C = 1;                // block shape (b200amg_bsr_op_create); n_row counts BLOCK rows then
    Sell sellJ;                      // diagonal-free copy for Jacobi on a big long-row operator, built at the first such call
    bool triedJ = false;
};

// Block operator behind the same handle: n_brow block rows of R x C blocks (row-major), nnz = stored blocks.  The
// row-partitioned multi-GPU cycle applies local block operators to [owned | halo] vectors through b200amg_op_apply
// (MATVEC / RESIDUAL / ADD: scipy bsr_matvec arithmetic; JACOBI and JACOBI_BSR1: amg_core.bsr_jacobi from x to y).
extern "C" int b200amg_bsr_op_create(int n_brow, int R, int C, const int *Ap, const int *Aj, const double *Ax,
                                     void **handle_out, void *stream)
{
    if (n_brow < 0 || !handle_out || R < 1 || C < 1 || R > 8 || C > 8) return set_err(B200AMG_EINVAL, "bsr_op_create: bad arguments");
    OpHandle *h = new OpHandle{n_brow, -1, Ap, Aj, Ax, Sell(), ValueDict(), RowPat()};
    h->R = R;
    h->C = C;
    h->nnz = n_brow > 0 ? nnz_lookup(Ap, n_brow, S(stream)) : 0;
    *handle_out = h;
    return 0;
}

extern "C" int b200amg_op_create(int n_row, const int *Ap, const int *Aj, const double *Ax, int layout,
                                 void **handle_out, void *stream)
{
    if (n_row < 0 || !handle_out || layout < 0 || layout > 2) return set_err(B200AMG_EINVAL, "op_create: bad arguments");
    OpHandle *h = new OpHandle{n_row, -1, Ap, Aj, Ax, Sell(), ValueDict(), RowPat()};
    if (n_row > 0) {
        int ends[2] = {0, 0};
        cudaError_t e = cudaMemcpyAsync(&ends[0], Ap, sizeof(int), cudaMemcpyDeviceToHost, S(stream));
        if (e == cudaSuccess) e = cudaMemcpyAsync(&ends[1], Ap + n_row, sizeof(int), cudaMemcpyDeviceToHost, S(stream));
        if (e == cudaSuccess) e = cudaStreamSynchronize(S(stream));
        if (e != cudaSuccess) { delete h; return set_err((int)e, "op_create: cannot read row pointers"); }
        h->nnz = (long long)ends[1] - ends[0];
        if (layout == 2 || (layout == 0 && sell_wanted(h->nnz, n_row))) {
            // value slots as 8-bit codes when the operator has <= 256 distinct values (B200AMG_VALUE_DICT=0: never)
            const char *ev = getenv("B200AMG_VALUE_DICT");
            ValueDict vd;
            int rc = (ev && ev[0] == '0') ? 0 : dict_build(h->nnz, Ax, &vd, S(stream));
            if (!rc) rc = sell_build(n_row, h->nnz, Ap, Aj, Ax, &h->sell, S(stream), nullptr, vd.codes, vd.dict);
            dict_free(&vd);
            if (rc) { delete h; return rc; }
            if (layout == 2 && !h->sell.ok()) { delete h; return set_err(B200AMG_ESTATE, "op_create: sliced-ELL layout declined (padding too large)"); }
        }
        const char *env = getenv("B200AMG_VALUE_DICT");
        const int dict_level = (env && env[0] >= '0' && env[0] <= '3') ? env[0] - '0' : 3;
        if (layout == 0 && !h->sell.ok() && (double)h->nnz / n_row <= 8.0 && dict_level) {
            // each layout is declined silently when the operator does not qualify (too many distinct row
            // patterns / (offset, value) pairs / values); the next one down is tried
            int rc = dict_level >= 3 ? rowpat_build(n_row, h->nnz, Ap, Aj, Ax, &h->rp, S(stream)) : 0;
            if (!rc && !h->rp.ok() && dict_level >= 2) rc = pattern_build(n_row, h->nnz, Ap, Aj, Ax, &h->vd, S(stream));
            if (!rc && !h->rp.ok() && !h->vd.ok()) rc = dict_build(h->nnz, Ax, &h->vd, S(stream));
            if (rc) { delete h; return rc; }
        }
    }
    *handle_out = h;
    return 0;
}

extern "C" int b200amg_op_apply(void *handle, int op, const double *x, const double *b, double *y, double omega,
                                void *stream)
{
    OpHandle *h = reinterpret_cast<OpHandle *>(handle);
    if (!h || op < 0 || op > OP_JACOBI_BSR1 || op == OP_SPMV_AB) return set_err(B200AMG_EINVAL, "op_apply: bad arguments");
    if (h->n_row == 0) return 0;
    if (h->R != 1 || h->C != 1) {
        switch (op) {
        case OP_MATVEC: return bsr_matvec(h->n_row, h->R, h->C, h->Ap, h->Aj, h->Ax, x, nullptr, y, 0, S(stream), h->nnz);
        case OP_RESIDUAL: return bsr_matvec(h->n_row, h->R, h->C, h->Ap, h->Aj, h->Ax, x, b, y, 1, S(stream), h->nnz);
        case OP_ADD: return bsr_matvec(h->n_row, h->R, h->C, h->Ap, h->Aj, h->Ax, x, b, y, 2, S(stream), h->nnz);
        default:
            if (h->R != h->C) return set_err(B200AMG_EINVAL, "op_apply: Jacobi needs square blocks");
            return bsr_jacobi(h->R, 0, h->n_row, h->Ap, h->Aj, h->Ax, x, y, b, omega, S(stream));
        }
    }
    StreamArgs a{h->Ap, h->Aj, h->Ax, x, b, y, h->nnz, 0, h->n_row, 1.0, 0.0, omega};
    h->vd.apply(a);
    if (h->rp.ok()) return rowpat_op(op, h->rp, a, S(stream));
    if (h->sell.ok() && op >= OP_JACOBI && !h->sell.cc && h->n_row >= 400000) {
        // the Jacobi sweeps of a big long-row operator stream a copy without the diagonal entries (vcycle.cu has the
        // story); built at the first such call outside stream capture (the build allocates and synchronises)
        if (!h->triedJ) {
            cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
            const char *ev = getenv("B200AMG_SELL_DIAGFREE");
            if (cudaStreamIsCapturing(S(stream), &cs) == cudaSuccess && cs == cudaStreamCaptureStatusNone) {
                h->triedJ = true;
                if (!(ev && ev[0] == '0')) {
                    const int rc = sell_build(h->n_row, h->nnz, h->Ap, h->Aj, h->Ax, &h->sellJ, S(stream), nullptr, nullptr, nullptr, true);
                    if (rc) return rc;
                }
            }
        }
        if (h->sellJ.ok()) return sell_op(op, h->sellJ, a, S(stream));
    }
    if (h->sell.ok()) return sell_op(op, h->sell, a, S(stream));
    return stream_op(op, a, S(stream));
}

extern "C" int b200amg_op_layout(void *handle)
{
    OpHandle *h = reinterpret_cast<OpHandle *>(handle);
    if (!h || h->R != 1 || h->C != 1) return 0;
    if (h->rp.ok()) return 5;
    if (h->sell.ok()) return h->sell.cc ? 6 : (h->sell.cd ? 7 : 2);
    return h->vd.ok() ? (h->vd.offs ? 4 : 3) : 1;
}

extern "C" int b200amg_op_destroy(void *handle)
{
    OpHandle *h = reinterpret_cast<OpHandle *>(handle);
    if (!h) return 0;
    sell_free(&h->sell);
    sell_free(&h->sellJ);
    dict_free(&h->vd);
    rowpat_free(&h->rp);
    delete h;
    return 0;
}
