// Cross-file declarations inside libb200amg.
#pragma once
#include "common.cuh"

namespace b200amg {

enum : int {
    OP_MATVEC = 0, OP_SPMV_AB, OP_RESIDUAL, OP_ADD, OP_JACOBI, OP_JACOBI_BSR1
};

struct StreamArgs {
    const int *Ap, *Aj;
    const double *Ax;
    const double *x;
    const double *b;
    double *y;
    long long nnz;         // stored entries of the rows in range, or -1 when the caller does not know
    int row_begin, row_end;
    double alpha, beta, omega;
    const unsigned char *codes = nullptr;   // optional value dictionary (engine-built): Ax[k] == dict[codes[k]]
    const double *dict = nullptr;           // 256 doubles
    // optional pattern dictionary (square operators): codes index (column offset, value) pairs,
    // Aj[k] == row + offs[codes[k]] as well -- a constant-coefficient stencil then costs 1 byte per entry
    const int *offs = nullptr;
};

// value dictionary of an operator with <= 256 distinct values (csr_pipe.cu DICT path); codes == nullptr: declined
struct ValueDict {
    unsigned char *codes = nullptr;
    double *dict = nullptr;
    int *offs = nullptr;             // pattern dictionary only: column offset (Aj - row) per code
    bool ok() const { return codes != nullptr; }
    void apply(StreamArgs &a) const { a.codes = codes; a.dict = dict; a.offs = offs; }
};
int dict_build(long long nnz, const double *Ax, ValueDict *out, cudaStream_t st);
// (column offset, value) pairs of a square operator; declined when there are more than 256 distinct pairs
int pattern_build(int n_row, long long nnz, const int *Ap, const int *Aj, const double *Ax, ValueDict *out,
                  cudaStream_t st);
void dict_free(ValueDict *d);

// ---- row-pattern layout (rowpat.cu): one byte per row = number of the row's (offset, value) sequence in a
// table of <= 255 patterns of <= 8 entries (constant-coefficient stencils); codes == nullptr: declined
struct __align__(16) RowPatEntry {
    double v;
    int off, len;                    // column offset (Aj - row); len = entries of the whole pattern
};
constexpr int RP_TILE_MAXSEG = 8;    // contiguous x segments a chunk of rows gathers from (rowpat_tile.cu)
constexpr int RP_TILE_LEN = 8;       // == RP_LEN
struct RowPat {
    unsigned char *codes = nullptr;  // [n]
    RowPatEntry *table = nullptr;    // [npat * 8]
    int npat = 0, n = 0, far = 0;    // far = largest column offset in the table
    int xlen = 0;                    // entries of x the operator may reference (> n on a row slab with halo columns)
    int common_len = 0;              // entries of the typical (middle) row
    // plan of the TMA-staged sweep (rowpat_tile.cu), made from the typical row's pattern; tile_ok == false: not eligible
    bool tile_ok = false;
    int tile_len = 0, nseg = 0, xs_elems = 0, tile_stages = 0;
    int seg_lo[RP_TILE_MAXSEG] = {}, seg_len[RP_TILE_MAXSEG] = {}, seg_base[RP_TILE_MAXSEG] = {};
    int sbase[RP_TILE_LEN] = {0, 0, 0, 0, 0, 0, 0, 0};
    double cval[RP_TILE_LEN] = {0, 0, 0, 0, 0, 0, 0, 0};
    unsigned short *tmask = nullptr; // [256] device: presence mask of every pattern relative to the typical one
    bool ok() const { return codes != nullptr; }
};
int rowpat_build(int n_row, long long nnz, const int *Ap, const int *Aj, const double *Ax, RowPat *out, cudaStream_t st);
void rowpat_free(RowPat *r);
int rowpat_op(int op, const RowPat &r, const StreamArgs &a, cudaStream_t st);
// TMA-staged sweep (rowpat_tile.cu): -1 = not eligible (rowpat.cu's kernel runs), else the launch status
int rowpat_tile_op(int op, const RowPat &r, const StreamArgs &a, cudaStream_t st);

int rows_per_cta_for(long long nnz, int nrows);
// persistent TMA pipeline (csr_pipe.cu): returns 1 when it took the launch (*rc = status), 0 when
// the case is not eligible (unknown nnz, unaligned arrays) and csr_stream must run instead
int pipe_op(int op, const StreamArgs &a, long long nnz, cudaStream_t st, int *rc);
// ---- sliced-ELL copy of a long-row CSR operator (sell.cu): slices of 32 rows, column-major inside a
// slice, so that thread-per-row loads of (Aj, Ax) coalesce and neighbouring rows gather neighbouring x
struct Sell {
    int n = 0;                       // rows
    long long padded = 0;            // stored slots (32 * sum of slice widths)
    long long *slice_off = nullptr;  // [n_slices + 1] slot offset of every slice
    int *cj = nullptr;               // [padded]   32-bit columns, or
    unsigned short *cd = nullptr;    // [padded]   16-bit column steps: column of slot k = first[row] + cd[0..k] (sorted rows)
    int *first = nullptr;            // [n]        first stored column of every row (with cd)
    double *dg = nullptr;            // [n]        copies for smoothers: the diagonal entries are NOT among the slots,
                                     //            dg[slot row] = the last stored one (0.0: none)
    double *cv = nullptr;            // [padded]   (fp64 values) or
    unsigned char *cc = nullptr;     // [padded]   8-bit value codes into dict[256]
    double *dict = nullptr;
    int *row_len = nullptr;          // [n]
    bool ok() const { return cj != nullptr || cd != nullptr; }
};
int sell_build(int n, long long nnz, const int *Ap, const int *Aj, const double *Ax, Sell *out, cudaStream_t st,
               const int *perm = nullptr, const unsigned char *codes = nullptr, const double *dict = nullptr,
               bool drop_diag = false);
// one colour [p0, p1) of a colour-grouped copy (slot row r = matrix row Id[r]); x updated in place
int sell_gs_color(bool bsrf, const Sell &s, const int *Id, int p0, int p1, double *x, const double *b, cudaStream_t st);
void sell_free(Sell *s);
// same operations / arithmetic as stream_op, on the sliced-ELL copy
int sell_op(int op, const Sell &s, const StreamArgs &a, cudaStream_t st);
bool sell_wanted(long long nnz, int nrows);

// nnz of rows [0, n_row) for the stand-alone entry points: one 4-byte read per (Ap, n_row), cached
long long nnz_lookup(const int *Ap, int n_row, cudaStream_t st);

int stream_op(int op, const StreamArgs &a, cudaStream_t st);
int gs_rows(bool bsrf, const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
            const int *Id, int p0, int p1, cudaStream_t st);
// a whole colour-ordered smoothing step of a small level in ONE launch (small_level.cu): -1 = not eligible
int gs_level_cta(int bs, bool bsrf, int n, long long stored_values, const int *Ap, const int *Aj, const double *Ax, const int *Id,
                 const int *cptr_dev, int K, double *x, const double *b, int iters, int sweep, bool x_is_zero, cudaStream_t st);
int gs_sequential(bool bsrf, const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
                  const int *Id, int start, int stop, int step, cudaStream_t st);

// nblocks / blocks_per_row (stored blocks; < 0: unknown) pick the schedule: warp per block row for long rows
int bsr_matvec(int n_brow, int R, int C, const int *Ap, const int *Aj, const double *Ax, const double *x,
               const double *b, double *y, int mode, cudaStream_t st, long long nblocks = -1);
int bsr_jacobi(int bs, int nb0, int nb1, const int *Ap, const int *Aj, const double *Ax, const double *xin,
               double *xout, const double *b, double omega, cudaStream_t st);
int bsr_gs_color(int bs, const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
                 const int *Id, int p0, int p1, bool backward, cudaStream_t st, double blocks_per_row = -1.0,
                 bool perm = false, bool static_op = false);    // static_op: no kernel in flight writes (Ap, Aj, Ax, Id)
// colour-grouped copy of a BSR operator for colour-ordered block Gauss-Seidel (row p = block row Id[p]); rp == nullptr: none
struct BsrColor {
    int *rp = nullptr, *cj = nullptr;
    double *cx = nullptr;
    bool ok() const { return rp != nullptr; }
};
int bsr_color_build(int nb, int bs, long long nblocks, const int *Ap, const int *Aj, const double *Ax, const int *Id,
                    BsrColor *out, cudaStream_t st);
void bsr_color_free(BsrColor *c);
int bsr_gs_sequential(int bs, const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
                      const int *Id, int start, int stop, int step, cudaStream_t st);

// block relaxation with stored diagonal-block inverses (block_relax.cu); block_jacobi_pp reads xin, writes xout
int block_jacobi_pp(int bs, int nb, const int *Ap, const int *Aj, const double *Ax, const double *xin, double *xout,
                    const double *b, const double *Dinv, double omega, cudaStream_t st);
int block_gs_color(int bs, const int *Ap, const int *Aj, const double *Ax, double *x, const double *b, const double *Dinv,
                   const int *Id, int p0, int p1, cudaStream_t st);
int block_gs_sequential(int bs, const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
                        const double *Dinv, const int *Id, int start, int stop, int step, cudaStream_t st);

int block_diag_inverse_gj(int n_brow, int bs, const int *Ap, const int *Aj, const double *Ax, double *Dinv, cudaStream_t st);
// amg_core.pinv_array on the device (pinv.cu, compiled without FMA contraction), n <= 6
int pinv_array(double *AA, int m, int n, int transA, cudaStream_t st);

int dot(long long n, const double *x, const double *y, double *out, double *work, cudaStream_t st);
int dot_noalloc(long long n, const double *x, const double *y, double *out, cudaStream_t st);   // scratch inside the library
int pcg_scalars(double *st, int which, cudaStream_t s);
int pcg_update_xr(long long n, const double *st, const double *p, const double *Ap, double *x, double *r, int update_r,
                  cudaStream_t s);
int pcg_update_p(long long n, const double *st, const double *z, double *p, cudaStream_t s);
int axpy(long long n, double a, const double *x, double *y, cudaStream_t st);
int scal(long long n, double a, const double *x, double *y, cudaStream_t st);
int extract_diag(int nb, int bs, const int *Ap, const int *Aj, const double *Ax, double *diag, cudaStream_t st);
int scale_diag(long long n, double omega, const double *b, const double *d, double *x, int bsrf, cudaStream_t st);
int dense_matvec(int n, const double *M, const double *x, double *y, cudaStream_t st);

}  // namespace b200amg
