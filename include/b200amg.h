/*
 * b200amg.h -- C-ABI of libb200amg.so: the B200 (sm_100a) drop-in for the solve-phase V-cycle
 * hot path of PyAMG (reference: rsmedleystevenson/pyamg, paths below relative to its root).
 *
 * Conventions
 *   - every array pointer is a DEVICE pointer (fp64 values, int32 indices) unless the
 *     parameter name ends in _host; the caller owns all buffers, nothing is retained except
 *     by the b200amg_vcycle_* handle (which only stores the pointers it is given);
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - return value: 0 on success, otherwise a cudaError_t value (>0) or a negative
 *     B200AMG_E* code; b200amg_last_error() gives the text;
 *   - kernels mutate x (and temp) in place and return nothing else, exactly like the
 *     reference's amg_core functions (pyamg/amg_core/amg_core.i:27-92, INPLACE_ARRAY1).
 *   - fp64 results are BIT-IDENTICAL to the reference's serial loops wherever the parallel
 *     schedule is order-free (SpMV, residual, prolongation, Jacobi, colour-ordered
 *     Gauss-Seidel, Galerkin product): every row is accumulated in CSR storage order with
 *     separately rounded multiply and add, like the reference's non-FMA x86-64 build.
 */
#ifndef B200AMG_H
#define B200AMG_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200AMG_EINVAL  (-1)   /* bad argument (negative size, unsupported blocksize, ...) */
#define B200AMG_ENOMEM  (-2)   /* caller-provided workspace too small */
#define B200AMG_ESTATE  (-3)   /* handle used in the wrong state */

const char *b200amg_last_error(void);
int b200amg_version(void);
/* number of kernel launches issued by this library since load (bench.py's gpu_launches) */
long long b200amg_launch_count(void);

/* ---- SpMV family: scipy.sparse `A * x` = _sparsetools csr_matvec / bsr_matvec
 *      call sites pyamg/multilevel.py:496 (b - A*x), :498 (R*residual), :544 (x += P*coarse_x),
 *      pyamg/util/linalg.py:109-112 (residual_norm) ---------------------------------------- */
/* y = A x */
int b200amg_csr_matvec(int n_row, int n_col, const int *Ap, const int *Aj, const double *Ax,
                       const double *x, double *y, void *stream);
/* y = alpha * (A x) + beta * y   (generic helper; not a reference call site) */
int b200amg_csr_spmv(int n_row, int n_col, const int *Ap, const int *Aj, const double *Ax,
                     const double *x, double *y, double alpha, double beta, void *stream);
/* r = b - A x                      (multilevel.py:496 fused with its numpy subtraction) */
int b200amg_csr_residual(int n_row, const int *Ap, const int *Aj, const double *Ax,
                         const double *x, const double *b, double *r, void *stream);
/* x += A xc                        (multilevel.py:544: x += P * coarse_x) */
int b200amg_csr_matvec_add(int n_row, const int *Ap, const int *Aj, const double *Ax,
                           const double *xc, double *x, void *stream);
/* BSR with R x C row-major blocks (scipy bsr_matvec); mode 0: y = A x, 1: y = b - A x,
 * 2: y += A x.  R, C in 1..8. */
int b200amg_bsr_matvec(int n_brow, int n_bcol, int R, int C, const int *Ap, const int *Aj,
                       const double *Ax, const double *x, const double *b, double *y, int mode,
                       void *stream);

/* ---- smoothers: amg_core signatures (pyamg/amg_core/relaxation.h) --------------------- */
/* amg_core.jacobi, relaxation.h:201-239.  omega is passed by value (the reference passes a
 * length-1 array, relaxation.py:412).  temp must hold n doubles. Only the contiguous forward
 * range the reference's Python wrapper uses (row_step == 1) is accepted. */
int b200amg_jacobi(const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
                   double *temp, int row_start, int row_stop, int row_step, double omega,
                   void *stream);
/* amg_core.bsr_jacobi, relaxation.h:267-360 (blocksize 1..8, row_step == 1) */
int b200amg_bsr_jacobi(const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
                       double *temp, int row_start, int row_stop, int row_step, int blocksize,
                       double omega, void *stream);
/* amg_core.gauss_seidel_indexed, relaxation.h:394-426.
 * color_ptr_host (HOST array of K+1 offsets into Id) declares that positions
 * [color_ptr[c], color_ptr[c+1]) of Id are mutually independent rows: each colour is then
 * relaxed in parallel, colours in sweep order -- bit-identical to the sequential sweep.
 * With color_ptr_host == NULL the sweep is executed strictly sequentially on the GPU
 * (one thread block), so arbitrary index lists with repeats keep reference semantics. */
int b200amg_gauss_seidel_indexed(const int *Ap, const int *Aj, const double *Ax, double *x,
                                 const double *b, const int *Id, int row_start, int row_stop,
                                 int row_step, const int *color_ptr_host, int K, void *stream);
/* amg_core.gauss_seidel, relaxation.h:33-62: lexicographic sweep, executed sequentially on
 * the GPU (one thread block; exact, slow) -- the parallel forms are the indexed / level ones. */
int b200amg_gauss_seidel(const int *Ap, const int *Aj, const double *Ax, double *x,
                         const double *b, int row_start, int row_stop, int row_step,
                         void *stream);
/* amg_core.bsr_gauss_seidel, relaxation.h:89-173 (blocksize 1..8).  Id/color_ptr_host as
 * above but over BLOCK rows; Id == NULL: the plain (row_start,row_stop,row_step) range,
 * executed sequentially. */
int b200amg_bsr_gauss_seidel(const int *Ap, const int *Aj, const double *Ax, double *x,
                             const double *b, int row_start, int row_stop, int row_step,
                             int blocksize, const int *Id, const int *color_ptr_host, int K,
                             void *stream);

/* ---- vector kernels: pyamg/util/linalg.py:17-54 (norm), :109-112 (residual_norm) ------ */
/* out_dev[0] = sum_i x_i*y_i  (deterministic two-stage reduction; work: >= 1024 doubles) */
int b200amg_dot(long long n, const double *x, const double *y, double *out_dev, double *work,
                void *stream);
/* b200amg_dot without any caller-provided or allocated scratch.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate (code under CUDA graph capture, serving loops, harnesses that
 * forbid cudaMalloc).  The partial sums live in a file-scope `static __device__ __align__(256)` array of 1024 doubles
 * (8 KB, held by every GPU that loads the library until exit); the call allocates nothing (no cudaMalloc / cudaFree /
 * allocating library routine).  Limit: n <= 268 435 456 (2^28) elements, larger n returns B200AMG_EINVAL with a message.
 * Same kernels and reduction order as b200amg_dot: the result is bit-identical to it.  All calls on a device share the
 * array (nothing needs clearing: every cell read is written first by the same call), so calls that OVERLAP on the GPU
 * (different streams without ordering) race -- serialise them. */
int b200amg_dot_noalloc(long long n, const double *x, const double *y, double *out_dev, void *stream);
/* ---- preconditioned CG with its scalars on the device (pyamg/krylov/_cg.py:109-170; multilevel.py:381-422 accel='cg').
 * `state`: >= 16 + capacity doubles on the device: [0] rz [1] pAp [2] rz_new [3] alpha [4] beta [5] tol (absolute, in
 * the preconditioned norm sqrt(<r, z>)) [6] flag [7] iterations done [8] maxiter [9] capacity of the history
 * [16 + k] sqrt(rz) after iteration k.  The caller writes [0], [5], [7] = 0, [8], [9], [16] and flag = 0, then enqueues
 * per iteration: Ap = A p; b200amg_dot(Ap, p) -> state + 1; b200amg_pcg_scalars(state, 0); b200amg_pcg_update_xr;
 * [r = b - A x on the iterations that recompute it]; z = M r; b200amg_dot(r, z) -> state + 2;
 * b200amg_pcg_scalars(state, 1); b200amg_pcg_update_p.  flag: 0 running, 1 converged, 2 maxiter, -1 / -2 / -3 the
 * reference's three abort conditions (indefinite matrix, indefinite / singular preconditioner).  After flag != 0
 * every kernel is a no-op, so iterations may be enqueued ahead of the host's look at the flag. */
int b200amg_pcg_scalars(double *state, int which, void *stream);
/* x += alpha p; update_r != 0: r -= alpha Ap (separately rounded, numpy's `x += alpha * p`) */
int b200amg_pcg_update_xr(long long n, const double *state, const double *p, const double *Ap, double *x, double *r,
                          int update_r, void *stream);
/* p = beta p + z */
int b200amg_pcg_update_p(long long n, const double *state, const double *z, double *p, void *stream);
/* y = a*x + y ; y = x ; x = 0 */
int b200amg_axpy(long long n, double a, const double *x, double *y, void *stream);
/* y = a*x */
int b200amg_scal(long long n, double a, const double *x, double *y, void *stream);
/* one Jacobi sweep from a zero initial guess == diagonal scaling (bit-identical):
 * x = omega*(b/diag) [CSR arithmetic] or (omega*b)/diag [bsr_flavor != 0]; rows with diag==0 -> 0 */
int b200amg_scale_diag(long long n, double omega, const double *b, const double *diag, double *x,
                       int bsr_flavor, void *stream);
/* diag[i*bs+k] = A_ii[k][k] of the LAST stored diagonal block of block row i (0 if absent) */
int b200amg_extract_diag(int n_brow, int blocksize, const int *Ap, const int *Aj, const double *Ax,
                         double *diag, void *stream);
/* y = M x, dense row-major M (n x n): coarse_grid_solver('pinv2'), multilevel.py:608-612 */
int b200amg_dense_matvec(int n, const double *M, const double *x, double *y, void *stream);

/* ---- operator handles: a level operator bound once to its best device layout ---------------
 * layout 0: automatic (sliced-ELL SELL-32 copy for 12..400 entries per row, else the caller's CSR),
 * 1: CSR kernels only, 2: force the sliced-ELL copy.  The handle keeps the caller's pointers
 * (not the arrays) and owns only its SELL copy.  op: 0  y = A x          (scipy csr_matvec)
 *                                                    2  y = b - A x      (multilevel.py:496)
 *                                                    3  y += A x         (multilevel.py:544)
 *                                                    4  y = jacobi(x)    out of place, amg_core.jacobi arithmetic
 *                                                    5  y = jacobi(x)    out of place, amg_core.bsr_jacobi(bs=1) arithmetic
 * Results are bit-identical across layouts.  Used by the row-partitioned multi-GPU cycle
 * (pyamg_b200/distributed.py), where every rank applies LOCAL operators to [owned | halo] vectors. */
/* (The first Jacobi sweep through a handle over >= 400 000 long rows builds a second copy without the diagonal
 * entries -- it allocates and synchronises once, never during stream capture; B200AMG_SELL_DIAGFREE=0 disables it.) */
int b200amg_op_create(int n_row, const int *Ap, const int *Aj, const double *Ax, int layout,
                      void **handle_out, void *stream);
int b200amg_op_apply(void *handle, int op, const double *x, const double *b, double *y, double omega,
                     void *stream);
/* The same handle over a BLOCK operator (n_brow block rows of R x C row-major blocks, 1..8): b200amg_op_apply then
 * runs scipy's bsr_matvec arithmetic (MATVEC / RESIDUAL / ADD; `_sparsetools/bsr.h` bsr_matvec) or, for square
 * blocks, one amg_core.bsr_jacobi sweep from x to y (pyamg/amg_core/relaxation.h:267-360; both JACOBI op codes).
 * Used for the block levels of the row-partitioned multi-GPU cycle; b200amg_op_layout returns 0. */
int b200amg_bsr_op_create(int n_brow, int R, int C, const int *Ap, const int *Aj, const double *Ax,
                          void **handle_out, void *stream);
int b200amg_op_layout(void *handle);     /* 1 CSR, 2 sliced-ELL, 3 CSR + 8-bit value codes, 4 CSR rows + 8-bit
                                           * (column offset, value) pattern codes, 5 one row-pattern
                                           * byte per row (stencil operators), 6 sliced-ELL with
                                           * 8-bit value codes, 7 sliced-ELL with 16-bit column
                                           * steps (sorted rows, >= 400 000 rows) */
int b200amg_op_destroy(void *handle);

/* ---- multi-GPU halo refresh over NVLink peer memory (pyamg_b200/distributed.py) -------
 * Level vectors live in a symmetric allocation (the same buffer mapped on every GPU of the node).
 * push: scatter v[send_idx[j]] to (double*)bases[dst_rank[j]] + dst_off[j] on the peers, then raise this
 *       rank's epoch in the flag cell sig_ptrs[i] of each of the n_nbr neighbours (system-scope release);
 * wait: spin (bounded by max_spins polls per neighbour; state[1] != 0 afterwards = timed out) until every
 *       neighbour's flag cell flag_ptrs[i] reached the next epoch.
 * state: 2 device uint32 per call site, zero-initialised (epoch, scratch) -- epochs live in memory so the
 * pair is CUDA-graph replay safe.  No reference counterpart: the reference is single-process. */
int b200amg_halo_push(const double *v, int n_send, const int *send_idx, const int *dst_rank,
                      const long long *dst_off, const unsigned long long *bases,
                      const unsigned long long *sig_ptrs, int n_nbr, unsigned *state, void *stream);
int b200amg_halo_wait(const unsigned long long *flag_ptrs, int n_nbr, unsigned *state,
                      long long max_spins, void *stream);
/* push + wait as ONE kernel node (the last CTA of the push signals the neighbours, then waits for theirs): what the
 * distributed cycle launches per halo refresh; push_state / wait_state as for the two separate calls. */
int b200amg_halo_exchange(const double *v, int n_send, const int *send_idx, const int *dst_rank,
                          const long long *dst_off, const unsigned long long *bases,
                          const unsigned long long *sig_ptrs, const unsigned long long *flag_ptrs, int n_nbr,
                          unsigned *push_state, unsigned *wait_state, long long max_spins, void *stream);

/* ---- setup-phase: Galerkin product and transpose (scipy csr_matmat / csr_tocsc;
 *      pyamg/aggregation/aggregation.py:429, pyamg/classical/classical.py:174,187) -------- */
/* Row-wise C = A B, exact zeros dropped (scipy drops them too); values bit-identical to scipy's
 * csr_matmat (same accumulation order, no FMA).  scipy_order == 0: rows SORTED by column;
 * scipy_order != 0: every row in scipy's own output order (reverse first touch), so that a
 * chained product (R*A)*P accumulates in exactly the reference's order.
 *  1. b200amg_spgemm_count : ub[i] = number of products of row i;
 *  2. b200amg_exclusive_scan_i64(ub) -> ubp[n_row+1] (scratch offsets, ubp[n_row] = total);
 *  3. b200amg_spgemm_numeric: every row into the scratch Tj/Tx (both sized
 *     ubp[n_row]) at offset ubp[i], cnt[i] = kept entries.  Tk (u64[ubp[n_row]]) is only
 *     needed when some row has more than 8192 products (else NULL).  work: n_row + 128 ints.
 *     Synchronises the stream once (row binning).  Rows of up to 8192 products in scipy's order are
 *     accumulated without a sort (spgemm_hash.cu; B200AMG_SPGEMM_HASH=0 switches that off), every other
 *     row by expand-sort-compress; the results are the same to the bit.
 *  4. b200amg_exclusive_scan_i32(cnt) -> Cp[n_row+1];
 *  5. b200amg_spgemm_compact: copy the kept entries to Cj/Cx. */
int b200amg_spgemm_count(int n_row, const int *Ap, const int *Aj, const int *Bp, long long *ub,
                         void *stream);
int b200amg_spgemm_numeric(int n_row, const int *Ap, const int *Aj, const double *Ax,
                           const int *Bp, const int *Bj, const double *Bx, const long long *ub,
                           const long long *ubp, unsigned long long *Tk, int *Tj, double *Tx,
                           int *cnt, int *work, int scipy_order, void *stream);
int b200amg_spgemm_compact(int n_row, const long long *ubp, const int *Tj, const double *Tx,
                           const int *Cp, int *Cj, double *Cx, void *stream);
/* out[0..n] = exclusive prefix sums of in[0..n) (out[n] = total);
 * scratch: b200amg_scan_scratch_elems(n) elements of the output type */
int b200amg_exclusive_scan_i64(long long n, const long long *in, long long *out, long long *scratch,
                               void *stream);
int b200amg_exclusive_scan_i32(long long n, const int *in, int *out, int *scratch, void *stream);
long long b200amg_scan_scratch_elems(long long n);
/* B = A^T in CSR; every row of B ascending in the original row index (like csr_tocsc).
 * work: b200amg_csr_transpose_work_bytes(...) bytes.  Synchronises the stream once. */
long long b200amg_csr_transpose_work_bytes(int n_row, int n_col, long long nnz);
int b200amg_csr_transpose(int n_row, int n_col, long long nnz, const int *Ap, const int *Aj,
                          const double *Ax, int *Bp, int *Bi, double *Bx, void *work, void *stream);

/* ---- scalar smoothed-aggregation prolongator pieces (setup; bit-identical op order) ------- */
/* out[k] = (Ax[k] * v[row]) [* c]: scipy csr_scale_rows (pyamg/util/utils.py:133-200) followed
 * by the scalar factor of pyamg/aggregation/smooth.py:164 when use_c != 0 */
int b200amg_csr_scale_rows(int n_row, const int *Ap, const double *Ax, const double *v, double c,
                           int use_c, double *out, void *stream);
/* amg_core.fit_candidates for ONE candidate and scalar unknowns (smoothed_aggregation.h:322-448):
 * (Mp, Mi, Mb) = members of each aggregate (ascending node, candidate value);
 * Bc[j] = ||b_j||, node_val[i] = b_i / ||b_agg(i)|| */
int b200amg_tentative_scalar(int n_agg, const int *Mp, const int *Mi, const double *Mb, double tol,
                             double *Bc, double *node_val, void *stream);
int b200amg_flag_nonneg(int n, const int *agg, int *flag, void *stream);
int b200amg_tentative_fill(int n, const int *agg, const double *node_val, const int *Tp, int *Tj,
                           double *Tx, void *stream);
/* C = A - B for sorted duplicate-free CSR rows, zero results not stored (scipy csr_binop_csr);
 * pass 0 writes cnt[i]; pass 1 fills Cj/Cx at Cp[i] */
int b200amg_csr_sub_sorted(int n_row, const int *Ap, const int *Aj, const double *Ax, const int *Bp,
                           const int *Bj, const double *Bx, int pass, int *cnt, const int *Cp,
                           int *Cj, double *Cx, void *stream);

/* the same for rows in any order (scipy csr_binop_csr_general): output in reverse first-touch
 * order over (A's row, then B's row) -- the storage order the reference's P = T - U ends up with */
int b200amg_csr_sub_general(int n_row, const int *Ap, const int *Aj, const double *Ax, const int *Bp,
                            const int *Bj, const double *Bx, int pass, int *cnt, const int *Cp,
                            int *Cj, double *Cx, void *stream);

/* ---- symmetric strength of connection (csrc/strength.cu): amg_core.symmetric_strength_of_connection,
 * pyamg/amg_core/smoothed_aggregation.h:48-99, for theta > 0 (theta == 0 keeps A's pattern) ------------- */
/* one value per block = sum of its squared entries in numpy's sum(axis=1) order (pyamg/strength.py:300-305) */
int b200amg_block_sqnorm(long long n_blocks, int R, int C, const double *Ax, double *out, void *stream);
/* diag[i] = |sum of the stored (i, i) entries| (smoothed_aggregation.h:62-71) */
int b200amg_abs_diag(int n_row, const int *Ap, const int *Aj, const double *Ax, double *diag, void *stream);
/* strong entries: the diagonal, and a_ij with a_ij^2 >= theta^2 |a_ii| |a_jj|; pass 0 writes cnt[i], pass 1 fills
 * Sj/Sx at Sp[i] in storage order */
int b200amg_symmetric_strength(int n_row, double theta, const int *Ap, const int *Aj, const double *Ax,
                               const double *diag, int pass, int *cnt, const int *Sp, int *Sj, double *Sx,
                               void *stream);

/* ---- block (BSR) smoothed-aggregation setup (csrc/bsr_setup.cu) -- BASELINE config 4 -------------- */
/* scipy bsr_scale_rows (pyamg/util/utils.py:133-200): out[blk][r][c] = Ax[blk][r][c] * v[R*i + r], then
 * times the scalar c of pyamg/aggregation/smooth.py:164 when use_c != 0 */
int b200amg_bsr_scale_rows(int n_brow, int R, int C, const int *Ap, const double *Ax, const double *v,
                           double c, int use_c, double *out, void *stream);
/* out block q = in block perm[q] (R x C, row-major), transposed to C x R when transpose_blocks != 0.
 * perm = the VALUE array of b200amg_csr_transpose run over block numbers 0..n_blocks-1 stored as doubles:
 * scipy bsr_transpose (P.H, pyamg/aggregation/aggregation.py:400-401) and bsr sort_indices
 * (pyamg/util/utils.py:571) */
int b200amg_bsr_gather_blocks(long long n_blocks, int R, int C, const double *perm, const double *in,
                              double *out, int transpose_blocks, void *stream);
/* amg_core.fit_candidates (pyamg/amg_core/smoothed_aggregation.h:322-448) for K1 unknowns per node and K2
 * candidates: (Mp, Mi) = members of each aggregate, ascending (AggOp in CSC form); B[node][K1][K2];
 * Qx[member][K1][K2] (CSC member order) receives the orthonormalised candidates, Rx[agg][K2][K2] the
 * coarse candidates */
int b200amg_fit_candidates(int n_agg, int K1, int K2, const int *Mp, const int *Mi, const double *B,
                           double tol, double *Qx, double *Rx, void *stream);
/* Q.T.tobsr() of pyamg/aggregation/tentative.py:158-161: block of member p goes to block Tp[Mi[p]] of T */
int b200amg_tentative_scatter(long long n_members, int K1, int K2, const int *Mi, const int *Tp,
                              const double *Qx, double *Tx, void *stream);
/* C = A - B on R x C blocks, all-zero result blocks not stored (scipy bsr_binop_bsr): general != 0 =
 * rows in any order, output in reverse first-touch order; general == 0 = canonical merge of sorted rows.
 * pass 0 writes cnt[i]; pass 1 fills Cj/Cx at Cp[i] */
int b200amg_bsr_sub(int n_brow, int R, int C, const int *Ap, const int *Aj, const double *Ax,
                    const int *Bp, const int *Bj, const double *Bx, int general, int pass, int *cnt,
                    const int *Cp, int *Cj, double *Cx, void *stream);
/* scipy bsr_matmat (the block R*A*P of pyamg/aggregation/aggregation.py:429): the block pattern is the
 * scalar SpGEMM of the block patterns, column-sorted (Sj) and in csr_matmat order (Lj) over one row
 * pointer Cp; slot_map turns the pair into scipy's first-touch block order (Cj) + the slot of every
 * sorted entry; numeric accumulates every block in visiting order (Cx is cleared inside) */
int b200amg_bsr_slot_map(int n_brow, const int *Cp, const int *Sj, const int *Lj, int *slot, int *Cj,
                         void *stream);
int b200amg_bsr_matmat_numeric(int n_brow, int R, int N, int C, const int *Ap, const int *Aj,
                               const double *Ax, const int *Bp, const int *Bj, const double *Bx,
                               const int *Cp, const int *Sj, const int *slot, long long n_out_blocks,
                               double *Cx, void *stream);

/* ---- block relaxation with stored diagonal-block inverses (csrc/block_relax.cu) --------------------- */
/* amg_core.block_jacobi (pyamg/amg_core/relaxation.h:661-727): temp <- x on the block rows of the sweep, then
 * x_i = (1 - omega) temp_i + omega Dinv_i (b_i - sum_{j != i} A_ij temp_j).  Dinv[n_brow][bs][bs] as the
 * reference's Tx.  omega by value (the reference passes a length-1 array, relaxation.py:894). */
int b200amg_block_jacobi(const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
                         const double *Dinv, double *temp, int row_start, int row_stop, int row_step,
                         double omega, int blocksize, void *stream);
/* amg_core.block_gauss_seidel (relaxation.h:753-806): x_i = Dinv_i (b_i - sum_{j != i} A_ij x_j) in place.
 * color_ptr_host == NULL: block rows row_start, row_start + row_step, ... (positions of Id when Id != NULL)
 * one after the other -- the reference's order, sequential on the device.  With (Id, color_ptr_host, K):
 * Id holds the block rows grouped by colour and the sweep covers whole colours, forward or backward, one
 * launch per colour (same contract as b200amg_bsr_gauss_seidel). */
int b200amg_block_gauss_seidel(const int *Ap, const int *Aj, const double *Ax, double *x, const double *b,
                               const double *Dinv, int row_start, int row_stop, int row_step, int blocksize,
                               const int *Id, const int *color_ptr_host, int K, void *stream);
/* amg_core.pinv_array (pyamg/amg_core/linalg.h:888-954 over svd_jacobi :455-780): every n x n block of AA
 * (m blocks; row-major when transA != 0) is replaced by its Moore-Penrose pseudo-inverse, operation for operation
 * as the reference computes it (n <= 6: above that the reference's Python layer switches to LAPACK). */
int b200amg_pinv_array(double *AA, int m, int n, int transA, void *stream);
/* inverses of the diagonal blocks, Dinv[n_brow][bs][bs] (pyamg/util/utils.py:601-690 get_block_diag(inv_flag=True)):
 * the stored diagonal block (zeros when absent) through pinv_array for bs < 7 -- the bits of the reference.  For
 * bs = 7, 8 the reference calls LAPACK gelss (no bit contract): Gauss-Jordan with partial pivoting, equal to
 * rounding, a numerically singular block giving a zero block instead of its pseudo-inverse. */
int b200amg_block_diag_inverse(int n_brow, int blocksize, const int *Ap, const int *Aj, const double *Ax,
                               double *Dinv, void *stream);
int b200amg_block_diag_inverse_gj(int n_brow, int blocksize, const int *Ap, const int *Aj, const double *Ax,
                                  double *Dinv, void *stream);   /* the Gauss-Jordan path at any block size */

/* ---- Ruge-Stuben setup (csrc/rs_setup.cu): pyamg/amg_core/ruge_stuben.h:45-93, :496-595 -------- */
/* amg_core.classical_strength_of_connection fused with the two things the reference's glue does to S:
 * flags & 1: drop the diagonal (the graph of the C/F splitting, pyamg/util/utils.py:1815-1827);
 * flags & 2: do not store exact zeros, flags & 4: rows in reverse storage order -- C = pattern(S) .* A as
 * scipy's csr_elmul_csr returns it for canonical (2) / non-canonical (2|4) A (classical/interpolate.py:55-57).
 * pass 0 writes cnt[i]; pass 1 fills Cj/Cx at Cp[i] (Cx = entries of A). */
int b200amg_rs_strength(int n_row, double theta, const int *Ap, const int *Aj, const double *Ax, int flags,
                        int pass, int *cnt, const int *Cp, int *Cj, double *Cx, void *stream);
/* flag_dev[0] = 1 when some row is not strictly increasing in column (scipy csr_has_canonical_format) */
int b200amg_csr_not_canonical(int n_row, const int *Ap, const int *Aj, int *flag_dev, void *stream);
/* amg_core.rs_direct_interpolation_pass1: entries per row of P (the caller scans them into Pp) */
int b200amg_rs_direct_interpolation_pass1(int n_nodes, const int *Sp, const int *Sj, const int *splitting,
                                          int *cnt, void *stream);
/* amg_core.rs_direct_interpolation_pass2; cmap = exclusive scan of splitting (coarse index of a C point,
 * the reference's `map`, ruge_stuben.h:589-595) */
int b200amg_rs_direct_interpolation_pass2(int n_nodes, const int *Ap, const int *Aj, const double *Ax,
                                          const int *Sp, const int *Sj, const double *Sx, const int *splitting,
                                          const int *cmap, const int *Pp, int *Pj, double *Px, void *stream);

/* ---- host-side, inherently sequential setup steps whose INTEGER result must be identical
 *      to the reference (arrays are HOST pointers) ------------------------------------------ */
/* amg_core.vertex_coloring_mis, pyamg/amg_core/graph.h:171-188: returns K, color[i] in [0,K) */
int b200amg_vertex_coloring_mis_host(int n, const int *Ap_host, const int *Aj_host, int *color_host);
/* amg_core.vertex_coloring_jones_plassmann / vertex_coloring_LDF (pyamg/amg_core/graph.h:242-335; methods 'JP' and
 * 'LDF' of pyamg/graph.py:84-125): weights = one random number per vertex (the reference draws sp.rand(N)).
 * Host arrays; return the number of colours. */
int b200amg_vertex_coloring_jones_plassmann_host(int n, const int *Ap, const int *Aj, int *color,
                                                 const double *weights);
int b200amg_vertex_coloring_ldf_host(int n, const int *Ap, const int *Aj, int *color, const double *weights);
/* Dependency levels ("wavefronts") of the lexicographic sweep of amg_core.gauss_seidel / bsr_gauss_seidel /
 * block_gauss_seidel (pyamg/amg_core/relaxation.h:33-62, :89-173, :753-806) over the rows (block rows) of the pattern
 * (Ap, Aj): level[i] = the parallel step at which row i sees exactly the operands of the sequential sweep.  Rows grouped
 * by level (ascending inside a level) + the level offsets are an (Id, color_ptr_host) pair for
 * b200amg_gauss_seidel_indexed / b200amg_bsr_gauss_seidel / b200amg_block_gauss_seidel: the lexicographic sweep, forward
 * or backward, bit for bit, one launch per level instead of one thread.  Host arrays; returns the number of levels. */
int b200amg_gs_wavefronts_host(int n, const int *Ap_host, const int *Aj_host, int *level_host);
/* amg_core.standard_aggregation, pyamg/amg_core/smoothed_aggregation.h:121-221: returns the
 * number of aggregates; agg[i] = aggregate of node i or -1; roots[k] = seed node of k */
int b200amg_standard_aggregation_host(int n, const int *Ap_host, const int *Aj_host, int *agg_host,
                                      int *roots_host);
/* amg_core.naive_aggregation (smoothed_aggregation.h:244-277, aggregate='naive'); host arrays, returns n_agg */
int b200amg_naive_aggregation_host(int n, const int *Ap, const int *Aj, int *agg, int *roots);
/* amg_core.rs_cf_splitting, pyamg/amg_core/ruge_stuben.h:157-297: S (no diagonal) and T = S^T as host CSR;
 * splitting[i] = 1 (C point) or 0 (F point) */
int b200amg_rs_cf_splitting_host(int n_nodes, const int *Sp_host, const int *Sj_host, const int *Tp_host,
                                 const int *Tj_host, int *splitting_host);

/* ---- whole-cycle engine: multilevel_solver.__solve, pyamg/multilevel.py:473-548 ------- */
typedef struct b200amg_level_desc {
    /* A: CSR (bs == 1) or BSR with bs x bs blocks; flavor 0: CSR arithmetic
     * (amg_core.jacobi / gauss_seidel*), 1: BSR arithmetic (amg_core.bsr_*), which differs in
     * association even for bs == 1 (relaxation.h:216-238 vs :313-350). */
    int n;            /* scalar unknowns on this level */
    int bs;           /* block size of A */
    int flavor;
    const int *Ap, *Aj; const double *Ax;
    /* P: (n x nc) and R: (nc x n), stored CSR or BSR(pr x pc) */
    int nc;
    int p_br, p_bc;   /* P block shape (R has the transposed shape) */
    const int *Pp, *Pj; const double *Px;
    const int *Rp, *Rj; const double *Rx;
    /* smoothers: kind 0 none, 1 jacobi(omega, iterations), 2 gauss_seidel_indexed over
     * colour-grouped rows (Id, color_ptr on HOST), 3 lexicographic gauss_seidel (sequential on
     * the device, exact), 4 polynomial (coefficients below), 5-7 block smoothers (Dinv below);
     * sweep 0 fwd / 1 bwd / 2 symmetric */
    int pre_kind, post_kind;
    double pre_omega, post_omega;
    int pre_iters, post_iters;
    int pre_sweep, post_sweep;
    const int *Id;              /* device: rows (block rows if bs>1) grouped by colour */
    const int *color_ptr_host;  /* host: K+1 offsets */
    int K;
    /* kind 4: polynomial smoother (pyamg/relaxation/relaxation.py:593-668; what 'chebyshev' resolves to,
     * smoothing.py:438-449): per iteration r = b - A x, h = c[0] r, h = c[k] r + A h (k = 1..ncoef-1), x += h */
    int pre_ncoef, post_ncoef;  /* 1..8 */
    double pre_coef[8], post_coef[8];
    /* kinds 5 block_jacobi(omega, iterations), 6 block_gauss_seidel over colour-grouped block rows (Id,
     * color_ptr as for kind 2), 7 lexicographic block_gauss_seidel (sequential on the device, exact):
     * pyamg/amg_core/relaxation.h:661-806 with the inverses of the diagonal blocks, [n / bs][bs][bs] */
    const double *Dinv;
    /* kind 2 on a BSR(1 x 1) level: nonzero = CSR arithmetic x_i = (b_i - sum) / a_ii -- the reference's
     * gauss_seidel_indexed converts its matrix to CSR first (relaxation.py:713) -- zero = the arithmetic of the level's
     * format (bsr_gauss_seidel for BSR: what the lexicographic gauss_seidel dispatches to, relaxation.py:351-354) */
    int gs_csr;
} b200amg_level_desc;
/* sizeof(b200amg_level_desc) as compiled: lets a binding check its mirror of the struct */
int b200amg_level_desc_size(void);

/* Build a cycle engine over nlev-1 described levels + the dense coarse operator
 * (coarse_inv: nc_last x nc_last row-major pseudo-inverse on the device, or NULL when the
 * coarsest matrix has no entries -> zero correction, multilevel.py:699-701).
 * work: device scratch of b200amg_vcycle_work_bytes(...) bytes. */
long long b200amg_vcycle_work_bytes(int nlev, const b200amg_level_desc *levels_host);
int b200amg_vcycle_create(int nlev, const b200amg_level_desc *levels_host, const double *coarse_inv,
                          void *work, long long work_bytes, void **handle_out);
/* one cycle: x <- cycle(x, b); cycle_kind 0 = V, 1 = W, 2 = F.  use_graph != 0 replays a CUDA
 * graph captured on first use for these (x, b) pointers. */
int b200amg_vcycle_run(void *handle, double *x, const double *b, int cycle_kind, int use_graph,
                       void *stream);
/* x <- cycle(0, b): one cycle from a zero initial guess = the preconditioner application of
 * multilevel_solver.aspreconditioner (pyamg/multilevel.py:274-314).  x need not be cleared; the first pre-smoothing
 * Jacobi sweep from the known-zero guess is the bit-identical diagonal scaling (one pass over A_0 less). */
int b200amg_vcycle_run_zero(void *handle, double *x, const double *b, int cycle_kind, int use_graph,
                            void *stream);
/* y = A_level x (op 0) or y = b - A_level x (op 2) in the layout the engine streams that operator in (A * p and
 * b - A * x of pyamg/krylov/_cg.py around the cycle) */
int b200amg_vcycle_apply_op(void *handle, int level, int op, const double *x, const double *b, double *y,
                           void *stream);
/* One UN-graphed cycle with a CUDA event after every operation: fills (host arrays, up to
 * max_ops entries) the operation code, its level and its device time in ms, in execution
 * order; *n_ops_host = entries written.  Synchronises the stream.  bench.py uses it for the
 * per-kernel share of a cycle and the roofline of the dominant kernel. */
#define B200AMG_OP_PRESMOOTH  0
#define B200AMG_OP_RESIDUAL   1
#define B200AMG_OP_RESTRICT   2
#define B200AMG_OP_COARSE     3
#define B200AMG_OP_PROLONG    4
#define B200AMG_OP_POSTSMOOTH 5
#define B200AMG_OP_COPYBACK   6
int b200amg_vcycle_profile(void *handle, double *x, const double *b, int cycle_kind, int max_ops,
                           int *op_code_host, int *op_level_host, float *op_ms_host,
                           int *n_ops_host, void *stream);
/* layout the engine streams an operator of `level` in (which: 0 A, 1 P, 2 R, 3 the colour-grouped copy of A that
 * colour-ordered Gauss-Seidel sweeps, 4 the diagonal-free copy of A that the Jacobi sweeps of a level of >= 400 000
 * rows stream): the b200amg_op_layout codes, 0 for block operators / no such copy.
 * Informational: every layout gives the same bits. */
int b200amg_vcycle_layout(void *handle, int level, int which);
/* rnorm2_dev[0] = ||b - A0 x||^2 on the finest level (residual_norm of multilevel.py:450,461) */
int b200amg_vcycle_residual_norm2(void *handle, const double *x, const double *b,
                                  double *rnorm2_dev, void *stream);
int b200amg_vcycle_destroy(void *handle);

#ifdef __cplusplus
}
#endif
#endif /* B200AMG_H */
