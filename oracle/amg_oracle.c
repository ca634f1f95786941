/*
 * amg_oracle.c -- TEST INFRASTRUCTURE ONLY.  CPU restatement ("port") of the reference's
 * solve-phase V-cycle hot path, fp64 / int32, plain serial C.
 *
 * Nothing in the product (pyamg_b200/) may import, link or call this file; it is the
 * checker used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks every function here against
 *   (a) the golden vectors of the reference's own tests
 *       (pyamg/relaxation/tests/test_relaxation.py:139-402) and
 *   (b) the reference's compiled kernels oracle/_ref/_amg_core.so (built by
 *       oracle/build_ref.sh from /root/reference/pyamg/amg_core/amg_core_wrap.cxx) and the
 *       installed scipy 1.18.1 _sparsetools, bit for bit on seeded inputs.
 *
 * Each function cites the reference lines it follows (paths relative to /root/reference).
 * The fp64 operation ORDER of the reference is kept exactly (row-internal storage order,
 * the b - sum vs (b - t1) - t2 association of the CSR vs BSR smoothers, ...), and the file
 * is compiled with -ffp-contract=off so no FMA changes roundings.
 */
#include <stdlib.h>
#include <string.h>
#include <math.h>

typedef int I;
typedef double T;

/* ---- scipy _sparsetools csr_matvec: y += A x  (call sites pyamg/multilevel.py:496,498,544) */
void orc_csr_matvec(I n_row, const I *Ap, const I *Aj, const T *Ax, const T *x, T *y)
{
    for (I i = 0; i < n_row; i++) {
        T s = y[i];
        for (I p = Ap[i]; p < Ap[i + 1]; p++)
            s += Ax[p] * x[Aj[p]];
        y[i] = s;
    }
}

/* ---- scipy _sparsetools bsr_matvec: y += A x with R x C row-major blocks.
 * R == C == 1 goes through csr_matvec exactly as scipy does. */
void orc_bsr_matvec(I n_brow, I R, I C, const I *Ap, const I *Aj, const T *Ax, const T *x, T *y)
{
    if (R == 1 && C == 1) { orc_csr_matvec(n_brow, Ap, Aj, Ax, x, y); return; }
    const I RC = R * C;
    for (I i = 0; i < n_brow; i++) {
        T *yb = y + (size_t)R * i;
        for (I p = Ap[i]; p < Ap[i + 1]; p++) {
            const T *blk = Ax + (size_t)RC * p;
            const T *xb = x + (size_t)C * Aj[p];
            for (I r = 0; r < R; r++) {
                T s = yb[r];
                for (I c = 0; c < C; c++)
                    s += blk[r * C + c] * xb[c];
                yb[r] = s;
            }
        }
    }
}

/* ---- amg_core.jacobi, pyamg/amg_core/relaxation.h:201-239 */
void orc_jacobi(const I *Ap, const I *Aj, const T *Ax, T *x, const T *b, T *temp,
                I row_start, I row_stop, I row_step, T omega)
{
    for (I i = row_start; i != row_stop; i += row_step)
        temp[i] = x[i];
    for (I i = row_start; i != row_stop; i += row_step) {
        T offd = 0, d = 0;
        for (I p = Ap[i]; p < Ap[i + 1]; p++) {
            I j = Aj[p];
            if (j == i) d = Ax[p];            /* last stored diagonal wins */
            else offd += Ax[p] * temp[j];
        }
        if (d != 0.0)
            x[i] = (1.0 - omega) * temp[i] + omega * ((b[i] - offd) / d);
    }
}

/* ---- amg_core.gauss_seidel, relaxation.h:33-62 */
void orc_gauss_seidel(const I *Ap, const I *Aj, const T *Ax, T *x, const T *b,
                      I row_start, I row_stop, I row_step)
{
    for (I i = row_start; i != row_stop; i += row_step) {
        T offd = 0, d = 0;
        for (I p = Ap[i]; p < Ap[i + 1]; p++) {
            I j = Aj[p];
            if (j == i) d = Ax[p];
            else offd += Ax[p] * x[j];
        }
        if (d != 0.0)
            x[i] = (b[i] - offd) / d;
    }
}

/* ---- amg_core.gauss_seidel_indexed, relaxation.h:394-426 (positions of Id, repeats allowed) */
void orc_gauss_seidel_indexed(const I *Ap, const I *Aj, const T *Ax, T *x, const T *b,
                              const I *Id, I row_start, I row_stop, I row_step)
{
    for (I k = row_start; k != row_stop; k += row_step) {
        I i = Id[k];
        T offd = 0, d = 0;
        for (I p = Ap[i]; p < Ap[i + 1]; p++) {
            I j = Aj[p];
            if (j == i) d = Ax[p];
            else offd += Ax[p] * x[j];
        }
        if (d != 0.0)
            x[i] = (b[i] - offd) / d;
    }
}

/* dense row-major block times vector started from zero: linalg.h:359-398 as used by the BSR
 * smoothers (overwrite='T', Strans='F'): out[r] = ((0 + a_r0 v_0) + a_r1 v_1) + ... */
static void blk_matvec0(const T *blk, const T *v, T *out, I bs)
{
    for (I r = 0; r < bs; r++) {
        T s = 0;
        for (I c = 0; c < bs; c++)
            s += blk[r * bs + c] * v[c];
        out[r] = s;
    }
}

/* ---- amg_core.bsr_gauss_seidel, relaxation.h:89-173 */
void orc_bsr_gauss_seidel(const I *Ap, const I *Aj, const T *Ax, T *x, const T *b,
                          I row_start, I row_stop, I row_step, I bs)
{
    const I B2 = bs * bs;
    T *acc = (T *)malloc(sizeof(T) * bs), *loc = (T *)malloc(sizeof(T) * bs);
    const I k0 = row_step < 0 ? bs - 1 : 0, k1 = row_step < 0 ? -1 : bs, dk = row_step < 0 ? -1 : 1;
    for (I i = row_start; i != row_stop; i += row_step) {
        long dptr = -1;
        for (I k = 0; k < bs; k++) acc[k] = b[(size_t)i * bs + k];
        for (I p = Ap[i]; p < Ap[i + 1]; p++) {
            I j = Aj[p];
            if (j == i) { dptr = (long)p * B2; continue; }
            blk_matvec0(Ax + (size_t)p * B2, x + (size_t)j * bs, loc, bs);
            for (I m = 0; m < bs; m++) acc[m] -= loc[m];
        }
        if (dptr < 0) continue;               /* no stored diagonal block: row untouched */
        for (I k = k0; k != k1; k += dk) {
            T d = 1.0;
            for (I kk = k0; kk != k1; kk += dk) {
                if (kk == k) d = Ax[dptr + k * bs + kk];
                else acc[k] -= Ax[dptr + k * bs + kk] * x[(size_t)i * bs + kk];
            }
            if (d != 0.0) x[(size_t)i * bs + k] = acc[k] / d;
        }
    }
    free(acc); free(loc);
}

/* ---- block rows relaxed in the order of an index list: bsr_gauss_seidel (relaxation.h:89-173) called for ONE block
 * row at a time, (i, i+1, +1) going forward through Id and (i, i-1, -1) going backward -- the sequential definition of
 * the block-multicolour sweep (the reference has no such entry point; its gauss_seidel_indexed is CSR-only,
 * relaxation.py:671-741).  Same arithmetic per row as orc_bsr_gauss_seidel. */
void orc_bsr_gauss_seidel_indexed(const I *Ap, const I *Aj, const T *Ax, T *x, const T *b, const I *Id,
                                  I start, I stop, I step, I bs)
{
    const I dir = step < 0 ? -1 : 1;
    for (I p = start; p != stop; p += step)
        orc_bsr_gauss_seidel(Ap, Aj, Ax, x, b, Id[p], Id[p] + dir, dir, bs);
}

/* ---- amg_core.block_gauss_seidel, relaxation.h:753-806: Tx = inverses of the diagonal blocks */
void orc_block_gauss_seidel(const I *Ap, const I *Aj, const T *Ax, T *x, const T *b, const T *Tx,
                            I row_start, I row_stop, I row_step, I bs)
{
    const I B2 = bs * bs;
    T *rsum = (T *)malloc(sizeof(T) * bs), *v = (T *)malloc(sizeof(T) * bs);
    for (I i = row_start; i != row_stop; i += row_step) {
        for (I k = 0; k < bs; k++) rsum[k] = 0.0;
        for (I p = Ap[i]; p < Ap[i + 1]; p++) {
            I j = Aj[p];
            if (j == i) continue;
            blk_matvec0(Ax + (size_t)p * B2, x + (size_t)j * bs, v, bs);
            for (I k = 0; k < bs; k++) rsum[k] += v[k];
        }
        for (I k = 0; k < bs; k++) rsum[k] = b[(size_t)i * bs + k] - rsum[k];
        blk_matvec0(Tx + (size_t)i * B2, rsum, x + (size_t)i * bs, bs);
    }
    free(rsum); free(v);
}

/* block rows in the order of an index list, one block_gauss_seidel call per row (cf. orc_bsr_gauss_seidel_indexed) */
void orc_block_gauss_seidel_indexed(const I *Ap, const I *Aj, const T *Ax, T *x, const T *b, const T *Tx, const I *Id,
                                    I start, I stop, I step, I bs)
{
    const I dir = step < 0 ? -1 : 1;
    for (I p = start; p != stop; p += step)
        orc_block_gauss_seidel(Ap, Aj, Ax, x, b, Tx, Id[p], Id[p] + dir, dir, bs);
}

/* ---- amg_core.block_jacobi, relaxation.h:661-727 */
void orc_block_jacobi(const I *Ap, const I *Aj, const T *Ax, T *x, const T *b, const T *Tx, T *temp,
                      I row_start, I row_stop, I row_step, T omega, I bs)
{
    const I B2 = bs * bs;
    T *rsum = (T *)malloc(sizeof(T) * bs), *v = (T *)malloc(sizeof(T) * bs);
    for (I i = row_start * bs; i != row_stop * bs; i += row_step * bs)
        for (I k = 0; k < bs; k++) temp[i + k] = x[i + k];
    for (I i = row_start; i != row_stop; i += row_step) {
        for (I k = 0; k < bs; k++) rsum[k] = 0.0;
        for (I p = Ap[i]; p < Ap[i + 1]; p++) {
            I j = Aj[p];
            if (j == i) continue;
            blk_matvec0(Ax + (size_t)p * B2, temp + (size_t)j * bs, v, bs);
            for (I k = 0; k < bs; k++) rsum[k] += v[k];
        }
        for (I k = 0; k < bs; k++) rsum[k] = b[(size_t)i * bs + k] - rsum[k];
        blk_matvec0(Tx + (size_t)i * B2, rsum, v, bs);
        for (I k = 0; k < bs; k++)
            x[(size_t)i * bs + k] = (1.0 - omega) * temp[(size_t)i * bs + k] + omega * v[k];
    }
    free(rsum); free(v);
}

/* ---- amg_core.bsr_jacobi, relaxation.h:267-360.  The temp copy (:301-303) takes the
 * |row_stop-row_start|*bs LEADING entries; the reference's loop never terminates for a
 * negative row_step (i += -1 from 0), and its Python wrapper only ever passes (0, N, 1)
 * (relaxation.py:403-404), so the copy is restated for the forward case only. */
void orc_bsr_jacobi(const I *Ap, const I *Aj, const T *Ax, T *x, const T *b, T *temp,
                    I row_start, I row_stop, I row_step, I bs, T omega)
{
    const I B2 = bs * bs;
    T *acc = (T *)malloc(sizeof(T) * bs), *loc = (T *)malloc(sizeof(T) * bs);
    const I k0 = row_step < 0 ? bs - 1 : 0, k1 = row_step < 0 ? -1 : bs, dk = row_step < 0 ? -1 : 1;
    for (I i = 0; i < abs(row_stop - row_start) * bs; i++)
        temp[i] = x[i];
    for (I i = row_start; i != row_stop; i += row_step) {
        long dptr = -1;
        for (I k = 0; k < bs; k++) acc[k] = b[(size_t)i * bs + k];
        for (I p = Ap[i]; p < Ap[i + 1]; p++) {
            I j = Aj[p];
            if (j == i) { dptr = (long)p * B2; continue; }
            blk_matvec0(Ax + (size_t)p * B2, temp + (size_t)j * bs, loc, bs);
            for (I m = 0; m < bs; m++) acc[m] -= loc[m];
        }
        if (dptr < 0) continue;
        for (I k = k0; k != k1; k += dk) {
            T d = 1.0;
            for (I kk = k0; kk != k1; kk += dk) {
                if (kk == k) d = Ax[dptr + k * bs + kk];
                else acc[k] -= Ax[dptr + k * bs + kk] * temp[(size_t)i * bs + kk];
            }
            if (d != 0.0)
                x[(size_t)i * bs + k] = (1.0 - omega) * temp[(size_t)i * bs + k] + omega * acc[k] / d;
        }
    }
    free(acc); free(loc);
}

/* ---- amg_core.vertex_coloring_mis, graph.h:171-188 with the greedy lexicographic MIS of
 * graph.h:34-61.  State encoding per round K: active = -1-K, in set = K, excluded = -2-K. */
I orc_vertex_coloring_mis(I n, const I *Ap, const I *Aj, I *color)
{
    for (I i = 0; i < n; i++) color[i] = -1;
    I done = 0, K = 0;
    while (done < n) {
        const I active = -1 - K, out = -2 - K;
        for (I i = 0; i < n; i++) {
            if (color[i] != active) continue;
            color[i] = K; done++;
            for (I p = Ap[i]; p < Ap[i + 1]; p++)
                if (color[Aj[p]] == active) color[Aj[p]] = out;
        }
        K++;
    }
    return K;
}

/* ======================= setup-phase pieces (Galerkin product and its inputs) ============= */

/* ---- scipy _sparsetools csr_matmat_maxnnz: structural nnz of C = A B (upper bound used to
 * allocate; call sites pyamg/aggregation/aggregation.py:429, pyamg/classical/classical.py:187) */
long orc_csr_matmat_maxnnz(I n_row, I n_col, const I *Ap, const I *Aj, const I *Bp, const I *Bj)
{
    I *mask = (I *)malloc(sizeof(I) * (n_col > 0 ? n_col : 1));
    for (I k = 0; k < n_col; k++) mask[k] = -1;
    long nnz = 0;
    for (I i = 0; i < n_row; i++)
        for (I p = Ap[i]; p < Ap[i + 1]; p++) {
            I j = Aj[p];
            for (I q = Bp[j]; q < Bp[j + 1]; q++)
                if (mask[Bj[q]] != i) { mask[Bj[q]] = i; nnz++; }
        }
    free(mask);
    return nnz;
}

/* ---- scipy _sparsetools csr_matmat (Gustavson row-wise product).  Per row: a dense
 * accumulator + a linked list threaded through the touched columns; columns come out in
 * REVERSE first-touch order, entries whose accumulated sum is exactly 0.0 are dropped.
 * Returns nnz written; Cp has n_row+1 entries. */
long orc_csr_matmat(I n_row, I n_col, const I *Ap, const I *Aj, const T *Ax,
                    const I *Bp, const I *Bj, const T *Bx, I *Cp, I *Cj, T *Cx)
{
    I *link = (I *)malloc(sizeof(I) * (n_col > 0 ? n_col : 1));
    T *acc = (T *)calloc(n_col > 0 ? n_col : 1, sizeof(T));
    for (I k = 0; k < n_col; k++) link[k] = -1;
    long nnz = 0;
    Cp[0] = 0;
    for (I i = 0; i < n_row; i++) {
        I head = -2, len = 0;
        for (I p = Ap[i]; p < Ap[i + 1]; p++) {
            I j = Aj[p];
            T a = Ax[p];
            for (I q = Bp[j]; q < Bp[j + 1]; q++) {
                I k = Bj[q];
                acc[k] += a * Bx[q];
                if (link[k] == -1) { link[k] = head; head = k; len++; }
            }
        }
        for (I t = 0; t < len; t++) {
            if (acc[head] != 0.0) { Cj[nnz] = head; Cx[nnz] = acc[head]; nnz++; }
            I nxt = link[head];
            link[head] = -1; acc[head] = 0;
            head = nxt;
        }
        Cp[i + 1] = (I)nnz;
    }
    free(link); free(acc);
    return nnz;
}

/* ---- scipy _sparsetools csr_tocsc (explicit transpose: P.T.tocsr(), classical.py:174;
 * P.H for SA, aggregation.py:400-401).  Rows inside each output row stay ascending. */
void orc_csr_tocsc(I n_row, I n_col, const I *Ap, const I *Aj, const T *Ax, I *Bp, I *Bi, T *Bx)
{
    const I nnz = Ap[n_row];
    for (I k = 0; k <= n_col; k++) Bp[k] = 0;
    for (I p = 0; p < nnz; p++) Bp[Aj[p] + 1]++;
    for (I k = 0; k < n_col; k++) Bp[k + 1] += Bp[k];
    I *fill = (I *)malloc(sizeof(I) * (n_col > 0 ? n_col : 1));
    memcpy(fill, Bp, sizeof(I) * n_col);
    for (I i = 0; i < n_row; i++)
        for (I p = Ap[i]; p < Ap[i + 1]; p++) {
            I dst = fill[Aj[p]]++;
            Bi[dst] = i; Bx[dst] = Ax[p];
        }
    free(fill);
}

/* ---- amg_core.symmetric_strength_of_connection, smoothed_aggregation.h:48-99 (real T):
 * keep a_ij when i==j or a_ij^2 >= theta^2 |a_ii| |a_jj| ; duplicates on the diagonal summed. */
void orc_symmetric_strength(I n, T theta, const I *Ap, const I *Aj, const T *Ax, I *Sp, I *Sj, T *Sx)
{
    T *dg = (T *)malloc(sizeof(T) * (n > 0 ? n : 1));
    for (I i = 0; i < n; i++) {
        T d = 0;
        for (I p = Ap[i]; p < Ap[i + 1]; p++) if (Aj[p] == i) d += Ax[p];
        dg[i] = fabs(d);
    }
    I nnz = 0; Sp[0] = 0;
    for (I i = 0; i < n; i++) {
        const T lim = theta * theta * dg[i];
        for (I p = Ap[i]; p < Ap[i + 1]; p++) {
            I j = Aj[p]; T a = Ax[p];
            if (j == i || a * a >= lim * dg[j]) { Sj[nnz] = j; Sx[nnz] = a; nnz++; }
        }
        Sp[i + 1] = nnz;
    }
    free(dg);
}

/* ---- amg_core.standard_aggregation, smoothed_aggregation.h:121-221 (three greedy passes).
 * agg[i] = aggregate of node i (-1: isolated); cpts[] = root nodes.  Returns #aggregates. */
I orc_standard_aggregation(I n, const I *Ap, const I *Aj, I *agg, I *cpts)
{
    for (I i = 0; i < n; i++) agg[i] = 0;
    I next = 1;
    for (I i = 0; i < n; i++) {                       /* pass 1: seed where all nbrs are free */
        if (agg[i]) continue;
        int nbr = 0, taken = 0;
        for (I p = Ap[i]; p < Ap[i + 1]; p++) {
            I j = Aj[p];
            if (j != i) { nbr = 1; if (agg[j]) { taken = 1; break; } }
        }
        if (!nbr) agg[i] = -n;
        else if (!taken) {
            agg[i] = next; cpts[next - 1] = i;
            for (I p = Ap[i]; p < Ap[i + 1]; p++) agg[Aj[p]] = next;
            next++;
        }
    }
    for (I i = 0; i < n; i++) {                       /* pass 2: attach to a neighbouring seed aggregate */
        if (agg[i]) continue;
        for (I p = Ap[i]; p < Ap[i + 1]; p++) {
            I a = agg[Aj[p]];
            if (a > 0) { agg[i] = -a; break; }
        }
    }
    next--;
    for (I i = 0; i < n; i++) {                       /* pass 3: renumber from 0, sweep up leftovers */
        I a = agg[i];
        if (a != 0) {
            agg[i] = a > 0 ? a - 1 : (a == -n ? -1 : -a - 1);
            continue;
        }
        agg[i] = next; cpts[next] = i;
        for (I p = Ap[i]; p < Ap[i + 1]; p++) if (agg[Aj[p]] == 0) agg[Aj[p]] = next;
        next++;
    }
    return next;
}

/* ---- amg_core.fit_candidates (real), smoothed_aggregation.h:322-448: per aggregate j
 * (CSC of AggOp: Ap/Ai) gather the K1 x K2 blocks of B, modified Gram-Schmidt QR of the
 * stacked block column, Q into Qx (nnz x K1 x K2), R into Rx (n_col x K2 x K2). */
void orc_fit_candidates(I n_col, I K1, I K2, const I *Ap, const I *Ai, T *Qx, const T *B, T *Rx, T tol)
{
    const I BS = K1 * K2;
    memset(Rx, 0, sizeof(T) * (size_t)n_col * K2 * K2);
    for (I j = 0; j < n_col; j++)
        for (I p = Ap[j]; p < Ap[j + 1]; p++)
            memcpy(Qx + (size_t)BS * p, B + (size_t)BS * Ai[p], sizeof(T) * BS);
    for (I j = 0; j < n_col; j++) {
        T *q0 = Qx + (size_t)BS * Ap[j], *q1 = Qx + (size_t)BS * Ap[j + 1];
        T *Rj = Rx + (size_t)j * K2 * K2;
        for (I c = 0; c < K2; c++) {
            T nrm = 0;
            for (T *q = q0 + c; q < q1; q += K2) nrm += (*q) * (*q);
            const T thresh = tol * sqrt(nrm);
            for (I a = 0; a < c; a++) {
                T d = 0;
                for (T *qa = q0 + a, *qc = q0 + c; qa < q1; qa += K2, qc += K2) d += (*qa) * (*qc);
                for (T *qa = q0 + a, *qc = q0 + c; qa < q1; qa += K2, qc += K2) *qc -= d * (*qa);
                Rj[K2 * a + c] = d;
            }
            nrm = 0;
            for (T *q = q0 + c; q < q1; q += K2) nrm += (*q) * (*q);
            nrm = sqrt(nrm);
            T scale;
            if (nrm > thresh) { scale = 1.0 / nrm; Rj[K2 * c + c] = nrm; }
            else { scale = 0; Rj[K2 * c + c] = 0; }
            for (T *q = q0 + c; q < q1; q += K2) *q *= scale;
        }
    }
}

/* ==== Ruge-Stuben setup pieces, restated from pyamg/amg_core/ruge_stuben.h ======== */

/* ---- amg_core.classical_strength_of_connection, ruge_stuben.h:45-93 (real T): off-diagonal a_ij is
 * strong when |a_ij| >= theta * max_{k != i} |a_ik| (the running maximum starts at DBL_MIN, the smallest
 * positive normal, :57); stored diagonal entries are always kept; order inside a row unchanged. */
void orc_classical_strength(I n, T theta, const I *Ap, const I *Aj, const T *Ax, I *Sp, I *Sj, T *Sx)
{
    I out = 0;
    Sp[0] = 0;
    for (I i = 0; i < n; i++) {
        T big = 2.2250738585072014e-308;
        for (I p = Ap[i]; p < Ap[i + 1]; p++)
            if (Aj[p] != i && fabs(Ax[p]) > big) big = fabs(Ax[p]);
        const T lim = theta * big;
        for (I p = Ap[i]; p < Ap[i + 1]; p++)
            if (Aj[p] == i || fabs(Ax[p]) >= lim) { Sj[out] = Aj[p]; Sx[out] = Ax[p]; out++; }
        Sp[i + 1] = out;
    }
}

/* ---- amg_core.rs_cf_splitting, ruge_stuben.h:157-297.  First-pass Ruge-Stuben coarsening: nodes
 * are kept in an array ordered by their measure lambda (= column count of S, i.e. row length of
 * T = S^T), one contiguous bucket per measure value; the last node of the array is taken as the next
 * C point, its undecided dependants become F points, and measures of their / its undecided
 * neighbours go up / down by one, which moves the node to the far / near end of its bucket and
 * shifts the bucket boundary.  The tie-breaking between equal measures IS the bucket mechanics, so
 * they are reproduced exactly, including: bucket starts are only initialised for measures < n
 * (:180-184), nodes with measure >= n-1 are never raised (:241), measure 0 is never lowered (:270).
 * split[i]: 1 = C, 0 = F. */
static void rs_swap(I *seq, I *where, I a, I b)
{
    const I na = seq[a], nb = seq[b];
    seq[a] = nb; seq[b] = na;
    where[nb] = a; where[na] = b;
}

void orc_rs_cf_splitting(I n, const I *Sp, const I *Sj, const I *Tp, const I *Tj, I *split)
{
    I *meas = (I *)calloc((size_t)n + 1, sizeof(I));
    I *first = (I *)calloc((size_t)n + 1, sizeof(I));
    I *fill = (I *)calloc((size_t)n + 1, sizeof(I));
    I *seq = (I *)malloc(sizeof(I) * ((size_t)n + 1));
    I *where = (I *)malloc(sizeof(I) * ((size_t)n + 1));
    for (I i = 0; i < n; i++) { meas[i] = Tp[i + 1] - Tp[i]; fill[meas[i]]++; }
    for (I v = 0, run = 0; v < n; v++) { first[v] = run; run += fill[v]; fill[v] = 0; }
    for (I i = 0; i < n; i++) {
        const I at = first[meas[i]] + fill[meas[i]]++;
        seq[at] = i; where[i] = at;
    }
    for (I i = 0; i < n; i++)
        split[i] = (meas[i] == 0 || (meas[i] == 1 && Tj[Tp[i]] == i)) ? 0 : 2;
    for (I top = n - 1; top >= 0; top--) {
        const I c = seq[top];
        fill[meas[c]]--;
        if (split[c] == 0) continue;
        split[c] = 1;
        for (I p = Tp[c]; p < Tp[c + 1]; p++) {
            const I f = Tj[p];
            if (split[f] != 2) continue;
            split[f] = 0;
            for (I q = Sp[f]; q < Sp[f + 1]; q++) {
                const I k = Sj[q];
                if (split[k] != 2 || meas[k] >= n - 1) continue;
                const I m = meas[k], to = first[m] + fill[m] - 1;
                rs_swap(seq, where, where[k], to);
                fill[m]--; fill[m + 1]++;
                first[m + 1] = to;
                meas[k] = m + 1;
            }
        }
        for (I p = Sp[c]; p < Sp[c + 1]; p++) {
            const I j = Sj[p];
            if (split[j] != 2 || meas[j] == 0) continue;
            const I m = meas[j], to = first[m];
            rs_swap(seq, where, where[j], to);
            fill[m]--; fill[m - 1]++;
            first[m]++;
            first[m - 1] = first[m] - fill[m - 1];
            meas[j] = m - 1;
        }
    }
    free(meas); free(first); free(fill); free(seq); free(where);
}

/* ---- amg_core.rs_direct_interpolation_pass1, ruge_stuben.h:496-515: row pointer of P.  A C row has
 * one entry; an F row one entry per strong C neighbour other than itself. */
void orc_rs_direct_interpolation_pass1(I n, const I *Sp, const I *Sj, const I *split, I *Pp)
{
    I run = 0;
    Pp[0] = 0;
    for (I i = 0; i < n; i++) {
        if (split[i] == 1) run++;
        else
            for (I p = Sp[i]; p < Sp[i + 1]; p++) run += (split[Sj[p]] == 1 && Sj[p] != i);
        Pp[i + 1] = run;
    }
}

/* ---- amg_core.rs_direct_interpolation_pass2, ruge_stuben.h:518-595: direct interpolation weights
 * w_ij = -(sum of all negative a_ik / sum of strong-C negative a_ik) a_ij / a_ii for negative a_ij, the
 * same with the positive sums for positive a_ij; without positive strong C neighbours the positive
 * off-diagonals are lumped into the diagonal (:569-572).  Sums run in storage order; columns are
 * renumbered to coarse indices at the end (:589-595). */
void orc_rs_direct_interpolation_pass2(I n, const I *Ap, const I *Aj, const T *Ax, const I *Sp, const I *Sj,
                                       const T *Sx, const I *split, const I *Pp, I *Pj, T *Px)
{
    for (I i = 0; i < n; i++) {
        if (split[i] == 1) { Pj[Pp[i]] = i; Px[Pp[i]] = 1; continue; }
        T sneg = 0, spos = 0, aneg = 0, apos = 0, d = 0;
        for (I p = Sp[i]; p < Sp[i + 1]; p++)
            if (split[Sj[p]] == 1 && Sj[p] != i) { if (Sx[p] < 0) sneg += Sx[p]; else spos += Sx[p]; }
        for (I p = Ap[i]; p < Ap[i + 1]; p++) {
            if (Aj[p] == i) d += Ax[p];
            else if (Ax[p] < 0) aneg += Ax[p];
            else apos += Ax[p];
        }
        const T alpha = aneg / sneg;
        T beta = apos / spos;
        if (spos == 0) { d += apos; beta = 0; }
        const T cn = -alpha / d, cp = -beta / d;
        I at = Pp[i];
        for (I p = Sp[i]; p < Sp[i + 1]; p++)
            if (split[Sj[p]] == 1 && Sj[p] != i) {
                Pj[at] = Sj[p];
                Px[at] = (Sx[p] < 0 ? cn : cp) * Sx[p];
                at++;
            }
    }
    I *renum = (I *)malloc(sizeof(I) * ((size_t)n + 1));
    for (I i = 0, run = 0; i < n; i++) { renum[i] = run; run += split[i]; }
    for (I p = 0; p < Pp[n]; p++) Pj[p] = renum[Pj[p]];
    free(renum);
}

/* ---- amg_core.pinv_array, linalg.h:888-954, on top of svd_jacobi, linalg.h:455-780 (one-sided Jacobi SVD of a
 * column-major m x n matrix, real case).  get_block_diag(inv_flag=True) calls it with TransA = 'T' for blocks smaller
 * than 7 (pyamg/util/utils.py:670-677): the inverses of the diagonal blocks the block smoothers use. */
static T vec_dot(const T *x, const T *y, I n) { T s = 0.0; for (I i = 0; i < n; i++) s += x[i] * y[i]; return s; }
static T vec_norm(const T *x, I n) { return sqrt(vec_dot(x, x, n)); }

static I svd_jacobi(const T *A, T *U, T *V, T *S, I m, I n)
{
    const T eps = 2.220446049250313e-16;
    if (m < n) return -1;
    if (n == 1 && m == 1) {
        T normA = fabs(A[0]);
        V[0] = 1.0; S[0] = normA;
        U[0] = (normA == 0.0) ? 1.0 : A[0] / normA;
        return 0;
    }
    I i, j, k, nsq = n * n, count = 1, sweep = 0;
    const I sweepmax = 15 * n > 30 ? 15 * n : 30;
    const T tolerance = sqrt((T)m) * eps;
    for (i = 0; i < nsq; i++) V[i] = 0.0;
    for (i = 0; i < nsq; i += n + 1) V[i] = 1.0;
    for (i = 0; i < m * n; i++) U[i] = A[i];
    for (j = 0; j < n; j++) S[j] = eps * vec_norm(U + j * m, m);
    while (count > 0 && sweep <= sweepmax) {
        count = n * (n - 1) / 2;
        I jm = 0, jn = 0;
        for (j = 0; j < n - 1; j++) {
            I km = (j + 1) * m, kn = (j + 1) * n;
            for (k = j + 1; k < n; k++) {
                T a = vec_norm(U + jm, m), b = vec_norm(U + km, m);
                T d = vec_dot(U + jm, U + km, m);
                T norm_d = fabs(d);
                T abserr_a = S[j], abserr_b = S[k];
                int sorted = (a >= b), orthog = (norm_d <= tolerance * a * b);
                int noisya = (a < abserr_a), noisyb = (b < abserr_b);
                if (sorted && (orthog || noisya || noisyb)) {
                    count--;
                } else if (!sorted || (norm_d == 0.0 && a == b)) {
                    S[j] = abserr_b; S[k] = abserr_a;
                    for (i = 0; i < m; i++) { T uj = U[jm + i], uk = U[km + i]; U[jm + i] = -uk; U[km + i] = uj; }
                    for (i = 0; i < n; i++) { T vj = V[jn + i], vk = V[kn + i]; V[jn + i] = -vk; V[kn + i] = vj; }
                } else {
                    T tau = (b * b - a * a) / (2.0 * norm_d);
                    T t = (tau < 0.0 ? -1.0 : 1.0) / (fabs(tau) + sqrt(1.0 + tau * tau));
                    T c = 1.0 / (sqrt(1.0 + t * t));
                    T sn = d * (t * c / norm_d);
                    T nsn = -sn;
                    T norm_sin = fabs(sn);
                    S[j] = fabs(c) * abserr_a + norm_sin * abserr_b;
                    S[k] = norm_sin * abserr_a + fabs(c) * abserr_b;
                    for (i = 0; i < m; i++) {
                        T uj = U[jm + i], uk = U[km + i];
                        U[jm + i] = uj * c + nsn * uk;
                        U[km + i] = sn * uj + uk * c;
                    }
                    for (i = 0; i < n; i++) {
                        T vj = V[jn + i], vk = V[kn + i];
                        V[jn + i] = vj * c + nsn * vk;
                        V[kn + i] = sn * vj + vk * c;
                    }
                }
                km += m; kn += n;
            }
            jm += m; jn += n;
        }
        sweep++;
    }
    T sigma_tol = 0.0;
    I iszero = n;
    for (j = 0; j < n; j++) {
        T cur = vec_norm(U + j * m, m);
        if (j == 0) { T alpha = 50.0 / sqrt(sqrt(eps)); sigma_tol = alpha * cur * eps; }
        if (cur <= sigma_tol) {
            iszero--;
            S[j] = 0.0;
            for (i = 0; i < m; i++) U[j * m + i] = 0.0;
        } else {
            S[j] = cur;
            for (i = 0; i < m; i++) U[j * m + i] = U[j * m + i] / cur;
        }
    }
    if (iszero == 0) {
        for (i = 0; i < nsq; i++) V[i] = 0.0;
        for (i = 0; i < nsq; i += n + 1) V[i] = 1.0;
        for (i = 0; i < n * m; i += m + 1) U[i] = 1.0;
        return 0;
    }
    return count > 0 ? 1 : 0;
}

void orc_pinv_array(T *AA, I m, I n, I transA)
{
    const I nsq = n * n;
    T *Tran = (T *)malloc(sizeof(T) * nsq), *U = (T *)malloc(sizeof(T) * nsq), *V = (T *)malloc(sizeof(T) * nsq);
    T *SinvUh = (T *)malloc(sizeof(T) * nsq), *S = (T *)malloc(sizeof(T) * n);
    for (I blk = 0; blk < m; blk++) {
        T *A = AA + (size_t)blk * nsq;
        if (transA) {
            for (I r = 0; r < n; r++) for (I c = 0; c < n; c++) Tran[c * n + r] = A[r * n + c];
            svd_jacobi(Tran, U, V, S, n, n);
        } else {
            svd_jacobi(A, U, V, S, n, n);
        }
        for (I j = 0; j < n; j++) if (S[j] != 0.0) S[j] = 1.0 / S[j];
        I counter = 0;
        for (I j = 0; j < n; j++) {                 /* Sinv(k) * U(j,k) -> SinvUh(k,j), column major */
            I uo = j;
            for (I k = 0; k < n; k++) { SinvUh[counter] = U[uo] * S[k]; counter++; uo += n; }
        }
        for (I r = 0; r < n; r++) for (I c = 0; c < n; c++) Tran[c * n + r] = V[r * n + c];
        /* gemm(Tran row-major n x n, SinvUh column-major, out row-major, overwrite): out[i][j] = sum_k Tran[i][k] SinvUh[k + j n] */
        for (I i = 0; i < n; i++)
            for (I j = 0; j < n; j++) {
                T sum = 0.0;
                for (I k = 0; k < n; k++) sum += Tran[i * n + k] * SinvUh[j * n + k];
                A[i * n + j] = sum;
            }
    }
    free(Tran); free(U); free(V); free(SinvUh); free(S);
}

/* ---- amg_core.vertex_coloring_jones_plassmann / vertex_coloring_LDF, graph.h:242-335, over one sweep of
 * maximal_independent_set_parallel (:90-160, run serially in index order with in-place state) and
 * vertex_coloring_first_fit (:200-222).  State in x: -1 uncoloured, -2 excluded this round, >= 0 colour. */
static I mis_sweep(I n, const I *Ap, const I *Aj, I K, I *x, const T *y)
{
    I found = 0;
    for (I i = 0; i < n; i++) {
        if (x[i] != -1) continue;
        const T yi = y[i];
        I p;
        for (p = Ap[i]; p < Ap[i + 1]; p++) {
            const I j = Aj[p];
            if (x[j] == K) { x[i] = -2; break; }
            if (x[j] == -1) {
                if (y[j] > yi) break;
                if (y[j] == yi && j > i) break;
            }
        }
        if (p == Ap[i + 1]) {
            for (p = Ap[i]; p < Ap[i + 1]; p++) if (x[Aj[p]] == -1) x[Aj[p]] = -2;
            found++;
            x[i] = K;
        }
    }
    return found;
}

static void first_fit(I n, const I *Ap, const I *Aj, I *x, I K)
{
    char *used = (char *)malloc((size_t)K + 1);
    for (I i = 0; i < n; i++) {
        if (x[i] != K) continue;
        for (I c = 0; c <= K; c++) used[c] = 0;
        for (I p = Ap[i]; p < Ap[i + 1]; p++) {
            const I j = Aj[p];
            if (j == i || x[j] < 0) continue;
            used[x[j]] = 1;
        }
        I c = 0;
        while (c < K && used[c]) c++;
        x[i] = c;
    }
    free(used);
}

I orc_vertex_coloring_jp(I n, const I *Ap, const I *Aj, I *x, T *z)
{
    for (I i = 0; i < n; i++) { x[i] = -1; z[i] += Ap[i + 1] - Ap[i]; }
    I done = 0, K = 0, best = -1;
    while (done < n) {
        done += mis_sweep(n, Ap, Aj, K, x, z);
        for (I i = 0; i < n; i++) if (x[i] == -2) x[i] = -1;
        first_fit(n, Ap, Aj, x, K);
        K++;
    }
    for (I i = 0; i < n; i++) if (x[i] > best) best = x[i];
    return best;
}

I orc_vertex_coloring_ldf(I n, const I *Ap, const I *Aj, I *x, const T *y)
{
    T *w = (T *)malloc(sizeof(T) * (n > 0 ? n : 1));
    for (I i = 0; i < n; i++) x[i] = -1;
    I done = 0, K = 0, best = -1;
    while (done < n) {
        for (I i = 0; i < n; i++) {
            if (x[i] != -1) continue;
            I free_nb = 0;
            for (I p = Ap[i]; p < Ap[i + 1]; p++) if (x[Aj[p]] == -1 && Aj[p] != i) free_nb++;
            w[i] = y[i] + free_nb;
        }
        done += mis_sweep(n, Ap, Aj, K, x, w);
        for (I i = 0; i < n; i++) if (x[i] == -2) x[i] = -1;
        first_fit(n, Ap, Aj, x, K);
        K++;
    }
    for (I i = 0; i < n; i++) if (x[i] > best) best = x[i];
    free(w);
    return best;
}
