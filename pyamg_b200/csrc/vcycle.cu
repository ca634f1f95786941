// Whole-cycle engine: multilevel_solver.__solve (pyamg/multilevel.py:473-548) as a fixed
// sequence of kernel launches over caller-owned level arrays, replayed as a CUDA graph.
//
//   presmooth -> r = b - A x -> b_c = R r -> recurse / dense coarse solve -> x += P x_c -> postsmooth
//
// Level vectors (r, b_c, x_c, Jacobi ping-pong buffer, extracted diagonal) live in one
// caller-provided work buffer.  Jacobi sweeps ping-pong between x and the spare buffer instead
// of the reference's temp copy (same values); a pre-smoothing Jacobi sweep from the zero
// initial guess of a coarse level is the bit-identical diagonal scaling (blas1.cu).
#include <stdlib.h>
#include <vector>
#include "internal.cuh"

namespace b200amg {

static bool diag_free_copies()
{
    const char *e = getenv("B200AMG_SELL_DIAGFREE");       // read per engine build (tests switch it)
    return !(e && e[0] == '0');
}

struct LevelRt {
    b200amg_level_desc d;
    double *r, *bc, *xc, *tmp, *diag;
    double *poly = nullptr;                   // third work vector of the polynomial smoother
    long long nzA = -1, nzP = -1, nzR = -1;   // stored entries (tile sizing of the stream kernels)
    Sell sA, sR, sP;                          // engine-owned sliced-ELL copies of long-row operators
    Sell sGS;                                 // colour-grouped copy of A for colour-ordered Gauss-Seidel
    Sell sJ;                                  // copy of A WITHOUT its diagonal entries for the Jacobi sweeps of a big level
    BsrColor bGS;                             // the same for a block operator (bsr.cu)
    RowPat rpA;                               // one byte per row of a stencil operator (rowpat.cu)
    ValueDict dA, dP, dR;                     // 8-bit value codes of operators with <= 256 distinct values
    std::vector<int> color_ptr;
    int *cptr_dev = nullptr;                  // the same offsets on the device (small_level.cu)
};

struct GraphKey {
    double *x; const double *b; int kind;   // kind: cycle kind | 4 when the cycle starts from a zero guess
    cudaGraphExec_t exec;
    long long launches;   // kernel nodes in the graph (for b200amg_launch_count)
};

// per-operation event timing of one un-graphed cycle (b200amg_vcycle_profile)
struct Prof {
    std::vector<cudaEvent_t> ev;
    std::vector<int> code, level;
    bool on = false;
};

struct Engine {
    Prof prof;
    std::vector<LevelRt> lev;
    const double *coarse_inv;
    int nc_last;
    double *dot_work, *r0;
    std::vector<GraphKey> graphs;
    bool diag_ready;
};

static void free_level(LevelRt &L)
{
    sell_free(&L.sA); sell_free(&L.sR); sell_free(&L.sP); sell_free(&L.sGS); sell_free(&L.sJ);
    dict_free(&L.dA); dict_free(&L.dP); dict_free(&L.dR);
    rowpat_free(&L.rpA);
    bsr_color_free(&L.bGS);
    if (L.cptr_dev) cudaFree(L.cptr_dev);
    L.cptr_dev = nullptr;
}

// B200AMG_VALUE_DICT=0 switches the value dictionary off (A/B measurements)
// B200AMG_VALUE_DICT: 0 = operators are streamed as stored, 1 = value dictionary only,
// 2 = (offset, value) pattern dictionary for A where it applies, 3 / unset = row-pattern layout for A where
// it applies; each level falls back to the one below when the operator does not qualify
static int use_value_dict()
{
    const char *e = getenv("B200AMG_VALUE_DICT");
    return (e && e[0] >= '0' && e[0] <= '3') ? e[0] - '0' : 3;
}

// level vectors are padded by an odd number of 2 KB pages so that same-index entries of different
// vectors (all 2^k-sized on structured grids) do not land on the same DRAM channel / L2 slice
static size_t al(size_t n_doubles) { return ((n_doubles + 31) & ~(size_t)31) + 9 * 256; }

long long work_doubles(int nlev, const b200amg_level_desc *L)
{
    size_t tot = al(1024) ;
    for (int k = 0; k < nlev - 1; ++k)
        tot += al(L[k].n) * (3 + (L[k].pre_kind == 4 || L[k].post_kind == 4)) /* r, tmp, diag [, poly] */ +
               al(L[k].nc) * 2 /* bc, xc */;
    return (long long)tot;
}

// ---- level operators ---------------------------------------------------------------------
static int apply_A(const LevelRt &L, int op_scalar, int bsr_mode, const double *x, const double *b, double *y,
                   cudaStream_t st)
{
    const b200amg_level_desc &d = L.d;
    if (d.bs == 1) {
        StreamArgs a{d.Ap, d.Aj, d.Ax, x, b, y, L.nzA, 0, d.n, 1.0, 0.0, 0.0};
        L.dA.apply(a);
        if (L.rpA.ok()) return rowpat_op(op_scalar, L.rpA, a, st);
        if (L.sA.ok()) return sell_op(op_scalar, L.sA, a, st);
        return stream_op(op_scalar, a, st);
    }
    return bsr_matvec(d.n / d.bs, d.bs, d.bs, d.Ap, d.Aj, d.Ax, x, b, y, bsr_mode, st, L.nzA);
}

static int restrict_r(const LevelRt &L, const double *r, double *bc, cudaStream_t st)
{
    const b200amg_level_desc &d = L.d;
    if (d.p_br == 1 && d.p_bc == 1) {
        StreamArgs a{d.Rp, d.Rj, d.Rx, r, nullptr, bc, L.nzR, 0, d.nc, 1.0, 0.0, 0.0};
        L.dR.apply(a);
        if (L.sR.ok()) return sell_op(OP_MATVEC, L.sR, a, st);
        return stream_op(OP_MATVEC, a, st);
    }
    return bsr_matvec(d.nc / d.p_bc, d.p_bc, d.p_br, d.Rp, d.Rj, d.Rx, r, nullptr, bc, 0, st, L.nzR);
}

// x_out = x_in + P x_c (x_in == nullptr: in place)
static int prolong_add(const LevelRt &L, const double *xc, const double *x_in, double *x, cudaStream_t st)
{
    const b200amg_level_desc &d = L.d;
    if (d.p_br == 1 && d.p_bc == 1) {
        StreamArgs a{d.Pp, d.Pj, d.Px, xc, x_in, x, L.nzP, 0, d.n, 1.0, 0.0, 0.0};
        L.dP.apply(a);
        if (L.sP.ok()) return sell_op(OP_ADD, L.sP, a, st);
        return stream_op(OP_ADD, a, st);
    }
    return bsr_matvec(d.n / d.p_br, d.p_br, d.p_bc, d.Pp, d.Pj, d.Px, xc, x_in, x, 2, st, L.nzP);
}

// one smoothing stage; *cur is the buffer holding the iterate (x or the spare), swapped by Jacobi
static int smooth(const LevelRt &L, bool post, int kind, double omega, int iters, int sweep, double **cur, double **spare,
                  const double *b, bool x_is_zero, cudaStream_t st)
{
    const b200amg_level_desc &d = L.d;
    if (kind == 0) {
        if (x_is_zero) B200_CUDA(cudaMemsetAsync(*cur, 0, sizeof(double) * d.n, st));
        return 0;
    }
    if (kind == 1) {                                   // Jacobi
        for (int it = 0; it < iters; ++it) {
            if (it == 0 && x_is_zero && d.bs == 1) {
                B200_TRY(scale_diag(d.n, omega, b, L.diag, *cur, d.flavor, st));
                continue;
            }
            if (it == 0 && x_is_zero) B200_CUDA(cudaMemsetAsync(*cur, 0, sizeof(double) * d.n, st));
            if (d.bs == 1) {
                StreamArgs a{d.Ap, d.Aj, d.Ax, *cur, b, *spare, L.nzA, 0, d.n, 1.0, 0.0, omega};
                L.dA.apply(a);
                if (L.rpA.ok()) B200_TRY(rowpat_op(d.flavor ? OP_JACOBI_BSR1 : OP_JACOBI, L.rpA, a, st));
                else if (L.sJ.ok()) B200_TRY(sell_op(d.flavor ? OP_JACOBI_BSR1 : OP_JACOBI, L.sJ, a, st));
                else if (L.sA.ok()) B200_TRY(sell_op(d.flavor ? OP_JACOBI_BSR1 : OP_JACOBI, L.sA, a, st));
                else B200_TRY(stream_op(d.flavor ? OP_JACOBI_BSR1 : OP_JACOBI, a, st));
            } else {
                B200_TRY(bsr_jacobi(d.bs, 0, d.n / d.bs, d.Ap, d.Aj, d.Ax, *cur, *spare, b, omega, st));
            }
            double *t = *cur; *cur = *spare; *spare = t;
        }
        if (iters == 0 && x_is_zero) B200_CUDA(cudaMemsetAsync(*cur, 0, sizeof(double) * d.n, st));
        return 0;
    }
    if (kind == 2) {                                   // colour-ordered Gauss-Seidel (indexed)
        const int K = d.K;
        {   // small level: the whole step in one launch, iterate in shared memory (same colour order, same bits)
            const int rc = gs_level_cta(d.bs, d.flavor != 0 && !d.gs_csr, d.n, L.nzA * d.bs * d.bs, d.Ap, d.Aj, d.Ax, d.Id,
                                        L.cptr_dev, K, *cur, b, iters, sweep, x_is_zero, st);
            if (rc >= 0) return rc;
        }
        if (x_is_zero) B200_CUDA(cudaMemsetAsync(*cur, 0, sizeof(double) * d.n, st));
        for (int it = 0; it < iters; ++it) {
            for (int pass = 0; pass < 2; ++pass) {
                const bool bwd = (pass == 1);
                if ((sweep == 0 && bwd) || (sweep == 1 && !bwd)) continue;
                for (int q = 0; q < K; ++q) {
                    const int c = bwd ? K - 1 - q : q;
                    const int p0 = L.color_ptr[c], p1 = L.color_ptr[c + 1];
                    if (d.bs == 1 && L.sGS.ok())
                        B200_TRY(sell_gs_color(d.flavor != 0 && !d.gs_csr, L.sGS, d.Id, p0, p1, *cur, b, st));
                    else if (d.bs == 1)
                        B200_TRY(gs_rows(d.flavor != 0 && !d.gs_csr, d.Ap, d.Aj, d.Ax, *cur, b, d.Id, p0, p1, st));
                    else if (L.bGS.ok())
                        B200_TRY(bsr_gs_color(d.bs, L.bGS.rp, L.bGS.cj, L.bGS.cx, *cur, b, d.Id, p0, p1, bwd, st,
                                              (double)L.nzA / (double)(d.n / d.bs), true, true));
                    else
                        B200_TRY(bsr_gs_color(d.bs, d.Ap, d.Aj, d.Ax, *cur, b, d.Id, p0, p1, bwd, st,
                                              (double)L.nzA / (double)(d.n / d.bs), false, true));
                }
            }
        }
        return 0;
    }
    if (kind == 3) {                                   // lexicographic Gauss-Seidel: exact, sequential
        if (x_is_zero) B200_CUDA(cudaMemsetAsync(*cur, 0, sizeof(double) * d.n, st));
        const int nb = d.n / d.bs;
        for (int it = 0; it < iters; ++it) {
            for (int pass = 0; pass < 2; ++pass) {
                const bool bwd = (pass == 1);
                if ((sweep == 0 && bwd) || (sweep == 1 && !bwd)) continue;
                const int r0 = bwd ? nb - 1 : 0, r1 = bwd ? -1 : nb, dr = bwd ? -1 : 1;
                if (d.bs == 1)
                    B200_TRY(gs_sequential(d.flavor != 0, d.Ap, d.Aj, d.Ax, *cur, b, nullptr, r0, r1, dr, st));
                else
                    B200_TRY(bsr_gs_sequential(d.bs, d.Ap, d.Aj, d.Ax, *cur, b, nullptr, r0, r1, dr, st));
            }
        }
        return 0;
    }
    if (kind == 4) {                                   // polynomial (Chebyshev): relaxation.py:593-668
        const double *coef = post ? d.post_coef : d.pre_coef;
        const int ncoef = post ? d.post_ncoef : d.pre_ncoef;
        double *r = L.r, *h = *spare, *t = L.poly;
        if (x_is_zero) B200_CUDA(cudaMemsetAsync(*cur, 0, sizeof(double) * d.n, st));
        for (int it = 0; it < iters; ++it) {
            // "norm(x) == 0 -> residual = b" of the reference is the same bits as b - A*0
            if (it == 0 && x_is_zero) B200_CUDA(cudaMemcpyAsync(r, b, sizeof(double) * d.n, cudaMemcpyDeviceToDevice, st));
            else B200_TRY(apply_A(L, OP_RESIDUAL, 1, *cur, b, r, st));
            B200_TRY(scal(d.n, coef[0], r, h, st));
            for (int c = 1; c < ncoef; ++c) {          // h = c*residual + A*h
                B200_TRY(scal(d.n, coef[c], r, t, st));
                B200_TRY(apply_A(L, OP_ADD, 2, h, nullptr, t, st));
                double *sw = h; h = t; t = sw;
            }
            B200_TRY(axpy(d.n, 1.0, h, *cur, st));     // x += h
        }
        return 0;
    }
    if (kind == 5) {                                   // block Jacobi: ping-pong like kind 1
        const int nb = d.n / d.bs;
        if (x_is_zero) B200_CUDA(cudaMemsetAsync(*cur, 0, sizeof(double) * d.n, st));
        for (int it = 0; it < iters; ++it) {
            B200_TRY(block_jacobi_pp(d.bs, nb, d.Ap, d.Aj, d.Ax, *cur, *spare, b, d.Dinv, omega, st));
            double *t = *cur; *cur = *spare; *spare = t;
        }
        return 0;
    }
    if (kind == 6 || kind == 7) {                      // block Gauss-Seidel: colour-ordered / lexicographic
        if (x_is_zero) B200_CUDA(cudaMemsetAsync(*cur, 0, sizeof(double) * d.n, st));
        const int nb = d.n / d.bs, K = d.K;
        for (int it = 0; it < iters; ++it) {
            for (int pass = 0; pass < 2; ++pass) {
                const bool bwd = (pass == 1);
                if ((sweep == 0 && bwd) || (sweep == 1 && !bwd)) continue;
                if (kind == 7) {
                    B200_TRY(block_gs_sequential(d.bs, d.Ap, d.Aj, d.Ax, *cur, b, d.Dinv, nullptr, bwd ? nb - 1 : 0,
                                                 bwd ? -1 : nb, bwd ? -1 : 1, st));
                    continue;
                }
                for (int q = 0; q < K; ++q) {
                    const int c = bwd ? K - 1 - q : q;
                    B200_TRY(block_gs_color(d.bs, d.Ap, d.Aj, d.Ax, *cur, b, d.Dinv, d.Id, L.color_ptr[c], L.color_ptr[c + 1], st));
                }
            }
        }
        return 0;
    }
    return set_err(B200AMG_EINVAL, "smooth: unknown smoother kind");
}

static int mark(Engine &E, int code, int level, cudaStream_t st)
{
    if (!E.prof.on) return 0;
    cudaEvent_t e;
    B200_CUDA(cudaEventCreate(&e));
    B200_CUDA(cudaEventRecord(e, st));
    E.prof.ev.push_back(e);
    E.prof.code.push_back(code);
    E.prof.level.push_back(level);
    return 0;
}

static int cycle(Engine &E, int k, double *x, const double *b, int kind, bool x_is_zero, cudaStream_t st)
{
    LevelRt &L = E.lev[k];
    const b200amg_level_desc &d = L.d;
    const int nlev = (int)E.lev.size() + 1;
    double *cur = x, *spare = L.tmp;
    B200_TRY(smooth(L, false, d.pre_kind, d.pre_omega, d.pre_iters, d.pre_sweep, &cur, &spare, b, x_is_zero, st));
    B200_TRY(mark(E, B200AMG_OP_PRESMOOTH, k, st));
    B200_TRY(apply_A(L, OP_RESIDUAL, 1, cur, b, L.r, st));
    B200_TRY(mark(E, B200AMG_OP_RESIDUAL, k, st));
    B200_TRY(restrict_r(L, L.r, L.bc, st));
    B200_TRY(mark(E, B200AMG_OP_RESTRICT, k, st));
    if (k == nlev - 2) {
        if (E.coarse_inv) B200_TRY(dense_matvec(d.nc, E.coarse_inv, L.bc, L.xc, st));
        else B200_CUDA(cudaMemsetAsync(L.xc, 0, sizeof(double) * d.nc, st));
        B200_TRY(mark(E, B200AMG_OP_COARSE, k + 1, st));
    } else if (kind == 0) {
        B200_TRY(cycle(E, k + 1, L.xc, L.bc, 0, true, st));
    } else if (kind == 1) {
        B200_TRY(cycle(E, k + 1, L.xc, L.bc, 1, true, st));
        B200_TRY(cycle(E, k + 1, L.xc, L.bc, 1, false, st));
    } else {
        B200_TRY(cycle(E, k + 1, L.xc, L.bc, 2, true, st));
        B200_TRY(cycle(E, k + 1, L.xc, L.bc, 0, false, st));
    }
    // The iterate must end in x.  Jacobi sweeps ping-pong between x and the spare buffer; when the sweeps still to
    // come would leave the result in the spare buffer, the correction is written OUT OF PLACE (other buffer =
    // iterate + P x_c, same bits) -- one more swap, and the D2D copy back to x disappears.
    const int swaps_to_come = (d.post_kind == 1 || d.post_kind == 5) ? d.post_iters : 0;
    if (((cur == x) ? 0 : 1) != (swaps_to_come & 1)) {
        B200_TRY(prolong_add(L, L.xc, cur, spare, st));
        double *t = cur; cur = spare; spare = t;
    } else {
        B200_TRY(prolong_add(L, L.xc, nullptr, cur, st));
    }
    B200_TRY(mark(E, B200AMG_OP_PROLONG, k, st));
    B200_TRY(smooth(L, true, d.post_kind, d.post_omega, d.post_iters, d.post_sweep, &cur, &spare, b, false, st));
    B200_TRY(mark(E, B200AMG_OP_POSTSMOOTH, k, st));
    if (cur != x) {
        B200_CUDA(cudaMemcpyAsync(x, cur, sizeof(double) * d.n, cudaMemcpyDeviceToDevice, st));
        B200_TRY(mark(E, B200AMG_OP_COPYBACK, k, st));
    }
    return 0;
}

static int ensure_diag(Engine &E, cudaStream_t st)
{
    if (E.diag_ready) return 0;
    for (auto &L : E.lev)
        B200_TRY(extract_diag(L.d.n / L.d.bs, L.d.bs, L.d.Ap, L.d.Aj, L.d.Ax, L.diag, st));
    E.diag_ready = true;
    return 0;
}

}  // namespace b200amg

using namespace b200amg;

extern "C" long long b200amg_vcycle_work_bytes(int nlev, const b200amg_level_desc *levels_host)
{
    if (nlev < 1 || !levels_host) return B200AMG_EINVAL;
    return work_doubles(nlev, levels_host) * (long long)sizeof(double);
}

extern "C" int b200amg_vcycle_create(int nlev, const b200amg_level_desc *Ld, const double *coarse_inv, void *work,
                                     long long work_bytes, void **handle_out)
{
    if (nlev < 2 || !Ld || !handle_out) return set_err(B200AMG_EINVAL, "vcycle_create: need >= 2 levels");
    if (work_bytes < b200amg_vcycle_work_bytes(nlev, Ld)) return set_err(B200AMG_ENOMEM, "vcycle_create: work too small");
    Engine *E = new Engine();
    double *w = reinterpret_cast<double *>(work);
    E->dot_work = w; w += al(1024);
    E->coarse_inv = coarse_inv;
    E->diag_ready = false;
    for (int k = 0; k < nlev - 1; ++k) {
        LevelRt L;
        L.d = Ld[k];
        if (L.d.bs < 1 || L.d.bs > 8 || L.d.n % L.d.bs) { delete E; return set_err(B200AMG_EINVAL, "vcycle_create: bad blocksize"); }
        L.r = w; w += al(L.d.n);
        L.tmp = w; w += al(L.d.n);
        L.diag = w; w += al(L.d.n);
        if (L.d.pre_kind == 4 || L.d.post_kind == 4) { L.poly = w; w += al(L.d.n); }
        for (int side = 0; side < 2; ++side) {
            const int kind = side ? L.d.post_kind : L.d.pre_kind, nc = side ? L.d.post_ncoef : L.d.pre_ncoef;
            if (kind == 4 && (nc < 1 || nc > 8)) { delete E; return set_err(B200AMG_EINVAL, "vcycle_create: polynomial smoother needs 1..8 coefficients"); }
        }
        L.bc = w; w += al(L.d.nc);
        L.xc = w; w += al(L.d.nc);
        for (int side = 0; side < 2; ++side) {
            const int kind = side ? L.d.post_kind : L.d.pre_kind;
            if (kind >= 5 && kind <= 7 && !L.d.Dinv) { delete E; return set_err(B200AMG_EINVAL, "vcycle_create: block smoother without Dinv"); }
        }
        if (L.d.pre_kind == 2 || L.d.post_kind == 2 || L.d.pre_kind == 6 || L.d.post_kind == 6) {
            if (!L.d.color_ptr_host || !L.d.Id || L.d.K < 0) { delete E; return set_err(B200AMG_EINVAL, "vcycle_create: colour ordering missing"); }
            L.color_ptr.assign(L.d.color_ptr_host, L.d.color_ptr_host + L.d.K + 1);
            L.d.color_ptr_host = nullptr;
            if (cudaMalloc(&L.cptr_dev, sizeof(int) * (L.d.K + 1)) != cudaSuccess ||
                cudaMemcpy(L.cptr_dev, L.color_ptr.data(), sizeof(int) * (L.d.K + 1), cudaMemcpyHostToDevice) != cudaSuccess) {
                cudaGetLastError();
                if (L.cptr_dev) cudaFree(L.cptr_dev);
                L.cptr_dev = nullptr;                    // the per-colour launches do not need it
            }
        }
        // nonzero counts (one small D2H read per matrix, at create time only) size the CTAs
        auto last = [](const int *p, int n, int *out) {
            return p ? cudaMemcpy(out, p + n, sizeof(int), cudaMemcpyDeviceToHost) : cudaSuccess;
        };
        int nzA = 0, nzP = 0, nzR = 0;
        if (last(L.d.Ap, L.d.n / L.d.bs, &nzA) != cudaSuccess || last(L.d.Pp, L.d.n / L.d.p_br, &nzP) != cudaSuccess ||
            last(L.d.Rp, L.d.nc / L.d.p_bc, &nzR) != cudaSuccess) {
            delete E;
            return set_err(B200AMG_EINVAL, "vcycle_create: cannot read row pointers");
        }
        L.nzA = nzA; L.nzP = nzP; L.nzR = nzR;
        // long-row scalar operators get a sliced-ELL copy (sell.cu); a declined / failed build keeps CSR
        int src = 0;
        // (with 8-bit value codes in the value slots when the operator has <= 256 distinct values: R_0 of a
        // stencil problem has ~85)
        auto sell_with_codes = [&](int n, long long nz, const int *Xp, const int *Xj, const double *Xx, Sell *out) -> int {
            ValueDict vd;
            int rc = use_value_dict() ? dict_build(nz, Xx, &vd, nullptr) : 0;
            if (!rc) rc = sell_build(n, nz, Xp, Xj, Xx, out, nullptr, nullptr, vd.codes, vd.dict);
            dict_free(&vd);
            return rc;
        };
        if (L.d.bs == 1 && sell_wanted(nzA, L.d.n)) src = sell_with_codes(L.d.n, nzA, L.d.Ap, L.d.Aj, L.d.Ax, &L.sA);
        // Jacobi on a big uncoded level: a second copy without the diagonal entries (the sweep never adds them to the row
        // sum; the others keep their order: same bits).  The loop loses the per-slot diagonal test and with it half of its
        // registers -- the smoothing pass then streams like the residual pass.  B200AMG_SELL_DIAGFREE=0: A/B switch.
        if (!src && L.sA.ok() && !L.sA.cc && L.d.n >= 400000 && (L.d.pre_kind == 1 || L.d.post_kind == 1) && diag_free_copies())
            src = sell_build(L.d.n, nzA, L.d.Ap, L.d.Aj, L.d.Ax, &L.sJ, nullptr, nullptr, nullptr, nullptr, true);
        if (!src && L.d.p_br == 1 && L.d.p_bc == 1) {
            if (sell_wanted(nzR, L.d.nc)) src = sell_with_codes(L.d.nc, nzR, L.d.Rp, L.d.Rj, L.d.Rx, &L.sR);
            if (!src && sell_wanted(nzP, L.d.n)) src = sell_with_codes(L.d.n, nzP, L.d.Pp, L.d.Pj, L.d.Px, &L.sP);
        }
        // short-row operators streamed by the TMA pipeline: try an 8-bit value dictionary (dict.cu)
        // (A first as a pattern dictionary -- offset and value in one code -- then values only)
        if (!src && L.d.bs == 1 && !L.sA.ok() && nzA > 0 && (double)nzA / L.d.n <= 8.0 && use_value_dict()) {
            if (use_value_dict() >= 3) src = rowpat_build(L.d.n, nzA, L.d.Ap, L.d.Aj, L.d.Ax, &L.rpA, nullptr);
            if (!src && !L.rpA.ok() && use_value_dict() >= 2)
                src = pattern_build(L.d.n, nzA, L.d.Ap, L.d.Aj, L.d.Ax, &L.dA, nullptr);
            if (!src && !L.rpA.ok() && !L.dA.ok()) src = dict_build(nzA, L.d.Ax, &L.dA, nullptr);
        }
        if (!src && L.d.p_br == 1 && L.d.p_bc == 1 && use_value_dict()) {
            if (!L.sP.ok() && nzP > 0 && (double)nzP / L.d.n <= 8.0) src = dict_build(nzP, L.d.Px, &L.dP, nullptr);
            if (!src && !L.sR.ok() && nzR > 0 && (double)nzR / L.d.nc <= 8.0) src = dict_build(nzR, L.d.Rx, &L.dR, nullptr);
        }
        // colour-ordered Gauss-Seidel: colour-grouped SELL copy of A when the index list is a permutation
        if (!src && L.d.bs == 1 && (L.d.pre_kind == 2 || L.d.post_kind == 2) && L.d.K > 0 &&
            L.color_ptr[0] == 0 && L.color_ptr[L.d.K] == L.d.n && L.d.n >= 128)
            src = sell_build(L.d.n, nzA, L.d.Ap, L.d.Aj, L.d.Ax, &L.sGS, nullptr, L.d.Id, nullptr, nullptr, diag_free_copies());
        // ... and the block form: rows of a colour contiguous (B200AMG_BSR_COLOR_COPY=0: A/B switch)
        if (!src && L.d.bs > 1 && (L.d.pre_kind == 2 || L.d.post_kind == 2) && L.d.K > 0 && L.color_ptr[0] == 0 &&
            L.color_ptr[L.d.K] == L.d.n / L.d.bs && L.d.n / L.d.bs >= 200000) {   // large levels only: no gain on small ones (run 7)
            const char *ev = getenv("B200AMG_BSR_COLOR_COPY");
            if (!(ev && ev[0] == '0')) src = bsr_color_build(L.d.n / L.d.bs, L.d.bs, nzA, L.d.Ap, L.d.Aj, L.d.Ax, L.d.Id, &L.bGS, nullptr);
        }
        if (src) {
            free_level(L);
            for (auto &Lp : E->lev) free_level(Lp);
            delete E;
            return src;
        }
        E->lev.push_back(L);
    }
    E->nc_last = Ld[nlev - 2].nc;
    E->r0 = E->lev[0].r;
    *handle_out = E;
    return 0;
}

static int run_cycle(Engine *E, double *x, const double *b, int cycle_kind, int use_graph, bool from_zero, cudaStream_t st)
{
    B200_TRY(ensure_diag(*E, st));
    if (!use_graph || st == nullptr) return cycle(*E, 0, x, b, cycle_kind, from_zero, st);
    const int key_kind = cycle_kind | (from_zero ? 4 : 0);
    for (auto &g : E->graphs)
        if (g.x == x && g.b == b && g.kind == key_kind) {
            B200_CUDA(cudaGraphLaunch(g.exec, st));
            g_launches += g.launches;
            return 0;
        }
    cudaGraph_t graph;
    const long long before = g_launches;
    B200_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    int rc = cycle(*E, 0, x, b, cycle_kind, from_zero, st);
    cudaError_t ce = cudaStreamEndCapture(st, &graph);
    if (rc) return rc;
    if (ce != cudaSuccess) return set_err((int)ce, "vcycle_run: stream capture failed");
    GraphKey key{x, b, key_kind, nullptr, g_launches - before};
    B200_CUDA(cudaGraphInstantiate(&key.exec, graph, 0));
    B200_CUDA(cudaGraphDestroy(graph));
    if (E->graphs.size() >= 12) { cudaGraphExecDestroy(E->graphs.front().exec); E->graphs.erase(E->graphs.begin()); }
    E->graphs.push_back(key);
    B200_CUDA(cudaGraphLaunch(key.exec, st));
    return 0;
}

extern "C" int b200amg_vcycle_run(void *handle, double *x, const double *b, int cycle_kind, int use_graph, void *stream)
{
    Engine *E = reinterpret_cast<Engine *>(handle);
    if (!E || cycle_kind < 0 || cycle_kind > 2) return set_err(B200AMG_EINVAL, "vcycle_run: bad arguments");
    return run_cycle(E, x, b, cycle_kind, use_graph, false, reinterpret_cast<cudaStream_t>(stream));
}

// x <- cycle(0, b): the preconditioner application (multilevel.py:274-314 calls solve(b, x0 = zeros, maxiter = 1)).
// x need not be cleared by the caller; the first pre-smoothing sweep from the known-zero guess is the bit-identical
// diagonal scaling / a memset, like on the coarse levels of every cycle -- one pass over A_0 less than b200amg_vcycle_run
extern "C" int b200amg_vcycle_run_zero(void *handle, double *x, const double *b, int cycle_kind, int use_graph, void *stream)
{
    Engine *E = reinterpret_cast<Engine *>(handle);
    if (!E || cycle_kind < 0 || cycle_kind > 2) return set_err(B200AMG_EINVAL, "vcycle_run_zero: bad arguments");
    return run_cycle(E, x, b, cycle_kind, use_graph, true, reinterpret_cast<cudaStream_t>(stream));
}

// y = A_level x (op 0) or y = b - A_level x (op 2) in the layout the engine streams that operator in: what a Krylov
// method around the cycle needs (pyamg/krylov/_cg.py: A*p, b - A*x) without a second copy of A in plain CSR traffic
extern "C" int b200amg_vcycle_apply_op(void *handle, int level, int op, const double *x, const double *b, double *y,
                                      void *stream)
{
    Engine *E = reinterpret_cast<Engine *>(handle);
    if (!E || level < 0 || level >= (int)E->lev.size() || (op != 0 && op != 2))
        return set_err(B200AMG_EINVAL, "vcycle_apply_op: bad arguments");
    return apply_A(E->lev[level], op == 0 ? OP_MATVEC : OP_RESIDUAL, op == 0 ? 0 : 1, x, b, y,
                   reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int b200amg_vcycle_profile(void *handle, double *x, const double *b, int cycle_kind, int max_ops,
                                      int *op_code_host, int *op_level_host, float *op_ms_host, int *n_ops_host,
                                      void *stream)
{
    Engine *E = reinterpret_cast<Engine *>(handle);
    if (!E || cycle_kind < 0 || cycle_kind > 2 || !n_ops_host) return set_err(B200AMG_EINVAL, "vcycle_profile: bad arguments");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    B200_TRY(ensure_diag(*E, st));
    Prof &P = E->prof;
    P.on = true;
    int rc = mark(*E, -1, 0, st);
    if (!rc) rc = cycle(*E, 0, x, b, cycle_kind, false, st);
    P.on = false;
    cudaError_t ce = cudaStreamSynchronize(st);
    int n = 0;
    if (!rc && ce == cudaSuccess) {
        for (size_t i = 1; i < P.ev.size() && n < max_ops; ++i, ++n) {
            float ms = 0.f;
            cudaEventElapsedTime(&ms, P.ev[i - 1], P.ev[i]);
            op_code_host[n] = P.code[i];
            op_level_host[n] = P.level[i];
            op_ms_host[n] = ms;
        }
    }
    for (auto e : P.ev) cudaEventDestroy(e);
    P.ev.clear(); P.code.clear(); P.level.clear();
    *n_ops_host = n;
    if (rc) return rc;
    if (ce != cudaSuccess) return set_err((int)ce, "vcycle_profile: stream synchronize failed");
    return 0;
}

extern "C" int b200amg_vcycle_residual_norm2(void *handle, const double *x, const double *b, double *out, void *stream)
{
    Engine *E = reinterpret_cast<Engine *>(handle);
    if (!E) return set_err(B200AMG_EINVAL, "vcycle_residual_norm2: null handle");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    LevelRt &L = E->lev[0];
    B200_TRY(apply_A(L, OP_RESIDUAL, 1, x, b, L.r, st));
    return dot(L.d.n, L.r, L.r, out, E->dot_work, st);
}

// layout the engine streams an operator in: 1 CSR as given, 2 sliced-ELL copy, 3 CSR + 8-bit value codes,
// 4 CSR row pointers + 8-bit (offset, value) pattern codes, 5 one row-pattern byte per row,
// 6 sliced-ELL copy with 8-bit value codes, 7 sliced-ELL copy with 16-bit column steps;
// 0 = block operator / no such level.  which = 3: the colour-grouped copy of A that colour-ordered Gauss-Seidel
// sweeps (2 or 7; 0 when the level has none); which = 4: the diagonal-free copy of A for the Jacobi sweeps of a big level
extern "C" int b200amg_vcycle_layout(void *handle, int level, int which)
{
    Engine *E = reinterpret_cast<Engine *>(handle);
    if (!E || level < 0 || level >= (int)E->lev.size() || which < 0 || which > 4) return 0;
    const LevelRt &L = E->lev[level];
    if (which == 3) return L.sGS.ok() ? (L.sGS.cd ? 7 : 2) : 0;
    if (which == 4) return L.sJ.ok() ? (L.sJ.cd ? 7 : 2) : 0;      // the diagonal-free copy the Jacobi sweeps of a big level stream
    if (which == 0 ? L.d.bs != 1 : (L.d.p_br != 1 || L.d.p_bc != 1)) return 0;
    const Sell &s = which == 0 ? L.sA : which == 1 ? L.sP : L.sR;
    const ValueDict &d = which == 0 ? L.dA : which == 1 ? L.dP : L.dR;
    if (which == 0 && L.rpA.ok()) return 5;
    if (s.ok()) return s.cc ? 6 : (s.cd ? 7 : 2);
    return d.ok() ? (d.offs ? 4 : 3) : 1;
}

extern "C" int b200amg_vcycle_destroy(void *handle)
{
    Engine *E = reinterpret_cast<Engine *>(handle);
    if (!E) return 0;
    for (auto &g : E->graphs) cudaGraphExecDestroy(g.exec);
    for (auto &L : E->lev) free_level(L);
    delete E;
    return 0;
}

extern "C" int b200amg_level_desc_size(void) { return (int)sizeof(b200amg_level_desc); }
