// Ruge-Stuben (classical AMG) setup pieces on the device.  Row-parallel restatements of
//   amg_core.classical_strength_of_connection    pyamg/amg_core/ruge_stuben.h:45-93
//   amg_core.rs_direct_interpolation_pass1/2     pyamg/amg_core/ruge_stuben.h:496-595
// with the fp64 operation order of the serial reference (sums in storage order, separately rounded),
// so that P -- and through the Galerkin product every coarse operator -- equals the reference's bit
// for bit.  The C/F splitting itself is sequential and runs on the host (host_graph.cu).
#include "internal.cuh"

namespace b200amg {

static inline cudaStream_t S(void *s) { return reinterpret_cast<cudaStream_t>(s); }

// Strong entries of row i: the stored diagonal, and off-diagonals with |a_ij| >= theta * max_{k != i} |a_ik|
// (the maximum starts at DBL_MIN like the reference's numeric_limits<F>::min()).  One kernel serves the two
// uses the reference's glue makes of S:
//   * the graph the C/F splitting runs on = S without its diagonal (pyamg/util/utils.py:1815-1827):
//     RS_DROP_DIAG;
//   * C = pattern(S) .* A for interpolation (pyamg/classical/interpolate.py:55-57): scipy's csr_elmul_csr
//     does not store exact zeros (RS_DROP_ZERO) and, when A's rows are not canonical (coarse levels: rows
//     come unsorted out of csr_matmat), takes csr_binop_csr_general, whose rows come out in REVERSE
//     storage order (RS_REVERSE).
// pass 0: cnt[i]; pass 1: entries at Cp[i].
enum : int { RS_DROP_DIAG = 1, RS_DROP_ZERO = 2, RS_REVERSE = 4 };

__global__ void __launch_bounds__(256) rs_strength_kernel(int n, double theta, const int *__restrict__ Ap,
                                                          const int *__restrict__ Aj, const double *__restrict__ Ax,
                                                          int flags, int pass, int *__restrict__ cnt,
                                                          const int *__restrict__ Cp, int *__restrict__ Cj,
                                                          double *__restrict__ Cx)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const int lo = Ap[i], hi = Ap[i + 1];
    double big = 2.2250738585072014e-308;
    for (int p = lo; p < hi; ++p)
        if (Aj[p] != i) big = fmax(big, fabs(Ax[p]));
    const double lim = mul_rn(theta, big);
    int out = 0;
    for (int p = lo; p < hi; ++p) {
        const double a = Ax[p];
        const bool diag = Aj[p] == i;
        bool keep = diag ? !(flags & RS_DROP_DIAG) : fabs(a) >= lim;
        if ((flags & RS_DROP_ZERO) && a == 0.0) keep = false;
        out += keep;
    }
    if (pass == 0) { cnt[i] = out; return; }
    const int base = Cp[i];
    int k = 0;
    for (int p = lo; p < hi; ++p) {
        const double a = Ax[p];
        const bool diag = Aj[p] == i;
        bool keep = diag ? !(flags & RS_DROP_DIAG) : fabs(a) >= lim;
        if ((flags & RS_DROP_ZERO) && a == 0.0) keep = false;
        if (!keep) continue;
        const int at = base + ((flags & RS_REVERSE) ? out - 1 - k : k);
        Cj[at] = Aj[p];
        Cx[at] = a;
        ++k;
    }
}

// flag[0] = 1 when some row is not strictly increasing in column (scipy csr_has_canonical_format is false)
__global__ void __launch_bounds__(256) csr_not_canonical_kernel(int n, const int *__restrict__ Ap,
                                                                const int *__restrict__ Aj, int *flag)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const int lo = Ap[i], hi = Ap[i + 1];
    bool bad = lo > hi;
    for (int p = lo + 1; p < hi; ++p) bad |= Aj[p - 1] >= Aj[p];
    if (bad) flag[0] = 1;
}

// entries of row i of P: 1 for a C point, one per strong C neighbour (other than i) for an F point
__global__ void __launch_bounds__(256) rs_interp_count_kernel(int n, const int *__restrict__ Sp, const int *__restrict__ Sj,
                                                              const int *__restrict__ split, int *__restrict__ cnt)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    if (split[i] == 1) { cnt[i] = 1; return; }
    int c = 0;
    for (int p = Sp[i]; p < Sp[i + 1]; ++p) c += (split[Sj[p]] == 1 && Sj[p] != i);
    cnt[i] = c;
}

// direct interpolation weights, thread = row; cmap[i] = number of C points before i (coarse index of a C point)
__global__ void __launch_bounds__(256) rs_interp_fill_kernel(int n, const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                             const double *__restrict__ Ax, const int *__restrict__ Sp,
                                                             const int *__restrict__ Sj, const double *__restrict__ Sx,
                                                             const int *__restrict__ split, const int *__restrict__ cmap,
                                                             const int *__restrict__ Pp, int *__restrict__ Pj,
                                                             double *__restrict__ Px)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    int at = Pp[i];
    if (split[i] == 1) { Pj[at] = cmap[i]; Px[at] = 1.0; return; }
    double sneg = 0.0, spos = 0.0, aneg = 0.0, apos = 0.0, d = 0.0;
    for (int p = Sp[i]; p < Sp[i + 1]; ++p) {
        const int j = Sj[p];
        if (split[j] != 1 || j == i) continue;
        const double s = Sx[p];
        if (s < 0) sneg = add_rn(sneg, s); else spos = add_rn(spos, s);
    }
    for (int p = Ap[i]; p < Ap[i + 1]; ++p) {
        const double a = Ax[p];
        if (Aj[p] == i) d = add_rn(d, a);
        else if (a < 0) aneg = add_rn(aneg, a);
        else apos = add_rn(apos, a);
    }
    const double alpha = div_rn(aneg, sneg);
    double beta = div_rn(apos, spos);
    if (spos == 0) { d = add_rn(d, apos); beta = 0.0; }
    const double cn = div_rn(-alpha, d), cp = div_rn(-beta, d);
    for (int p = Sp[i]; p < Sp[i + 1]; ++p) {
        const int j = Sj[p];
        if (split[j] != 1 || j == i) continue;
        const double s = Sx[p];
        Pj[at] = cmap[j];
        Px[at] = mul_rn(s < 0 ? cn : cp, s);
        ++at;
    }
}

}  // namespace b200amg

using namespace b200amg;

extern "C" int b200amg_rs_strength(int n_row, double theta, const int *Ap, const int *Aj, const double *Ax, int flags,
                                   int pass, int *cnt, const int *Cp, int *Cj, double *Cx, void *stream)
{
    if (n_row <= 0) return 0;
    if (theta < 0.0 || theta > 1.0) return set_err(B200AMG_EINVAL, "rs_strength: expected theta in [0,1]");
    B200_LAUNCH(rs_strength_kernel, (n_row + 255) / 256, 256, 0, S(stream), n_row, theta, Ap, Aj, Ax, flags, pass, cnt, Cp, Cj, Cx);
    return check_launch("rs_strength_kernel");
}

extern "C" int b200amg_csr_not_canonical(int n_row, const int *Ap, const int *Aj, int *flag, void *stream)
{
    B200_CUDA(cudaMemsetAsync(flag, 0, sizeof(int), S(stream)));
    if (n_row <= 0) return 0;
    B200_LAUNCH(csr_not_canonical_kernel, (n_row + 255) / 256, 256, 0, S(stream), n_row, Ap, Aj, flag);
    return check_launch("csr_not_canonical_kernel");
}

extern "C" int b200amg_rs_direct_interpolation_pass1(int n_nodes, const int *Sp, const int *Sj, const int *splitting,
                                                     int *cnt, void *stream)
{
    if (n_nodes <= 0) return 0;
    B200_LAUNCH(rs_interp_count_kernel, (n_nodes + 255) / 256, 256, 0, S(stream), n_nodes, Sp, Sj, splitting, cnt);
    return check_launch("rs_interp_count_kernel");
}

extern "C" int b200amg_rs_direct_interpolation_pass2(int n_nodes, const int *Ap, const int *Aj, const double *Ax,
                                                     const int *Sp, const int *Sj, const double *Sx, const int *splitting,
                                                     const int *cmap, const int *Pp, int *Pj, double *Px, void *stream)
{
    if (n_nodes <= 0) return 0;
    B200_LAUNCH(rs_interp_fill_kernel, (n_nodes + 255) / 256, 256, 0, S(stream), n_nodes, Ap, Aj, Ax, Sp, Sj, Sx, splitting, cmap,
                                                                        Pp, Pj, Px);
    return check_launch("rs_interp_fill_kernel");
}
