// TEST INFRASTRUCTURE ONLY: C entry points over the sliced-ELL layout of pyamg_b200/csrc/sell.cu when that file is compiled
// for the host (tests/test_emulated_kernels.py): build a copy (32-bit columns / 16-bit column steps with escape slots /
// without the diagonal entries, optionally colour-grouped), apply an operation or sweep the colours, free it.  The scan
// that the build uses lives in scan.cu in the product; here it is a serial host loop.
#include "internal.cuh"
#include "setup.cuh"

namespace b200amg {

template <typename TI, typename TO>
int exclusive_scan(long long n, const TI *in, TO *out, TO *, cudaStream_t)
{
    TO run = 0;
    for (long long i = 0; i < n; ++i) {
        const TO v = (TO)in[i];
        out[i] = run;
        run += v;
    }
    out[n] = run;
    return 0;
}
template int exclusive_scan<long long, long long>(long long, const long long *, long long *, long long *, cudaStream_t);
long long scan_scratch_elems(long long n) { return n + 1; }

}  // namespace b200amg

using namespace b200amg;

// op: the engine's op codes (0 y = A x, 2 y = b - A x, 3 y += A x, 4 Jacobi, 5 Jacobi with bsr(1x1) arithmetic).
// layout_out: 2 / 7 (16-bit steps) or 0 when the build declined; escapes_out: rows flagged as holding escape slots.
extern "C" int emu_sell_apply(int n, long long nnz, const int *Ap, const int *Aj, const double *Ax, int drop_diag, int op,
                              const double *x, const double *b, double *y, double omega, int *layout_out, int *escapes_out)
{
    Sell s;
    int rc = sell_build(n, nnz, Ap, Aj, Ax, &s, nullptr, nullptr, nullptr, nullptr, drop_diag != 0);
    if (rc) return rc;
    *layout_out = s.ok() ? (s.cd ? 7 : 2) : 0;
    *escapes_out = 0;
    if (!s.ok()) return 0;
    for (int i = 0; i < n; ++i) *escapes_out += s.row_len[i] < 0 ? 1 : 0;
    StreamArgs a{Ap, Aj, Ax, x, b, y, nnz, 0, n, 1.0, 0.0, omega};
    rc = sell_op(op, s, a, nullptr);
    sell_free(&s);
    return rc;
}

// colour-ordered sweeps over a colour-grouped copy: passes forward (dir >= 0) then/or backward like the engine does
extern "C" int emu_sell_gs(int n, long long nnz, const int *Ap, const int *Aj, const double *Ax, const int *Id, const int *cptr,
                           int K, int bsrf, int drop_diag, int sweep, double *x, const double *b, int *layout_out)
{
    Sell s;
    int rc = sell_build(n, nnz, Ap, Aj, Ax, &s, nullptr, Id, nullptr, nullptr, drop_diag != 0);
    if (rc) return rc;
    *layout_out = s.ok() ? (s.cd ? 7 : 2) : 0;
    if (!s.ok()) return 0;
    for (int pass = 0; pass < 2 && !rc; ++pass) {
        const bool bwd = pass == 1;
        if ((sweep == 0 && bwd) || (sweep == 1 && !bwd)) continue;
        for (int q = 0; q < K && !rc; ++q) {
            const int c = bwd ? K - 1 - q : q;
            rc = sell_gs_color(bsrf != 0, s, Id, cptr[c], cptr[c + 1], x, b, nullptr);
        }
    }
    sell_free(&s);
    return rc;
}
