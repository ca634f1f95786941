"""Import a hierarchy built elsewhere (e.g. by the reference PyAMG on the host: `ml.levels`
with scipy A/P/R) into a device multilevel_solver.  Formats are kept (CSR vs BSR decides which
amg_core arithmetic the smoothers use, SURVEY.md 2 'Observed formats'), column order inside rows
is kept (it is the fp summation order of the reference)."""
from .multilevel import multilevel_solver
from .smoothing import change_smoothers
from .sparse import DevMat


def import_hierarchy(levels, presmoother=None, postsmoother=None, coarse_solver='pinv2'):
    """levels: objects with .A and (all but the last) .P, .R as scipy csr/bsr matrices or DevMat;
    optional per-level .pre / .post resolved smoother specs (used when presmoother is None)."""
    out = []
    for src in levels:
        lvl = multilevel_solver.level()
        for nm in ('A', 'P', 'R'):
            M = getattr(src, nm, None)
            if M is not None:
                setattr(lvl, nm, M if isinstance(M, DevMat) else DevMat.from_scipy(M))
        out.append(lvl)
    ml = multilevel_solver(out, coarse_solver=coarse_solver)
    if presmoother is None:
        presmoother = [_norho(getattr(s, 'pre', None)) for s in levels[:-1]]
        postsmoother = [_norho(getattr(s, 'post', None)) for s in levels[:-1]]
    change_smoothers(ml, presmoother, postsmoother)
    return ml


def _norho(spec):
    """imported specs carry RESOLVED parameters (omega already divided by rho)"""
    if isinstance(spec, tuple) and spec[0] in ('jacobi', 'block_jacobi'):
        kw = dict(spec[1])
        kw.setdefault('withrho', False)
        return (spec[0], kw)
    return spec
