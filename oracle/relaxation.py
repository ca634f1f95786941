"""TEST INFRASTRUCTURE ONLY -- restatement of the reference's smoother front-ends
(pyamg/relaxation/relaxation.py: make_system :21-105, gauss_seidel :280-354, jacobi :357-427,
gauss_seidel_indexed :671-741) over ``oracle.core`` (reference .so if built, else the C port),
plus the multicolour composition of SURVEY.md 3.5 (pyamg/graph.py:84-125 + gauss_seidel_indexed).
"""
import numpy as np
import scipy.sparse as sparse

import oracle as _o


def make_system(A, x, b, formats=None):
    """relaxation.py:21-105 -- validate / normalise (A, x, b)."""
    if formats is None:
        pass
    elif formats == ['csr']:
        if sparse.issparse(A) and A.format == 'csr':
            pass
        elif sparse.issparse(A) and A.format == 'bsr':
            A = A.tocsr()
        else:
            A = sparse.csr_matrix(A)
    elif not (sparse.issparse(A) and A.format in formats):
        A = sparse.csr_matrix(A).asformat(formats[0])
    if not isinstance(x, np.ndarray):
        raise ValueError('expected numpy array for argument x')
    if not isinstance(b, np.ndarray):
        raise ValueError('expected numpy array for argument b')
    M, N = A.shape
    if M != N:
        raise ValueError('expected square matrix')
    if x.shape not in [(M,), (M, 1)]:
        raise ValueError('x has invalid dimensions')
    if b.shape not in [(M,), (M, 1)]:
        raise ValueError('b has invalid dimensions')
    if A.dtype != x.dtype or A.dtype != b.dtype:
        raise TypeError('arguments A, x, and b must have the same dtype')
    if not x.flags.carray:
        raise ValueError('x must be contiguous in memory')
    return A, np.ravel(x), np.ravel(b)


def _sweep_range(sweep, n):
    if sweep == 'forward':
        return 0, n, 1
    if sweep == 'backward':
        return n - 1, -1, -1
    raise ValueError("valid sweep directions are 'forward', 'backward', and 'symmetric'")


def gauss_seidel(A, x, b, iterations=1, sweep='forward', core=None):
    core = core or _o.core
    A, x, b = make_system(A, x, b, formats=['csr', 'bsr'])
    if A.format == 'csr':
        bs = 1
    else:
        R, C = A.blocksize
        if R != C:
            raise ValueError('BSR blocks must be square')
        bs = R
    if sweep == 'symmetric':
        for _ in range(iterations):
            gauss_seidel(A, x, b, 1, 'forward', core)
            gauss_seidel(A, x, b, 1, 'backward', core)
        return
    r0, r1, dr = _sweep_range(sweep, len(x) // bs)
    for _ in range(iterations):
        if A.format == 'csr':
            core.gauss_seidel(A.indptr, A.indices, A.data, x, b, r0, r1, dr)
        else:
            core.bsr_gauss_seidel(A.indptr, A.indices, np.ravel(A.data), x, b, r0, r1, dr, bs)


def jacobi(A, x, b, iterations=1, omega=1.0, core=None):
    core = core or _o.core
    A, x, b = make_system(A, x, b, formats=['csr', 'bsr'])
    n = A.shape[0]
    if n <= 0:
        return
    temp = np.empty_like(x)
    om = np.array([omega], dtype=A.dtype)
    for _ in range(iterations):
        if A.format == 'csr':
            core.jacobi(A.indptr, A.indices, A.data, x, b, temp, 0, n, 1, om)
        else:
            R, C = A.blocksize
            if R != C:
                raise ValueError('BSR blocks must be square')
            core.bsr_jacobi(A.indptr, A.indices, np.ravel(A.data), x, b, temp,
                            0, n // R, 1, R, om)


def gauss_seidel_indexed(A, x, b, indices, iterations=1, sweep='forward', core=None):
    core = core or _o.core
    A, x, b = make_system(A, x, b, formats=['csr'])
    indices = np.asarray(indices, dtype='intc')
    if sweep == 'symmetric':
        for _ in range(iterations):
            gauss_seidel_indexed(A, x, b, indices, 1, 'forward', core)
            gauss_seidel_indexed(A, x, b, indices, 1, 'backward', core)
        return
    r0, r1, dr = _sweep_range(sweep, len(indices))
    for _ in range(iterations):
        core.gauss_seidel_indexed(A.indptr, A.indices, A.data, x, b, indices, r0, r1, dr)


def bsr_gauss_seidel_indexed(A, x, b, indices, iterations=1, sweep='forward', core=None):
    """Block rows of a BSR matrix relaxed in the order of `indices` (block-row numbers), each through the reference's
    bsr_gauss_seidel restricted to that one row -- the sequential definition of the block-multicolour sweep
    (tests/golden/make_golden.py::sa_elasticity).  A core without the fused loop (the reference's own _amg_core) is
    driven one row per call."""
    core = core or _o.core
    A, x, b = make_system(A, x, b, formats=['bsr'])
    bs = A.blocksize[0]
    if A.blocksize[0] != A.blocksize[1]:
        raise ValueError('BSR blocks must be square')
    indices = np.asarray(indices, dtype='intc')
    data = np.ravel(A.data)
    fused = getattr(core, 'bsr_gauss_seidel_indexed', None)
    for _ in range(iterations):
        for direction in (('forward', 'backward') if sweep == 'symmetric' else (sweep,)):
            r0, r1, dr = _sweep_range(direction, len(indices))
            if fused is not None:
                fused(A.indptr, A.indices, data, x, b, indices, r0, r1, dr, bs)
            else:
                for p in range(r0, r1, dr):
                    i = int(indices[p])
                    core.bsr_gauss_seidel(A.indptr, A.indices, data, x, b, i, i + dr, dr, bs)


def _block_system(A, x, b, blocksize, Dinv):
    A, x, b = make_system(A, x, b, formats=['csr', 'bsr'])
    A = A.tobsr(blocksize=(blocksize, blocksize))
    if Dinv.shape[0] != A.shape[0] // blocksize:
        raise ValueError('Dinv and A have incompatible dimensions')
    if Dinv.shape[1] != blocksize or Dinv.shape[2] != blocksize:
        raise ValueError('Dinv and blocksize are incompatible')
    return A, x, b, np.ravel(np.ascontiguousarray(Dinv, dtype=np.float64))


def block_jacobi(A, x, b, Dinv, blocksize, iterations=1, omega=1.0, core=None):
    """pyamg/relaxation/relaxation.py:430-492 with the inverses of the diagonal blocks GIVEN (the reference computes
    them with LAPACK gelss, utils.py:601-690 -> util/linalg.py:568-649)"""
    core = core or _o.core
    A, x, b, Tx = _block_system(A, x, b, blocksize, Dinv)
    nb = A.shape[0] // blocksize
    temp = np.empty_like(x)
    om = np.array([omega], dtype=A.dtype)
    for _ in range(iterations):
        core.block_jacobi(A.indptr, A.indices, np.ravel(A.data), x, b, Tx, temp, 0, nb, 1, om, blocksize)


def block_gauss_seidel(A, x, b, Dinv, blocksize, iterations=1, sweep='forward', indices=None, core=None):
    """pyamg/relaxation/relaxation.py:495-592 (Dinv given).  `indices`: block rows relaxed in that order instead, one
    reference call per block row -- the sequential definition of the block-multicolour sweep."""
    core = core or _o.core
    A, x, b, Tx = _block_system(A, x, b, blocksize, Dinv)
    data = np.ravel(A.data)
    nb = A.shape[0] // blocksize
    if indices is not None:
        indices = np.asarray(indices, dtype='intc')
    fused = getattr(core, 'block_gauss_seidel_indexed', None)
    for _ in range(iterations):
        for direction in (('forward', 'backward') if sweep == 'symmetric' else (sweep,)):
            r0, r1, dr = _sweep_range(direction, nb if indices is None else len(indices))
            if indices is None:
                core.block_gauss_seidel(A.indptr, A.indices, data, x, b, Tx, r0, r1, dr, blocksize)
            elif fused is not None:
                fused(A.indptr, A.indices, data, x, b, Tx, indices, r0, r1, dr, blocksize)
            else:
                for p in range(r0, r1, dr):
                    i = int(indices[p])
                    core.block_gauss_seidel(A.indptr, A.indices, data, x, b, Tx, i, i + dr, dr, blocksize)


def polynomial(A, x, b, coefficients, iterations=1, core=None):
    """pyamg/relaxation/relaxation.py:593-668 (what 'chebyshev' resolves to, smoothing.py:438-449): numpy / scipy
    expressions exactly as the reference writes them."""
    A, x, b = make_system(A, x, b, formats=None)
    for _ in range(iterations):
        if np.sqrt(np.inner(x.conj(), x).real) == 0:
            residual = b
        else:
            residual = (b - A * x)
        h = coefficients[0] * residual
        for c in coefficients[1:]:
            h = c * residual + A * h
        x += h


def vertex_coloring_mis(G, core=None):
    """pyamg/graph.py:84-125 with method='MIS' (-> graph.h:171-188)."""
    core = core or _o.core
    G = sparse.csr_matrix(G)
    colors = np.empty(G.shape[0], dtype='intc')
    core.vertex_coloring_mis(G.shape[0], np.ascontiguousarray(G.indptr, 'intc'),
                             np.ascontiguousarray(G.indices, 'intc'), colors)
    return colors


def color_ordering(colors):
    """Rows grouped by colour, ascending inside a colour.
    Returns (indices int32[n], color_ptr int32[K+1])."""
    K = int(colors.max()) + 1 if len(colors) else 0
    idx = np.argsort(colors, kind='stable').astype('intc')
    ptr = np.zeros(K + 1, dtype='intc')
    np.cumsum(np.bincount(colors, minlength=K), out=ptr[1:])
    return idx, ptr
