"""Synthetic problem generators of the BASELINE configs, built directly in HBM
(pyamg/gallery/laplacian.py:14-69 `poisson`, via stencil.py:12-146 `stencil_grid`;
diffusion.py:19-138 `diffusion_stencil_2d`).  Output equals the reference generators' CSR
arrays (C-order grid numbering, ascending columns, boundary entries absent) -- checked
against them in tests/.  Index construction uses torch tensor ops (setup plumbing only)."""
import numpy as np
import torch

from .sparse import DevMat, _dev


def stencil_grid(S, grid, device=None):
    """CSR matrix of stencil S (odd-sized N-D array) on a regular grid, Dirichlet boundaries
    (stencil.py:12-146): zero stencil entries and out-of-grid neighbours produce no entry."""
    dev = device or _dev()
    S = np.asarray(S, dtype=np.float64)
    grid = tuple(int(g) for g in grid)
    if S.ndim != len(grid) or any(s % 2 != 1 for s in S.shape):
        raise ValueError('stencil / grid mismatch')
    N = int(np.prod(grid))
    strides = [int(np.prod(grid[d + 1:])) for d in range(len(grid))]
    centre = [s // 2 for s in S.shape]
    offs = []
    for idx in np.ndindex(*S.shape):
        if S[idx] != 0:
            rel = [i - c for i, c in zip(idx, centre)]
            offs.append((sum(r * st for r, st in zip(rel, strides)), rel, float(S[idx])))
    offs.sort(key=lambda t: t[0])                 # ascending column inside every row
    coords = []
    lin = torch.arange(N, device=dev, dtype=torch.int64)
    rem = lin
    for d, st in enumerate(strides):
        coords.append(torch.div(rem, st, rounding_mode='floor'))
        rem = rem - coords[-1] * st
    valid = torch.empty((N, len(offs)), dtype=torch.bool, device=dev)
    for q, (_, rel, _) in enumerate(offs):
        ok = torch.ones(N, dtype=torch.bool, device=dev)
        for d, r in enumerate(rel):
            if r:
                ok &= (coords[d] + r >= 0) & (coords[d] + r < grid[d])
        valid[:, q] = ok
    del coords, rem
    counts = valid.sum(dim=1, dtype=torch.int32)
    indptr = torch.zeros(N + 1, dtype=torch.int32, device=dev)
    indptr[1:] = torch.cumsum(counts, 0).to(torch.int32)
    off_t = torch.tensor([o[0] for o in offs], dtype=torch.int64, device=dev)
    val_t = torch.tensor([o[2] for o in offs], dtype=torch.float64, device=dev)
    mask = valid.reshape(-1)
    cols = (lin[:, None] + off_t[None, :]).reshape(-1)[mask].to(torch.int32)
    vals = val_t[None, :].expand(N, -1).reshape(-1)[mask].contiguous()
    return DevMat(indptr, cols.contiguous(), vals, (N, N), (1, 1), 'csr', True)


def poisson(grid, device=None):
    """laplacian.py:14-69: 2*D on the diagonal, -1 to the 2*D axis neighbours."""
    grid = tuple(int(g) for g in grid)
    D = len(grid)
    S = np.zeros((3,) * D)
    for d in range(D):
        for s in (0, 2):
            idx = [1] * D
            idx[d] = s
            S[tuple(idx)] = -1
    S[(1,) * D] = 2 * D
    return stencil_grid(S, grid, device)


def diffusion_stencil_2d(epsilon=1.0, theta=0.0, type='FE'):
    """diffusion.py:19-138: rotated anisotropic diffusion, Q1 finite elements ('FE') or finite
    differences ('FD'), 3x3 stencil."""
    eps, th = float(epsilon), float(theta)
    cc, ss, cs = np.cos(th) ** 2, np.sin(th) ** 2, np.cos(th) * np.sin(th)
    if type == 'FE':
        # Q1 element stencil of -(eps c^2+s^2)u_xx - 2(eps-1)cs u_xy - (c^2+eps s^2)u_yy, times 6
        corner_p = (-1 * eps - 1) * cc + (-1 * eps - 1) * ss + (3 * eps - 3) * cs      # NW / SE
        corner_m = (-1 * eps - 1) * cc + (-1 * eps - 1) * ss + (-3 * eps + 3) * cs     # NE / SW
        north = (2 * eps - 4) * cc + (-4 * eps + 2) * ss
        west = (-4 * eps + 2) * cc + (2 * eps - 4) * ss
        centre = (8 * eps + 8) * cc + (8 * eps + 8) * ss
        scale = 6.0
    elif type == 'FD':
        corner_p = 0.5 * (eps - 1) * cs
        corner_m = -corner_p
        north = -(eps * ss + cc)
        west = -(eps * cc + ss)
        centre = 2.0 * (eps + 1)
        scale = 1.0
    else:
        raise ValueError('type must be FE or FD')
    st = np.array([[corner_p, north, corner_m],
                   [west, centre, west],
                   [corner_m, north, corner_p]]) / scale
    return st


# ---- BASELINE config 4: P1 tetrahedral linear elasticity on a structured grid -------------------------------------
_TET_RING = ((1, 3), (3, 2), (2, 6), (6, 4), (4, 5), (5, 1))     # the 6 tetrahedra (c0, cp, cq, c7) around the diagonal c0-c7


def _p13d_local(verts, lame, mu):
    """local stiffness matrix of a P1 tetrahedron (pyamg/gallery/elasticity.py:340-357)"""
    A = np.vstack((np.ones((1, 4)), verts.T))
    grad = np.linalg.inv(A)[:, 1:]
    R = np.zeros((6, 12))
    R[[0, 3, 4], 0::3] = grad.T
    R[[3, 1, 5], 1::3] = grad.T
    R[[4, 5, 2], 2::3] = grad.T
    C = np.zeros((6, 6))
    C[0:3, 0:3] = lame + 2 * mu * np.eye(3)
    C[3:6, 3:6] = mu * np.eye(3)
    return np.linalg.det(A) / 6 * np.dot(np.dot(R.T, C), R)


def linear_elasticity_tets(nv, E=1e5, nu=0.3, device=None):
    """P1 linear elasticity (pyamg/gallery/elasticity.py:215-322, `linear_elasticity_p1`) on nv^3 vertices of the unit
    cube, every cube cut into 6 tetrahedra around its main diagonal, the face x = 0 clamped by leaving its vertices out:
    -> (A as BSR with 3 x 3 blocks, B = the 6 rigid-body modes as an (n, 6) tensor).

    All cubes are congruent, so the operator is a 15-point block stencil whose blocks are sums of the local stiffness
    blocks of the tetrahedra that exist around a vertex; it is assembled per stencil offset instead of per element
    (12.3 M elements at nv = 128).  Same pattern as the reference generator on the same mesh, values equal up to the
    rounding of the duplicate sums (checked against a reference-built fixture in tests/test_cabi.py)."""
    dev = device or _dev()
    nv = int(nv)
    if nv < 3:
        raise ValueError('need at least 3 vertices per side')
    lame = E * nu / ((1 + nu) * (1 - 2 * nu))
    mu = E / (2 + 2 * nu)
    h = 1.0 / (nv - 1)
    corner = [np.array([a, b, d]) for d in (0, 1) for b in (0, 1) for a in (0, 1)]        # c[a + 2b + 4d] -> (x, y, z)
    # contrib[d] = list of (offset of the row vertex inside its cube, 3x3 block) for the stencil offset d
    contrib = {}
    for p, q in _TET_RING:
        loc = [corner[0], corner[p], corner[q], corner[7]]
        K = _p13d_local(np.array(loc, dtype=np.float64) * h, lame, mu)
        for a in range(4):
            for b in range(4):
                d = tuple(int(v) for v in (loc[b] - loc[a]))
                contrib.setdefault(d, []).append((tuple(int(v) for v in loc[a]), K[3 * a:3 * a + 3, 3 * b:3 * b + 3]))
    nx, ny, nz = nv - 1, nv, nv                       # free vertices: original x index 1..nv-1
    N = nx * ny * nz
    stride = (1, nx, nx * ny)
    offs = sorted(contrib, key=lambda d: d[0] * stride[0] + d[1] * stride[1] + d[2] * stride[2])
    lin = torch.arange(N, device=dev, dtype=torch.int64)
    ck = torch.div(lin, nx * ny, rounding_mode='floor')
    cj = torch.div(lin - ck * nx * ny, nx, rounding_mode='floor')
    ci = lin - ck * nx * ny - cj * nx
    co = (ci + 1, cj, ck)                             # original vertex coordinates
    valid = torch.zeros((N, len(offs)), dtype=torch.bool, device=dev)
    data = torch.zeros((N, len(offs), 9), dtype=torch.float64, device=dev)
    for q, d in enumerate(offs):
        for oa, blk in contrib[d]:
            ok = torch.ones(N, dtype=torch.bool, device=dev)
            for dim in range(3):                      # the cube (vertex - oa) must exist: 0 <= c <= nv - 2
                c = co[dim] - oa[dim]
                ok &= (c >= 0) & (c <= nv - 2)
            ok &= (co[0] + d[0] >= 1)                 # neighbour not on the clamped face
            valid[:, q] |= ok
            data[:, q, :] += ok[:, None].to(torch.float64) * torch.from_numpy(np.ascontiguousarray(blk).ravel()).to(dev)[None, :]
    counts = valid.sum(dim=1, dtype=torch.int32)
    indptr = torch.zeros(N + 1, dtype=torch.int32, device=dev)
    indptr[1:] = torch.cumsum(counts, 0).to(torch.int32)
    off_t = torch.tensor([d[0] * stride[0] + d[1] * stride[1] + d[2] * stride[2] for d in offs], dtype=torch.int64, device=dev)
    mask = valid.reshape(-1)
    cols = (lin[:, None] + off_t[None, :]).reshape(-1)[mask].to(torch.int32).contiguous()
    vals = data.reshape(-1, 9)[mask].reshape(-1).contiguous()
    A = DevMat(indptr, cols, vals, (3 * N, 3 * N), (3, 3), 'bsr', True)
    x, y, z = (c.to(torch.float64) / float(nv - 1) for c in co)
    B = torch.zeros((N, 3, 6), dtype=torch.float64, device=dev)
    B[:, 0, 0] = 1
    B[:, 1, 1] = 1
    B[:, 2, 2] = 1
    B[:, 0, 3], B[:, 1, 3] = -y, x                    # rotations (-y, x, 0), (-z, 0, x), (0, -z, y)
    B[:, 0, 4], B[:, 2, 4] = -z, x
    B[:, 1, 5], B[:, 2, 5] = -z, y
    return A, B.reshape(3 * N, 6)
