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
