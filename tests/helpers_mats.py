import numpy as np
import scipy.sparse as sp


def lap1d(n):
    return sp.spdiags([-np.ones(n), 2 * np.ones(n), -np.ones(n)], [-1, 0, 1], n, n, format='csr')


def poisson3d(m):
    I = sp.eye(m); T = lap1d(m)
    return sp.csr_matrix(sp.kron(sp.kron(T, I), I) + sp.kron(sp.kron(I, T), I) + sp.kron(sp.kron(I, I), T))
