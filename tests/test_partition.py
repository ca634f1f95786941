"""Index plumbing of the row-partitioned cycle (pyamg_b200/partition.py) on its own: slab cutting with column
renumbering, extra (halo-point) rows in the same numbering, halo plans and their block expansion.  Pure torch on the
CPU; the multi-process tests (test_distributed_cpu.py) run the same functions under gloo."""
import importlib

import numpy as np
import pytest
import scipy.sparse as sp
import torch

part = importlib.import_module('pyamg_b200.partition')


def t3(A):
    return (torch.from_numpy(A.indptr.astype(np.int32)), torch.from_numpy(A.indices.astype(np.int32)),
            torch.from_numpy(np.ravel(A.data).astype(np.float64)))


def back_to_global(ip, ix, dx, n_own_cols, col0, halo, shape, blockvals=1):
    """undo the [owned | halo] column numbering"""
    ix = ix.numpy().astype(np.int64)
    h = halo.numpy()
    cols = np.where(ix < n_own_cols, ix + col0, h[np.maximum(ix - n_own_cols, 0)] if h.size else 0)
    if blockvals == 1:
        return sp.csr_matrix((dx.numpy(), cols, ip.numpy()), shape=shape)
    r = int(round(np.sqrt(blockvals)))
    return sp.bsr_matrix((dx.numpy().reshape(-1, r, r), cols, ip.numpy()), shape=(shape[0] * r, shape[1] * r))


@pytest.mark.parametrize('seed', [0, 1])
def test_localize_keeps_rows_and_order(seed):
    rng = np.random.RandomState(seed)
    A = sp.random(40, 30, 0.25, random_state=rng, format='csr')
    A.sort_indices()
    ip, ix, dx = t3(A)
    lp, lx, ld, halo = part.localize(ip, ix, dx, 8, 19, 5, 14)
    assert torch.all(halo[1:] > halo[:-1]) and not torch.any((halo >= 5) & (halo < 14))
    B = back_to_global(lp, lx, ld, 9, 5, halo, (11, 30))
    assert (B != A[8:19]).nnz == 0
    assert np.array_equal(ld.numpy(), A.data[A.indptr[8]:A.indptr[19]])            # stored order untouched


def test_localize_with_halo_point_rows_shares_the_numbering():
    """the rows of P for a rank's halo points travel in the numbering of its slab rows (distributed.py: the halo of x is
    corrected locally after prolongation); their columns join the halo list"""
    rng = np.random.RandomState(2)
    A = sp.random(50, 24, 0.3, random_state=rng, format='csr')
    A.sort_indices()
    ip, ix, dx = t3(A)
    rows = torch.tensor([44, 3, 21, 20], dtype=torch.int64)
    lp, lx, ld, halo, (ep, ex, ed) = part.localize(ip, ix, dx, 10, 20, 6, 12, 1, rows)
    assert (back_to_global(lp, lx, ld, 6, 6, halo, (10, 24)) != A[10:20]).nnz == 0
    assert (back_to_global(ep, ex, ed, 6, 6, halo, (4, 24)) != A[[44, 3, 21, 20]]).nnz == 0
    alone = part.localize(ip, ix, dx, 10, 20, 6, 12)[3]
    assert set(alone.tolist()) <= set(halo.tolist())
    gp, gj, gx = part.gather_rows(ip, ix, dx, torch.tensor([], dtype=torch.int64))
    assert gp.tolist() == [0] and gj.numel() == 0 and gx.numel() == 0


def test_block_rows_carry_their_values():
    rng = np.random.RandomState(3)
    pat = sp.random(12, 12, 0.4, random_state=rng, format='csr')
    pat.sort_indices()
    data = rng.rand(pat.nnz, 2, 2)
    A = sp.bsr_matrix((data, pat.indices, pat.indptr), shape=(24, 24))
    ip, ix, dx = t3(A)
    lp, lx, ld, halo, (ep, ex, ed) = part.localize(ip, ix, dx, 4, 9, 4, 9, 4, torch.tensor([1, 10], dtype=torch.int64))
    B = back_to_global(lp, lx, ld, 5, 4, halo, (5, 12), 4)
    assert (sp.csr_matrix(B) != sp.csr_matrix(A)[8:18]).nnz == 0
    E = back_to_global(ep, ex, ed, 5, 4, halo, (2, 12), 4)
    assert (sp.csr_matrix(E) != sp.csr_matrix(A)[[2, 3, 20, 21]]).nnz == 0


def test_plans_and_block_expansion():
    split = [0, 4, 9, 12]
    halos = [torch.tensor([4, 5, 11]), torch.tensor([0, 3, 9]), torch.tensor([8])]
    plans = part.build_plans(halos, split)
    # rank 1 sends its local entries 0, 1 (global 4, 5) to rank 0 and local 4 (global 8) to rank 2
    assert plans[1].send_idx.tolist() == [0, 1, 4] and plans[1].send_counts == [2, 0, 1]
    assert plans[0].recv_counts == [0, 2, 1] and plans[0].n_halo == 3
    for p in plans:
        q = part.HaloPlan.unpack(p.pack())
        assert q.send_idx.tolist() == p.send_idx.tolist() and q.recv_counts == p.recv_counts
    e = plans[1].expanded(3)
    assert e.n_own == 15 and e.n_halo == 9 and e.send_counts == [6, 0, 3]
    assert e.send_idx.tolist() == [0, 1, 2, 3, 4, 5, 12, 13, 14]
    assert plans[1].expanded(1) is plans[1]
