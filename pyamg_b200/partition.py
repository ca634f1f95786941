"""Row-slab partition of a multigrid hierarchy over G ranks: index plumbing only.

Every partitioned level is cut into G contiguous row slabs (rank p owns rows
[split[p], split[p+1]) of A_l and P_l, and the rows of R_l of its coarse slab).  A local operator
keeps its rows exactly as stored (same entries, same order: the fp summation order of the reference
is untouched) with columns renumbered into the rank's local vector layout

        [ owned entries (global - split[p]) | halo entries, ascending global index ]

so the single-GPU kernels run unchanged on `[owned | halo]` vectors.  Because ownership ranges are
contiguous and the halo is sorted, the halo region is the concatenation, by owner rank, of what each
peer must send: one all-to-all (grouped NCCL send/recv over NVLink) with precomputed counts refreshes it.

Everything here is torch tensor arithmetic and therefore device-agnostic: rank 0 runs it on its GPU
for the real thing, the CPU tests run the same code on gloo.  No floating-point arithmetic happens here.
"""
import torch


def split_points(n, G):
    """contiguous, near-equal slabs: split[p] = floor(n p / G)"""
    return [(n * p) // G for p in range(G + 1)]


def owner_counts(sorted_cols, split):
    """how many entries of an ascending global index list fall into each owner's range"""
    bounds = torch.tensor(split, dtype=sorted_cols.dtype, device=sorted_cols.device)
    pos = torch.searchsorted(sorted_cols, bounds)
    return (pos[1:] - pos[:-1]).tolist()


def gather_rows(indptr, indices, data, rows, blockvals=1):
    """the rows `rows` (int64 tensor of row numbers, any order) of a CSR/BSR matrix, entries in stored order
    -> (indptr int64 starting at 0, GLOBAL column indices int64, data)"""
    rows = rows.to(torch.int64)
    ptr = indptr.to(torch.int64)
    starts, lens = ptr[rows], ptr[rows + 1] - ptr[rows]
    ip = torch.zeros(rows.numel() + 1, dtype=torch.int64, device=indptr.device)
    if rows.numel():
        ip[1:] = torch.cumsum(lens, 0)
    total = int(ip[-1])
    pos = torch.arange(total, dtype=torch.int64, device=indptr.device)
    src = pos - torch.repeat_interleave(ip[:-1], lens) + torch.repeat_interleave(starts, lens)
    vals = data.reshape(-1, blockvals)[src].reshape(-1) if total else data[:0]
    return ip, indices.to(torch.int64)[src], vals.contiguous().clone()


def localize(indptr, indices, data, row0, row1, col0, col1, blockvals=1, extra_rows=None):
    """Rows [row0, row1) of a CSR/BSR matrix with columns renumbered for the rank owning the
    column range [col0, col1).  -> (indptr int32, indices int32, data, halo_cols int64 ascending).
    For a BSR operator rows / columns count BLOCKS and every stored block carries `blockvals` = R * C values.
    extra_rows (int64 tensor of row numbers outside the slab): ALSO return those rows as a second operator
    (indptr, indices, data) in the SAME column numbering -- its columns join the halo list."""
    s, e = int(indptr[row0]), int(indptr[row1])
    ip = (indptr[row0:row1 + 1] - indptr[row0]).to(torch.int32).contiguous()
    cols = indices[s:e].to(torch.int64)
    own = (cols >= col0) & (cols < col1)
    far = cols[~own]
    if extra_rows is not None:
        ip2, cols2, dx2 = gather_rows(indptr, indices, data, extra_rows, blockvals)
        own2 = (cols2 >= col0) & (cols2 < col1)
        far = torch.cat([far, cols2[~own2]])
    halo_cols = torch.unique(far)                              # sorted ascending

    def renumber(c, o):
        loc = torch.empty_like(c)
        loc[o] = c[o] - col0
        if halo_cols.numel():
            loc[~o] = (col1 - col0) + torch.searchsorted(halo_cols, c[~o])
        return loc.to(torch.int32).contiguous()
    out = (ip, renumber(cols, own), data[s * blockvals:e * blockvals].contiguous().clone(), halo_cols)
    if extra_rows is not None:
        out = out + ((ip2.to(torch.int32).contiguous(), renumber(cols2, own2), dx2),)
    return out


class HaloPlan:
    """what one rank sends / receives to refresh the halo of one vector space"""

    def __init__(self, n_own, n_halo, send_idx, send_counts, recv_counts):
        self.n_own, self.n_halo = int(n_own), int(n_halo)
        self.send_idx = send_idx                  # int64 local indices into the owned part, grouped by peer
        self.send_counts = [int(c) for c in send_counts]
        self.recv_counts = [int(c) for c in recv_counts]

    def to(self, device):
        self.send_idx = self.send_idx.to(device)
        return self

    def expanded(self, bs):
        """the plan of the SCALAR vector space when this plan counts blocks of bs unknowns: block q is the scalars
        q * bs .. q * bs + bs - 1, in the owned part and in the halo alike"""
        if bs == 1:
            return self
        k = torch.arange(bs, dtype=torch.int64, device=self.send_idx.device)
        idx = (self.send_idx.to(torch.int64)[:, None] * bs + k[None, :]).reshape(-1)
        return HaloPlan(self.n_own * bs, self.n_halo * bs, idx, [c * bs for c in self.send_counts],
                        [c * bs for c in self.recv_counts])

    def pack(self):
        return {'n_own': self.n_own, 'n_halo': self.n_halo, 'send_idx': self.send_idx.cpu(),
                'send_counts': self.send_counts, 'recv_counts': self.recv_counts}

    @classmethod
    def unpack(cls, d):
        return cls(d['n_own'], d['n_halo'], d['send_idx'], d['send_counts'], d['recv_counts'])


def build_plans(halo_lists, split):
    """halo_lists[p] = ascending global indices rank p needs (none inside its own range).
    -> [HaloPlan for every rank]"""
    G = len(split) - 1
    recv = [owner_counts(h, split) for h in halo_lists]                     # recv[p][q]: p gets from q
    plans = []
    for q in range(G):
        parts, counts = [], []
        for p in range(G):
            h = halo_lists[p]
            lo = sum(recv[p][:q])
            seg = h[lo:lo + recv[p][q]] - split[q]                         # q's local indices p wants
            parts.append(seg)
            counts.append(recv[p][q])
        send_idx = torch.cat(parts) if parts else torch.empty(0, dtype=torch.int64)
        plans.append(HaloPlan(split[q + 1] - split[q], halo_lists[q].numel(), send_idx.to(torch.int64),
                              counts, recv[q]))
    return plans


def partition_operator(indptr, indices, data, row_split, col_split):
    """-> per rank (ip, ix, dx) and the HaloPlans of the operator's COLUMN space"""
    G = len(row_split) - 1
    locs, halos = [], []
    for p in range(G):
        ip, ix, dx, h = localize(indptr, indices, data, row_split[p], row_split[p + 1], col_split[p], col_split[p + 1])
        locs.append((ip, ix, dx))
        halos.append(h)
    return locs, build_plans(halos, col_split)
