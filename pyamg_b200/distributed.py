"""Row-partitioned V-cycle over G GPUs of one node (one process per GPU, torch.distributed).

* Levels with at least `agglomerate_below` rows are cut into contiguous row slabs (partition.py);
  every rank applies its LOCAL operators (same rows, same storage order -> same bits as the
  single-GPU kernels) to `[owned | halo]` vectors; the halo is refreshed by ONE all-to-all
  (grouped NCCL send/recv over NVLink) before each operator application:
      x before the pre-smoothing sweep, x before the residual, r before the restriction,
      x_c before the prolongation, x before the post-smoothing sweep.
* Levels below the threshold are agglomerated on rank 0: b_c is gathered there, the lower part of the
  cycle runs in the ordinary single-GPU engine (multilevel_solver over levels[ka:], one CUDA-graph
  replay), x_c is scattered back.
* Norms: local dot + all-reduce.
Jacobi parity across ranks is exact (row arithmetic is rank-independent); norms differ in summation
order only.  The hierarchy is built on rank 0 (setup across GPUs is out of the first milestone)
and its slabs are shipped to their owners.

`backend` supplies the per-rank compute (default: libb200amg through the C-ABI; there is no CPU
fallback -- the CPU/gloo tests inject a checker backend from tests/).
"""
import ctypes

import numpy as np
import torch
import torch.distributed as dist

from . import partition as part

OP_MATVEC, OP_RESIDUAL, OP_ADD, OP_JACOBI, OP_JACOBI_BSR1 = 0, 2, 3, 4, 5


class DeviceBackend:
    """per-rank compute through libb200amg (operator handles: pipeline / sliced-ELL / stream kernels)"""

    def __init__(self):
        from . import _lib
        from .sparse import _dev
        self._lib, self.device = _lib, _dev()

    def stream(self):
        return torch.cuda.current_stream().cuda_stream

    class Op:
        """n_row rows (BLOCK rows of R x C blocks when R * C > 1) of a local operator"""

        def __init__(self, be, ip, ix, dx, n_row, R=1, C=1):
            self.be, self.ip, self.ix, self.dx, self.n_row, self.R, self.C = be, ip, ix, dx, int(n_row), int(R), int(C)
            self.h = ctypes.c_void_p()
            if self.R * self.C > 1:
                be._lib.call('b200amg_bsr_op_create', self.n_row, self.R, self.C, ip.data_ptr(), ix.data_ptr(), dx.data_ptr(),
                             ctypes.byref(self.h), be.stream())
            else:
                be._lib.call('b200amg_op_create', self.n_row, ip.data_ptr(), ix.data_ptr(), dx.data_ptr(), 0,
                             ctypes.byref(self.h), be.stream())

        def apply(self, op, x, b, y, omega=0.0):
            self.be._lib.call('b200amg_op_apply', self.h, op, x.data_ptr(), 0 if b is None else b.data_ptr(),
                              y.data_ptr(), float(omega), self.be.stream())

        def layout(self):
            """b200amg_op_layout code: 1 CSR, 2 sliced-ELL, 3 value codes, 4 pattern codes"""
            return int(self.be._lib.load().b200amg_op_layout(self.h))

        def diagonal(self):
            d = torch.empty(self.n_row * self.R, dtype=torch.float64, device=self.dx.device)
            self.be._lib.call('b200amg_extract_diag', self.n_row, self.R, self.ip.data_ptr(), self.ix.data_ptr(),
                              self.dx.data_ptr(), d.data_ptr(), self.be.stream())
            return d

        def __del__(self):
            try:
                self.be._lib.load().b200amg_op_destroy(self.h)
            except Exception:
                pass

    def make_op(self, ip, ix, dx, n_row, R=1, C=1):
        return DeviceBackend.Op(self, ip, ix, dx, n_row, R, C)

    def scale_diag(self, omega, b, diag, x, flavor):
        self._lib.call('b200amg_scale_diag', diag.numel(), float(omega), b.data_ptr(), diag.data_ptr(), x.data_ptr(),
                       int(flavor), self.stream())

    def gs_color(self, op, x, b, Id, p0, p1, flavor, backward=False):
        """rows Id[p0:p1] (one colour of the GLOBAL colouring, local numbering) relaxed in parallel, x in place:
        amg_core.gauss_seidel_indexed / bsr_gauss_seidel arithmetic (relaxation.h:394-426, :89-173).  `backward` matters
        for block rows only: the point sweep inside the diagonal block runs in the direction of the pass (:150-163)"""
        if p1 <= p0:
            return
        cp = (ctypes.c_int * 2)(int(p0), int(p1))
        if flavor or op.R > 1:                  # block rows: amg_core.bsr_gauss_seidel over the listed BLOCK rows
            r0, r1, dr = (int(p1) - 1, int(p0) - 1, -1) if backward else (int(p0), int(p1), 1)
            self._lib.call('b200amg_bsr_gauss_seidel', op.ip.data_ptr(), op.ix.data_ptr(), op.dx.data_ptr(), x.data_ptr(),
                           b.data_ptr(), r0, r1, dr, op.R, Id.data_ptr(), cp, 1, self.stream())
        else:
            self._lib.call('b200amg_gauss_seidel_indexed', op.ip.data_ptr(), op.ix.data_ptr(), op.dx.data_ptr(), x.data_ptr(),
                           b.data_ptr(), Id.data_ptr(), int(p0), int(p1), 1, cp, 1, self.stream())

    def dot(self, x, y, n):
        if not hasattr(self, '_dw'):
            self._dw = torch.empty(1024, dtype=torch.float64, device=self.device)
        out = torch.empty(1, dtype=torch.float64, device=self.device)
        self._lib.call('b200amg_dot', int(n), x.data_ptr(), y.data_ptr(), out.data_ptr(), self._dw.data_ptr(), self.stream())
        return out


class _Level:
    pass


def _send_tensors(tensors, dst, group):
    for t in tensors:
        dist.send(t.contiguous(), dst, group=group)


def _recv_tensors(specs, src, device, group):
    out = []
    for n, dt in specs:
        t = torch.empty(n, dtype=dt, device=device)
        dist.recv(t, src, group=group)
        out.append(t)
    return out


def describe(ml):
    """rank-0 description of a device multilevel_solver for DistributedSolver: per non-coarsest level
    the raw CSR arrays of A, P, R (torch tensors), sizes, arithmetic flavor and resolved smoothers"""
    out = []
    for L in ml.levels[:-1]:
        if L.A.blocksize[0] != L.A.blocksize[1] or L.P.blocksize[0] != L.A.blocksize[0] or L.R.blocksize[1] != L.A.blocksize[0] \
                or L.R.blocksize[0] != L.P.blocksize[1]:
            raise NotImplementedError('distributed cycle: block shapes of A, P, R do not chain')
        pre, post = L.presmoother, L.postsmoother
        for sm in (pre, post):
            if sm.kind not in (0, 1, 2):
                raise NotImplementedError('distributed cycle: Jacobi, colour-ordered Gauss-Seidel or None smoothers only')
        d = {'n': L.A.shape[0], 'nc': L.P.shape[1], 'flavor': 0 if L.A.format == 'csr' else 1,
             'bs': int(L.A.blocksize[0]), 'bsc': int(L.P.blocksize[1]),     # unknowns per block row: this level / the next
             'pre': (pre.kind, pre.omega, pre.iterations, pre.sweep), 'post': (post.kind, post.omega, post.iterations, post.sweep)}
        ordered = [sm for sm in (pre, post) if sm.kind == 2]
        if ordered:                                     # one GLOBAL colour ordering per level (multilevel._build_engine checks it)
            d['Id'] = ordered[0].indices
            d['color_ptr'] = [int(v) for v in ordered[0].color_ptr]
            d['gs_csr'] = int(ordered[0].name in ('gauss_seidel_indexed', 'multicolor_gauss_seidel'))
        for nm in ('A', 'P', 'R'):
            M = getattr(L, nm)
            d[nm] = (M.indptr, M.indices, M.data)
        out.append(d)
    return out


class DistributedSolver:
    """Collective constructor.  On rank 0: `levels` = describe(ml) (or an equivalent list) and
    `coarse_factory(ka)` -> object with cycle_from_zero(b_c) -> x_c running levels[ka:] on rank 0;
    both None on the other ranks."""

    @classmethod
    def from_ml(cls, ml, **kw):
        """ml: device multilevel_solver on rank 0, None elsewhere"""
        if ml is None:
            return cls(None, None, **kw)
        from .multilevel import multilevel_solver

        def coarse(ka):
            sub = multilevel_solver(ml.levels[ka:], coarse_solver=ml.coarse_solver_spec)
            return sub
        return cls(describe(ml), coarse, **kw)

    def __init__(self, levels, coarse_factory, group=None, agglomerate_below=None, backend=None, device=None,
                 halo='auto', min_rows_per_rank=16384):
        self.group = group
        self.rank, self.G = dist.get_rank(group), dist.get_world_size(group)
        self.be = backend if backend is not None else DeviceBackend()
        self.device = device if device is not None else self.be.device
        self.exchanges = 0
        self.halo_mode = halo
        import os
        self._fused_halo = os.environ.get('B200AMG_HALO_FUSED', '1') != '0'    # one kernel node per refresh (A/B switch)
        self._local_halo_prolong = os.environ.get('B200AMG_HALO_PROLONG', '1') != '0'   # halo of x corrected locally (A/B switch)
        self._p2p = None
        self._graph = None
        self._prof = None
        meta = [None]
        if self.rank == 0:
            nlev = len(levels) + 1
            ka = nlev - 1
            # a level stays partitioned while every rank keeps a slab worth launching kernels for: the threshold
            # scales with the number of ranks (at 2 GPUs the 89 k-row level 2 of 2 x 256^3 stays partitioned instead
            # of making rank 0 run it alone while rank 1 idles); `agglomerate_below` overrides it with a global size
            limit = agglomerate_below if agglomerate_below is not None else min_rows_per_rank * self.G
            for k in range(1, nlev - 1):
                if levels[k]['n'] < limit:
                    ka = k
                    break
            info = [{q: levels[k].get(q, 1) if q in ('bs', 'bsc') else levels[k][q] for q in ('n', 'nc', 'flavor', 'pre', 'post', 'bs', 'bsc')}
                    for k in range(ka)]
            for k in range(ka):
                info[k]['K'] = len(levels[k]['color_ptr']) - 1 if 'color_ptr' in levels[k] else 0
                info[k]['gs_csr'] = int(levels[k].get('gs_csr', 1))
            meta[0] = {'ka': ka, 'nlev': nlev, 'levels': info}
        dist.broadcast_object_list(meta, src=0, group=group)
        meta = meta[0]
        self.ka, self.nlev = meta['ka'], meta['nlev']
        self.levels = []
        for k in range(self.ka):
            self.levels.append(self._distribute_level(k, meta['levels'][k], levels[k] if self.rank == 0 else None))
        # agglomerated lower part of the cycle: ordinary single-GPU engine on rank 0
        self.coarse = None
        if self.rank == 0:
            self.coarse = coarse_factory(self.ka)
            self.nc_global = meta['levels'][-1]['nc']
        self._alloc()

    # ---- setup ---------------------------------------------------------------------------------
    def _distribute_level(self, k, info, src):
        G, rank, dev = self.G, self.rank, self.device
        n, nc = info['n'], info['nc']
        # block levels are cut between BLOCK rows (a block row is relaxed as a unit); vectors, halo plans and every
        # size below count scalars.  The coarse split of level k is the fine split of level k + 1 (bsc_k == bs_{k+1}).
        bs, bsc = info.get('bs', 1), info.get('bsc', 1)
        spb, spcb = part.split_points(n // bs, G), part.split_points(nc // bsc, G)
        sp, spc = [v * bs for v in spb], [v * bsc for v in spcb]
        L = _Level()
        L.info, L.split, L.csplit = info, sp, spc
        L.n_own, L.nc_own = sp[rank + 1] - sp[rank], spc[rank + 1] - spc[rank]
        shapes = {'A': (spb, spb, bs, bs), 'P': (spb, spcb, bs, bsc), 'R': (spcb, spb, bsc, bs)}
        # Correcting the halo of x locally after prolongation (instead of refreshing it) is safe only when the post-smoother
        # does not refresh x IN PLACE: a neighbour that runs ahead would push its next colour's values into this rank's
        # halo while this rank's own correction (a read-modify-write of the same entries) is still to come.  An
        # out-of-place sweep (Jacobi: reads x, writes the other buffer; the next push goes to THAT buffer) cannot collide:
        # every later push into this buffer's halo is ordered behind an exchange this rank takes part in after its
        # correction.  (The peer-memory protocol has no "halo consumed" signal: it relies on consecutive refreshes
        # alternating between buffers.)  Without post-smoothing nobody reads the halo of x: nothing to do.
        local_halo = L.local_halo = bool(self._local_halo_prolong and info['post'][0] == 1 and info['post'][2] >= 1)
        for nm in ('A', 'P', 'R'):
            rs, cs, bR, bC = shapes[nm]
            header = [None]
            if rank == 0:
                Mp, Mj, Mx = src[nm]
                # P also ships its rows for every rank's A-halo points (extra): after x += P x_c the halo of x is
                # corrected locally with the same row arithmetic -- the refresh of x before post-smoothing goes away
                extra = halosA if (nm == 'P' and local_halo) else [None] * G
                halos = [part.localize(Mp, Mj, Mx, rs[p], rs[p + 1], cs[p], cs[p + 1], bR * bC, extra[p])[3] for p in range(G)]
                if nm == 'A':
                    halosA = halos
                plans = part.build_plans(halos, cs)
                header[0] = [pl.pack() for pl in plans]
                if nm == 'P':
                    # what every rank's x_c looks like, [owned | halo], as indices into the GLOBAL coarse vector (scalars):
                    # when this is the last partitioned level, rank 0 holds all of x_c after the coarse solve and ships
                    # slab AND halo in one scatter -- one halo refresh per cycle less
                    k = torch.arange(bC, dtype=torch.int64, device=halos[0].device)
                    L.xc_layout = [torch.cat([torch.arange(spc[p], spc[p + 1], dtype=torch.int64, device=halos[p].device),
                                              (halos[p].to(torch.int64)[:, None] * bC + k[None, :]).reshape(-1)])
                                   for p in range(G)]
            dist.broadcast_object_list(header, src=0, group=self.group)
            plan = part.HaloPlan.unpack(header[0][rank]).expanded(bC).to(dev)
            with_halo_rows = nm == 'P' and local_halo
            if rank == 0:
                mine = None
                for p in range(G):
                    got = part.localize(Mp, Mj, Mx, rs[p], rs[p + 1], cs[p], cs[p + 1], bR * bC, extra[p])
                    ops = [got[:3]] + ([got[4]] if with_halo_rows else [])
                    if p == 0:
                        mine = ops
                    else:
                        for (ip, ix, dx) in ops:
                            sizes = torch.tensor([ip.numel(), ix.numel()], dtype=torch.int64, device=dev)
                            dist.send(sizes, p, group=self.group)
                            _send_tensors((ip, ix, dx), p, self.group)
                    del got, ops
                ops = mine
            else:
                ops = []
                for _ in range(2 if with_halo_rows else 1):
                    sizes = torch.empty(2, dtype=torch.int64, device=dev)
                    dist.recv(sizes, 0, group=self.group)
                    n_ip, n_ix = [int(v) for v in sizes.tolist()]
                    ops.append(_recv_tensors([(n_ip, torch.int32), (n_ix, torch.int32), (n_ix * bR * bC, torch.float64)], 0, dev,
                                             self.group))
            ip, ix, dx = ops[0]
            setattr(L, nm, self.be.make_op(ip, ix, dx, ip.numel() - 1, bR, bC))
            setattr(L, 'plan' + nm, plan)
            if with_halo_rows:
                ip, ix, dx = ops[1]
                L.Ph = self.be.make_op(ip, ix, dx, ip.numel() - 1, bR, bC) if ip.numel() > 1 else None
        L.diag = L.A.diagonal()
        # colour-ordered Gauss-Seidel: the GLOBAL colouring cut by slab -- every rank keeps, colour by colour, the rows of
        # its slab in local numbering (ascending inside a colour like the global list), so a colour step relaxes the
        # same rows with the same operands as the single-GPU sweep once the halo is fresh
        L.Id, L.cptr = None, None
        if info.get('K', 0) > 0:
            parts = [None] * G
            if rank == 0:
                Id = torch.as_tensor(src['Id']).to(torch.int64).cpu()
                cp = src['color_ptr']
                for p in range(G):
                    segs, ptr = [], [0]
                    for c in range(len(cp) - 1):
                        seg = Id[cp[c]:cp[c + 1]]                   # (block) rows of this colour
                        seg = seg[(seg >= spb[p]) & (seg < spb[p + 1])] - spb[p]
                        segs.append(seg)
                        ptr.append(ptr[-1] + int(seg.numel()))
                    parts[p] = (torch.cat(segs).to(torch.int32) if segs else torch.zeros(0, dtype=torch.int32), ptr)
            got = [None]
            dist.scatter_object_list(got, parts if rank == 0 else None, src=0, group=self.group)
            L.Id, L.cptr = got[0][0].to(dev).contiguous(), got[0][1]
        return L

    def _alloc(self):
        dev, f64 = self.device, torch.float64
        K = len(self.levels)
        # vectors that receive halos: X_k / XS_k (iterate + Jacobi spare), R_k, and X_K (coarse vector of
        # the last partitioned level).  X_{k+1} doubles as x_c of level k.
        sizes = {}
        for k, L in enumerate(self.levels):
            halo_x = max(L.planA.n_halo, self.levels[k - 1].planP.n_halo if k > 0 else 0)
            sizes[('x', k)] = sizes[('xs', k)] = L.n_own + halo_x
            sizes[('r', k)] = L.n_own + L.planR.n_halo
        Ll = self.levels[-1]
        sizes[('x', K)] = Ll.nc_own + Ll.planP.n_halo
        self._buf_sizes = sizes
        use_p2p = self.halo_mode in ('auto', 'p2p') and self.G > 1 and isinstance(self.be, DeviceBackend)
        bufs = None
        if use_p2p:
            try:
                bufs = self._alloc_symmetric(sizes)
            except Exception as exc:                      # noqa: BLE001
                if self.halo_mode == 'p2p':
                    raise
                import warnings
                warnings.warn('symmetric memory unavailable (%s): halo refresh falls back to NCCL all-to-all' % (exc,))
                bufs = None
        if bufs is None:
            bufs = {key: torch.zeros(max(n, 1), dtype=f64, device=dev) for key, n in sizes.items()}
        self._bufs = bufs
        for k, L in enumerate(self.levels):
            L.x, L.xs, L.r = bufs[('x', k)], bufs[('xs', k)], bufs[('r', k)]
            L.xc = bufs[('x', k + 1)]
            L.bc = torch.zeros(max(L.nc_own, 1), dtype=f64, device=dev)
            L.b = torch.zeros(max(L.n_own, 1), dtype=f64, device=dev) if k == 0 else self.levels[k - 1].bc
        if self.rank == 0:
            self.cb = torch.zeros(self.nc_global, dtype=f64, device=dev)
        if self._p2p is not None:
            self._build_p2p_tables()

    # ---- NVLink peer-memory halo refresh (csrc/halo.cu) ---------------------------------------------
    def _alloc_symmetric(self, sizes):
        import torch.distributed._symmetric_memory as symm
        grp = self.group if self.group is not None else dist.group.WORLD
        al = lambda n: (max(n, 1) + 31) & ~31
        offs, tot = {}, 0
        for key in sorted(sizes):
            offs[key] = tot
            tot += al(sizes[key])
        t = torch.tensor([tot], dtype=torch.int64, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        arena = symm.empty(int(t.item()), dtype=torch.float64, device=self.device)
        arena.zero_()
        nslots = 3 * len(self.levels)
        flags = symm.empty(nslots * self.G + 32, dtype=torch.int32, device=self.device)
        flags.zero_()
        torch.cuda.synchronize()
        ha = symm.rendezvous(arena, grp)
        hf = symm.rendezvous(flags, grp)
        self._p2p = {'arena': arena, 'flags': flags, 'ha': ha, 'hf': hf, 'offs': offs,
                     'arena_ptrs': [int(p) for p in ha.buffer_ptrs], 'flag_ptrs': [int(p) for p in hf.buffer_ptrs]}
        torch.cuda.synchronize()
        dist.barrier(group=self.group)
        return {key: arena[offs[key]:offs[key] + max(sizes[key], 1)] for key in sizes}

    def _build_p2p_tables(self):
        """per plan: element-wise destinations on the peers; per (plan, target buffer): peer base table"""
        P, G, me, dev = self._p2p, self.G, self.rank, self.device
        K = len(self.levels)
        mine = {'offs': {'%s%d' % key: v for key, v in P['offs'].items()},
                'n_own': [L.n_own for L in self.levels] + [self.levels[-1].nc_own],
                'recv': {}}
        plans = {}
        for k, L in enumerate(self.levels):
            plans[('A', k)] = (L.planA, k, [('x', k), ('xs', k)])
            plans[('R', k)] = (L.planR, k, [('r', k)])
            plans[('P', k)] = (L.planP, k + 1, [('x', k + 1)] + ([('xs', k + 1)] if k + 1 < K else []))
        for (kind, k), (pl, _, _) in plans.items():
            mine['recv']['%s%d' % (kind, k)] = pl.recv_counts
        allmeta = [None] * G
        dist.all_gather_object(allmeta, mine, group=self.group)
        nbr = set()
        for (kind, k), (pl, _, _) in plans.items():
            for q in range(G):
                if q != me and (pl.send_counts[q] > 0 or pl.recv_counts[q] > 0):
                    nbr.add(q)
        nbr = sorted(nbr)
        P['nbr'] = nbr
        u64 = lambda v: torch.tensor(v, dtype=torch.int64, device=dev)
        P['plans'] = {}
        for slot, ((kind, k), (pl, vec_level, targets)) in enumerate(sorted(plans.items())):
            name = '%s%d' % (kind, k)
            dst_rank, dst_off = [], []
            for p in range(G):
                cnt = pl.send_counts[p]
                if cnt == 0:
                    continue
                rc = allmeta[p]['recv'][name]                       # what p receives, by source rank
                base = allmeta[p]['n_own'][vec_level] + sum(rc[:me])
                dst_rank.append(torch.full((cnt,), p, dtype=torch.int32))
                dst_off.append(torch.arange(base, base + cnt, dtype=torch.int64))
            n_send = int(pl.send_idx.numel())
            ent = {'n_send': n_send,
                   'send_idx': pl.send_idx.to(torch.int32).contiguous(),
                   'dst_rank': (torch.cat(dst_rank) if dst_rank else torch.zeros(0, dtype=torch.int32)).to(dev),
                   'dst_off': (torch.cat(dst_off) if dst_off else torch.zeros(0, dtype=torch.int64)).to(dev),
                   'sig': u64([P['flag_ptrs'][q] + 4 * (slot * G + me) for q in nbr] or [0]),
                   'flg': u64([P['flag_ptrs'][me] + 4 * (slot * G + q) for q in nbr] or [0]),
                   'push_state': torch.zeros(2, dtype=torch.int32, device=dev),
                   'wait_state': torch.zeros(2, dtype=torch.int32, device=dev),
                   'bases': {}}
            for tkey in targets:
                tname = '%s%d' % tkey
                ent['bases'][self._bufs[tkey].data_ptr()] = u64([P['arena_ptrs'][p] + 8 * allmeta[p]['offs'][tname] for p in range(G)])
            P['plans'][id(pl)] = ent
        torch.cuda.synchronize()
        dist.barrier(group=self.group)

    def halo_errors(self):
        """number of halo waits that timed out (0 on a healthy run)"""
        if self._p2p is None:
            return 0
        return int(sum(int(e['wait_state'][1].item()) for e in self._p2p['plans'].values()))

    # ---- communication -------------------------------------------------------------------------
    def _mark(self, label):
        if self._prof is not None:
            e = torch.cuda.Event(enable_timing=True)
            e.record(torch.cuda.current_stream())
            self._prof.append((label, e))

    def profile_cycle(self):
        """one eager cycle with a CUDA event after every step -> [(label, ms)] on this rank"""
        self._prof = []
        self._mark('start')
        self._cycle_eager()
        torch.cuda.synchronize()
        ev, self._prof = self._prof, None
        return [(ev[i][0], ev[i - 1][1].elapsed_time(ev[i][1])) for i in range(1, len(ev))]

    def exchange(self, plan, v):
        """refresh v[n_own : n_own + n_halo] from the owners (v = [owned | halo])"""
        if self.G == 1:
            return                                        # (a rank with nothing to send or receive still joins the collective)
        self._mark('compute')
        if self._p2p is not None:
            ent = self._p2p['plans'][id(plan)]
            lib, st, nn = self.be._lib, self.be.stream(), len(self._p2p['nbr'])
            if self._fused_halo:
                lib.call('b200amg_halo_exchange', v.data_ptr(), ent['n_send'], ent['send_idx'].data_ptr(),
                         ent['dst_rank'].data_ptr(), ent['dst_off'].data_ptr(), ent['bases'][v.data_ptr()].data_ptr(),
                         ent['sig'].data_ptr(), ent['flg'].data_ptr(), nn, ent['push_state'].data_ptr(),
                         ent['wait_state'].data_ptr(), 1 << 24, st)
            else:
                lib.call('b200amg_halo_push', v.data_ptr(), ent['n_send'], ent['send_idx'].data_ptr(), ent['dst_rank'].data_ptr(),
                         ent['dst_off'].data_ptr(), ent['bases'][v.data_ptr()].data_ptr(), ent['sig'].data_ptr(), nn,
                         ent['push_state'].data_ptr(), st)
                lib.call('b200amg_halo_wait', ent['flg'].data_ptr(), nn, ent['wait_state'].data_ptr(), 1 << 24, st)
            self.exchanges += 1
            self._mark('exchange')
            return
        send = v[:plan.n_own].index_select(0, plan.send_idx)
        recv = v[plan.n_own:plan.n_own + plan.n_halo]
        dist.all_to_all_single(recv, send, output_split_sizes=plan.recv_counts, input_split_sizes=plan.send_counts,
                               group=self.group)
        self.exchanges += 1
        self._mark('exchange')

    def _gather_coarse(self, L):
        spc = L.csplit
        sizes = [spc[p + 1] - spc[p] for p in range(self.G)]
        out = self.cb if self.rank == 0 else torch.empty(0, dtype=torch.float64, device=self.device)
        dist.all_to_all_single(out, L.bc[:L.nc_own], output_split_sizes=sizes if self.rank == 0 else [0] * self.G,
                               input_split_sizes=[L.nc_own] + [0] * (self.G - 1), group=self.group)

    def _scatter_coarse(self, L, xc_global):
        """x_c of the agglomerated levels back to the slabs -- owned part AND halo of every rank ([owned | halo] is the
        layout of L.xc), so the prolongation needs no refresh of x_c after it"""
        mine = L.nc_own + L.planP.n_halo
        if self.rank == 0:
            if not hasattr(self, '_scat'):
                idx = torch.cat(L.xc_layout).to(self.device)
                self._scat = (idx, [int(t.numel()) for t in L.xc_layout],
                              torch.empty(max(int(idx.numel()), 1), dtype=torch.float64, device=self.device))
            idx, sizes, buf = self._scat
            src = torch.index_select(xc_global, 0, idx, out=buf[:idx.numel()])
        else:
            sizes, src = [0] * self.G, torch.empty(0, dtype=torch.float64, device=self.device)
        dist.all_to_all_single(L.xc[:mine], src, output_split_sizes=[mine] + [0] * (self.G - 1),
                               input_split_sizes=sizes, group=self.group)

    # ---- the cycle (multilevel.py:473-548 on slabs) ----------------------------------------------
    def _smooth(self, L, spec, x, xs, b, x_is_zero, halo_fresh=False):
        kind, omega, iters = spec[:3]
        if kind == 0 or iters == 0:
            if x_is_zero:
                x[:L.n_own].zero_()
            return x, xs
        if kind == 2:
            # colour-ordered Gauss-Seidel over the GLOBAL colours (relaxation.h:394-426 semantics): one halo refresh
            # per colour, then this rank's rows of the colour in parallel; bit-identical to the single-GPU sweep
            sweep = spec[3] if len(spec) > 3 else 'forward'
            if x_is_zero:
                x[:L.n_own].zero_()
            K = len(L.cptr) - 1
            for _ in range(iters):
                for backward in (False, True):
                    if (sweep == 'forward' and backward) or (sweep == 'backward' and not backward):
                        continue
                    for q in range(K):
                        c = K - 1 - q if backward else q
                        if halo_fresh:
                            halo_fresh = False                    # (the caller has just made the halo of x current)
                        else:
                            self.exchange(L.planA, x)
                        self.be.gs_color(L.A, x, b, L.Id, L.cptr[c], L.cptr[c + 1], L.info['flavor'] and not L.info.get('gs_csr', 1), backward)
            return x, xs
        op = OP_JACOBI_BSR1 if L.info['flavor'] else OP_JACOBI
        for it in range(iters):
            if it == 0 and x_is_zero:
                self.be.scale_diag(omega, b, L.diag, x, L.info['flavor'])     # bit-identical zero-guess sweep
                continue
            if halo_fresh and it == 0:
                pass                                              # (the caller has just made the halo of x current)
            else:
                self.exchange(L.planA, x)
            L.A.apply(op, x, b, xs, omega)
            x, xs = xs, x
        return x, xs

    def _cycle(self, k, x, b, x_is_zero):
        L = self.levels[k]
        xs = L.xs if L.xs is not x else L.x
        x, xs = self._smooth(L, L.info['pre'], x, xs, b, x_is_zero)
        self.exchange(L.planA, x)
        L.A.apply(OP_RESIDUAL, x, b, L.r)
        self.exchange(L.planR, L.r)
        L.R.apply(OP_MATVEC, L.r, None, L.bc)
        xc = L.xc
        scattered_with_halo = False
        if k + 1 < len(self.levels):
            xc, _ = self._cycle(k + 1, L.xc, L.bc, True)          # the iterate may end in the child's spare buffer
        else:
            self._mark('compute')
            self._gather_coarse(L)
            self._mark('gather')
            xg = self.coarse.cycle_from_zero(self.cb) if self.rank == 0 else None
            self._mark('coarse_rank0')
            self._scatter_coarse(L, xg)
            scattered_with_halo = True
            self._mark('scatter')
        if not scattered_with_halo:
            self.exchange(L.planP, xc)
        L.P.apply(OP_ADD, xc, None, x)
        fresh = False
        if L.local_halo:
            # the halo of x was refreshed before the residual and nothing has touched x since: the correction of the halo
            # points is P's rows for them times the same x_c -- what their owners compute, bit for bit
            if getattr(L, 'Ph', None) is not None:
                L.Ph.apply(OP_ADD, xc, None, x[L.n_own:])
            fresh = True
        x, xs = self._smooth(L, L.info['post'], x, xs, b, False, halo_fresh=fresh)
        return x, xs

    def _cycle_eager(self):
        L = self.levels[0]
        L.x, L.xs = self._cycle(0, L.x, L.b, False)        # ping-pong: no copy back
        self._mark('compute')

    def cycle(self):
        """one V-cycle on the resident (x, b) slabs (a CUDA-graph replay after capture())"""
        if self._graph is not None:
            self._graph.replay()
        else:
            self._cycle_eager()

    def release_graph(self):
        """drop the captured cycle (required before the process group is destroyed)"""
        if self._graph is not None:
            torch.cuda.synchronize()
            self._graph.reset()
            self._graph = None

    def capture(self, warmup=2):
        """Capture one whole distributed cycle -- slab kernels, halo all-to-alls (NCCL), the rank-0 coarse
        engine -- into ONE CUDA graph per rank: replays cost one launch instead of ~60 Python-driven
        ones.  Collective (every rank must call it); device backend only."""
        L = self.levels[0]
        x0 = L.x.clone()
        for _ in range(warmup):                             # engines / communicators / allocator warm
            self._cycle_eager()
        torch.cuda.synchronize()
        dist.barrier(group=self.group)
        home = L.x
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._cycle_eager()
            if L.x is not home:                             # odd number of sweeps: keep the iterate in place
                home[:L.n_own].copy_(L.x[:L.n_own])
                L.x, L.xs = home, L.x
        self._graph = g
        L.x.copy_(x0)
        torch.cuda.synchronize()
        dist.barrier(group=self.group)

    # ---- data movement / norms ---------------------------------------------------------------------
    def load(self, b_global=None, x0_global=None):
        """scatter b (and x0, default 0) from rank 0 to the slabs"""
        L = self.levels[0]
        sp = L.split
        sizes = [sp[p + 1] - sp[p] for p in range(self.G)]
        for dst, src in ((L.b, b_global), (L.x, x0_global)):
            if src is None and dst is L.x:
                L.x.zero_()
                continue
            if self.rank == 0:
                if getattr(self, '_stage', None) is None:
                    self._stage = torch.empty(sp[-1], dtype=torch.float64, device=self.device)
                self._stage.copy_(torch.as_tensor(src, dtype=torch.float64).reshape(-1), non_blocking=True)
                s = self._stage
            else:
                s = torch.empty(0, dtype=torch.float64, device=self.device)
            dist.all_to_all_single(dst[:L.n_own], s, output_split_sizes=[L.n_own] + [0] * (self.G - 1),
                                   input_split_sizes=sizes if self.rank == 0 else [0] * self.G, group=self.group)

    def load_local(self, b_slab, x0_slab=None):
        """every rank loads ITS OWN slab of b (and x0, default 0) -- a host (pinned) or device vector of n_own entries --
        through its own PCIe link: the distributed form of solve(b_host); no collective, asynchronous"""
        L = self.levels[0]
        L.b[:L.n_own].copy_(torch.as_tensor(b_slab).reshape(-1), non_blocking=True)
        if x0_slab is None:
            L.x[:L.n_own].zero_()
        else:
            L.x[:L.n_own].copy_(torch.as_tensor(x0_slab).reshape(-1), non_blocking=True)

    def store_local(self, out):
        """this rank's slab of x -> `out` (host pinned or device vector of n_own entries); asynchronous"""
        L = self.levels[0]
        out.copy_(L.x[:L.n_own], non_blocking=True)
        return out

    def slab(self):
        """(first row, one past the last row) of this rank's slab of the finest level"""
        sp = self.levels[0].split
        return int(sp[self.rank]), int(sp[self.rank + 1])

    def check_halo(self):
        """raise if a peer-memory halo wait timed out since the last check (a stalled peer would otherwise give
        silently wrong iterates); one small D2H read per call site"""
        n = self.halo_errors()
        if n:
            raise RuntimeError('distributed cycle: %d halo wait(s) timed out -- a peer rank stalled; the iterate is not valid' % n)

    def gather_x(self):
        L = self.levels[0]
        sp = L.split
        sizes = [sp[p + 1] - sp[p] for p in range(self.G)]
        if getattr(self, '_xg', None) is None:
            self._xg = torch.empty(sp[-1] if self.rank == 0 else 0, dtype=torch.float64, device=self.device)
        out = self._xg
        dist.all_to_all_single(out, L.x[:L.n_own], output_split_sizes=sizes if self.rank == 0 else [0] * self.G,
                               input_split_sizes=[L.n_own] + [0] * (self.G - 1), group=self.group)
        return out if self.rank == 0 else None

    def residual_norm(self):
        L = self.levels[0]
        self.exchange(L.planA, L.x)
        L.A.apply(OP_RESIDUAL, L.x, L.b, L.r)
        s = self.be.dot(L.r, L.r, L.n_own)
        dist.all_reduce(s, group=self.group)
        return float(np.sqrt(s.item()))

    def norm_b(self):
        L = self.levels[0]
        s = self.be.dot(L.b, L.b, L.n_own)
        dist.all_reduce(s, group=self.group)
        return float(np.sqrt(s.item()))

    def solve(self, b=None, x0=None, tol=1e-5, maxiter=100, residuals=None):
        """multilevel_solver.solve semantics (multilevel.py:316-471) on the partitioned problem.
        b / x0 are global arrays on rank 0 (ignored elsewhere); returns global x on rank 0, None elsewhere."""
        self.load(b, x0)
        normb = self.norm_b()
        if normb != 0:
            tol = tol * normb
        res = [self.residual_norm()]
        while len(res) <= maxiter and res[-1] > tol:
            self.cycle()
            res.append(self.residual_norm())
        if residuals is not None:
            residuals[:] = res
        self.check_halo()
        x = self.gather_x()
        return x


class _Participant:
    """What a rank WITHOUT the hierarchy holds in place of the multilevel_solver: its solve() joins the collective
    that rank 0 enters through `ml.solve(b, distributed=True)`, with the same keyword arguments."""

    def __init__(self, **kw):
        self._kw, self._S = kw, None

    def solver(self):
        if self._S is None:
            self._S = DistributedSolver.from_ml(None, **self._kw)
        return self._S

    def solve(self, b=None, x0=None, tol=1e-5, maxiter=100, cycle='V', residuals=None, **ignored):
        if str(cycle).upper() != 'V':
            raise NotImplementedError('distributed solve: V-cycles only')
        return self.solver().solve(None, None, tol=tol, maxiter=maxiter, residuals=residuals)


def participant(**kw):
    """the stand-in for `ml` on the ranks that do not hold the hierarchy (rank != 0):
        ml = b200.smoothed_aggregation_solver(A, ...) if rank == 0 else b200.distributed.participant()
        x = ml.solve(b if rank == 0 else None, tol=1e-8, distributed=True)       # collective; x on rank 0"""
    return _Participant(**kw)
