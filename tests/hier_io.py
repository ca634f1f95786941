"""Hierarchy <-> .npz helpers for the golden fixtures (tests only).

A fixture stores, per level k: A (and P, R on non-coarsest levels) as
``L{k}_{A|P|R}_{indptr,indices,data,shape,blocksize}`` (blocksize (1,1) + format flag for CSR),
the resolved smoother parameters, plus b, the reference residual history and final x.
"""
import json

import numpy as np
import scipy.sparse as sp


def pack_matrix(out, key, M):
    out[key + '_indptr'] = np.asarray(M.indptr, np.int32)
    out[key + '_indices'] = np.asarray(M.indices, np.int32)
    out[key + '_data'] = np.asarray(M.data, np.float64)
    out[key + '_shape'] = np.asarray(M.shape, np.int64)
    out[key + '_fmt'] = np.asarray([1 if M.format == 'bsr' else 0] +
                                   (list(M.blocksize) if M.format == 'bsr' else [1, 1]), np.int64)


def unpack_matrix(z, key):
    shape = tuple(int(s) for s in z[key + '_shape'])
    isbsr, R, C = (int(v) for v in z[key + '_fmt'])
    if isbsr:
        return sp.bsr_matrix((z[key + '_data'].reshape(-1, R, C), z[key + '_indices'], z[key + '_indptr']),
                             shape=shape, blocksize=(R, C))
    return sp.csr_matrix((z[key + '_data'], z[key + '_indices'], z[key + '_indptr']), shape=shape)


def save(path, levels, smoothers, extra):
    """levels: objects with A (and P, R); smoothers: list (per non-coarsest level) of
    (pre_spec, post_spec) with spec = (name, {params: scalars or int arrays})."""
    out = {}
    meta = {'nlevels': len(levels), 'smoothers': []}
    for k, lvl in enumerate(levels):
        pack_matrix(out, 'L%d_A' % k, lvl.A)
        if k < len(levels) - 1:
            pack_matrix(out, 'L%d_P' % k, lvl.P)
            pack_matrix(out, 'L%d_R' % k, lvl.R)
            entry = []
            for which, (name, kw) in zip(('pre', 'post'), smoothers[k]):
                kwj = {}
                for pk, pv in kw.items():
                    if isinstance(pv, np.ndarray):
                        akey = 'L%d_%s_%s' % (k, which, pk)
                        out[akey] = pv
                        kwj[pk] = {'array': akey}
                    else:
                        kwj[pk] = pv
                entry.append([name, kwj])
            meta['smoothers'].append(entry)
    for k, v in extra.items():
        out['X_' + k] = np.asarray(v)
    out['meta'] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
    np.savez_compressed(path, **out)


class Lvl:
    pass


def load(path):
    """-> (levels with A,P,R,pre,post ; extra dict)."""
    z = np.load(path)
    meta = json.loads(bytes(z['meta']).decode())
    levels = []
    for k in range(meta['nlevels']):
        lvl = Lvl()
        lvl.A = unpack_matrix(z, 'L%d_A' % k)
        if k < meta['nlevels'] - 1:
            lvl.P = unpack_matrix(z, 'L%d_P' % k)
            lvl.R = unpack_matrix(z, 'L%d_R' % k)
            specs = []
            for name, kwj in meta['smoothers'][k]:
                kw = {pk: (z[pv['array']] if isinstance(pv, dict) else pv) for pk, pv in kwj.items()}
                specs.append((name, kw))
            lvl.pre, lvl.post = specs
        levels.append(lvl)
    extra = {k[2:]: z[k] for k in z.files if k.startswith('X_')}
    return levels, extra
