"""A/B of the Galerkin product R A P on the device: the sorting rows (B200AMG_SPGEMM_HASH=0) against the sort-free
rows (=1: rows of up to 512 products, =3: up to 8192, the longer ones staged in global memory) on the levels of the config-2 hierarchy.  Every mode's result
(indptr, indices in scipy's order, values) must equal mode 0's to the bit.  Prints one JSON line.
    python tools/spgemm_ab.py [--grid 256]"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..'))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import pyamg_b200 as amg  # noqa: E402
from pyamg_b200 import sparse  # noqa: E402


MODES = ('0', '1', '3')


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--grid', type=int, default=256)
    args = ap.parse_args()
    torch.cuda.set_device(0)
    out = {'grid': args.grid, 'levels': []}
    A = amg.gallery.poisson((args.grid,) * 3)
    setup = {}
    ml = None
    for mode in ('1', '0', '3'):                       # (the first build also pays the process warm-up)
        os.environ['B200AMG_SPGEMM_HASH'] = mode
        del ml
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        np.random.seed(0)
        ml = amg.smoothed_aggregation_solver(A, max_coarse=500)
        torch.cuda.synchronize()
        setup[mode] = round(time.perf_counter() - t0, 3)
    out['setup_s'] = setup
    for k, L in enumerate(ml.levels[:-1]):
        Rm, Am, Pm = L.R, L.A, L.P
        rec = {'level': k, 'rows_R': int(Rm.shape[0]), 'nnz_R': int(Rm.nnz), 'nnz_A': int(Am.nnz)}
        want = None
        for mode in MODES:
            os.environ['B200AMG_SPGEMM_HASH'] = mode
            best = None
            for _ in range(2):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                RA = sparse.matmat(Rm, Am, True)
                torch.cuda.synchronize()
                t1 = time.perf_counter()
                Ac = sparse.matmat(RA, Pm, True)
                torch.cuda.synchronize()
                t2 = time.perf_counter()
                if best is None or t2 - t0 < best[0]:
                    best = (t2 - t0, t1 - t0, t2 - t1)
            got = (RA.indptr, RA.indices, RA.data, Ac.indptr, Ac.indices, Ac.data)
            if want is None:
                want = got
                same = True
            else:
                same = all(torch.equal(g, w) for g, w in zip(got, want))
            rec['mode' + mode] = {'s': round(best[0], 4), 's_RA': round(best[1], 4), 's_RAP': round(best[2], 4),
                                  'bit_identical_to_mode0': bool(same)}
            del RA, Ac
        del want, got
        out['levels'].append(rec)
    print(json.dumps(out))


if __name__ == '__main__':
    main()
