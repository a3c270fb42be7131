"""Per-solve overhead of C2 (TV denoising 1024^2, 200 updates, on-chip resident kernel): host time of each phase of
PDS.iterate() with a device sync after each, then whole solves, then a cProfile of 20 solves."""
import cProfile
import os
import pstats
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from pycsou_b200.func.loss import SquaredL2Loss
from pycsou_b200.func.penalty import L21Norm, tv_groups
from pycsou_b200.linop.diff import Gradient
from pycsou_b200.opt.proxalgs import PDS
from pycsou_b200.util import synth


def main(dtype, iters=200):
    tdt = torch.float32 if dtype == 'f32' else torch.float64
    ndt = np.float32 if dtype == 'f32' else np.float64
    shape = (1024, 1024)
    N = shape[0] * shape[1]
    _, y = synth.tv_denoise_problem(shape, 0.1, 0)
    K = Gradient(shape=shape, kind='forward', dtype=ndt.__name__)
    K.lipschitz_cst = K.diff_lipschitz_cst = float(np.sqrt(8.))
    F = (1 / 2) * SquaredL2Loss(dim=N, data=torch.from_numpy(y.astype(ndt)).cuda())
    s = PDS(dim=N, F=F, H=0.1 * L21Norm(dim=2 * N, groups=tv_groups(N)), K=K, rho=0.9, max_iter=iters - 1,
            min_iter=iters - 1, accuracy_threshold=0, verbose=None, dtype=tdt, output='torch')
    s.iterate(); s.reset(); s.iterate()
    acc = []

    def timed(label, fn):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = fn()
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        acc.append(f'{label} host {1e3 * (t1 - t0):.3f} / synced {1e3 * (time.perf_counter() - t0):.3f} ms')
        return r
    timed('reset', s.reset)
    timed('make_plan', s._make_plan)
    plan = s._plan
    st = timed('init_state', s._init_state)
    timed('alloc_diag', s._alloc_diag)
    st = timed(f'run_many({iters})', lambda: plan.run_many(st[0], st[1], iters, s._diag_dev[0:iters]))
    timed('diag D2H', lambda: s._diag_host[:iters].copy_(s._diag_dev[:iters]))
    timed('release', lambda: plan.release(*st))
    timed('build_diagnostics', lambda: s._build_diagnostics(iters))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(10):
        s.reset(); s.iterate()
    torch.cuda.synchronize()
    acc.append(f'10 x (reset + iterate): wall {1e3 * (time.perf_counter() - t0) / 10:.3f} ms per solve')
    print(f'[{dtype}] ' + '; '.join(acc), flush=True)
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(20):
        s.reset(); s.iterate()
    torch.cuda.synchronize()
    pr.disable()
    pstats.Stats(pr).sort_stats('cumulative').print_stats(22)


for d in sys.argv[1:] or ['f64']:
    main(d)
