"""Per-solve overhead of the single-GPU path: host time of each phase of PDS.iterate() with a device sync after each."""
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench


def main(dtype):
    tdt = torch.float32 if dtype == 'f32' else torch.float64
    psf, y64 = bench.make_problem(4096, 'sep', 0)
    y = torch.from_numpy(y64.astype(np.float32 if dtype == 'f32' else np.float64)).cuda()
    s = bench.make_solver(y, psf, 4096, tdt, 100, 'torch')
    s.iterate()
    s.reset()
    s.iterate()
    for rep in range(2):
        acc = []

        def timed(label, fn):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            r = fn()
            t1 = time.perf_counter()
            torch.cuda.synchronize()
            acc.append(f'{label} host {1e3 * (t1 - t0):.2f} / synced {1e3 * (time.perf_counter() - t0):.2f} ms')
            return r
        s.reset()
        timed('make_plan', s._make_plan)
        plan = s._plan
        timed('begin_solve', plan.begin_solve)
        st = timed('init_state', s._init_state)
        timed('alloc_diag', s._alloc_diag)
        st = timed('run_many(100)', lambda: plan.run_many(st[0], st[1], 100, s._diag_dev[0:100]))
        timed('diag D2H', lambda: s._diag_host[:100].copy_(s._diag_dev[:100]))
        timed('release', lambda: plan.release(*st))
        timed('build_diagnostics', lambda: s._build_diagnostics(100))
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            s.reset()
            s.iterate()
        e1.record()
        torch.cuda.synchronize()
        acc.append(f'3 x (reset + iterate): wall {1e3 * (time.perf_counter() - t0) / 3:.2f} ms, device {e0.elapsed_time(e1) / 3:.2f} ms per solve')
        print(f'[{dtype}] rep {rep}: ' + '; '.join(acc), flush=True)


for d in sys.argv[1:] or ['f64', 'f32']:
    main(d)
