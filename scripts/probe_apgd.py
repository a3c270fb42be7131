"""Scratch probe: time of the one-launch APGD solve (kernel only and through iterate())."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tests.test_gpu_apgd_run import _build
for dtype in ('float64', 'float32'):
    kw = dict(max_iter=499, min_iter=499, accuracy_threshold=0)
    s, _, _ = _build(4096, 65, 'l1', 'CD', dtype=dtype, **kw)
    s.iterate(); s.reset(); s.iterate()
    plan = s._plan
    x, aux, t = s._init_state()
    diag = torch.zeros(502, 4, dtype=torch.float64, device='cuda')
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); plan.run(x, aux, 1, 0, diag); e1.record(); torch.cuda.synchronize()
    t0 = time.perf_counter(); s.reset(); s.iterate(); torch.cuda.synchronize(); t1 = time.perf_counter()
    print(f'{dtype}: kernel {e0.elapsed_time(e1):.3f} ms for 500 updates ({2 * e0.elapsed_time(e1):.2f} us/update); iterate() {1e3 * (t1 - t0):.2f} ms')
