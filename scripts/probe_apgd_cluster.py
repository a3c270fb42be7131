"""Scratch probe: the cluster APGD kernel against the single-block one (PCS_APGD_CLUSTER=0): iterates bitwise, update counts,
diagnostics, and the time of 500 updates."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tests.test_gpu_apgd_run import _build

def solve(flag, *a, **kw):
    os.environ['PCS_APGD_CLUSTER'] = flag
    s, _, _ = _build(*a, **kw)
    est, _, diag = s.iterate()
    return s, est, diag

ok = True
for n, nh, g, acc in [(4096, 65, 'l1', 'CD'), (4096, 65, 'l1', 'BT'), (5000, 64, 'l1', 'CD'), (8192, 129, 'l1', 'CD'), (2048, 33, 'nonneg', None),
                      (2049, 9, 'none', 'BT'), (8191, 255, 'l1', 'CD'), (3000, 1, 'l1', 'CD'), (6000, 200, 'nonneg', 'CD')]:
    for dtype in ('float64', 'float32'):
        kw = dict(max_iter=300, min_iter=5, accuracy_threshold=3e-3, dtype=dtype)
        s0, e0, d0 = solve('0', n, nh, g, acc, **kw)
        s1, e1, d1 = solve('1', n, nh, g, acc, **kw)
        err = max(np.linalg.norm(e0[k].astype(float) - e1[k].astype(float)) / np.linalg.norm(e0[k].astype(float)) for k in ('iterand', 'past_aux'))
        same = err < (1e-12 if dtype == 'float64' else 1e-4) and s0.iter == s1.iter and e0['past_t'] == e1['past_t']
        r0, r1 = np.asarray(d0['Relative Improvement'], float)[1:], np.asarray(d1['Relative Improvement'], float)[1:]
        close = len(r0) == len(r1) and np.allclose(r0, r1, rtol=1e-9 if dtype == 'float64' else 1e-2)
        ok &= same and close
        print(f'n={n} nh={nh} {g} {acc} {dtype}: updates {s0.iter}/{s1.iter} rel diff {err:.1e} ok {same} diag close {close}')
print('ALL OK' if ok else 'MISMATCH')
for flag, nt in (("0", "-"), ("1", "128")):
    os.environ['PCS_APGD_CLUSTER'] = flag
    for dtype in ('float64', 'float32'):
        kw = dict(max_iter=499, min_iter=499, accuracy_threshold=0)
        s, _, _ = _build(4096, 65, 'l1', 'CD', dtype=dtype, **kw)
        s.iterate(); s.reset(); s.iterate()
        plan = s._plan
        x, aux, t = s._init_state()
        diag = torch.zeros(502, 4, dtype=torch.float64, device='cuda')
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(5):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); plan.run(x, aux, 1, 0, diag); b.record(); torch.cuda.synchronize()
            best = min(best, a.elapsed_time(b))
        t0 = time.perf_counter(); s.reset(); s.iterate(); torch.cuda.synchronize(); t1 = time.perf_counter()
        print(f'cluster={flag} nt={nt} {dtype}: kernel {best:.3f} ms for 500 updates ({2 * best:.2f} us/update); iterate() {1e3 * (t1 - t0):.2f} ms')
# stepping one update per launch (the iterates() generator): launch + set-up cost of the two kernels
for flag in ('0', '1'):
    os.environ['PCS_APGD_CLUSTER'] = flag
    kw = dict(max_iter=499, min_iter=499, accuracy_threshold=0)
    s, _, _ = _build(4096, 65, 'l1', 'CD', **kw)
    s.iterate()
    plan = s._plan
    x, aux, t = s._init_state()
    diag = torch.zeros(4, 4, dtype=torch.float64, device='cuda')
    for _ in range(3): plan.run(x, aux, 1.0, 0, diag, single=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for k in range(200): x, aux, tt, it = plan.run(x, aux, 1.0, k, diag, single=True)
    torch.cuda.synchronize()
    print(f'cluster={flag}: one update per launch (with the read-back of n_iter): {1e6 * (time.perf_counter() - t0) / 200:.1f} us per update')
