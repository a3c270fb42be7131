"""Scratch probe: single-block against cluster APGD kernel below n = 2048 (PCS_APGD_CLUSTER_MIN_N=256)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tests.test_gpu_apgd_run import _build
from pycsou_b200.opt._fused import apgd_last_launch
os.environ['PCS_APGD_CLUSTER_MIN_N'] = '256'
for n, nh in ((512, 33), (768, 33), (1024, 33), (1024, 65), (1536, 65), (2047, 65)):
    for dtype in ('float64', 'float32'):
        res = {}
        for flag in ('0', '1'):
            os.environ['PCS_APGD_CLUSTER'] = flag
            kw = dict(max_iter=499, min_iter=499, accuracy_threshold=0)
            s, _, _ = _build(n, nh, 'l1', 'CD', dtype=dtype, **kw)
            est, _, _ = s.iterate()
            plan = s._plan
            x, aux, t = s._init_state()
            diag = torch.zeros(502, 4, dtype=torch.float64, device='cuda')
            torch.cuda.synchronize()
            best = 1e9
            for _ in range(5):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(); plan.run(x, aux, 1, 0, diag); b.record(); torch.cuda.synchronize()
                best = min(best, a.elapsed_time(b))
            res[flag] = (best, est['iterand'].astype(float), apgd_last_launch()['kernel'])
        err = np.linalg.norm(res['0'][1] - res['1'][1]) / np.linalg.norm(res['0'][1])
        print(f'n={n} nh={nh} {dtype}: single block {2 * res["0"][0]:.2f} us/update, cluster {2 * res["1"][0]:.2f} us/update ({res["1"][2]}), rel diff {err:.1e}')
