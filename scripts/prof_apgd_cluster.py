"""One 500-update C1 solve (N = 4096, 65 taps, fp64 then fp32) for ncu: `ncu -k regex:apgd_cluster -c 2 --set full ...`."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tests.test_gpu_apgd_run import _build
for dtype in ('float64', 'float32'):
    s, _, _ = _build(4096, 65, 'l1', 'CD', dtype=dtype, max_iter=499, min_iter=499, accuracy_threshold=0)
    s.iterate()
    torch.cuda.synchronize()
    print(dtype, s.iter)
