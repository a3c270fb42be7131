"""Scratch probe: cycles per phase of the cluster APGD kernel (library built with PCS_NVCC_EXTRA=-DAPC_PHASE_TIMING into
another path and loaded through PCS_LIB_PATH; the kernel prints from two blocks)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tests.test_gpu_apgd_run import _build
for dtype in ('float64', 'float32'):
    s, _, _ = _build(4096, 65, 'l1', 'CD', dtype=dtype, max_iter=499, min_iter=499, accuracy_threshold=0)
    print(dtype, flush=True)
    s.iterate(); torch.cuda.synchronize()
