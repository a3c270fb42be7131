"""Stand-alone kernels at 4096^2 (N = 16.7 M): CUDA-event time and GB/s against the algorithmic bytes of each.
    python scripts/probe_ops.py [f32|f64|both] [reps]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from pycsou_b200 import _kernels as K


def timeit(fn, n):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return 1e3 * e0.elapsed_time(e1) / n


def main(dtype, reps):
    tdt = torch.float32 if dtype == 'f32' else torch.float64
    s = 4 if dtype == 'f32' else 8
    H = W = 4096
    N = H * W
    x, y, w = (torch.randn(N, dtype=tdt, device='cuda') for _ in range(3))
    z, z2 = torch.randn(2 * N, dtype=tdt, device='cuda'), torch.empty(2 * N, dtype=tdt, device='cuda')
    o = torch.empty(N, dtype=tdt, device='cuda')
    mask = torch.rand(N, device='cuda') < 0.3
    idx = torch.nonzero(mask).reshape(-1)
    m = idx.numel()
    ym = torch.empty(m, dtype=tdt, device='cuda')
    rows = [
        ('lincomb2 (axpby)', lambda: K.lincomb2(o, 1.0, x, -0.3, y), 3 * N * s),
        ('lincomb3', lambda: K.lincomb3(o, 1.0, x, -0.3, y, 0.2, w), 4 * N * s),
        ('prox_l1', lambda: K.prox_l1(o, x, 0.3), 2 * N * s),
        ('proj_nonneg', lambda: K.proj_nonneg(o, x), 2 * N * s),
        ('prox_l21 planes (2N)', lambda: K.prox_l21(z2, z, 0.3, N, 2), 4 * N * s),
        ('sumsq', lambda: K.sumsq(x), N * s),
        ('diffsumsq', lambda: K.diffsumsq(x, y), 2 * N * s),
        ('sumgroupnorm (2N)', lambda: K.sumgroupnorm(z, N, 2), 2 * N * s),
        ('firstderiv axis0 centered', lambda: K.firstderiv(o, x, 1, H, W, 'centered', True, 1.0), 2 * N * s),
        ('firstderiv axis1 centered', lambda: K.firstderiv(o, x, H, W, 1, 'centered', True, 1.0), 2 * N * s),
        ('firstderiv axis0 adj', lambda: K.firstderiv(o, x, 1, H, W, 'centered', True, 1.0, adjoint=True), 2 * N * s),
        ('secondderiv axis1', lambda: K.secondderiv(o, x, H, W, 1, True, 1.0), 2 * N * s),
        ('laplacian2d', lambda: K.laplacian2d(o, x, H, W, 1.0, 1.0, 1.0, 1.0, True), 2 * N * s),
        ('laplacian2d adj', lambda: K.laplacian2d(o, x, H, W, 1.0, 1.0, 1.0, 1.0, True, adjoint=True), 2 * N * s),
        ('take (Masking fwd, 30 %)', lambda: K.take(ym, x, idx, 1, N, 1), (m + m) * s + 8 * m),
        ('put (Masking adj, 30 %)', lambda: K.put(o, ym, idx, 1, N, 1), (N + m) * s + 8 * m),
    ]
    for name, fn, nbytes in rows:
        us = timeit(fn, reps)
        print(f'[{dtype}] {name:28s}: {us:7.1f} us  {nbytes / us / 1e3:7.1f} GB/s')


which = sys.argv[1] if len(sys.argv) > 1 else 'both'
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 30
for d in (('f32', 'f64') if which == 'both' else (which,)):
    main(d, reps)
