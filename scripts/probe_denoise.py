"""Scratch probe: fused kernel without the data-term convolution (TV denoise, HL = 0 variant) at 4096^2."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pycsou_b200.func.loss import SquaredL2Loss
from pycsou_b200.func.penalty import L21Norm, NonNegativeOrthant, tv_groups
from pycsou_b200.linop.diff import Gradient
from pycsou_b200.opt.proxalgs import PDS

def run(dtype, size=4096):
    tdt = torch.float32 if dtype == 'f32' else torch.float64
    N = size * size
    y = torch.rand(N, dtype=tdt, device='cuda')
    K = Gradient(shape=(size, size), kind='forward'); K.lipschitz_cst = K.diff_lipschitz_cst = float(np.sqrt(8.))
    s = PDS(dim=N, F=(1 / 2) * SquaredL2Loss(dim=N, data=y), G=NonNegativeOrthant(dim=N), H=0.1 * L21Norm(dim=2 * N, groups=tv_groups(N)),
            K=K, beta=1.0, max_iter=19, min_iter=19, accuracy_threshold=0, verbose=None, dtype=tdt, output='torch')
    s.iterate()
    plan = s._plan
    x, z = s._init_state()
    drow = torch.zeros(4, dtype=torch.float64, device='cuda')
    for _ in range(10): x, z = plan.step(x, z, drow)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 200
    e0.record()
    for _ in range(n): x, z = plan.step(x, z, drow)
    e1.record(); torch.cuda.synchronize()
    us = 1e3 * e0.elapsed_time(e1) / n
    isz = 4 if dtype == 'f32' else 8
    print(f'{dtype} denoise {size}^2: {us:.1f} us/iter, {7 * N * isz / us / 1e3:.0f} GB/s = {7 * N * isz / us / 1e3 / 6459 * 100:.1f}% of roofline')
run('f32'); run('f64')
