"""Kernel-only timing of the fused Condat-Vu update: CUDA events around 200 back-to-back plan.step calls (best of 3).

    python scripts/probe_kernel.py [f32|f64|both] [size]
"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench


def run(dtype, size):
    tdt = torch.float32 if dtype == 'f32' else torch.float64
    psf, y64 = bench.make_problem(size, 'sep', 0)
    y = torch.from_numpy(y64.astype(np.float32 if dtype == 'f32' else np.float64)).cuda()
    s = bench.make_solver(y, psf, size, tdt, 10, 'torch')
    s.iterate()
    plan = s._plan
    x, z = s._init_state()
    drow = torch.zeros(4, dtype=torch.float64, device='cuda')
    for _ in range(20):
        x, z = plan.step(x, z, drow)
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 200
        e0.record()
        for _ in range(n):
            x, z = plan.step(x, z, drow)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, 1e3 * e0.elapsed_time(e1) / n)
    from pycsou_b200.opt._fused import last_launch
    L = last_launch()
    print(f'{dtype}: {best:.1f} us/iter ({L["kernel"]}, {L["blocks"]} blocks) checksum {float(x.double().sum()):.6f}')


which = sys.argv[1] if len(sys.argv) > 1 else 'both'
size = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
for d in (('f32', 'f64') if which == 'both' else (which,)):
    run(d, size)
