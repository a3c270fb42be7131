"""Stand-alone Gradient kernels and the two-kernel Condat-Vu update for centred differences at 4096^2: CUDA-event times,
effective GB/s against the algorithmic bytes (grad fwd / adj: 3 N s; PDS update: 7 N s; the two kernels move 11 N s).

    python scripts/probe_grad.py [f32|f64|both] [size] [reps]
"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from pycsou_b200 import _kernels as K


def timeit(fn, n=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return 1e3 * e0.elapsed_time(e1) / n


def main(dtype, size, reps):
    from pycsou_b200.func.loss import SquaredL2Loss
    from pycsou_b200.func.penalty import L21Norm, tv_groups
    from pycsou_b200.linop.diff import Gradient
    from pycsou_b200.opt.proxalgs import PDS
    from pycsou_b200.util import synth
    tdt = torch.float32 if dtype == 'f32' else torch.float64
    isz = 4 if dtype == 'f32' else 8
    H = W = size
    N = H * W
    x = torch.randn(N, dtype=tdt, device='cuda')
    z = torch.randn(2 * N, dtype=tdt, device='cuda')
    o2, o1 = torch.empty(2 * N, dtype=tdt, device='cuda'), torch.empty(N, dtype=tdt, device='cuda')
    for kind in ('forward', 'centered'):
        us = timeit(lambda: K.grad2d(o2, x, H, W, kind, True, 1., 1.), reps)
        print(f'[{dtype}] grad2d_fwd {kind:9s}: {us:7.1f} us  {3 * N * isz / us / 1e3:7.1f} GB/s (3 N s)')
        us = timeit(lambda: K.grad2d(o1, z, H, W, kind, True, 1., 1., adjoint=True), reps)
        print(f'[{dtype}] grad2d_adj {kind:9s}: {us:7.1f} us  {3 * N * isz / us / 1e3:7.1f} GB/s (3 N s)')
    _, y = synth.tv_denoise_problem((H, W), 0.1, 0)
    yd = torch.from_numpy(y.astype(np.float32 if dtype == 'f32' else np.float64)).cuda()
    for kind, fused in (('centered', True), ('centered', False), ('forward', True)):
        Kop = Gradient(shape=(H, W), kind=kind, dtype='float32' if dtype == 'f32' else 'float64')
        Kop.lipschitz_cst = Kop.diff_lipschitz_cst = float(np.sqrt(8.))
        s = PDS(dim=N, F=(1 / 2) * SquaredL2Loss(dim=N, data=yd), H=0.1 * L21Norm(2 * N, tv_groups(N)), K=Kop, rho=0.9,
                max_iter=9, min_iter=9, accuracy_threshold=0, verbose=None, dtype=tdt, output='torch', fused=fused)
        s.iterate()
        st = [s._init_state()]
        drow = torch.zeros(4, dtype=torch.float64, device='cuda')

        def step():
            st[0] = s._plan.step(*st[0], drow) if s._plan is not None else s._generic_step(*st[0], drow)
        us = timeit(step, reps)
        name = type(s._plan).__name__ if s._plan is not None else 'op-by-op'
        print(f'[{dtype}] PDS TV denoise {kind:9s} {name:12s}: {us:7.1f} us/update  {7 * N * isz / us / 1e3:7.1f} GB/s (7 N s)')


which = sys.argv[1] if len(sys.argv) > 1 else 'both'
size = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 50
for d in (('f32', 'f64') if which == 'both' else (which,)):
    main(d, size, reps)
