"""Scratch probe: per-phase cycles of the on-chip resident kernel (needs a -DPDS_RES_TIMING build), 1024^2, 200 updates."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pycsou_b200.func.loss import SquaredL2Loss
from pycsou_b200.func.penalty import L21Norm, tv_groups
from pycsou_b200.linop.diff import Gradient
from pycsou_b200.opt.proxalgs import PDS
from pycsou_b200.util import synth
NAMES = ['interior rows (phase 1)', 'counter wait', 'halo loads + rows 0, Rc', 'barrier 1', 'phase 2', 'warp sums', 'barrier 2', '(loop tail)']
for dtype in sys.argv[1:] or ['f64', 'f32']:
    tdt = torch.float32 if dtype == 'f32' else torch.float64
    ndt = np.float32 if dtype == 'f32' else np.float64
    shape, iters = (1024, 1024), 200
    N = shape[0] * shape[1]
    _, y = synth.tv_denoise_problem(shape, 0.1, 0)
    K = Gradient(shape=shape, kind='forward', dtype=ndt.__name__); K.lipschitz_cst = K.diff_lipschitz_cst = float(np.sqrt(8.))
    F = (1 / 2) * SquaredL2Loss(dim=N, data=torch.from_numpy(y.astype(ndt)).cuda())
    s = PDS(dim=N, F=F, H=0.1 * L21Norm(dim=2 * N, groups=tv_groups(N)), K=K, rho=0.9, max_iter=iters - 1, min_iter=iters - 1,
            accuracy_threshold=0, verbose=None, dtype=tdt, output='torch')
    s.iterate(); s.reset(); s.iterate()
    torch.cuda.synchronize()
    ws = s._plan.ws.view(torch.float64).cpu().numpy() if s._plan.ws.dtype != torch.float64 else s._plan.ws.cpu().numpy()
    rt = ws[240000:240000 + 8 * 147].reshape(147, 8) / iters
    print(f'== {dtype}: cycles per update, mean over blocks (min .. max); total {rt.sum(1).mean():.0f} cycles = {rt.sum(1).mean() / 1.9e3:.2f} us at 1.9 GHz')
    for k, n in enumerate(NAMES):
        print(f'   {n:28s} {rt[:, k].mean():8.0f}  ({rt[:, k].min():.0f} .. {rt[:, k].max():.0f})')
