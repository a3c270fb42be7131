"""Scratch probe: phase cycle counts of the fused kernel in TV-denoise mode (needs a -DPDS_PHASE_TIMING build)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pycsou_b200.func.loss import SquaredL2Loss
from pycsou_b200.func.penalty import L21Norm, NonNegativeOrthant, tv_groups
from pycsou_b200.linop.diff import Gradient
from pycsou_b200.opt.proxalgs import PDS
NAMES = ['wait_x', 'hpass', 'sync1(+hcorr)', 'colpass', 'wait_step', 'primal', 'sync2', 'dual', 'sync3']
def run(dtype, size=4096):
    tdt = torch.float32 if dtype == 'f32' else torch.float64
    N = size * size
    y = torch.rand(N, dtype=tdt, device='cuda')
    K = Gradient(shape=(size, size), kind='forward'); K.lipschitz_cst = K.diff_lipschitz_cst = float(np.sqrt(8.))
    s = PDS(dim=N, F=(1 / 2) * SquaredL2Loss(dim=N, data=y), G=NonNegativeOrthant(dim=N), H=0.1 * L21Norm(dim=2 * N, groups=tv_groups(N)),
            K=K, beta=1.0, max_iter=9, min_iter=9, accuracy_threshold=0, verbose=None, dtype=tdt, output='torch')
    s.iterate()
    torch.cuda.synchronize()
    ws = s._plan.ws.cpu().numpy()
    pt = ws[8192:8192 + 16 * 296].reshape(296, 16)
    pt = pt[pt[:, 15] > 0]
    tot = pt[:, 15]
    print(f'== {dtype} denoise: {len(pt)} blocks, total cycles min {tot.min():.0f} mean {tot.mean():.0f} max {tot.max():.0f}')
    m = pt[:, :9].mean(0)
    for k, n in enumerate(NAMES):
        print(f'   {n:14s} mean {m[k]:9.0f} cyc  {100 * m[k] / tot.mean():5.1f}%')
size = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
run('f32', size); run('f64', size)
