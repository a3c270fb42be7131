"""Scratch probe: C2 (1024^2 TV denoising, 200 updates) and C3 through PDS.iterate(), with one library call per update
(PCS_PDS_BATCH=0) and with the updates up to min_iter enqueued by one call."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pycsou_b200.func.loss import SquaredL2Loss
from pycsou_b200.func.penalty import L21Norm, tv_groups
from pycsou_b200.linop.diff import Gradient
from pycsou_b200.opt.proxalgs import PDS
from pycsou_b200.util import synth

def run(dtype, size=1024, iters=200):
    tdt = torch.float32 if dtype == 'f32' else torch.float64
    N = size * size
    _, y = synth.tv_denoise_problem((size, size), 0.1, 0)
    y = torch.from_numpy(np.asarray(y).reshape(-1)).to('cuda', tdt)
    K = Gradient(shape=(size, size), kind='forward'); K.lipschitz_cst = K.diff_lipschitz_cst = float(np.sqrt(8.))
    s = PDS(dim=N, F=(1 / 2) * SquaredL2Loss(dim=N, data=y), G=None, H=0.1 * L21Norm(dim=2 * N, groups=tv_groups(N)), K=K,
            beta=1.0, rho=0.9, max_iter=iters - 1, min_iter=iters - 1, accuracy_threshold=0, verbose=None, dtype=tdt, output='torch')
    out = {}
    for batch in ('0', '1'):
        os.environ['PCS_PDS_BATCH'] = batch
        s.reset(); est, _, _ = s.iterate(); s.reset()
        torch.cuda.synchronize(); t0 = time.perf_counter()
        est, _, d = s.iterate()
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
        out[batch] = est['primal_variable'].clone()
        isz = 4 if dtype == 'f32' else 8
        print(f'{dtype} denoise {size}^2 x{iters}: batch={batch}: {1e6 * dt / iters:.1f} us/update wall = {iters / dt:.0f} it/s, '
              f'{7 * N * isz * iters / dt / 1e9:.0f} GB/s effective (L2-resident)' )
    assert torch.equal(out['0'], out['1'])
run('f64'); run('f32'); run('f32', 4096, 50)
