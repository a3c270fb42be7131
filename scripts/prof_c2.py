"""Profile target: 12 fused updates of the 1024^2 TV-denoising problem (C2), fp64 then fp32."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pycsou_b200.func.loss import SquaredL2Loss
from pycsou_b200.func.penalty import L21Norm, tv_groups
from pycsou_b200.linop.diff import Gradient
from pycsou_b200.opt.proxalgs import PDS
from pycsou_b200.util import synth
size = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
for tdt in (torch.float64, torch.float32):
    N = size * size
    _, y = synth.tv_denoise_problem((size, size), 0.1, 0)
    y = torch.from_numpy(np.asarray(y).reshape(-1)).to('cuda', tdt)
    K = Gradient(shape=(size, size), kind='forward'); K.lipschitz_cst = K.diff_lipschitz_cst = float(np.sqrt(8.))
    s = PDS(dim=N, F=(1 / 2) * SquaredL2Loss(dim=N, data=y), G=None, H=0.1 * L21Norm(dim=2 * N, groups=tv_groups(N)), K=K,
            beta=1.0, rho=0.9, max_iter=11, min_iter=11, accuracy_threshold=0, verbose=None, dtype=tdt, output='torch')
    s.iterate()
torch.cuda.synchronize()
