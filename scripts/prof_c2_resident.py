"""One fp64 C2 solve (1024^2, 200 updates, resident kernel) for ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pycsou_b200.func.loss import SquaredL2Loss
from pycsou_b200.func.penalty import L21Norm, tv_groups
from pycsou_b200.linop.diff import Gradient
from pycsou_b200.opt.proxalgs import PDS
from pycsou_b200.util import synth
dtype = sys.argv[1] if len(sys.argv) > 1 else 'f64'
tdt = torch.float32 if dtype == 'f32' else torch.float64
ndt = np.float32 if dtype == 'f32' else np.float64
shape, iters = (1024, 1024), 200
N = shape[0] * shape[1]
_, y = synth.tv_denoise_problem(shape, 0.1, 0)
K = Gradient(shape=shape, kind='forward', dtype=ndt.__name__); K.lipschitz_cst = K.diff_lipschitz_cst = float(np.sqrt(8.))
F = (1 / 2) * SquaredL2Loss(dim=N, data=torch.from_numpy(y.astype(ndt)).cuda())
s = PDS(dim=N, F=F, H=0.1 * L21Norm(dim=2 * N, groups=tv_groups(N)), K=K, rho=0.9, max_iter=iters - 1, min_iter=iters - 1,
        accuracy_threshold=0, verbose=None, dtype=tdt, output='torch')
for _ in range(3):
    s.reset(); s.iterate()
torch.cuda.synchronize()
print('done')
