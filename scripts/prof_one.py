"""Profile target: a few fused PDS updates of the 4096^2 deconvolution problem (run under ncu -k regex:pds_tv)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
dtype = sys.argv[1] if len(sys.argv) > 1 else 'f32'
tdt = torch.float32 if dtype == 'f32' else torch.float64
psf, y64 = bench.make_problem(4096, 'sep', 0)
y = torch.from_numpy(y64.astype(np.float32 if dtype == 'f32' else np.float64)).cuda()
s = bench.make_solver(y, psf, 4096, tdt, 12, 'torch')
s.iterate()
torch.cuda.synchronize()
print('done', type(s._plan).__name__)
