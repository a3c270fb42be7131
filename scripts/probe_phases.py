"""Scratch probe: per-block phase cycle counts of the fused PDS kernel (needs a -DPDS_PHASE_TIMING build)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench

NAMES = ['wait_x', 'hpass', 'sync1(+hcorr)', 'colpass', 'wait_step', 'primal', 'sync2', 'dual', 'sync3', 'issue_step', 'issue_x']
def run(dtype):
    tdt = torch.float32 if dtype == 'f32' else torch.float64
    psf, y64 = bench.make_problem(4096, 'sep', 0)
    y = torch.from_numpy(y64.astype(np.float32 if dtype == 'f32' else np.float64)).cuda()
    s = bench.make_solver(y, psf, 4096, tdt, 20, 'torch')
    s.iterate()
    plan = s._plan
    x, z = s._state
    drow = torch.zeros(4, dtype=torch.float64, device='cuda')
    for _ in range(5):
        x, z = plan.step(x, z, drow)
    torch.cuda.synchronize()
    ws = plan.ws.cpu().numpy()
    nb = 296
    pt = ws[8192:8192 + 16 * nb].reshape(nb, 16)
    tot = pt[:, 15]
    print(f'== {dtype}: block total cycles: min {tot.min():.0f} mean {tot.mean():.0f} max {tot.max():.0f}  (max/mean {tot.max()/tot.mean():.3f})')
    order = np.argsort(tot)
    print('   slowest blocks:', order[-8:], tot[order[-8:]].astype(int))
    print('   fastest blocks:', order[:8], tot[order[:8]].astype(int))
    m = pt[:, :11].mean(0)
    for k, n in enumerate(NAMES):
        print(f'   {n:14s} mean {m[k]:9.0f} cyc  {100 * m[k] / tot.mean():5.1f}%   (max over blocks {pt[:, k].max():9.0f})')
    print(f'   unaccounted (prologue, finish) {100 * (1 - m.sum() / tot.mean()):.1f}%')
run('f32'); run('f64')
