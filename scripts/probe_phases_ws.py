"""Scratch probe: per-block phase cycle counts of the warp-specialised PDS kernel (needs a -DPDS_PHASE_TIMING build)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench

PW = ['wait_step', 'wait_g', 'primal', 'bar1', 'dual', 'bar2', 'issue_step']
CV = ['wait_x', 'hpass', 'bar', 'hcorr+issue_x', 'colpass', 'wait_b', 'store_g+arrive']
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
    for names, off in ((PW, 0), (CV, 8)):
        tot = pt[:, off + 7]
        print(f'== {dtype} {"pointwise" if off == 0 else "conv"} warps: total cycles min {tot.min():.0f} mean {tot.mean():.0f} max {tot.max():.0f}')
        m = pt[:, off:off + 7].mean(0)
        for k, n in enumerate(names):
            print(f'   {n:16s} mean {m[k]:9.0f} cyc  {100 * m[k] / tot.mean():5.1f}%')
run('f32'); run('f64')

def dist(dtype):
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
    tot = pt[:, 7]
    print(f'== {dtype}: per-block total (pointwise thread), kilo-cycles, block index 0..295 in rows of 9 (= one strip per row for the hi strips)')
    for r in range(0, nb, 9):
        print(f'   {r:3d}: ' + ' '.join(f'{v / 1e3:5.0f}' for v in tot[r:r + 9]))
    pair = tot[:148] + tot[148:]
    print('   SM-pair sums (block b + block b+148): min %.0f mean %.0f max %.0f' % (pair.min() / 1e3, pair.mean() / 1e3, pair.max() / 1e3))
    for k, n in enumerate(PW):
        print(f'   {n:12s} slowest-16-block mean {pt[np.argsort(tot)[-16:], k].mean():9.0f}   fastest-16 mean {pt[np.argsort(tot)[:16], k].mean():9.0f}')
dist('f32')
