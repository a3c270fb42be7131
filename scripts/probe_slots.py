"""Scratch probe: which of the two CTAs of an SM runs faster (needs a -DPDS_PHASE_TIMING build, even split)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench

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
    info = ws[8192 + 16 * nb:8192 + 16 * nb + 4 * nb].reshape(nb, 4)     # per blockIdx: smid, slot, start clock, bid
    tot_by_bid = pt[:, 7]
    smid, slot, start, bid = info[:, 0].astype(int), info[:, 1].astype(int), info[:, 2], info[:, 3].astype(int)
    tot = tot_by_bid[bid]
    n_first_fast = n_slot0_fast = n_lowidx_fast = npairs = 0
    gaps = []
    for sm in np.unique(smid):
        m = np.where(smid == sm)[0]
        if len(m) != 2:
            print('   SM', sm, 'has', len(m), 'CTAs'); continue
        a, b = m                                    # a < b in blockIdx
        fast = a if tot[a] < tot[b] else b
        other = b if fast == a else a
        npairs += 1
        n_lowidx_fast += fast == a
        n_slot0_fast += slot[fast] == 0
        n_first_fast += start[fast] < start[other]
        gaps.append(abs(start[a] - start[b]))
    print(f'== {dtype}: {npairs} SM pairs; faster CTA is the lower blockIdx in {n_lowidx_fast}, is slot 0 in {n_slot0_fast}, '
          f'started first (clock64) in {n_first_fast}; start gap median {np.median(gaps):.0f} cycles, max {np.max(gaps):.0f}')
    print('   block totals: fast-of-pair mean %.0f, slow-of-pair mean %.0f' % (
        np.mean([min(tot[np.where(smid == sm)[0]]) for sm in np.unique(smid) if (smid == sm).sum() == 2]),
        np.mean([max(tot[np.where(smid == sm)[0]]) for sm in np.unique(smid) if (smid == sm).sum() == 2])))
run('f32'); run('f64')
