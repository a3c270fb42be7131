"""Host-side logic of the slab decomposition (pycsou_b200/opt/slab.py) on CPU: world_size 2 and 3 over gloo.

The layout arithmetic and ``gather_halo`` are the product code; the per-slab update is done by the numpy oracle on the
halo-extended slab (the fused CUDA kernel needs a GPU).  The test shows that halos of 14 rows above / 15 below for x
and y and 1 row for z reproduce the undecomposed Condat-Vu iteration (reference opt/proxalgs.py:343-355) on the rows
a rank owns -- i.e. that the halo depths, the row ranges and the exchange pattern the kernel relies on are right.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import np_oracle as O
from pycsou_b200.opt.slab import HALO_DN, HALO_UP, SlabComm, SlabLayout, gather_halo
from pycsou_b200.util import synth

LAM, ITERS = 0.05, 6


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _problem(Hg, W):
    _, psf, y = synth.tv_deconv_problem((Hg, W), None, 0.02, 3)
    return psf, y.reshape(Hg, W)


def _global_solution(Hg, W, psf, y, iters):
    N = Hg * W
    F = O.HalfSquaredL2Loss(y.reshape(-1), O.Convolve2D(N, psf, (Hg, W)), G_lipschitz=1.0)
    tau, sigma, rho = O.pds_step_sizes(1.0, np.sqrt(8.))
    x, z = np.zeros(N), np.zeros(2 * N)
    H, K, G = LAM * O.TVL21Norm(2 * N, N), O.Gradient((Hg, W), kind='forward'), O.NonNegativeOrthant(N)
    for _ in range(iters):
        x, z = O.pds_update(x, z, F, G, H, K, tau, sigma, 0.9)
    return x.reshape(Hg, W), z.reshape(2, Hg, W)


def _worker(rank, world, port, Hg, W, out):
    os.environ['MASTER_ADDR'], os.environ['MASTER_PORT'] = '127.0.0.1', str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        comm = SlabComm()
        lay = comm.layout(Hg, W)
        psf, y = _problem(Hg, W)
        y_own = torch.from_numpy(y[lay.r0:lay.r1].copy())
        y_ext = gather_halo(y_own, HALO_UP, HALO_DN, rank, world).numpy()
        assert y_ext.shape[0] == lay.H_loc
        assert np.array_equal(y_ext, y[lay.grow0:lay.grow0 + lay.H_loc])          # the halo rows are the right ones
        n_ext = lay.H_loc
        Ne = n_ext * W
        F = O.HalfSquaredL2Loss(y_ext.reshape(-1), O.Convolve2D(Ne, psf, (n_ext, W)), G_lipschitz=1.0)
        H, K, G = LAM * O.TVL21Norm(2 * Ne, Ne), O.Gradient((n_ext, W), kind='forward'), O.NonNegativeOrthant(Ne)
        tau, sigma, rho = O.pds_step_sizes(1.0, np.sqrt(8.))
        x_own = torch.zeros(lay.n_own, W, dtype=torch.float64)
        z_own = torch.zeros(2, lay.n_own, W, dtype=torch.float64)
        diag = torch.zeros(ITERS, 4, dtype=torch.float64)
        for it in range(ITERS):
            x_ext = gather_halo(x_own, HALO_UP, HALO_DN, rank, world).numpy()
            z_ext = np.zeros((2, n_ext, W))
            for p in (0, 1):
                zz = gather_halo(z_own[p], 1, 1, rank, world).numpy()              # z needs one row either side only
                lo = lay.row_begin - (1 if rank > 0 else 0)
                z_ext[p, lo:lo + zz.shape[0]] = zz
            xn, zn = O.pds_update(x_ext.reshape(-1), z_ext.reshape(-1), F, G, H, K, tau, sigma, 0.9)
            xn = torch.from_numpy(xn.reshape(n_ext, W)[lay.row_begin:lay.row_end].copy())
            zn = torch.from_numpy(zn.reshape(2, n_ext, W)[:, lay.row_begin:lay.row_end].copy())
            diag[it] = torch.tensor([float(((xn - x_own) ** 2).sum()), float((x_own ** 2).sum()),
                                     float(((zn - z_own) ** 2).sum()), float((z_own ** 2).sum())], dtype=torch.float64)
            x_own, z_own = xn, zn
        comm.all_reduce_sum(diag)
        np.savez(os.path.join(out, f'rank{rank}.npz'), x=x_own.numpy(), z=z_own.numpy(), diag=diag.numpy(),
                 rows=np.array([lay.r0, lay.r1]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('world,Hg,W', [(2, 70, 40), (3, 101, 36)])
def test_slab_iteration_matches_global(world, Hg, W, tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(world, port, Hg, W, str(tmp_path)), nprocs=world, join=True)
    psf, y = _problem(Hg, W)
    xg, zg = _global_solution(Hg, W, psf, y, ITERS)
    xg5, zg5 = _global_solution(Hg, W, psf, y, ITERS - 1)
    covered = 0
    for r in range(world):
        d = np.load(os.path.join(str(tmp_path), f'rank{r}.npz'))
        r0, r1 = d['rows']
        assert r0 == covered
        covered = r1
        assert np.linalg.norm(d['x'] - xg[r0:r1]) <= 1e-12 * np.linalg.norm(xg)
        assert np.linalg.norm(d['z'] - zg[:, r0:r1]) <= 1e-12 * np.linalg.norm(zg)
        # all-reduced norms of the last update = the undecomposed ones (update_diagnostics, opt/proxalgs.py:366-394)
        ref = np.array([((xg - xg5) ** 2).sum(), (xg5 ** 2).sum(), ((zg - zg5) ** 2).sum(), (zg5 ** 2).sum()])
        assert np.allclose(d['diag'][-1], ref, rtol=1e-10)
    assert covered == Hg


def test_layout_arithmetic():
    for Hg, world in [(4096, 8), (1000, 3), (64, 2), (32768, 8)]:
        lays = [SlabLayout(Hg, 128, world, r) for r in range(world)]
        assert lays[0].r0 == 0 and lays[-1].r1 == Hg
        for a, b in zip(lays, lays[1:]):
            assert a.r1 == b.r0 and abs(a.n_own - b.n_own) <= 1
            # what a rank pushes down is exactly the upper halo of the rank below, and vice versa
            assert a.push_counts()[1] == b.ht == HALO_UP and b.push_counts()[0] == a.hb == HALO_DN
            # the same image row, seen from both ranks
            assert a.grow0 + a.row_end == b.grow0 + b.row_begin
        assert lays[0].ht == 0 and lays[-1].hb == 0
        for l in lays:
            assert l.H_loc == l.ht + l.n_own + l.hb and l.grow0 + l.row_begin == l.r0
            assert l.neighbour(-1) is None or l.neighbour(-1).r1 == l.r0
    with pytest.raises(ValueError):
        SlabLayout(40, 16, 2, 0)
