"""GPU parity of the slab-decomposed fused Condat-Vu iteration (csrc/pcs_pds.cu + opt/slab.py) against the numpy oracle.

Ranks are separate processes (tests/slab_worker.py under torch.distributed.run).  On a box with >= 2 GPUs they run one
per GPU over NCCL, free-running: the halo rows travel through NVLink and the kernels of neighbouring ranks hand over
through the release/acquire flag words.  On a single-GPU box they share cuda:0 (gloo control plane, CUDA IPC between
the processes) in LOCKSTEP (SlabComm.lockstep: every update is completed on all ranks before the next is launched), so
that no kernel ever spins on a flag another time-sliced process has yet to write; the kernel path is the same: peer
stores into the neighbour's halo rows, flag words, all-reduced stop-test norms.
"""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from oracle import np_oracle as O
from pycsou_b200.util import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _run_ranks(world, out, extra):
    import torch
    backend = 'nccl' if torch.cuda.device_count() >= world else 'gloo'
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', f'--nproc-per-node={world}',
           '--master-addr', '127.0.0.1', '--master-port', str(_free_port()),
           os.path.join(ROOT, 'tests', 'slab_worker.py'), '--out', str(out), '--backend', backend] + extra
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return [np.load(os.path.join(str(out), f'rank{k}.npz')) for k in range(world)]


def _oracle(Hg, W, iters, data, threshold=0.0):
    N = Hg * W
    _, psf, y = synth.tv_deconv_problem((Hg, W), None, 0.02, 5)
    G = O.Convolve2D(N, psf, (Hg, W)) if data == 'conv' else None
    F = O.HalfSquaredL2Loss(y, G, G_lipschitz=1.0)
    kw = dict(max_iter=iters - 1, min_iter=iters - 1, accuracy_threshold=0) if threshold == 0 else \
        dict(max_iter=iters - 1, min_iter=2, accuracy_threshold=threshold)
    x, z, diag, n = O.pds_iterate(N, F=F, G=O.NonNegativeOrthant(N), H=0.05 * O.TVL21Norm(2 * N, N),
                                  K=O.Gradient((Hg, W), kind='forward'), K_lipschitz=np.sqrt(8.), beta=1.0, **kw)
    return x.reshape(Hg, W), z.reshape(2, Hg, W), diag, n


def _check(parts, xo, zo, tol):
    covered = 0
    for d in parts:
        r0, r1 = d['rows']
        assert r0 == covered
        covered = r1
        W = xo.shape[1]
        assert np.linalg.norm(d['x'].reshape(-1, W) - xo[r0:r1]) <= tol * np.linalg.norm(xo)
        assert np.linalg.norm(d['z'].reshape(2, -1, W) - zo[:, r0:r1]) <= tol * np.linalg.norm(zo)
    assert covered == xo.shape[0]


@pytest.mark.parametrize('world,rows,cols,data', [(1, 96, 200, 'conv'), (2, 192, 200, 'conv'), (3, 200, 136, 'conv'),
                                                  (2, 128, 260, 'identity')])
def test_slab_fp64_matches_oracle(world, rows, cols, data, tmp_path):
    iters = 12
    parts = _run_ranks(world, tmp_path, ['--rows', str(rows), '--cols', str(cols), '--iters', str(iters), '--data', data])
    xo, zo, diag, _ = _oracle(rows, cols, iters, data)
    _check(parts, xo, zo, 1e-10)
    for d in parts:
        # all-reduced relative improvements = the undecomposed ones; the second solve on the same plan repeats the first
        assert np.allclose(d['diag'][:, 1], np.asarray(diag['Relative Improvement (primal variable)']), rtol=1e-9)
        assert np.allclose(d['diag'][:, 2], np.asarray(diag['Relative Improvement (dual variable)']), rtol=1e-9)
        assert np.array_equal(d['x2'], d['x'])


def test_slab_stop_test_is_consistent_across_ranks(tmp_path):
    rows, cols, iters, thr = 192, 200, 60, 2e-2
    parts = _run_ranks(2, tmp_path, ['--rows', str(rows), '--cols', str(cols), '--iters', str(iters), '--threshold',
                                     str(thr)])
    xo, zo, _, n = _oracle(rows, cols, iters, 'conv', thr)
    assert 3 < n < iters
    for d in parts:
        assert int(d['n_iter']) == n
    _check(parts, xo, zo, 1e-10)


def test_slab_fp32_objective(tmp_path):
    rows, cols, iters = 192, 264, 20
    parts = _run_ranks(2, tmp_path, ['--rows', str(rows), '--cols', str(cols), '--iters', str(iters), '--dtype', 'f32'])
    xo, zo, _, _ = _oracle(rows, cols, iters, 'conv')
    _check(parts, xo, zo, 2e-5)


@pytest.mark.parametrize('dtype', ['f64', 'f32'])
def test_slab_full_size_equals_single_gpu(dtype, tmp_path):
    """Two 2048 x 4096 slabs of one 4096^2 TV-deconvolution (the benchmarked size and launch geometry of a slab) against
    the whole image on one GPU: the fp64 iterates must be IDENTICAL on every pixel, fp32 to rounding of the different
    strip / segment geometry (the kernels are deterministic; only the all-reduced norms differ in summation order)."""
    parts = _run_ranks(2, tmp_path, ['--rows', '4096', '--cols', '4096', '--iters', '6', '--dtype', dtype, '--check-single'])
    assert [tuple(d['rows']) for d in parts] == [(0, 2048), (2048, 4096)]
    for d in parts:
        assert str(d['plan_whole']) == 'PdsTvPlan' and int(d['n_iter']) == 6
        if dtype == 'f64':
            assert bool(d['x_equal']) and bool(d['z_equal']), (float(d['x_maxabs']), float(d['z_maxabs']))
        else:
            assert float(d['x_maxabs']) <= 2e-6 and float(d['z_maxabs']) <= 2e-6
        assert np.allclose(d['diag'][:, 1:], d['diag_ref'][:, 1:], rtol=1e-9 if dtype == 'f64' else 1e-4)


def test_slab_checkpoint_resume_and_iterates(tmp_path):
    """SURVEY 8(f)-4 for slab-decomposed solves: 7 updates, per-rank checkpoint files, 7 more on a new solver == 14 updates
    of the oracle; ``iterates(14)`` streams fresh per-rank arrays and ends on the same iterate."""
    rows, cols, iters = 192, 200, 14
    parts = _run_ranks(2, tmp_path, ['--rows', str(rows), '--cols', str(cols), '--iters', str(iters), '--resume'])
    xo, zo, _, _ = _oracle(rows, cols, iters, 'conv')
    _check(parts, xo, zo, 1e-10)
    for d in parts:
        assert bool(d['ck_exists']) and int(d['n_fresh']) >= 2
        assert np.array_equal(d['xi'], d['x']) and np.array_equal(d['zi'], d['z'])
