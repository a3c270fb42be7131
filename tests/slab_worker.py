"""One rank of a slab-decomposed Condat-Vu TV solve (run by tests/test_gpu_slab.py through torch.distributed.run).

    python -m torch.distributed.run --nproc-per-node 2 ... tests/slab_worker.py --out DIR --backend nccl|gloo ...

With ``--backend gloo`` every rank uses cuda:0: CUDA IPC works between processes sharing a device, so the fused halo
push / flag handshake of the kernel is exercised on a single-GPU box too (the ranks' kernels are time-sliced).
Saves this rank's rows of the iterates and the diagnostics to ``DIR/rank<r>.npz``.
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    ap = argparse.ArgumentParser()
    ap.add_argument('--out', required=True)
    ap.add_argument('--backend', default='nccl')
    ap.add_argument('--rows', type=int, default=192)
    ap.add_argument('--cols', type=int, default=200)
    ap.add_argument('--iters', type=int, default=12)
    ap.add_argument('--dtype', default='f64')
    ap.add_argument('--data', default='conv', choices=['conv', 'identity'])
    ap.add_argument('--threshold', type=float, default=0.0)
    ap.add_argument('--resume', action='store_true',
                    help='SURVEY 8(f)-4 in slab mode: k1 updates, per-rank checkpoint, k2 more from the checkpoint on a '
                         'new solver; and the same updates streamed through iterates(n).  Saves both results')
    ap.add_argument('--check-single', action='store_true',
                    help='large images: instead of saving the iterates, every rank also solves the WHOLE image alone '
                         '(single-GPU kernel path) and saves how its rows of the slab solve differ from it')
    a = ap.parse_args()
    rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
    dev = int(os.environ.get('LOCAL_RANK', '0')) if a.backend == 'nccl' else 0
    torch.cuda.set_device(dev)
    os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
    if a.backend == 'nccl':
        dist.init_process_group('nccl', device_id=torch.device('cuda', dev))
    else:
        dist.init_process_group('gloo')
    from pycsou_b200.func.loss import SquaredL2Loss
    from pycsou_b200.func.penalty import L21Norm, NonNegativeOrthant, tv_groups
    from pycsou_b200.linop.conv import Convolve2D
    from pycsou_b200.linop.diff import Gradient
    from pycsou_b200.opt.proxalgs import PDS
    from pycsou_b200.opt.slab import SlabComm
    from pycsou_b200.util import synth
    try:
        Hg, W = a.rows, a.cols
        N = Hg * W
        _, psf, y = synth.tv_deconv_problem((Hg, W), None, 0.02, 5)
        y = y.reshape(Hg, W)
        comm = SlabComm()
        lay = comm.layout(Hg, W)
        tdt = torch.float64 if a.dtype == 'f64' else torch.float32
        K = Gradient(shape=(Hg, W), kind='forward')
        K.lipschitz_cst = K.diff_lipschitz_cst = float(np.sqrt(8.))
        F = (1 / 2) * SquaredL2Loss(dim=N, data=lay.local(y[lay.r0:lay.r1].reshape(-1)))
        if a.data == 'conv':
            F = F * Convolve2D(size=N, filter=psf, shape=(Hg, W))
        H = 0.05 * L21Norm(dim=2 * N, groups=tv_groups(N))
        kw = dict(max_iter=a.iters - 1, min_iter=a.iters - 1, accuracy_threshold=0) if a.threshold == 0 else \
            dict(max_iter=a.iters - 1, min_iter=2, accuracy_threshold=a.threshold)
        pds = PDS(dim=N, F=F, G=NonNegativeOrthant(dim=N), H=H, K=K, beta=1.0, verbose=None, dtype=tdt, slab=comm,
                  output='numpy', **kw)
        if a.resume:
            k1, k2 = a.iters // 2, a.iters - a.iters // 2

            def mk(k):
                return PDS(dim=N, F=F, G=NonNegativeOrthant(dim=N), H=H, K=K, beta=1.0, verbose=None, dtype=tdt, slab=comm,
                           output='numpy', max_iter=k - 1, min_iter=k - 1, accuracy_threshold=0)
            s1 = mk(k1)
            s1.iterate()
            ck = os.path.join(a.out, 'ck.pt')
            s1.save_checkpoint(ck)
            s2 = mk(k2)
            assert s2.load_checkpoint(ck) == k1
            est, _, _ = s2.iterate()
            s3 = mk(a.iters)
            last, ptrs = None, set()
            for last in s3.iterates(a.iters):            # streamed: fresh arrays of this rank's rows every update
                ptrs.add(last['primal_variable'].ctypes.data)
            np.savez(os.path.join(a.out, f'rank{rank}.npz'), x=est['primal_variable'], z=est['dual_variable'],
                     rows=np.array([lay.r0, lay.r1]), xi=last['primal_variable'], zi=last['dual_variable'],
                     ck_exists=os.path.exists(f'{ck}.rank{rank}of{world}'), n_fresh=len(ptrs))
            s3._plan.close()
            return
        est, _, diag = pds.iterate()
        if a.check_single:
            Fw = (1 / 2) * SquaredL2Loss(dim=N, data=y.reshape(-1))
            if a.data == 'conv':
                Fw = Fw * Convolve2D(size=N, filter=psf, shape=(Hg, W))
            whole = PDS(dim=N, F=Fw, G=NonNegativeOrthant(dim=N), H=H, K=K, beta=1.0, verbose=None, dtype=tdt,
                        output='numpy', **kw)
            ref, _, dref = whole.iterate()
            xr = ref['primal_variable'].reshape(Hg, W)[lay.r0:lay.r1].reshape(-1)
            zr = ref['dual_variable'].reshape(2, Hg, W)[:, lay.r0:lay.r1].reshape(-1)
            np.savez(os.path.join(a.out, f'rank{rank}.npz'), rows=np.array([lay.r0, lay.r1]),
                     x_equal=np.array_equal(est['primal_variable'], xr), z_equal=np.array_equal(est['dual_variable'], zr),
                     x_maxabs=np.max(np.abs(est['primal_variable'] - xr)), z_maxabs=np.max(np.abs(est['dual_variable'] - zr)),
                     x_norm=np.linalg.norm(xr), diag=diag.to_numpy(dtype=float), diag_ref=dref.to_numpy(dtype=float),
                     plan_whole=type(whole._plan).__name__, n_iter=pds.iter)
            pds._plan.close()
            return
        est2 = None
        if a.threshold == 0:                      # a second solve on the same plan: epochs and buffers carry over
            pds.reset()
            est2, _, _ = pds.iterate()
        np.savez(os.path.join(a.out, f'rank{rank}.npz'), x=est['primal_variable'], z=est['dual_variable'],
                 rows=np.array([lay.r0, lay.r1]), diag=diag.to_numpy(dtype=float), n_iter=pds.iter,
                 x2=est2['primal_variable'] if est2 is not None else np.zeros(0))
        pds._plan.close()
    finally:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
