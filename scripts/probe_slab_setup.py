"""Per-solve set-up cost of the slab path, piece by piece (torchrun, 2+ ranks): wall ms with a device sync after each."""
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
import bench


def timed(label, fn, acc):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    r = fn()
    torch.cuda.synchronize()
    acc.append((label, 1e3 * (time.perf_counter() - t0)))
    return r


def main():
    rank, lr = int(os.environ['RANK']), int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(lr)
    dist.init_process_group('nccl', device_id=torch.device('cuda', lr))
    from pycsou_b200.opt.slab import SlabComm, release_arenas
    dtype = sys.argv[1] if len(sys.argv) > 1 else 'f32'
    tdt = torch.float32 if dtype == 'f32' else torch.float64
    psf, y64 = bench.make_problem(4096, 'sep', rank)
    y = torch.from_numpy(y64.astype(np.float32 if dtype == 'f32' else np.float64)).cuda()
    comm = SlabComm()
    s = bench.make_solver(y, psf, 4096, tdt, 100, 'torch', comm)
    s.iterate()
    for rep in range(2):
        acc = []
        s.reset()
        plan = s._plan
        timed('begin_solve (halo of y + A^T y)', plan.begin_solve, acc)
        st = timed('init_state (2 barriers + zero)', lambda: plan.init_state(None, None), acc)
        timed('barrier alone', comm.barrier, acc)
        timed('_alloc_diag', s._alloc_diag, acc)

        def run():
            x, z = st
            for it in range(100):
                x, z = plan.step(x, z, s._diag_dev[it])
            return x, z
        x, z = timed('100 steps', run, acc)
        timed('_reduce_diag (all-reduce of the norms)', lambda: s._reduce_diag(100), acc)
        timed('check', plan.check, acc)
        timed('extract (clone own rows)', lambda: plan.extract(x, z), acc)
        t0 = time.perf_counter()
        s.reset()
        s.iterate()
        torch.cuda.synchronize()
        acc.append(('whole iterate()', 1e3 * (time.perf_counter() - t0)))
        if rank == 0:
            print(f'[{dtype}] rep {rep}: ' + '; '.join(f'{k} {v:.2f} ms' for k, v in acc), flush=True)
    release_arenas()
    dist.destroy_process_group()


main()
