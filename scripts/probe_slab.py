"""Where does the per-step cost of the slab path go?  Run under torchrun with 2+ ranks.

For the world group (real neighbours) and for a one-rank group (same Python / kernel path, no neighbour) it times 200
back-to-back plan.step calls: host enqueue time per step, device time per step (CUDA events), on every rank.
"""
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
import bench


def loop(s, n=200):
    plan = s._plan
    x, z = plan.init_state(None, None)
    drow = torch.zeros(4, dtype=torch.float64, device='cuda')
    for _ in range(20):
        x, z = plan.step(x, z, drow)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(n):
        x, z = plan.step(x, z, drow)
    e1.record()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    plan.check()
    return 1e6 * (t1 - t0) / n, 1e3 * e0.elapsed_time(e1) / n


def main():
    rank, world, lr = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(lr)
    dist.init_process_group('nccl', device_id=torch.device('cuda', lr))
    from pycsou_b200.opt.slab import SlabComm, release_arenas
    dtype = sys.argv[1] if len(sys.argv) > 1 else 'f32'
    tdt = torch.float32 if dtype == 'f32' else torch.float64
    size = 4096
    psf, y64 = bench.make_problem(size, 'sep', rank)
    y = torch.from_numpy(y64.astype(np.float32 if dtype == 'f32' else np.float64)).cuda()
    solo = [dist.new_group([r]) for r in range(world)][rank]
    for name, comm in (('one-rank group (no neighbour)', SlabComm(solo)), ('world group', SlabComm())):
        s = bench.make_solver(y, psf, size, tdt, 10, 'torch', comm)
        s.iterate()
        host, dev = loop(s)
        print(f'[{dtype}] rank {rank} {name}: host enqueue {host:.1f} us/step, device {dev:.1f} us/step', flush=True)
        # whole solves through iterate(): 100 updates
        s2 = bench.make_solver(y, psf, size, tdt, 100, 'torch', comm)
        s2.iterate()
        s2.reset()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        s2.iterate()
        e1.record()
        torch.cuda.synchronize()
        print(f'[{dtype}] rank {rank} {name}: iterate() {1e3 * e0.elapsed_time(e1) / 100:.1f} us/update', flush=True)
        del s, s2
        release_arenas()
    dist.destroy_process_group()


main()
