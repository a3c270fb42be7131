"""Scratch probe: host enqueue time vs device time of the fused PDS step loop (launch-gap diagnosis)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench

def run(dtype):
    tdt = torch.float32 if dtype == 'f32' else torch.float64
    psf, y64 = bench.make_problem(4096, 'sep', 0)
    y = torch.from_numpy(y64.astype(np.float32 if dtype == 'f32' else np.float64)).cuda()
    s = bench.make_solver(y, psf, 4096, tdt, 100, 'torch')
    s.iterate(); s.reset(); s.iterate()
    plan = s._plan
    x, z = s._init_state()
    drow = torch.zeros(4, dtype=torch.float64, device='cuda')
    for _ in range(10):
        x, z = plan.step(x, z, drow)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 200
    t0 = time.perf_counter(); e0.record()
    for _ in range(n):
        x, z = plan.step(x, z, drow)
    e1.record(); t1 = time.perf_counter()
    torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f'{dtype}: raw plan.step loop: host enqueue {1e6*(t1-t0)/n:.1f} us/iter, device {1e3*e0.elapsed_time(e1)/n:.1f} us/iter, wall {1e6*(t2-t0)/n:.1f}')
    # full solver loop
    s.reset(); torch.cuda.synchronize()
    t0 = time.perf_counter(); e0.record(); s.iterate(); e1.record(); torch.cuda.synchronize(); t1 = time.perf_counter()
    print(f'{dtype}: solver.iterate(): device {1e3*e0.elapsed_time(e1)/100:.1f} us/iter, wall {1e6*(t1-t0)/100:.1f}')

run('f32'); run('f64')
