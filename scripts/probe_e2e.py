"""Scratch probe: where the end-to-end time of one c3 solve from host buffers goes (bench.py e2e leg)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench

def T():
    torch.cuda.synchronize(); return time.perf_counter()

size, iters = 4096, 100
psf, y64 = bench.make_problem(size, 'sep', 0)
y_host = torch.from_numpy(y64.astype(np.float32)).pin_memory()
for rep in range(4):
    t0 = T()
    s = bench.make_solver(y_host, psf, size, torch.float32, iters, 'numpy')
    t1 = T()
    s._make_plan()
    t2 = T()
    state = s._init_state()
    t3 = T()
    x, z = s._run_loop(state)
    t4 = T()
    s._plan.release(x, z)
    xo = s._as_output(x)
    t5 = T()
    zo = s._as_output(z)
    t6 = T()
    d = s._build_diagnostics(s.iter)
    t7 = T()
    print(f'rep {rep}: construct {1e3*(t1-t0):.2f} ms, plan (H2D y, A^T y) {1e3*(t2-t1):.2f}, init state {1e3*(t3-t2):.2f}, '
          f'loop {1e3*(t4-t3):.2f}, D2H x {1e3*(t5-t4):.2f}, D2H z {1e3*(t6-t5):.2f}, diagnostics {1e3*(t7-t6):.2f}, total {1e3*(t7-t0):.2f}')
    del xo, zo
# raw PCIe rates
h = torch.empty(4096 * 4096 * 2, dtype=torch.float32, pin_memory=True)
dd = torch.empty_like(h, device='cuda')
for _ in range(2):
    t0 = T(); dd.copy_(h, non_blocking=True); t1 = T(); h.copy_(dd, non_blocking=True); t2 = T()
print(f'pinned H2D {h.numel()*4/1e9/(t1-t0):.1f} GB/s, D2H {h.numel()*4/1e9/(t2-t1):.1f} GB/s')

def e2e_step():
    s = bench.make_solver(y_host, psf, size, torch.float32, iters, 'numpy')
    out, _, _ = s.iterate()
    return out
import gc
gc.collect(); gc.disable()      # results must be freed by reference counting alone (no solver <-> plan cycle)
for label, use_sampler in (('no sampler', False), ('nvidia-smi sampler running', True)):
    sampler = bench.ClockSampler(0).start() if use_sampler else None
    out = None
    ts = []
    for _ in range(8):
        t0 = T(); out = e2e_step(); ts.append(1e3 * (T() - t0))
    if sampler: sampler.stop()
    print(label, 'e2e steps (keeping the previous result alive):', ' '.join(f'{t:.1f}' for t in ts), 'ms')
print('objects found by the cycle collector afterwards:', gc.collect())
