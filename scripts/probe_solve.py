"""Scratch probe: where does a 100-update solve spend time besides the update kernels?"""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from pycsou_b200.opt import proxalgs as PA

def run(dtype):
    tdt = torch.float32 if dtype == 'f32' else torch.float64
    psf, y64 = bench.make_problem(4096, 'sep', 0)
    y = torch.from_numpy(y64.astype(np.float32 if dtype == 'f32' else np.float64)).cuda()
    s = bench.make_solver(y, psf, 4096, tdt, 100, 'torch')
    for _ in range(3):
        s.reset(); s.iterate()
    torch.cuda.synchronize()
    marks = []
    def mark(name):
        torch.cuda.synchronize(); marks.append((name, time.perf_counter()))
    orig_init, orig_loop, orig_build = s._init_state, s._run_loop, s._build_diagnostics
    def init():
        mark('pre-init'); r = orig_init(); mark('init'); return r
    def loop(st):
        r = orig_loop(st); mark('loop'); return r
    def build(n):
        r = orig_build(n); mark('diag'); return r
    s._init_state, s._run_loop, s._build_diagnostics = init, loop, build
    for rep in range(3):
        marks.clear()
        s.reset(); mark('start'); s.iterate(); mark('end')
        out = ' '.join(f'{b[0]}={1e3*(b[1]-a[1]):.3f}ms' for a, b in zip(marks, marks[1:]))
        print(dtype, out, f'total={1e3*(marks[-1][1]-marks[0][1]):.3f}ms')

run('f32'); run('f64')
