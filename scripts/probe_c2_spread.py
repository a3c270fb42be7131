"""Scratch probe: spread of the C2 (1024^2 TV denoising, resident kernel) time per update over repeated solves on one box."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
for rep in range(6):
    r = bench.c2_measure()
    print(rep, {t: round(r[t]['us_per_update'], 2) for t in ('f64', 'f32')}, r['f64']['kernel']['kernel'], flush=True)
