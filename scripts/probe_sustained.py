"""Does the fused kernel slow down under sustained load?  Time per update in consecutive chunks of 100 launches over
~N seconds, with nvidia-smi clocks / power sampled every 50 ms alongside.   python scripts/probe_sustained.py f64 [chunks]"""
import os
import subprocess
import sys
import threading
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench

dtype = sys.argv[1] if len(sys.argv) > 1 else 'f64'
chunks = int(sys.argv[2]) if len(sys.argv) > 2 else 40
tdt = torch.float32 if dtype == 'f32' else torch.float64
if os.environ.get('PCS_PROBE_NUM_SMS'):          # geometry built for fewer SMs than the device has (pcs_set_num_sms)
    from pycsou_b200 import _lib
    _lib.call_plain('pcs_set_num_sms', int(os.environ['PCS_PROBE_NUM_SMS']))
psf, y64 = bench.make_problem(4096, 'sep', 0)
y = torch.from_numpy(y64.astype(np.float32 if dtype == 'f32' else np.float64)).cuda()
s = bench.make_solver(y, psf, 4096, tdt, 10, 'torch')
s.iterate()
plan = s._plan
x, z = s._init_state()
drow = torch.zeros(4, dtype=torch.float64, device='cuda')
rows = []
proc = subprocess.Popen(['nvidia-smi', '-i', '0', '--query-gpu=clocks.sm,clocks.mem,power.draw,temperature.gpu,'
                         'clocks_event_reasons.sw_power_cap,clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown',
                         '--format=csv,noheader,nounits', '-lms', '50'], stdout=subprocess.PIPE, text=True)
t_start = time.perf_counter()
threading.Thread(target=lambda: [rows.append((time.perf_counter() - t_start, ln.strip())) for ln in proc.stdout], daemon=True).start()
time.sleep(1.0)
torch.cuda.synchronize()
evs = [torch.cuda.Event(enable_timing=True) for _ in range(chunks + 1)]
t0 = time.perf_counter() - t_start
evs[0].record()
for c in range(chunks):
    for _ in range(100):
        x, z = plan.step(x, z, drow)
    evs[c + 1].record()
torch.cuda.synchronize()
t1 = time.perf_counter() - t_start
time.sleep(0.3)
proc.terminate()
us = [10 * evs[c].elapsed_time(evs[c + 1]) for c in range(chunks)]
from pycsou_b200.opt._fused import last_launch
print(last_launch())
print(f'[{dtype}] us/update per chunk of 100: ' + ' '.join(f'{u:.1f}' for u in us))
print(f'[{dtype}] load from t={t0:.2f}s to t={t1:.2f}s; nvidia-smi (t, sm MHz, mem MHz, W, C, power_cap, hw_slow, sw_thermal):')
for t, ln in rows[::4]:
    print(f'   {t:6.2f}  {ln}')
