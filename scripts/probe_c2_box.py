"""Scratch probe: C2 time per update on this box, the resident kernel's phase cycles (-DPDS_RES_TIMING build given by
PCS_RES_TIMING_LIB, run in a child process) and the block -> SM map of a cooperative 147-block launch."""
import sys, os, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
r = bench.c2_measure()
print('C2 us/update:', {t: round(r[t]['us_per_update'], 2) for t in ('f64', 'f32')}, flush=True)
lib = os.environ.get('PCS_RES_TIMING_LIB')
if lib:
    env = dict(os.environ, PCS_LIB_PATH=lib)
    out = subprocess.run([sys.executable, os.path.join(os.path.dirname(__file__), 'probe_phases_resident.py'), 'f64'], env=env,
                         capture_output=True, text=True, timeout=200)
    print(out.stdout[-1500:], out.stderr[-500:] if out.returncode else '')
