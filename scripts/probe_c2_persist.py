"""C2 (TV denoising 1024^2, 200 updates per solve): on-chip resident kernel vs persistent multi-update kernel vs one launch
per update."""
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench

for res, mode in (('0', '0'), ('0', '1'), ('1', '1')):
    os.environ['PCS_PDS_RESIDENT'] = res
    os.environ['PCS_PDS_PERSIST'] = mode
    r = bench.c2_measure()
    for tag in ('f64', 'f32'):
        print(f"PCS_PDS_RESIDENT={res} PCS_PDS_PERSIST={mode} {tag}: {r[tag]['us_per_update']:.2f} us/update  {r[tag]['iter_per_s']:.0f} it/s  "
              f"frac(L2-resident) {r[tag]['roofline_frac_L2_resident']:.3f}  {r[tag]['kernel']['kernel']} x {r[tag]['kernel']['blocks']} blocks", flush=True)
