"""Stand-alone SEPARABLE Convolve2D at 4096^2: streaming kernel vs the tile kernel (PCS_CONV2D_SEP_TILE=1 in a child
process), forward and adjoint, bitwise comparison of the two.   python scripts/probe_conv2d_sep.py [K ...]"""
import os
import subprocess
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch


def run(K, size=4096):
    from pycsou_b200 import _kernels as Kn
    out = {}
    for dt in (torch.float32, torch.float64):
        g = torch.Generator(device='cpu').manual_seed(K)
        hc = torch.rand(K, generator=g, dtype=torch.float64).to(dt).cuda()
        hr = torch.rand(K, generator=g, dtype=torch.float64).to(dt).cuda()
        x = torch.rand(size * size + 0, generator=g, dtype=torch.float64).to(dt).cuda()
        sub = torch.rand(size * size, generator=g, dtype=torch.float64).to(dt).cuda()
        o = torch.empty_like(x)
        for adj in (False, True):
            for s_ in (None, sub):
                Kn.conv2d_sep(o, x, hc, hr, K // 2, K // 2, size, size, adjoint=adj, sub=s_)
                out[(str(dt), adj, s_ is not None)] = float(o.double().sum()), float(o.double().abs().max()), o[::4099].clone().cpu()
        for _ in range(3):
            Kn.conv2d_sep(o, x, hc, hr, K // 2, K // 2, size, size)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            Kn.conv2d_sep(o, x, hc, hr, K // 2, K // 2, size, size)
        e1.record()
        torch.cuda.synchronize()
        us = 1e3 * e0.elapsed_time(e1) / 20
        s = 4 if dt == torch.float32 else 8
        print(f'K={K} {dt}: {us:7.1f} us  {2 * size * size * s / us / 1e3:7.1f} GB/s (read + write once)', flush=True)
    return out


if __name__ == '__main__':
    if os.environ.get('PCS_PROBE_CHILD'):
        import pickle
        K = int(sys.argv[1])
        res = run(K)
        pickle.dump(res, open(f'/tmp/conv_sep_{K}_{os.environ.get("PCS_CONV2D_SEP_TILE", "0")}.pkl', 'wb'))
    else:
        import pickle
        for K in [int(a) for a in sys.argv[1:]] or [15, 7, 3]:
            for tile in ('0', '1'):
                print('--- tile kernel' if tile == '1' else '--- streaming kernel', flush=True)
                env = dict(os.environ, PCS_PROBE_CHILD='1', PCS_CONV2D_SEP_TILE=tile)
                subprocess.run([sys.executable, __file__, str(K)], env=env, check=True)
            a = pickle.load(open(f'/tmp/conv_sep_{K}_0.pkl', 'rb'))
            b = pickle.load(open(f'/tmp/conv_sep_{K}_1.pkl', 'rb'))
            same = all(torch.equal(a[k][2], b[k][2]) and a[k][0] == b[k][0] for k in a)
            print(f'K={K}: streaming == tile bitwise on {len(a)} cases: {same}', flush=True)
