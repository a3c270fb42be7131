"""Scratch probe: stand-alone non-separable 15x15 Convolve2D at 4096^2 (forward + adjoint), fp32 and fp64."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pycsou_b200.linop.conv import Convolve2D
from pycsou_b200.util import synth
size = 4096
N = size * size
psf = synth.rotated_gaussian_psf_2d(15)
for dt in (torch.float32, torch.float64):
    A = Convolve2D(size=N, filter=psf, shape=(size, size), dtype='float32' if dt == torch.float32 else 'float64')
    x = torch.rand(N, dtype=dt, device='cuda')
    for _ in range(3): y = A._apply(x); z = A._adjoint(y)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): y = A._apply(x); z = A._adjoint(y)
    e1.record(); torch.cuda.synchronize()
    us = 1e3 * e0.elapsed_time(e1) / 40
    print(f'{dt}: {us:.1f} us per 15x15 convolution = {225 * N / us / 1e6:.1f} TFMA/s ({"37.2" if dt == torch.float32 else "18.6"} peak)')
