r"""Element-wise proximal helpers on the device (mirror of ``pycsou/math/prox.py``).

``sign``, ``soft``, ``proj_l1_ball``, ``proj_l2_ball``, ``proj_linfty_ball``, ``proj_nonnegative_orthant``,
``proj_segment`` (reference ``math/prox.py:17-343``).
Unlike the reference the projections never mutate their argument (the reference's in-place behaviour is an aliasing
hazard); values are identical.
"""
from __future__ import annotations

from numbers import Number

import torch

from .. import _kernels as K
from .._array import like_input, to_device

__all__ = ['sign', 'soft', 'proj_l1_ball', 'proj_l2_ball', 'proj_linfty_ball', 'proj_nonnegative_orthant',
           'proj_segment']


def _edge(fn):
    def wrapped(x, *a, **k):
        return like_input(fn(to_device(x), *a, **k), x)
    wrapped.__name__, wrapped.__doc__ = fn.__name__, fn.__doc__
    wrapped.device = fn
    return wrapped


@_edge
def sign(x):
    """``conj(x)/|x|`` where ``x != 0``, else 0 (note the conjugate, reference ``math/prox.py:61-64``)."""
    if x.is_complex():
        mag = x.abs()
        return torch.where(mag != 0, x.conj() / mag.clamp_min(torch.finfo(mag.dtype).tiny), torch.zeros_like(x))
    return torch.sign(x)


@_edge
def soft(x, tau: Number):
    """Soft thresholding ``max(|x| - tau, 0) * sign(x)`` (reference ``math/prox.py:114``)."""
    if x.is_complex():
        mag = x.abs()
        return (mag - tau).clamp_min(0) * sign.device(x)
    # real: identical to x - clip(x, -tau, tau)
    return K.lincomb2(torch.empty_like(x), 1.0, x, -1.0, K.proj_segment(torch.empty_like(x), x, -tau, tau))


@_edge
def proj_l1_ball(x, radius: Number):
    """``x`` (the same object) if ``||x||_1 <= radius`` else ``soft(x, mu)`` with ``sum max(|x| - mu, 0) = radius``
    (reference ``math/prox.py:157-164``).  The reference brackets ``mu`` with ``scipy.optimize.brentq`` (absolute
    tolerance 2e-12); here monotone Newton steps on the device reach the exact root (``K.threshold_newton``), so the
    two agree to brentq's tolerance.  Real arrays only."""
    if x.is_complex():
        raise NotImplementedError('proj_l1_ball: real arrays only on the device')
    if float(K.sumabs(x).item()) <= radius:
        return x
    return K.prox_l1(torch.empty_like(x), x, K.threshold_newton(x, radius, 0.0))


@_edge
def proj_l2_ball(x, radius: Number):
    """``x`` if ``||x|| <= radius`` else ``radius * x / ||x||`` (reference ``math/prox.py:207-210``)."""
    return K.scale_by_norm(torch.empty_like(x), x, K.sumsq(x), radius, 1)


@_edge
def proj_linfty_ball(x, radius: Number):
    """Clip to ``[-radius, radius]`` (reference ``math/prox.py:253-256``; real-valued only)."""
    if x.is_complex():
        raise NotImplementedError('proj_linfty_ball uses ordered comparisons: real arrays only')
    return K.proj_segment(torch.empty_like(x), x, -radius, radius)


@_edge
def proj_nonnegative_orthant(x):
    """``max(Re x, 0)`` (reference ``math/prox.py:295-297``)."""
    x = x.real.contiguous() if x.is_complex() else x
    return K.proj_nonneg(torch.empty_like(x), x)


@_edge
def proj_segment(x, a: Number = 0, b: Number = 1):
    """``clip(Re x, a, b)`` (reference ``math/prox.py:340-343``)."""
    x = x.real.contiguous() if x.is_complex() else x
    return K.proj_segment(torch.empty_like(x), x, a, b)
