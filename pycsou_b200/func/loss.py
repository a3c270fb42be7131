r"""Loss functionals (mirror of the hot-path part of ``pycsou/func/loss.py``).

``ProximableLoss(f, data) = f(. - data)`` is ``ProxFuncPreComp(f, scale=1, shift=-data)`` (reference ``loss.py:69``);
``DifferentiableLoss(f, data) = f.shifter(-data)`` (``loss.py:120``).  ``data`` may be a numpy array or a tensor; it is
uploaded once, lazily, in the dtype of the first argument it meets.
"""
from __future__ import annotations

from numbers import Number

import numpy as np
import torch

from .. import _kernels as K
from .._array import to_device
from ..core.functional import DifferentiableFunctional, ProxFuncPreComp, ProximableFunctional
from .base import IndicatorFunctional
from .penalty import (L1Ball, L1Norm, L2Ball, L2Norm, LInftyBall, LInftyNorm, SquaredL1Norm, SquaredL2Norm)

__all__ = ['ProximableLoss', 'DifferentiableLoss', 'L2Loss', 'SquaredL2Loss', 'L2BallLoss', 'L1Loss', 'SquaredL1Loss',
           'L1BallLoss', 'LInftyLoss', 'LInftyBallLoss', 'ConsistencyLoss', 'KLDivergence']


def _neg(data):
    """``-data``; for arrays the sign stays pending until the upload (``_array.Negated``)."""
    from .._array import Negated, SlabRows
    if isinstance(data, Number):
        return -data
    if isinstance(data, Negated):
        return data.a
    return Negated(data if isinstance(data, (torch.Tensor, SlabRows)) else np.asarray(data))


def ProximableLoss(func: ProximableFunctional, data) -> ProximableFunctional:
    return ProxFuncPreComp(func, scale=1, shift=_neg(data))


def DifferentiableLoss(func: DifferentiableFunctional, data) -> DifferentiableFunctional:
    return func.shifter(shift=_neg(data))


def L2Loss(dim: int, data) -> ProximableFunctional:
    return ProximableLoss(L2Norm(dim=dim), data=data)


def SquaredL2Loss(dim: int, data) -> DifferentiableFunctional:
    r""":math:`\|y - x\|_2^2` (reference ``func/loss.py:165-219``)."""
    return DifferentiableLoss(SquaredL2Norm(dim=dim), data=data)


def L2BallLoss(dim: int, data, radius: Number = 1) -> ProximableFunctional:
    return ProximableLoss(L2Ball(dim=dim, radius=radius), data=data)


def L1Loss(dim: int, data) -> ProximableFunctional:
    return ProximableLoss(L1Norm(dim=dim), data=data)


def LInftyBallLoss(dim: int, data, radius: Number = 1) -> ProximableFunctional:
    return ProximableLoss(LInftyBall(dim=dim, radius=radius), data=data)


def SquaredL1Loss(dim: int, data, prox_computation='sort') -> ProximableFunctional:
    """``||y - x||_1^2`` (reference ``func/loss.py:329-368``)."""
    return ProximableLoss(SquaredL1Norm(dim=dim, prox_computation=prox_computation), data=data)


def L1BallLoss(dim: int, data, radius: Number = 1) -> ProximableFunctional:
    """Indicator of ``||y - x||_1 <= radius`` (reference ``func/loss.py:371-426``)."""
    return ProximableLoss(L1Ball(dim=dim, radius=radius), data=data)


def LInftyLoss(dim: int, data) -> ProximableFunctional:
    """``||y - x||_inf`` (reference ``func/loss.py:429-475``)."""
    return ProximableLoss(LInftyNorm(dim=dim), data=data)


def ConsistencyLoss(dim: int, data) -> IndicatorFunctional:
    """Indicator of ``x == y`` up to ``np.allclose``'s default tolerances; prox returns the data (reference
    ``func/loss.py:536-586``)."""
    def y_like(x):
        if isinstance(data, Number):
            return torch.full_like(x, data)
        return to_device(data).to(device=x.device, dtype=x.dtype).reshape(-1)

    def cond(x):
        yy = y_like(x)
        return bool(torch.all(torch.abs(x - yy) <= 1e-8 + 1e-5 * torch.abs(yy)).item())

    def proj(x):
        return y_like(x).clone()
    return IndicatorFunctional(dim=dim, condition_func=cond, projection_func=proj, kind='consistency')


class KLDivergence(ProximableFunctional):
    r"""Generalised Kullback-Leibler divergence :math:`\sum_i y_i\log(y_i/x_i) - y_i + x_i` (reference
    ``func/loss.py:597-682``); prox ``(x - tau + sqrt((x - tau)^2 + 4 tau y)) / 2``.  ``data`` is uploaded once."""

    def __init__(self, dim: int, data):
        super(KLDivergence, self).__init__(dim=dim, data=None, is_differentiable=False, is_linear=False)
        self.data = data
        self._data_dev = None

    def _d(self, x):
        if self._data_dev is None or self._data_dev.dtype != x.dtype or self._data_dev.device != x.device:
            d = self.data
            if isinstance(d, Number):
                d = np.full(self.dim, d)
            self._data_dev = to_device(d, x.dtype)
        return self._data_dev

    def _apply(self, x):
        return float(K.reduce_kl(x, self._d(x)).item())

    def _prox(self, x, tau):
        return K.prox_kl(torch.empty_like(x), x, self._d(x), tau)
