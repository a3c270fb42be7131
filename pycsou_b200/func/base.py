r"""Functional glue (mirror of ``pycsou/func/base.py``): separable stacks, indicators, null objects, ``LpNorm``."""
from __future__ import annotations

from numbers import Number
from typing import Callable

import numpy as np
import torch

from .. import _kernels as K
from .._array import to_device
from ..core.functional import DifferentiableFunctional, LinearFunctional, ProximableFunctional
from ..core.map import DiffMapHStack, MapHStack

__all__ = ['ProxFuncHStack', 'DiffFuncHStack', 'ExplicitLinearFunctional', 'IndicatorFunctional',
           'NullDifferentiableFunctional', 'NullProximableFunctional', 'LpNorm']


class ProxFuncHStack(ProximableFunctional, MapHStack):
    r"""Separable sum :math:`\sum_i f_i(x_i)` over concatenated blocks (reference ``func/base.py:21-89``)."""

    def __init__(self, *proxfuncs, n_jobs: int = 1, joblib_backend: str = 'loky'):
        MapHStack.__init__(self, *proxfuncs, n_jobs=n_jobs, joblib_backend=joblib_backend)
        self.proxfuncs = self.maps
        ProximableFunctional.__init__(self, dim=self.shape[1], data=None, is_differentiable=self.is_differentiable,
                                      is_linear=self.is_linear)

    def _apply(self, x):
        tot = 0.0
        for f, xi in zip(self.proxfuncs, self._split(x)):
            v = f._apply(xi)
            tot = tot + (float(v.item()) if isinstance(v, torch.Tensor) else v)
        return tot

    def _prox(self, x, tau):
        out = torch.empty_like(x)
        edges = [0] + [int(s) for s in self.sections]
        for i, (f, xi) in enumerate(zip(self.proxfuncs, self._split(x))):
            out[edges[i]:edges[i + 1]].copy_(f._prox(xi, tau))
        return out


class DiffFuncHStack(DifferentiableFunctional, DiffMapHStack):
    def __init__(self, *difffuncs, n_jobs: int = 1, joblib_backend: str = 'loky'):
        DiffMapHStack.__init__(self, *difffuncs, n_jobs=n_jobs, joblib_backend=joblib_backend)
        self.difffuncs = self.maps
        DifferentiableFunctional.__init__(self, dim=self.shape[1], data=None, is_linear=self.is_linear,
                                          lipschitz_cst=self.lipschitz_cst, diff_lipschitz_cst=self.diff_lipschitz_cst)

    def _apply(self, x):
        tot = 0.0
        for f, xi in zip(self.maps, self._split(x)):
            v = f._apply(xi)
            tot = tot + (float(v.item()) if isinstance(v, torch.Tensor) else v)
        return tot

    def _jacobianT(self, x):
        return torch.cat([f._jacobianT(xi) for f, xi in zip(self.maps, self._split(x))])


class ExplicitLinearFunctional(LinearFunctional):
    r""":math:`x\mapsto\langle v, x\rangle` (reference ``func/base.py:124-137``)."""

    def __init__(self, vec, dtype: type = np.float64):
        self.vec = np.asarray(vec.detach().cpu() if isinstance(vec, torch.Tensor) else vec).flatten().astype(dtype)
        self._vec_dev = None
        super(ExplicitLinearFunctional, self).__init__(dim=self.vec.size, dtype=dtype, is_explicit=True)

    def _vec_like(self, x):
        if self._vec_dev is None or self._vec_dev.dtype != x.dtype:
            self._vec_dev = to_device(self.vec, x.dtype)
        return self._vec_dev

    def _apply(self, x):
        return K.dot(self._vec_like(x), x)

    def _adjoint(self, y):
        v = self._vec_like(y)
        return K.scale(torch.empty_like(v), float(y.reshape(-1)[0].item()), v)


class IndicatorFunctional(ProximableFunctional):
    r"""Indicator of a convex set given by a membership test and a projection (reference ``func/base.py:140-168``).

    ``condition_func`` / ``projection_func`` receive device tensors.  The library's own indicator sets
    (``NonNegativeOrthant``, ``Segment``, ``L2Ball``, ...) additionally carry a ``kind`` tag so the solvers can fuse
    their projection into the primal update kernel.
    """

    def __init__(self, dim: int, condition_func: Callable, projection_func: Callable, kind=None, params=()):
        super(IndicatorFunctional, self).__init__(dim=dim, data=None, is_differentiable=False, is_linear=False)
        self.condition_func = condition_func
        self.projection_func = projection_func
        self.kind = kind
        self.params = tuple(params)

    def _apply(self, x, **kwargs):
        return 0 if self.condition_func(x, **kwargs) else np.inf

    def _prox(self, x, tau, **kwargs):
        return self.projection_func(x, **kwargs)


class NullDifferentiableFunctional(DifferentiableFunctional):
    def __init__(self, dim: int):
        super(NullDifferentiableFunctional, self).__init__(dim=dim, is_linear=True, lipschitz_cst=0,
                                                           diff_lipschitz_cst=0)

    def _apply(self, x):
        return 0

    def _jacobianT(self, x):
        return torch.zeros(self.dim, dtype=x.dtype, device=x.device)


class NullProximableFunctional(ProximableFunctional):
    def __init__(self, dim: int):
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            super(NullProximableFunctional, self).__init__(dim=dim, is_linear=True)

    def _apply(self, x):
        return 0

    def _prox(self, x, tau):
        return x


class LpNorm(ProximableFunctional):
    r""":math:`\ell_p` norm whose prox follows from the dual-ball projection via Moreau:
    ``x - tau * proj_lq_ball(x / tau, radius=1)`` (reference ``func/base.py:215-240``)."""

    def __init__(self, dim: int, proj_lq_ball: Callable):
        super(LpNorm, self).__init__(dim=dim, data=None, is_differentiable=False, is_linear=False)
        self.proj_lq_ball = proj_lq_ball

    def _prox(self, x, tau):
        proj = getattr(self.proj_lq_ball, 'device', self.proj_lq_ball)
        xs = K.scale(torch.empty_like(x), 1.0 / tau, x)
        return K.lincomb2(xs, 1.0, x, -tau, proj(xs, radius=1))
