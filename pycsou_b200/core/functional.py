r"""Functionals and the proximal calculus (mirror of ``pycsou/core/functional.py``, device-resident).

``ProximableFunctional.prox`` / ``fenchel_prox`` (Moreau identity, reference ``functional.py:207``) and the
automatic prox rules ``ProxFuncPostComp`` (:math:`\alpha f`), ``ProxFuncPreComp`` (:math:`f(\alpha x+b)`),
``ProxFuncAffineSum`` and ``ProxFuncPreCompUnitOp`` (:math:`f\circ U`) keep the reference's names, arguments
and formulas (``functional.py:253-313``).  The public methods accept numpy arrays at the edge; solvers call the
device-side ``_prox`` / ``_fenchel_prox`` directly.
"""
from __future__ import annotations

import warnings
from numbers import Number
from typing import Union

import numpy as np
import torch

from .. import _kernels as K
from .._array import is_array, like_input, to_device
from .linop import LinearOperator, UnitaryOperator
from .map import DifferentiableMap, Map, MapComp, MapSum


class Functional(Map):
    r"""Base class for functionals :math:`f:\mathbb{C}^N\to\mathbb{R}`."""

    def __init__(self, dim: int, data=None, is_differentiable: bool = False, is_linear: bool = False):
        Map.__init__(self, shape=(1, dim), is_differentiable=is_differentiable, is_linear=is_linear)
        self.data = data
        self.dim = dim


class DifferentiableFunctional(Functional, DifferentiableMap):
    def __init__(self, dim: int, data=None, is_linear: bool = False, lipschitz_cst: float = np.inf,
                 diff_lipschitz_cst: float = np.inf):
        Functional.__init__(self, dim=dim, data=data, is_differentiable=True, is_linear=is_linear)
        DifferentiableMap.__init__(self, shape=self.shape, is_linear=self.is_linear, lipschitz_cst=lipschitz_cst,
                                   diff_lipschitz_cst=diff_lipschitz_cst)


class LinearFunctional(Functional, LinearOperator):
    def __init__(self, dim: int, data=None, dtype: type = np.float64, is_explicit: bool = False,
                 is_dense: bool = False, is_sparse: bool = False, is_dask: bool = False):
        Functional.__init__(self, dim=dim, data=data, is_differentiable=True, is_linear=True)
        LinearOperator.__init__(self, shape=self.shape, dtype=dtype, is_explicit=is_explicit, is_dense=is_dense,
                                is_sparse=is_sparse, is_dask=is_dask, is_symmetric=False)


class ProximableFunctional(Functional):
    r"""Functional with a simple proximal operator (reference ``core/functional.py:90-250``)."""

    def __init__(self, dim: int, data=None, is_differentiable: bool = False, is_linear: bool = False):
        if is_differentiable == True or is_linear == True:
            warnings.warn(
                'For differentiable and/or linear maps, consider the dedicated classes DifferentiableMap and LinearOperator.')
        super(ProximableFunctional, self).__init__(dim=dim, data=data, is_differentiable=is_differentiable,
                                                   is_linear=is_linear)

    def prox(self, x, tau: Number):
        if type(self)._prox is ProximableFunctional._prox:
            raise NotImplementedError('ProximableFunctional subclasses must implement prox or _prox')
        return like_input(self._prox(to_device(x), tau), x)

    def _prox(self, x: torch.Tensor, tau: Number) -> torch.Tensor:
        return self.prox(x, tau)

    def fenchel_prox(self, z, sigma: Number):
        return like_input(self._fenchel_prox(to_device(z), sigma), z)

    def _fenchel_prox(self, z: torch.Tensor, sigma: Number) -> torch.Tensor:
        # z - sigma * prox_{f/sigma}(z/sigma)   (reference core/functional.py:207)
        zs = K.scale(torch.empty_like(z), 1.0 / sigma, z)
        p = self._prox(zs, 1.0 / sigma)
        return K.lincomb2(zs, 1.0, z, -sigma, p)

    def shifter(self, shift) -> 'ProxFuncPreComp':
        return ProxFuncPreComp(prox_func=self, scale=1, shift=shift)

    def __add__(self, other):
        if isinstance(other, LinearFunctional):
            return ProxFuncAffineSum(self, linear_part=other, intercept=0)
        if isinstance(other, Map):
            return MapSum(self, other)
        raise NotImplementedError

    def __mul__(self, other):
        if isinstance(other, Number) or is_array(other):
            return ProxFuncPreComp(self, scale=other, shift=0)
        if isinstance(other, UnitaryOperator):
            return ProxFuncPreCompUnitOp(self, other)
        if isinstance(other, Map):
            return MapComp(self, other)
        raise NotImplementedError

    def __rmul__(self, other):
        if isinstance(other, Number) and other > 0:
            return ProxFuncPostComp(self, scale=other, shift=0)
        if isinstance(other, Map):
            return MapComp(other, self)
        raise NotImplementedError


def _fval(v):
    return float(v.item()) if isinstance(v, torch.Tensor) else v


class ProxFuncPostComp(ProximableFunctional):
    r""":math:`\alpha f + c`: ``prox_{tau (alpha f)} = prox_f(., alpha tau)`` (reference ``functional.py:253-265``)."""

    def __init__(self, prox_func: ProximableFunctional, scale: Number, shift: Number):
        super(ProxFuncPostComp, self).__init__(dim=prox_func.dim, data=prox_func.data,
                                               is_differentiable=prox_func.is_differentiable)
        self.prox_func = prox_func
        self.scale = scale
        self.shift = shift

    def _apply(self, x):
        return self.scale * _fval(self.prox_func._apply(x)) + self.shift

    def _prox(self, x, tau):
        return self.prox_func._prox(x, tau * self.scale)


class ProxFuncAffineSum(ProximableFunctional):
    def __init__(self, prox_func: ProximableFunctional, linear_part: LinearFunctional, intercept: Number):
        if not isinstance(linear_part, LinearFunctional) or linear_part.dim != prox_func.dim:
            raise TypeError('Invalid affine sum.')
        super(ProxFuncAffineSum, self).__init__(dim=prox_func.dim, data=prox_func.data,
                                                is_differentiable=prox_func.is_differentiable)
        self.prox_func = prox_func
        self.linear_part = linear_part
        self.intercept = intercept

    def _apply(self, x):
        return _fval(self.prox_func._apply(x)) + _fval(self.linear_part._apply(x)) + self.intercept

    def _prox(self, x, tau):
        a = self.linear_part._vec_like(x)        # the explicit vector of the linear functional
        return self.prox_func._prox(K.lincomb2(torch.empty_like(x), 1.0, x, -tau, a), tau)


class ProxFuncPreComp(ProximableFunctional):
    r""":math:`f(\alpha x + b)` with scalar :math:`\alpha` (reference ``functional.py:286-299``)."""

    def __init__(self, prox_func: ProximableFunctional, scale, shift):
        super(ProxFuncPreComp, self).__init__(dim=prox_func.dim, data=prox_func.data,
                                              is_differentiable=prox_func.is_differentiable)
        self.prox_func = prox_func
        self.scale = scale
        self.shift = shift
        self._shift_dev = None
        if is_array(scale):
            raise NotImplementedError('array-valued pre-composition scale')

    def _shift_like(self, x):
        if self._shift_dev is None or self._shift_dev.dtype != x.dtype or self._shift_dev.device != x.device:
            s = self.shift
            if isinstance(s, Number):
                s = np.full(self.dim, s)
            self._shift_dev = to_device(s, x.dtype)
        return self._shift_dev

    def _affine(self, x):
        return K.lincomb2(torch.empty_like(x), self.scale, x, 1.0, self._shift_like(x))

    def _apply(self, x):
        return self.prox_func._apply(self._affine(x))

    def _prox(self, x, tau):
        # (prox_f(alpha x + b, alpha^2 tau) - b) / alpha
        p = self.prox_func._prox(self._affine(x), tau * (self.scale ** 2))
        return K.lincomb2(torch.empty_like(x), 1.0 / self.scale, p, -1.0 / self.scale, self._shift_like(x))


class ProxFuncPreCompUnitOp(ProximableFunctional):
    r""":math:`f\circ U` with unitary :math:`U`: ``U^H prox_f(U x)`` (reference ``functional.py:302-313``)."""

    def __init__(self, prox_func: ProximableFunctional, unitary_op: UnitaryOperator):
        super(ProxFuncPreCompUnitOp, self).__init__(dim=prox_func.dim, data=prox_func.data,
                                                    is_differentiable=prox_func.is_differentiable)
        self.prox_func = prox_func
        self.unitary_op = unitary_op

    def _apply(self, x):
        return self.prox_func._apply(self.unitary_op._apply(x))

    def _prox(self, x, tau):
        return self.unitary_op._adjoint(self.prox_func._prox(self.unitary_op._apply(x), tau))
