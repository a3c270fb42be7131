r"""Abstract maps and their arithmetic (mirror of ``pycsou/core/map.py``, device-resident).

Same class names, constructor arguments, ``shape=(out, in)`` convention and composition rules as the reference
(``core/map.py:20-1001``): ``+`` builds ``MapSum``/``DiffMapSum``, ``*`` / ``@`` with a map builds
``MapComp``/``DiffMapComp``, ``*`` with an array *evaluates* the map (``map.py:288-289``), a scalar is wrapped in
a ``HomothetyMap`` (``map.py:283-299``), Lipschitz constants propagate as in ``map.py:564-610``.

Difference by design: arrays are 1-D CUDA tensors.  Every map has a device-side ``_apply`` that the solvers call
directly; the public ``__call__`` accepts numpy arrays / CPU tensors (one H2D copy) and mirrors the input kind on
return.  User subclasses may override either ``__call__`` (reference style, written against the tensor API) or
``_apply``.
"""
from __future__ import annotations

from abc import ABC
from numbers import Number
from typing import Tuple, Union

import numpy as np
import torch

from .. import _kernels as K
from .._array import is_array, like_input, to_device
from ..util.misc import is_range_broadcastable, range_broadcast_shape


def _scale_any(a, alpha):
    """alpha * a for a tensor or number (device kernel for tensors)."""
    if isinstance(a, torch.Tensor):
        return K.scale(torch.empty_like(a), alpha, a)
    return a * alpha


def _add_any(a, b):
    if isinstance(a, torch.Tensor) and isinstance(b, torch.Tensor):
        return K.lincomb2(torch.empty_like(a), 1.0, a, 1.0, b)
    if isinstance(a, torch.Tensor) or isinstance(b, torch.Tensor):
        t, s = (a, b) if isinstance(a, torch.Tensor) else (b, a)
        if s == 0:
            return t
        raise NotImplementedError('tensor + nonzero scalar')
    return a + b


def jac_mul(a, b):
    """Product used by the chain rule ``map2.jacobianT(x) * map1.jacobianT(map2(x))`` (core/map.py:609-610).

    Covers the operand kinds the reference produces: LinearOperator * array (application, core/linop.py:460-461),
    array * scalar, scalar * scalar, LinearOperator * LinearOperator (composition).
    """
    from .linop import LinearOperator
    if isinstance(a, LinearOperator):
        if isinstance(b, torch.Tensor):
            return a._apply(b)
        return a * b
    if isinstance(b, LinearOperator):
        if isinstance(a, torch.Tensor):
            raise NotImplementedError('array * LinearOperator')
        return a * b
    if isinstance(a, torch.Tensor) and isinstance(b, torch.Tensor):
        return K.mul(torch.empty_like(a), a, b)
    if isinstance(a, torch.Tensor):
        return _scale_any(a, b)
    if isinstance(b, torch.Tensor):
        return _scale_any(b, a)
    return a * b


class Map(ABC):
    r"""Base class for multidimensional maps :math:`\mathbb{R}^M\to\mathbb{R}^N` (reference ``core/map.py:20-335``)."""

    def __init__(self, shape: Tuple[int, int], is_linear: bool = False, is_differentiable: bool = False):
        if len(shape) > 2:
            raise NotImplementedError(
                'Shapes of map objects must be tuples of length 2 (tensorial maps not supported).')
        self.shape = shape
        self.is_linear = is_linear
        self.is_functional = True if self.shape[0] == 1 else False
        self.is_differentiable = is_differentiable
        super(Map, self).__init__()

    # -- evaluation ------------------------------------------------------------------------------------------
    def _in_dtype(self, arg):
        return None

    def __call__(self, arg):
        if type(self)._apply is Map._apply:
            raise NotImplementedError('Map subclasses must implement __call__ or _apply')
        x = to_device(arg, self._in_dtype(arg))
        out = self._apply(x)
        if self.is_functional and isinstance(out, torch.Tensor) and out.numel() == 1:
            return float(out.item())
        return like_input(out, arg)

    def _apply(self, x: torch.Tensor):
        return self.__call__(x)

    def apply_along_axis(self, arr, axis: int = 0):
        a = arr if isinstance(arr, np.ndarray) else np.asarray(arr.detach().cpu())
        if a.shape[axis] != self.shape[1]:
            raise ValueError(
                f"Array size along specified axis and the map domain's dimension differ: {a.shape[axis]} != {self.shape[1]}.")
        return np.apply_along_axis(func1d=lambda v: np.asarray(self.__call__(np.ascontiguousarray(v))), axis=axis,
                                   arr=a)

    def shifter(self, shift) -> 'MapShifted':
        return MapShifted(map=self, shift=shift)

    # -- arithmetic ------------------------------------------------------------------------------------------
    def __add__(self, other: 'Map') -> 'MapSum':
        if isinstance(other, Map):
            return MapSum(self, other)
        raise NotImplementedError

    def __radd__(self, other: 'Map') -> 'MapSum':
        if isinstance(other, Map):
            return MapSum(other, self)
        raise NotImplementedError

    def _wrap_scalar(self, other, size):
        from ..linop.base import HomothetyMap
        return HomothetyMap(constant=other, size=size)

    def __mul__(self, other):
        if isinstance(other, Number):
            other = self._wrap_scalar(other, self.shape[1])
        if is_array(other):
            return self(other)
        if isinstance(other, Map):
            return MapComp(self, other)
        raise NotImplementedError

    def __rmul__(self, other):
        if isinstance(other, Number):
            other = self._wrap_scalar(other, self.shape[0])
        if isinstance(other, Map):
            return MapComp(other, self)
        raise NotImplementedError

    def __matmul__(self, other):
        return self.__mul__(other)

    def __neg__(self):
        return self.__mul__(-1)

    def __sub__(self, other):
        return self.__add__(other.__neg__())

    def __pow__(self, power: int):
        if type(power) is int:
            exp_map = self
            for _ in range(1, power):
                exp_map = self.__mul__(exp_map)
            return exp_map
        raise NotImplementedError

    def __truediv__(self, scalar: Number):
        if isinstance(scalar, Number):
            return self.__mul__(1 / scalar)
        raise NotImplementedError


def _shift_size(shift):
    return 1 if isinstance(shift, Number) else int(np.prod(tuple(shift.shape)))


class MapShifted(Map):
    """``x -> map(x + shift)`` (reference ``core/map.py:337-346``)."""

    def __init__(self, map: Map, shift):
        self.map = map
        self.shift = shift
        if _shift_size(shift) != map.shape[1]:
            raise TypeError('Invalid shift size.')
        self._shift_dev = None
        Map.__init__(self, shape=map.shape, is_linear=map.is_linear, is_differentiable=map.is_differentiable)

    def _shift_like(self, x):
        if self._shift_dev is None or self._shift_dev.dtype != x.dtype or self._shift_dev.device != x.device:
            s = self.shift
            if isinstance(s, Number):
                s = np.full(self.shape[1], s)
            self._shift_dev = to_device(s, x.dtype)
        return self._shift_dev

    def _shifted(self, x):
        return K.lincomb2(torch.empty_like(x), 1.0, x, 1.0, self._shift_like(x))

    def _apply(self, x):
        return self.map._apply(self._shifted(x))


class MapSum(Map):
    def __init__(self, map1: Map, map2: Map):
        if not is_range_broadcastable(map1.shape, map2.shape):
            raise ValueError('Cannot sum two maps with inconsistent range or domain sizes.')
        Map.__init__(self, shape=range_broadcast_shape(map1.shape, map2.shape),
                     is_linear=map1.is_linear & map2.is_linear,
                     is_differentiable=map1.is_differentiable & map2.is_differentiable)
        self.map1, self.map2 = map1, map2

    def _apply(self, x):
        a, b = self.map1._apply(x), self.map2._apply(x)
        if isinstance(a, torch.Tensor) and isinstance(b, torch.Tensor) and a.numel() == b.numel() and a.numel() > 1:
            return K.lincomb2(torch.empty_like(a), 1.0, a, 1.0, b)
        return a + b      # functional values (python floats / 1-element tensors)


class MapComp(Map):
    def __init__(self, map1: Map, map2: Map):
        if map1.shape[1] != map2.shape[0] and map2.shape[0] != 1 and map1.shape[1] != 1:
            raise ValueError('Cannot compose two maps with inconsistent range or domain sizes.')
        Map.__init__(self, shape=(map1.shape[0], map2.shape[1]),
                     is_linear=map1.is_linear & map2.is_linear,
                     is_differentiable=map1.is_differentiable & map2.is_differentiable)
        self.map1 = map1
        self.map2 = map2

    def _apply(self, x):
        return self.map1._apply(self.map2._apply(x))


class DifferentiableMap(Map):
    r"""Differentiable map with Lipschitz bookkeeping (reference ``core/map.py:390-561``)."""

    def __init__(self, shape: Tuple[int, int], is_linear: bool = False, lipschitz_cst: float = np.inf,
                 diff_lipschitz_cst: float = np.inf):
        super(DifferentiableMap, self).__init__(shape=shape, is_linear=is_linear, is_differentiable=True)
        self.lipschitz_cst = lipschitz_cst
        self.diff_lipschitz_cst = diff_lipschitz_cst

    def jacobianT(self, arg):
        """Transpose of the Jacobian at ``arg`` (public edge; device path is ``_jacobianT``)."""
        if type(self)._jacobianT is DifferentiableMap._jacobianT:
            raise NotImplementedError('DifferentiableMap subclasses must implement jacobianT or _jacobianT')
        x = to_device(arg, self._in_dtype(arg)) if is_array(arg) else arg
        return like_input(self._jacobianT(x), arg)

    def _jacobianT(self, x):
        return self.jacobianT(x)

    def gradient(self, arg):
        return self.jacobianT(arg)

    def _gradient(self, x: torch.Tensor):
        return self._jacobianT(x)

    def compute_lipschitz_cst(self):
        pass

    def compute_diff_lipschitz_cst(self):
        pass

    def shifter(self, shift) -> 'DiffMapShifted':
        return DiffMapShifted(map=self, shift=shift)

    def __add__(self, other):
        if isinstance(other, DifferentiableMap):
            return DiffMapSum(self, other)
        if isinstance(other, Map):
            return MapSum(self, other)
        raise NotImplementedError

    def __radd__(self, other):
        if isinstance(other, DifferentiableMap):
            return DiffMapSum(self, other)
        if isinstance(other, Map):
            return MapSum(self, other)
        raise NotImplementedError

    def __mul__(self, other):
        if isinstance(other, Number):
            other = self._wrap_scalar(other, self.shape[1])
        if is_array(other):
            return self(other)
        if isinstance(other, DifferentiableMap):
            return DiffMapComp(self, other)
        if isinstance(other, Map):
            return MapComp(self, other)
        raise NotImplementedError

    def __rmul__(self, other):
        if isinstance(other, Number):
            other = self._wrap_scalar(other, self.shape[0])
        if isinstance(other, DifferentiableMap):
            return DiffMapComp(other, self)
        if isinstance(other, Map):
            return MapComp(other, self)
        raise NotImplementedError


class DiffMapShifted(MapShifted, DifferentiableMap):
    def __init__(self, map: DifferentiableMap, shift):
        MapShifted.__init__(self, map=map, shift=shift)
        DifferentiableMap.__init__(self, shape=self.shape, is_linear=self.is_linear,
                                   lipschitz_cst=self.map.lipschitz_cst,
                                   diff_lipschitz_cst=self.map.diff_lipschitz_cst)

    def _jacobianT(self, x):
        return self.map._jacobianT(self._shifted(x))


class DiffMapSum(MapSum, DifferentiableMap):
    def __init__(self, map1: DifferentiableMap, map2: DifferentiableMap):
        MapSum.__init__(self, map1=map1, map2=map2)
        DifferentiableMap.__init__(self, shape=self.shape, is_linear=self.is_linear,
                                   lipschitz_cst=self.map1.lipschitz_cst + self.map2.lipschitz_cst,
                                   diff_lipschitz_cst=self.map1.diff_lipschitz_cst + self.map2.diff_lipschitz_cst)

    def _jacobianT(self, x):
        return _add_any(self.map1._jacobianT(x), self.map2._jacobianT(x))


class DiffMapComp(MapComp, DifferentiableMap):
    def __init__(self, map1: DifferentiableMap, map2: DifferentiableMap):
        from ..linop.base import HomothetyMap

        MapComp.__init__(self, map1=map1, map2=map2)
        lipschitz_cst = self.map2.lipschitz_cst * self.map1.lipschitz_cst
        if isinstance(map1, HomothetyMap):      # reference core/map.py:602-605
            diff_lipschitz_cst = self.map1.diff_lipschitz_cst * self.map2.diff_lipschitz_cst
        else:
            diff_lipschitz_cst = self.map1.diff_lipschitz_cst * self.map2.diff_lipschitz_cst * self.map2.lipschitz_cst
        DifferentiableMap.__init__(self, shape=self.shape, is_linear=self.is_linear,
                                   lipschitz_cst=lipschitz_cst, diff_lipschitz_cst=diff_lipschitz_cst)

    def _jacobianT(self, x):
        from ..linop.base import HomothetyMap
        inner = self.map2._apply(x) if not isinstance(self.map1, HomothetyMap) else None  # Homothety ignores its arg
        return jac_mul(self.map2._jacobianT(x), self.map1._jacobianT(inner))


class MapStack(Map):
    r"""Stack maps vertically (``axis=0``) or horizontally (``axis=1``) (reference ``core/map.py:613-751``).

    ``n_jobs`` / ``joblib_backend`` are accepted for signature compatibility; blocks run back-to-back on the
    current CUDA stream (the device, not a process pool, provides the parallelism).
    """

    def __init__(self, *maps: Map, axis: int, n_jobs: int = 1, joblib_backend: str = 'loky'):
        self.maps = list(maps)
        if np.abs(axis) > 1:
            raise ValueError('Axis must be one of {0, 1,-1}.')
        self.axis = int(axis)
        self.is_linear_list = [m.is_linear for m in self.maps]
        self.is_differentiable_list = [m.is_differentiable for m in self.maps]
        self.shapes = np.array([m.shape for m in self.maps])
        self.block_sizes = [m.shape[axis] for m in self.maps]
        self.sections = np.cumsum(self.block_sizes)
        self.n_jobs = n_jobs
        self.joblib_backend = joblib_backend
        if not self.is_valid_stack():
            raise ValueError('Inconsistent map shapes for  stacking.')
        Map.__init__(self, shape=self.get_shape(), is_linear=bool(np.all(self.is_linear_list)),
                     is_differentiable=bool(np.all(self.is_differentiable_list)))

    def _split(self, x):
        edges = [0] + [int(s) for s in self.sections]
        return [x[edges[i]:edges[i + 1]] for i in range(len(self.maps))]

    def _apply(self, x):
        if self.axis == 0:
            outs = [m._apply(x) for m in self.maps]
            outs = [o.reshape(-1) if isinstance(o, torch.Tensor) else
                    torch.full((1,), float(o), dtype=x.dtype, device=x.device) for o in outs]
            return torch.cat(outs)
        result = None
        for m, xi in zip(self.maps, self._split(x)):
            v = m._apply(xi)
            result = v if result is None else _add_any(result, v)
        return result

    def is_valid_stack(self) -> bool:
        col_sizes = [m.shape[1 - self.axis] for m in self.maps]
        return np.unique(col_sizes).size == 1

    def get_shape(self) -> Tuple[int, int]:
        sizes = [m.shape[self.axis] for m in self.maps]
        if self.axis == 0:
            return (int(np.sum(sizes)), self.maps[0].shape[1 - self.axis])
        return (self.maps[0].shape[1 - self.axis], int(np.sum(sizes)))


class MapVStack(MapStack):
    def __init__(self, *maps: Map, n_jobs: int = 1, joblib_backend: str = 'loky'):
        super(MapVStack, self).__init__(*maps, axis=0, n_jobs=n_jobs, joblib_backend=joblib_backend)


class MapHStack(MapStack):
    def __init__(self, *maps: Map, n_jobs: int = 1, joblib_backend: str = 'loky'):
        super(MapHStack, self).__init__(*maps, axis=1, n_jobs=n_jobs, joblib_backend=joblib_backend)


class DiffMapStack(MapStack, DifferentiableMap):
    def __init__(self, *diffmaps: DifferentiableMap, axis: int, n_jobs: int = 1, joblib_backend: str = 'loky'):
        MapStack.__init__(self, *diffmaps, axis=axis, n_jobs=n_jobs, joblib_backend=joblib_backend)
        if axis == 0:       # reference core/map.py:928-933
            lipschitz_cst = np.sqrt(np.sum([m.lipschitz_cst ** 2 for m in self.maps]))
            diff_lipschitz_cst = np.sqrt(np.sum([m.diff_lipschitz_cst ** 2 for m in self.maps]))
        else:
            lipschitz_cst = np.max([m.lipschitz_cst for m in self.maps])
            diff_lipschitz_cst = np.max([m.diff_lipschitz_cst for m in self.maps])
        DifferentiableMap.__init__(self, shape=self.shape, is_linear=self.is_linear, lipschitz_cst=lipschitz_cst,
                                   diff_lipschitz_cst=diff_lipschitz_cst)

    def _jacobianT(self, x):
        from ..linop.base import LinOpHStack, LinOpVStack
        from .linop import LinearOperator
        jl = []
        args = [x] * len(self.maps) if self.axis == 0 else self._split(x)
        for m, a in zip(self.maps, args):
            j = m._jacobianT(a)
            if not isinstance(j, LinearOperator):
                raise NotImplementedError('stacked Jacobians of non-linear maps')
            jl.append(j)
        stack = LinOpVStack if self.axis == 0 else LinOpHStack
        return stack(*jl, n_jobs=self.n_jobs, joblib_backend=self.joblib_backend)


class DiffMapVStack(DiffMapStack):
    def __init__(self, *diffmaps: DifferentiableMap, n_jobs: int = 1, joblib_backend: str = 'loky'):
        super(DiffMapVStack, self).__init__(*diffmaps, axis=0, n_jobs=n_jobs, joblib_backend=joblib_backend)


class DiffMapHStack(DiffMapStack):
    def __init__(self, *diffmaps: DifferentiableMap, n_jobs: int = 1, joblib_backend: str = 'loky'):
        super(DiffMapHStack, self).__init__(*diffmaps, axis=1, n_jobs=n_jobs, joblib_backend=joblib_backend)
