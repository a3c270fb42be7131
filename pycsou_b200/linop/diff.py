r"""Finite-difference operators (mirror of the hot-path part of ``pycsou/linop/diff.py``).

``FirstDerivative`` (default ``kind='forward'``, ``edge=True``; reference ``diff.py:24-130``), ``SecondDerivative``
(``diff.py:133-219``), ``Gradient`` (default ``kind='centered'``, ``edge=True``; ``diff.py:777-882``) and ``Laplacian``
(``diff.py:885-957``) keep the reference's signatures and defaults.  Where the reference wraps a pylops operator,
these factories return device-resident ``LinearOperator`` subclasses whose ``_apply`` / ``_adjoint`` launch the
stencil kernels of ``csrc/pcs_diff.cu``.  Directional and generalised derivatives and ``Integration1D`` (SURVEY.md
section 8(f) rank 3) are built on the same kernels.
"""
from __future__ import annotations

from numbers import Number
from typing import Optional, Tuple, Union

import numpy as np
import torch

from .. import _kernels as K
from .._array import to_device
from ..core.linop import LinearOperator

__all__ = ['FirstDerivative', 'SecondDerivative', 'Gradient', 'Laplacian', 'GeneralisedDerivative',
           'FirstDirectionalDerivative', 'SecondDirectionalDerivative', 'DirectionalGradient', 'DirectionalLaplacian',
           'GeneralisedLaplacian', 'Integration1D']

_KINDS = ('forward', 'centered', 'backward')


def _axis_view(size, shape, axis):
    """(outer, n, inner) view of a C-order array differentiated along ``axis``."""
    if shape is None:
        return 1, int(size), 1
    shape = tuple(int(s) for s in shape)
    if int(np.prod(shape)) != size:
        raise ValueError('product of dims must equal N!')
    axis = axis if axis >= 0 else len(shape) + axis
    if not 0 <= axis < len(shape):
        raise ValueError(f'axis {axis} is out of range for shape {shape}')
    return int(np.prod(shape[:axis], dtype=np.int64)), shape[axis], int(np.prod(shape[axis + 1:], dtype=np.int64))


class FirstDerivativeOperator(LinearOperator):
    """First derivative along one axis of an n-D array (device stencil; semantics of pylops v1.x FirstDerivative)."""

    def __init__(self, size, shape, axis, step, edge, dtype, kind):
        if kind not in _KINDS:
            raise NotImplementedError('kind must be forward, centered, or backward')
        self.outer, self.n, self.inner = _axis_view(size, shape, axis)
        self.step, self.edge, self.kind = float(step), bool(edge), kind
        self.array_shape = None if shape is None else tuple(shape)
        self.axis = axis
        super().__init__(shape=(size, size), dtype=np.dtype(dtype))

    def _lipschitz_upper_bound(self):
        # two entries of magnitude 1/h per row and per column (forward / backward): ||D|| <= 2/h
        return 2.0 / abs(self.step) if self.kind in ('forward', 'backward') else None

    def _apply(self, x):
        return K.firstderiv(torch.empty_like(x), x, self.outer, self.n, self.inner, self.kind, self.edge, self.step)

    def _adjoint(self, y):
        return K.firstderiv(torch.empty_like(y), y, self.outer, self.n, self.inner, self.kind, self.edge, self.step,
                            adjoint=True)


class SecondDerivativeOperator(LinearOperator):
    def __init__(self, size, shape, axis, step, edge, dtype):
        self.outer, self.n, self.inner = _axis_view(size, shape, axis)
        self.step, self.edge = float(step), bool(edge)
        self.array_shape = None if shape is None else tuple(shape)
        self.axis = axis
        super().__init__(shape=(size, size), dtype=np.dtype(dtype))

    def _apply(self, x):
        return K.secondderiv(torch.empty_like(x), x, self.outer, self.n, self.inner, self.edge, self.step)

    def _adjoint(self, y):
        return K.secondderiv(torch.empty_like(y), y, self.outer, self.n, self.inner, self.edge, self.step, adjoint=True)


class GradientOperator(LinearOperator):
    r"""Stack of per-axis first derivatives, output laid out ``[d/daxis0 ; d/daxis1 ; ...]`` (pylops ``Gradient``)."""

    def __init__(self, shape, step, edge, dtype, kind):
        if kind not in _KINDS:
            raise NotImplementedError('kind must be forward, centered, or backward')
        self.array_shape = tuple(int(s) for s in shape)
        nd = len(self.array_shape)
        self.steps = tuple([float(step)] * nd) if isinstance(step, Number) else tuple(float(s) for s in step)
        if len(self.steps) != nd:
            raise ValueError('one sampling step per axis is required')
        self.edge, self.kind = bool(edge), kind
        self.N = int(np.prod(self.array_shape))
        super().__init__(shape=(nd * self.N, self.N), dtype=np.dtype(dtype))

    def _lipschitz_upper_bound(self):
        # ||K||^2 <= sum over axes of ||D_axis||^2 <= sum 4 / h_axis^2 (forward / backward differences); sqrt(8) in 2-D
        if self.kind not in ('forward', 'backward'):
            return None
        return float(np.sqrt(sum(4.0 / (h * h) for h in self.steps)))

    def _apply(self, x):
        nd, N = len(self.array_shape), self.N
        out = torch.empty(nd * N, dtype=x.dtype, device=x.device)
        if nd == 2 and not x.is_complex():
            H, W = self.array_shape
            return K.grad2d(out, x, H, W, self.kind, self.edge, self.steps[0], self.steps[1])
        for ax in range(nd):
            o, n, i = _axis_view(N, self.array_shape, ax)
            K.firstderiv(out[ax * N:(ax + 1) * N], x, o, n, i, self.kind, self.edge, self.steps[ax])
        return out

    def _adjoint(self, z):
        nd, N = len(self.array_shape), self.N
        out = torch.empty(N, dtype=z.dtype, device=z.device)
        if nd == 2 and not z.is_complex():
            H, W = self.array_shape
            return K.grad2d(out, z, H, W, self.kind, self.edge, self.steps[0], self.steps[1], adjoint=True)
        for ax in range(nd):
            o, n, i = _axis_view(N, self.array_shape, ax)
            K.firstderiv(out, z[ax * N:(ax + 1) * N], o, n, i, self.kind, self.edge, self.steps[ax], adjoint=True,
                         accumulate=ax > 0)
        return out


class LaplacianOperator(LinearOperator):
    r""":math:`w_0\,\partial^2_{axis0} + w_1\,\partial^2_{axis1}` on a 2-D array (pylops ``Laplacian``)."""

    def __init__(self, shape, weights, step, edge, dtype):
        self.array_shape = tuple(int(s) for s in shape)
        if len(self.array_shape) != 2:
            raise NotImplementedError('Laplacian is implemented for 2-D arrays (the reference differentiates axes 0 and 1)')
        self.weights = (float(weights[0]), float(weights[1]))
        self.steps = (float(step[0]), float(step[1]))
        self.edge = bool(edge)
        N = int(np.prod(self.array_shape))
        super().__init__(shape=(N, N), dtype=np.dtype(dtype))

    def _run(self, x, adjoint):
        if x.is_complex():
            raise NotImplementedError('complex Laplacian')
        H, W = self.array_shape
        return K.laplacian2d(torch.empty_like(x), x, H, W, self.weights[0], self.weights[1], self.steps[0],
                             self.steps[1], self.edge, adjoint=adjoint)

    def _apply(self, x):
        return self._run(x, False)

    def _adjoint(self, y):
        return self._run(y, True)


def FirstDerivative(size: int, shape: Optional[tuple] = None, axis: int = 0, step: float = 1.0, edge: bool = True,
                    dtype: str = 'float64', kind: str = 'forward') -> LinearOperator:
    return FirstDerivativeOperator(size, shape, axis, step, edge, dtype, kind)


def SecondDerivative(size: int, shape: Optional[tuple] = None, axis: int = 0, step: float = 1.0, edge: bool = True,
                     dtype: str = 'float64') -> LinearOperator:
    return SecondDerivativeOperator(size, shape, axis, step, edge, dtype)


def Gradient(shape: tuple, step: Union[tuple, float] = 1., edge: bool = True, dtype: str = 'float64',
             kind: str = 'centered') -> LinearOperator:
    return GradientOperator(shape, step, edge, dtype, kind)


def Laplacian(shape: tuple, weights: Tuple[float] = (1, 1), step: Union[tuple, float] = 1., edge: bool = True,
              dtype: str = 'float64') -> LinearOperator:
    if isinstance(step, Number):
        step = [step] * len(shape)
    return LaplacianOperator(shape, weights, step, edge, dtype)


def _kill_edges(shape, order, axes, low=True, high=True):
    """Diagonal mask that zeroes the first / last ``order`` samples along ``axes`` (``diff.py:355-376, 1059-1067``)."""
    from .base import DiagonalOperator
    kill = np.ones(shape)
    for ax in axes:
        idx = [slice(None)] * kill.ndim
        if high:                                   # numpy's [-order:] (order 0 selects everything, as in the reference)
            idx[ax] = slice(-order, None)
            kill[tuple(idx)] = 0
        if low:
            idx[ax] = slice(None, order)
            kill[tuple(idx)] = 0
    return DiagonalOperator(kill.reshape(-1))


def GeneralisedDerivative(size: int, shape: Optional[tuple] = None, axis: int = 0, step: float = 1.0, edge: bool = True,
                          dtype: str = 'float64', kind_op='iterated', kind_diff='centered', **kwargs) -> LinearOperator:
    r"""Generalised derivative ``KillEdges * P(D)`` (reference ``diff.py:222-377``): ``kind_op`` = ``'iterated'``
    (:math:`D^N`), ``'sobolev'`` (:math:`(\alpha^2 I - D^2)^N`), ``'exponential'`` (:math:`(\alpha I + D)^N`) or
    ``'polynomial'`` (``coeffs``), composed from the stencil kernels op by op."""
    from .base import IdentityOperator, PolynomialLinearOperator
    D = FirstDerivative(size=size, shape=shape, axis=axis, step=step, edge=edge, dtype=dtype, kind=kind_diff)
    D.is_symmetric = False
    D2 = SecondDerivative(size=size, shape=shape, axis=axis, step=step, edge=edge, dtype=dtype)
    if kind_op == 'iterated':
        N = kwargs['order']
        Dgen, order = D ** N, N
    elif kind_op == 'sobolev':
        alpha, N = kwargs['constant'], kwargs['order']
        Dgen, order = ((alpha ** 2) * IdentityOperator(size=size) - D2) ** N, 2 * N
    elif kind_op == 'exponential':
        alpha, N = kwargs['constant'], kwargs['order']
        Dgen, order = (alpha * IdentityOperator(size=size) + D) ** N, N
    elif kind_op == 'polynomial':
        coeffs = kwargs['coeffs']
        Dgen, order = PolynomialLinearOperator(LinOp=D, coeffs=coeffs), len(coeffs) - 1
    else:
        raise NotImplementedError(
            'Supported generalised derivative types are: iterated, sobolev, exponential, polynomial.')
    kill = _kill_edges((Dgen.shape[0],) if shape is None else tuple(shape), order, (axis,),
                       low=kind_diff in ('backward', 'centered'), high=kind_diff in ('forward', 'centered'))
    return kill * Dgen


class FirstDirectionalDerivativeOperator(LinearOperator):
    r""":math:`d_v x = \sum_i v_i\,\partial_i x` (pylops ``FirstDirectionalDerivative`` = ``Sum * Diagonal(v) * Gradient``):
    the gradient kernels, one element-wise product with the direction field and a sum over the axes.  ``directions`` has
    shape ``(ndim,)`` (one direction everywhere) or ``(ndim, *shape)`` (a direction per pixel)."""

    def __init__(self, shape, directions, step, edge, dtype, kind):
        self.array_shape = tuple(int(s) for s in shape)
        nd, N = len(self.array_shape), int(np.prod(shape))
        v = np.asarray(directions, dtype=np.float64)
        if v.shape == (nd,):
            v = np.repeat(v, N)
        elif v.shape == (nd,) + self.array_shape:
            v = v.reshape(-1)
        else:
            raise ValueError('directions must have shape (ndim,) or (ndim, *shape)')
        self.directions, self.N, self.nd = v, N, nd
        self.grad = GradientOperator(shape, step, edge, dtype, kind)
        self._v = {}
        super().__init__(shape=(N, N), dtype=np.dtype(dtype))

    def _vdev(self, like):
        key = (like.dtype, like.device)
        if key not in self._v:
            self._v[key] = to_device(self.directions, like.dtype)
        return self._v[key]

    def _apply(self, x):
        if x.is_complex():
            raise NotImplementedError('complex directional derivative')
        g = self.grad._apply(x)
        K.mul(g, g, self._vdev(x))
        N = self.N
        out = K.lincomb2(torch.empty_like(x), 1.0, g[:N], 1.0, g[N:2 * N]) if self.nd > 1 else g[:N].clone()
        for ax in range(2, self.nd):
            K.lincomb2(out, 1.0, out, 1.0, g[ax * N:(ax + 1) * N])
        return out

    def _adjoint(self, y):
        if y.is_complex():
            raise NotImplementedError('complex directional derivative')
        v, N = self._vdev(y), self.N
        t = torch.empty(self.nd * N, dtype=y.dtype, device=y.device)
        for ax in range(self.nd):
            K.mul(t[ax * N:(ax + 1) * N], v[ax * N:(ax + 1) * N], y)
        return self.grad._adjoint(t)


class SecondDirectionalDerivativeOperator(LinearOperator):
    r""":math:`-d_v^T d_v` with the two outermost samples of every axis zeroed in the OUTPUT (pylops
    ``SecondDirectionalDerivative`` followed by the reference's edge-killing ``DiagonalOperator``, ``diff.py:596-606``)."""

    def __init__(self, shape, directions, step, edge, dtype):
        self.first = FirstDirectionalDerivativeOperator(shape, directions, step, edge, dtype, 'centered')
        kill = np.ones(shape)
        for axis in range(len(shape)):
            kill = np.swapaxes(kill, axis, 0)
            kill[-2:] = 0
            kill[:2] = 0
            kill = np.swapaxes(kill, 0, axis)
        self.kill_edges = kill.reshape(-1)
        self._k = {}
        N = self.first.N
        super().__init__(shape=(N, N), dtype=np.dtype(dtype))

    def _kdev(self, like):
        key = (like.dtype, like.device)
        if key not in self._k:
            self._k[key] = to_device(self.kill_edges, like.dtype)
        return self._k[key]

    def _apply(self, x):
        y = self.first._adjoint(self.first._apply(x))
        K.mul(y, y, self._kdev(x))
        return K.scale(y, -1.0, y)

    def _adjoint(self, z):
        t = K.mul(torch.empty_like(z), z, self._kdev(z))
        y = self.first._adjoint(self.first._apply(t))
        return K.scale(y, -1.0, y)


def FirstDirectionalDerivative(shape: tuple, directions, step=1., edge: bool = True, dtype: str = 'float64',
                               kind: str = 'centered') -> LinearOperator:
    return FirstDirectionalDerivativeOperator(shape, directions, step, edge, dtype, kind)


def SecondDirectionalDerivative(shape: tuple, directions, step=1., edge: bool = True,
                                dtype: str = 'float64') -> LinearOperator:
    return SecondDirectionalDerivativeOperator(shape, directions, step, edge, dtype)


def DirectionalGradient(first_directional_derivatives) -> LinearOperator:
    """Stack of first directional derivatives (reference ``diff.py:609-688``)."""
    from .base import LinOpVStack
    return LinOpVStack(*first_directional_derivatives)


def DirectionalLaplacian(second_directional_derivatives, weights=None) -> LinearOperator:
    """Weighted sum of second directional derivatives, exactly as the reference accumulates it (``diff.py:691-774``: the
    first weight is not applied)."""
    lap = second_directional_derivatives[0]
    if weights is None:
        weights = [1.] * len(second_directional_derivatives)
    elif len(weights) != len(second_directional_derivatives):
        raise ValueError('The number of weights and operators provided differ.')
    for i in range(1, len(second_directional_derivatives)):
        lap += weights[i] * second_directional_derivatives[i]
    return lap
def GeneralisedLaplacian(shape: Optional[tuple] = None, step: Union[tuple, float] = 1., edge: bool = True,
                         dtype: str = 'float64', kind='iterated', **kwargs) -> LinearOperator:
    r"""Generalised Laplacian ``KillEdges * P(\Delta)`` (reference ``diff.py:960-1068``): ``kind`` = ``'iterated'``,
    ``'sobolev'`` (:math:`(\alpha^2 I - \Delta)^N`) or ``'polynomial'``.  The Laplacian is flagged symmetric as in the
    reference (``:1039``), which only matters for the adjoint of the polynomial form."""
    from .base import IdentityOperator, PolynomialLinearOperator
    Delta = Laplacian(shape=shape, step=step, edge=edge, dtype=dtype)
    Delta.is_symmetric = True
    if kind == 'iterated':
        N = kwargs['order']
        Dgen, order = Delta ** N, 2 * N
    elif kind == 'sobolev':
        alpha, N = kwargs['constant'], kwargs['order']
        Dgen, order = ((alpha ** 2) * IdentityOperator(size=shape[0] * shape[1]) - Delta) ** N, 2 * N
    elif kind == 'polynomial':
        coeffs = kwargs['coeffs']
        Dgen, order = PolynomialLinearOperator(LinOp=Delta, coeffs=coeffs), 2 * (len(coeffs) - 1)
    else:
        raise NotImplementedError('Supported generalised derivative types are: iterated, sobolev, polynomial.')
    return _kill_edges(tuple(shape), order, range(len(shape))) * Dgen


class Integration1DOperator(LinearOperator):
    r"""Causal integration ``y[i] = step * sum_{j<=i} x[j]`` along ``axis`` (reference ``diff.py:1071-1137`` ->
    ``pylops.CausalIntegration(halfcurrent=False)``); the adjoint runs the sum from the far end."""

    def __init__(self, size, shape, axis, step, dtype):
        shape = (int(size),) if shape is None else tuple(int(v) for v in shape)
        if int(np.prod(shape)) != size:
            raise ValueError('product of dims must equal N!')
        self.array_shape, self.axis, self.step = shape, int(axis) % len(shape), float(step)
        self.view = _axis_view(int(size), shape, self.axis)
        super().__init__(shape=(size, size), dtype=np.dtype(dtype))

    def _apply(self, x):
        o, n, i = self.view
        return K.cumsum(torch.empty_like(x), x, o, n, i, self.step)

    def _adjoint(self, y):
        o, n, i = self.view
        return K.cumsum(torch.empty_like(y), y, o, n, i, self.step, adjoint=True)


def Integration1D(size: int, shape: Optional[tuple] = None, axis: int = 0, step: float = 1.,
                  dtype='float64') -> LinearOperator:
    return Integration1DOperator(size, shape, axis, step, dtype)
