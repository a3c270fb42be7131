"""numpy/scipy restatement of the pylops v1.x operators on the pycsou hot path.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

pylops is a third-party dependency of the reference (``requirements.txt:4``: ``pylops >= 1.9.2``;
the keyword API the reference uses -- ``N=, dims=, dir=, sampling=, nodir=, M=, iava=`` -- exists
only in pylops 1.x) and is NOT vendored under /root/reference, nor installable offline.  This module
restates the *published* v1.x semantics of exactly the operators the reference calls:

=====================  ==========================================  ==============================
stand-in               reference call site                         reference pin (doctest)
=====================  ==========================================  ==============================
FirstDerivative        pycsou/linop/diff.py:128-130                diff.py:72-78, core/map.py:418-423
SecondDerivative       pycsou/linop/diff.py:218-219                diff.py:210-211 (formula)
Gradient (VStack)      pycsou/linop/diff.py:882                    diff.py:814-820, base.py:213-233
Laplacian              pycsou/linop/diff.py:957                    diff.py:946-957 (formula)
Restriction            pycsou/linop/sampling.py:121                sampling.py:72-89
Convolve1D             pycsou/linop/conv.py:163                    conv.py:67-73
Convolve2D             pycsou/linop/conv.py:294                    conv.py:209-217
=====================  ==========================================  ==============================

Even-length filters: pylops 1.x releases differ in how they pad an even filter before calling
scipy 'same' convolution.  The stand-in implements the *documented intent* of the reference
(conv.py:159-162, 285-293: the filter is centred at index ``offset``):
``y[i] = sum_k h[k] x[i + offset - k]`` with zero extension, adjoint ``sum_k h[k] y[i + k - offset]``.
For odd filter sizes (all BASELINE configs) every 1.x release agrees with this.

Each operator exposes the scipy-LinearOperator-like surface pycsou's ``PyLopLinearOperator``
(linop/base.py:24-54) consumes: ``shape``, ``dtype``, ``explicit``, ``matvec``, ``rmatvec``.
"""
from __future__ import annotations

import numpy as np
import scipy.signal as _sps


class LinearOperator:
    """Minimal stand-in for ``pylops.LinearOperator`` (scipy LinearOperator subclass in pylops)."""

    def __init__(self, Op=None, explicit=False):
        self.explicit = explicit
        if Op is not None:
            self.Op = Op
            self.shape = Op.shape
            self.dtype = Op.dtype

    # scipy.sparse.linalg.LinearOperator.matvec semantics: accept (N,) or (N,1), no dtype cast.
    def matvec(self, x):
        x = np.asanyarray(x)
        M, N = self.shape
        if x.shape != (N,) and x.shape != (N, 1):
            raise ValueError('dimension mismatch')
        y = self._matvec(x.reshape(-1))
        return y.reshape(M) if x.ndim == 1 else y.reshape(M, 1)

    def rmatvec(self, x):
        x = np.asanyarray(x)
        M, N = self.shape
        if x.shape != (M,) and x.shape != (M, 1):
            raise ValueError('dimension mismatch')
        y = self._rmatvec(x.reshape(-1))
        return y.reshape(N) if x.ndim == 1 else y.reshape(N, 1)

    def _matvec(self, x):
        return self.Op.matvec(x)

    def _rmatvec(self, x):
        return self.Op.rmatvec(x)

    def __mul__(self, x):
        if np.isscalar(x):
            return _Scaled(self, x)
        return self.matvec(x)

    def __rmul__(self, a):
        return _Scaled(self, a)

    def __add__(self, other):
        return _Sum(self, other)


class _Scaled(LinearOperator):
    def __init__(self, A, alpha):
        self.A, self.alpha = A, alpha
        self.shape, self.dtype, self.explicit = A.shape, A.dtype, False

    def _matvec(self, x):
        return self.alpha * self.A.matvec(x)

    def _rmatvec(self, x):
        return np.conj(self.alpha) * self.A.rmatvec(x)


class _Sum(LinearOperator):
    def __init__(self, A, B):
        self.A, self.B = A, B
        self.shape, self.dtype, self.explicit = A.shape, A.dtype, False

    def _matvec(self, x):
        return self.A.matvec(x) + self.B.matvec(x)

    def _rmatvec(self, x):
        return self.A.rmatvec(x) + self.B.rmatvec(x)


def _to_first(a, dims, axis):
    a = np.reshape(a, dims)
    return np.swapaxes(a, axis, 0) if axis > 0 else a


def _from_first(a, axis):
    return (np.swapaxes(a, 0, axis) if axis > 0 else a).ravel()


class FirstDerivative(LinearOperator):
    """pylops v1.x ``FirstDerivative(N, dims, dir, sampling, edge, dtype, kind)``."""

    def __init__(self, N, dims=None, dir=0, sampling=1., edge=False, dtype='float64', kind='centered'):
        self.N = int(N)
        self.dims = (self.N,) if dims is None else tuple(dims)
        if int(np.prod(self.dims)) != self.N:
            raise ValueError('product of dims must equal N!')
        self.dir = dir if dir >= 0 else len(self.dims) + dir
        self.sampling, self.edge, self.kind = sampling, edge, kind
        if kind not in ('forward', 'centered', 'backward'):
            raise NotImplementedError('kind must be forward, centered, or backward')
        self.shape, self.dtype, self.explicit = (self.N, self.N), np.dtype(dtype), False

    def _matvec(self, x):
        x = _to_first(x, self.dims, self.dir)
        y = np.zeros(x.shape, self.dtype)
        h = self.sampling
        if self.kind == 'forward':
            y[:-1] = (x[1:] - x[:-1]) / h
        elif self.kind == 'backward':
            y[1:] = (x[1:] - x[:-1]) / h
        else:
            y[1:-1] = (0.5 * x[2:] - 0.5 * x[:-2]) / h
            if self.edge:
                y[0] = (x[1] - x[0]) / h
                y[-1] = (x[-1] - x[-2]) / h
        return _from_first(y, self.dir)

    def _rmatvec(self, x):
        x = _to_first(x, self.dims, self.dir)
        y = np.zeros(x.shape, self.dtype)
        h = self.sampling
        if self.kind == 'forward':
            y[:-1] -= x[:-1] / h
            y[1:] += x[:-1] / h
        elif self.kind == 'backward':
            y[:-1] -= x[1:] / h
            y[1:] += x[1:] / h
        else:
            y[:-2] -= (0.5 * x[1:-1]) / h
            y[2:] += (0.5 * x[1:-1]) / h
            if self.edge:
                y[0] -= x[0] / h
                y[1] += x[0] / h
                y[-2] -= x[-1] / h
                y[-1] += x[-1] / h
        return _from_first(y, self.dir)


class SecondDerivative(LinearOperator):
    """pylops v1.x ``SecondDerivative(N, dims, dir, sampling, edge, dtype)``."""

    def __init__(self, N, dims=None, dir=0, sampling=1., edge=False, dtype='float64'):
        self.N = int(N)
        self.dims = (self.N,) if dims is None else tuple(dims)
        if int(np.prod(self.dims)) != self.N:
            raise ValueError('product of dims must equal N!')
        self.dir = dir if dir >= 0 else len(self.dims) + dir
        self.sampling, self.edge = sampling, edge
        self.shape, self.dtype, self.explicit = (self.N, self.N), np.dtype(dtype), False

    def _matvec(self, x):
        x = _to_first(x, self.dims, self.dir)
        y = np.zeros(x.shape, self.dtype)
        h2 = self.sampling ** 2
        y[1:-1] = (x[2:] - 2 * x[1:-1] + x[0:-2]) / h2
        if self.edge:
            y[0] = (x[0] - 2 * x[1] + x[2]) / h2
            y[-1] = (x[-3] - 2 * x[-2] + x[-1]) / h2
        return _from_first(y, self.dir)

    def _rmatvec(self, x):
        x = _to_first(x, self.dims, self.dir)
        y = np.zeros(x.shape, self.dtype)
        h2 = self.sampling ** 2
        y[0:-2] += x[1:-1] / h2
        y[1:-1] -= 2 * x[1:-1] / h2
        y[2:] += x[1:-1] / h2
        if self.edge:
            y[0] += x[0] / h2
            y[1] -= 2 * x[0] / h2
            y[2] += x[0] / h2
            y[-3] += x[-1] / h2
            y[-2] -= 2 * x[-1] / h2
            y[-1] += x[-1] / h2
        return _from_first(y, self.dir)


class VStack(LinearOperator):
    """pylops ``VStack``: concatenated outputs; adjoint sums the per-block adjoints."""

    def __init__(self, ops, dtype=None):
        self.ops = list(ops)
        self.nops = [op.shape[0] for op in self.ops]
        self.shape = (int(np.sum(self.nops)), self.ops[0].shape[1])
        self.dtype = np.dtype(dtype) if dtype is not None else self.ops[0].dtype
        self.explicit = False
        self._edges = np.concatenate(([0], np.cumsum(self.nops)))

    def _matvec(self, x):
        return np.concatenate([op.matvec(x) for op in self.ops])

    def _rmatvec(self, x):
        y = np.zeros(self.shape[1], dtype=np.result_type(self.dtype, x.dtype))
        for i, op in enumerate(self.ops):
            y = y + op.rmatvec(x[self._edges[i]:self._edges[i + 1]])
        return y


def Gradient(dims, sampling=1, edge=False, dtype='float64', kind='centered'):
    """pylops v1.x ``Gradient``: VStack over axes 0..ndim-1 of FirstDerivative."""
    ndims = len(dims)
    if isinstance(sampling, (int, float)):
        sampling = [sampling] * ndims
    N = int(np.prod(dims))
    return VStack([FirstDerivative(N, dims=dims, dir=i, sampling=sampling[i], edge=edge, dtype=dtype, kind=kind)
                   for i in range(ndims)])


def Laplacian(dims, dirs=(0, 1), weights=(1, 1), sampling=(1, 1), edge=False, dtype='float64'):
    """pylops v1.x ``Laplacian``: w0*D2(dir0) + w1*D2(dir1)."""
    N = int(np.prod(dims))
    a = weights[0] * SecondDerivative(N, dims=dims, dir=dirs[0], sampling=sampling[0], edge=edge, dtype=dtype)
    b = weights[1] * SecondDerivative(N, dims=dims, dir=dirs[1], sampling=sampling[1], edge=edge, dtype=dtype)
    return a + b


class Restriction(LinearOperator):
    """pylops v1.x ``Restriction(M, iava, dims, dir, dtype, inplace)``."""

    def __init__(self, M, iava, dims=None, dir=0, dtype='float64', inplace=True):
        self.M, self.dir = int(M), dir
        self.iava = np.asarray(iava)
        if dims is None:
            self.N, self.dims, self.reshape = len(self.iava), (self.M,), False
        else:
            if int(np.prod(dims)) != self.M:
                raise ValueError('product of dims must equal M!')
            self.dims = tuple(dims)
            self.dimsd = list(dims)
            self.dimsd[dir] = len(self.iava)
            self.iavareshape = [1] * dir + [len(self.iava)] + [1] * (len(self.dims) - dir - 1)
            self.N, self.reshape = int(np.prod(self.dimsd)), True
        self.inplace = inplace
        self.shape, self.dtype, self.explicit = (self.N, self.M), np.dtype(dtype), False

    def _matvec(self, x):
        if not self.reshape:
            return x[self.iava]
        return np.take(np.reshape(x, self.dims), self.iava, axis=self.dir).ravel()

    def _rmatvec(self, x):
        y = np.zeros(self.dims, dtype=self.dtype)
        if not self.reshape:
            y[self.iava] = x
            return y
        x = np.reshape(x, self.dimsd)
        np.put_along_axis(y, np.reshape(self.iava, self.iavareshape), x, axis=self.dir)
        return y.ravel()


def _centre_pad(h, offset, axis_sizes):
    """Zero-pad ``h`` so that scipy's 'same' centring lands on ``offset`` along every axis."""
    pad = []
    for nh, off in zip(axis_sizes, offset):
        d = 2 * (nh // 2 - int(off))
        if nh % 2 == 0:
            d -= 1
        pad.append((d if d > 0 else 0, -d if d < 0 else 0))
    return np.pad(h, pad, mode='constant') if any(p != (0, 0) for p in pad) else h


class Convolve1D(LinearOperator):
    """pylops v1.x ``signalprocessing.Convolve1D(N, h, offset, dims, dir, dtype, method)``."""

    def __init__(self, N, h, offset=0, dims=None, dir=0, dtype='float64', method=None):
        h = np.asarray(h)
        if offset > len(h) - 1:
            raise ValueError('offset must be smaller than len(h) - 1')
        self.h = _centre_pad(h, (offset,), (len(h),))
        self.hstar = np.flip(self.h, axis=-1)
        self.dir = dir
        if dims is None:
            self.dims, self.reshape = (int(N),), False
            method = 'direct' if method is None else method
            if method not in ('direct', 'fft'):
                raise NotImplementedError('method must be direct or fft')
        else:
            if int(np.prod(dims)) != N:
                raise ValueError('product of dims must equal N!')
            self.dims, self.reshape = tuple(dims), True
            method = 'fft' if method is None else method
            if method not in ('fft', 'overlapadd'):
                raise NotImplementedError('method must be fft or overlapadd')
            shp = [1] * len(self.dims)
            shp[dir] = len(self.h)
            self.h = self.h.reshape(shp)
            self.hstar = self.hstar.reshape(shp)
        self.method = method
        self.shape, self.dtype, self.explicit = (int(N), int(N)), np.dtype(dtype), False

    def _apply(self, x, h):
        if not self.reshape:
            return _sps.convolve(x, h, mode='same', method=self.method)
        x = np.reshape(x, self.dims)
        f = _sps.fftconvolve if self.method == 'fft' else _sps.oaconvolve
        return f(x, h, mode='same', axes=self.dir).ravel()

    def _matvec(self, x):
        return self._apply(x, self.h)

    def _rmatvec(self, x):
        return self._apply(x, self.hstar)


class Convolve2D(LinearOperator):
    """pylops v1.x ``signalprocessing.Convolve2D(N, h, dims, offset, nodir, dtype, method)`` (2-D dims only)."""

    def __init__(self, N, h, dims, offset=(0, 0), nodir=None, dtype='float64', method='fft'):
        h = np.asarray(h)
        if h.ndim != 2:
            raise ValueError('h must be 2-dimensional')
        if nodir is not None or len(dims) != 2:
            raise NotImplementedError('stand-in supports 2-D dims with nodir=None (the only form pycsou uses)')
        if offset[0] > h.shape[0] - 1 or offset[1] > h.shape[1] - 1:
            raise ValueError('offset must be smaller than h.shape - 1')
        if int(np.prod(dims)) != N:
            raise ValueError('product of dims must equal N!')
        self.h = _centre_pad(h, offset, h.shape)
        self.dims, self.method = tuple(dims), method
        self.shape, self.dtype, self.explicit = (int(N), int(N)), np.dtype(dtype), False

    def _matvec(self, x):
        return _sps.convolve(np.reshape(x, self.dims), self.h, mode='same', method=self.method).ravel()

    def _rmatvec(self, x):
        return _sps.correlate(np.reshape(x, self.dims), self.h, mode='same', method=self.method).ravel()
