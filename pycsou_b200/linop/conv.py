r"""Convolution operators (mirror of ``pycsou/linop/conv.py:20-295``).

``Convolve1D(size, filter, reshape_dims, axis, dtype, method)`` and ``Convolve2D(size, filter, shape, dtype, method)``
keep the reference's signatures and its centring rule (filter centred at ``n//2`` for odd sizes, ``n//2 - 1`` for even
ones, ``conv.py:159-162, 285-293``): zero-padded "same" convolution forward, correlation (the exact transpose) in
adjoint mode.  The reference evaluates them with ``scipy.signal`` (direct or pocketfft); here compact filters run on
direct shared-memory stencil kernels, and a 2-D filter that is numerically rank-1 (e.g. any Gaussian) is detected by
SVD at construction and applied as a row pass + column pass (``csrc/pcs_conv.cu``).  Large non-separable 2-D filters
(more than 15 x 15 taps) take an FFT path: padded real transforms by cuFFT (where the reference calls pocketfft), the
spectral multiply and the crop + normalisation + residual ``A x - y`` fused into two kernels of this library.
``method`` is validated as in pylops but does not select the arithmetic path; results agree with either scipy method
to rounding.
"""
from __future__ import annotations

import os
from typing import Optional, Union

import numpy as np
import torch

from .. import _kernels as K
from .._array import to_device
from ..core.linop import LinearOperator

__all__ = ['Convolve1D', 'Convolve2D', 'MovingAverage1D', 'MovingAverage2D']

SEPARABLE_RTOL = 1e-13     # second singular value / first below this -> treated as exactly rank-1
FFT_MIN_TAPS = 15 * 15     # non-separable filters with more taps than this take the FFT path


def _centre(n: int) -> int:
    return n // 2 - 1 if n % 2 == 0 else n // 2


def _host(a):
    return a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)


class Convolve1DOperator(LinearOperator):
    def __init__(self, size, filter, reshape_dims, axis, dtype, method):
        h = _host(filter).reshape(-1)
        if np.iscomplexobj(h):
            raise NotImplementedError('the filter should be real-valued (reference conv.py docstring)')
        if reshape_dims is None:
            if method not in (None, 'direct', 'fft'):
                raise NotImplementedError('method must be direct or fft')
            self.outer, self.n, self.inner = 1, int(size), 1
        else:
            if method not in (None, 'fft', 'overlapadd'):
                raise NotImplementedError('method must be fft or overlapadd')
            dims = tuple(int(d) for d in reshape_dims)
            if int(np.prod(dims)) != size:
                raise ValueError('product of dims must equal N!')
            self.outer = int(np.prod(dims[:axis], dtype=np.int64))
            self.n, self.inner = dims[axis], int(np.prod(dims[axis + 1:], dtype=np.int64))
        self.filter = h.astype(np.float64)
        self.offset = _centre(h.size)
        self.method = method
        self._taps = {}
        super().__init__(shape=(size, size), dtype=np.dtype(dtype))

    def _h(self, like):
        key = (like.dtype, like.device)
        if key not in self._taps:
            self._taps[key] = to_device(self.filter, like.dtype)
        return self._taps[key]

    def _apply(self, x):
        return K.conv1d(torch.empty_like(x), x, self._h(x), self.offset, self.outer, self.n, self.inner)

    def _adjoint(self, y):
        return K.conv1d(torch.empty_like(y), y, self._h(y), self.offset, self.outer, self.n, self.inner, adjoint=True)


class Convolve2DOperator(LinearOperator):
    def __init__(self, size, filter, shape, dtype, method):
        h = _host(filter)
        if h.ndim != 2:
            raise ValueError('filter must be a 2D array.')
        if np.iscomplexobj(h):
            raise NotImplementedError('the filter should be real-valued')
        if len(shape) != 2 or int(np.prod(shape)) != size:
            raise ValueError('product of dims must equal N!')
        self.array_shape = (int(shape[0]), int(shape[1]))
        self.filter = h.astype(np.float64)
        self.offset = (_centre(h.shape[0]), _centre(h.shape[1]))
        self.method = method
        # rank-1 detection: h = s0 * u0 v0^T exactly (to rounding) -> separable kernels
        u, s, vt = np.linalg.svd(self.filter)
        self.is_separable = bool(s.size == 1 or s[1] <= SEPARABLE_RTOL * s[0])
        if self.is_separable:
            sc = np.sqrt(s[0])
            self.hcol, self.hrow = u[:, 0] * sc, vt[0] * sc
            if self.hcol.sum() < 0:
                self.hcol, self.hrow = -self.hcol, -self.hrow
        self._taps = {}
        # Large non-separable filters: direct evaluation costs kh*kw FMA per pixel (and the tile no longer fits in shared
        # memory beyond ~100 x 100 taps); they take the FFT path (padded real transforms by cuFFT, spectral multiply
        # and crop / normalisation / residual fused into two kernels of this library).  PCS_CONV2D_PATH=fft|direct forces one.
        forced = os.environ.get('PCS_CONV2D_PATH', '')
        self.use_fft = (not self.is_separable) and (forced == 'fft' or (forced != 'direct' and h.size > FFT_MIN_TAPS))
        if self.use_fft:
            from scipy.fft import next_fast_len
            H, W = self.array_shape
            self.padded = (int(next_fast_len(H + h.shape[0] - 1, real=True)), int(next_fast_len(W + h.shape[1] - 1, real=True)))
        self._spectrum = {}
        super().__init__(shape=(size, size), dtype=np.dtype(dtype))

    def _hf(self, like):
        """Half spectrum of the zero-padded filter (once per dtype / device)."""
        key = (like.dtype, like.device)
        if key not in self._spectrum:
            h = to_device(self.filter, like.dtype).reshape(self.filter.shape)
            self._spectrum[key] = torch.fft.rfft2(h, s=self.padded).contiguous()
        return self._spectrum[key]

    def _run_fft(self, x, adjoint, sub):
        H, W = self.array_shape
        P0, P1 = self.padded
        X = torch.fft.rfft2(x.view(H, W), s=(P0, P1)).contiguous()           # (no copy when cuFFT's layout is already dense)
        K.spectral_mul(X, self._hf(x), conj_b=adjoint)
        full = torch.fft.irfft2(X, s=(P0, P1), norm='forward')              # unnormalised: 1/(P0 P1) is applied by the crop
        # forward: y[i] = full[i + off]; adjoint (correlation): y[i] = c[i - off], indices wrapping into the zero padding
        s0, s1 = (-self.offset[0], -self.offset[1]) if adjoint else self.offset
        return K.crop_shift(torch.empty_like(x), full.contiguous().reshape(-1), H, W, P0, P1, s0, s1, 1.0 / (P0 * P1), sub=sub)

    def _dev(self, like):
        key = (like.dtype, like.device)
        if key not in self._taps:
            if self.is_separable:
                self._taps[key] = (to_device(self.hcol, like.dtype), to_device(self.hrow, like.dtype))
            else:
                self._taps[key] = (to_device(self.filter, like.dtype).reshape(self.filter.shape),)
        return self._taps[key]

    def _run(self, x, adjoint, sub=None):
        if x.is_complex():
            raise NotImplementedError('complex Convolve2D')
        H, W = self.array_shape
        if self.use_fft:
            return self._run_fft(x, adjoint, sub)
        t = self._dev(x)
        out = torch.empty_like(x)
        if self.is_separable:
            return K.conv2d_sep(out, x, t[0], t[1], self.offset[0], self.offset[1], H, W, adjoint=adjoint, sub=sub)
        K.conv2d_direct(out, x, t[0], self.offset[0], self.offset[1], H, W, adjoint=adjoint)
        if sub is not None:
            K.lincomb2(out, 1.0, out, -1.0, sub)
        return out

    def _apply(self, x):
        return self._run(x, False)

    def _adjoint(self, y):
        return self._run(y, True)

    def _residual(self, x, y):
        """A x - y in one pass (used by the solvers for the data-term gradient)."""
        return self._run(x, False, sub=y)


def Convolve1D(size: int, filter, reshape_dims: Optional[tuple] = None, axis: int = 0, dtype: type = 'float64',
               method: Optional[str] = None) -> LinearOperator:
    return Convolve1DOperator(size, filter, reshape_dims, axis, dtype, method)


def Convolve2D(size: int, filter, shape: tuple, dtype: type = 'float64', method: str = 'fft') -> LinearOperator:
    return Convolve2DOperator(size, filter, shape, dtype, method)


def _odd(n: int) -> int:
    n = int(n)
    return n + 1 if n % 2 == 0 else n       # pylops.Smoothing1D / Smoothing2D make the window odd


def MovingAverage1D(window_size: int, shape: tuple, axis: int = 0, dtype='float64') -> LinearOperator:
    r"""Moving average along ``axis`` (reference ``conv.py:298-350`` -> ``pylops.Smoothing1D``): a ``Convolve1D`` with
    the constant filter ``ones(n)/n`` (``n`` made odd), centred; self-adjoint away from the array ends."""
    n = _odd(window_size)
    shape = tuple(int(v) for v in np.atleast_1d(shape))
    size = int(np.prod(shape))
    return Convolve1DOperator(size, np.full(n, 1.0 / n), shape if len(shape) > 1 else None, axis, dtype, None)


def MovingAverage2D(window_shape: Union[tuple, list], shape: tuple, dtype='float64') -> LinearOperator:
    r"""2-D moving average (reference ``conv.py:351-418`` -> ``pylops.Smoothing2D``): a ``Convolve2D`` with the
    constant (rank-1, hence separable) filter ``ones(n0, n1) / (n0 n1)``, window sizes made odd."""
    n0, n1 = _odd(window_shape[0]), _odd(window_shape[1])
    size = int(np.prod(shape))
    return Convolve2DOperator(size, np.full((n0, n1), 1.0 / (n0 * n1)), tuple(shape), dtype, 'fft')
