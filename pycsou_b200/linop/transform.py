r"""Orthonormal 2-D discrete Fourier transform as a :class:`UnitaryOperator`.

The reference's ``pycsou/linop/transform.py`` is an empty file: it ships no DFT operator, and its
users write one against the ``UnitaryOperator`` API (``core/linop.py:574-615``), as the compressed-sensing MRI
configuration does (BASELINE.json configs[3]; generator ``tests/golden/gen_golden.py:_FFT2``).  ``DFT2`` is that
operator for device tensors: ``DFT2(shape)(x) = fft2(x.reshape(shape), norm='ortho').ravel()``, adjoint = inverse.

The transform itself is a plain library call (cuFFT through ``torch.fft``, exactly as the reference path would call
pocketfft); everything around it on the solver path (masking, proxes, updates, norms) runs on this package's kernels.
A leading batch dimension is supported: a flat vector of ``B * H * W`` samples is transformed as ``B`` images.
"""
from __future__ import annotations

import numpy as np
import torch

from ..core.linop import UnitaryOperator

__all__ = ['DFT2']


class DFT2(UnitaryOperator):
    def __init__(self, shape, dtype='complex128', batch: int = 1):
        if len(shape) != 2:
            raise ValueError('DFT2 needs a 2-D shape')
        self.array_shape = (int(shape[0]), int(shape[1]))
        self.batch = int(batch)
        dt = np.dtype(dtype)
        if dt.kind != 'c':
            raise ValueError('DFT2 works on complex vectors (complex64 / complex128)')
        super().__init__(size=self.batch * self.array_shape[0] * self.array_shape[1], dtype=dt)

    def _img(self, x: torch.Tensor) -> torch.Tensor:
        if not x.is_complex():
            x = x.to(torch.complex64 if x.dtype == torch.float32 else torch.complex128)
        return x.view(self.batch, *self.array_shape)

    def _apply(self, x):
        return torch.fft.fft2(self._img(x), norm='ortho').reshape(-1)

    def _adjoint(self, y):
        return torch.fft.ifft2(self._img(y), norm='ortho').reshape(-1)
