r"""Sampling operators (mirror of the hot-path part of ``pycsou/linop/sampling.py``).

``SubSampling`` (reference ``sampling.py:25-122``, pylops ``Restriction``), ``Masking`` (``sampling.py:125-196``) and
``DownSampling`` (``sampling.py:199-391``) keep the reference's signatures and public attributes
(``sampling_bool``, ``input_size``, ``nb_of_samples``, ``downsampling_mask``, ``output_shape`` ...).  Masks are
converted once into a sorted int64 index list on the device (``pcs_mask_count`` / ``pcs_mask_indices``); forward is a
gather, adjoint a zero-fill + scatter (``csrc/pcs_sampling.cu``).  Masks and indices are bit-exact by construction.
``Pooling`` (``sampling.py:394-536``) is a block-reduce kernel.  ``NNSampling``, ``GeneralisedVandermonde`` and
``MappedDistanceMatrix`` are out of scope and raise.
"""
from __future__ import annotations

from typing import Optional, Union

import numpy as np
import torch

from .. import _kernels as K
from .._array import device, torch_dtype
from ..core.linop import LinearOperator

__all__ = ['SubSampling', 'Masking', 'DownSampling', 'Pooling', 'NNSampling', 'GeneralisedVandermonde',
           'MappedDistanceMatrix']


class SubSamplingOperator(LinearOperator):
    """``y = x[idx]`` / ``take(X, idx, axis)``; adjoint = zeros then assignment (duplicates: last one wins)."""

    def __init__(self, size, sampling_indices, shape, axis, dtype, inplace):
        idx = np.asarray(sampling_indices.detach().cpu() if isinstance(sampling_indices, torch.Tensor)
                         else sampling_indices).reshape(-1).astype(np.int64)
        if shape is None:
            self.outer, self.n, self.inner = 1, int(size), 1
        else:
            dims = tuple(int(d) for d in shape)
            if int(np.prod(dims)) != size:
                raise ValueError('product of dims must equal M!')
            self.outer = int(np.prod(dims[:axis], dtype=np.int64))
            self.n, self.inner = dims[axis], int(np.prod(dims[axis + 1:], dtype=np.int64))
        if idx.size and (idx.min() < -self.n or idx.max() >= self.n):
            raise IndexError('sampling index out of range')
        idx = np.where(idx < 0, idx + self.n, idx)
        self.sampling_indices = idx
        # adjoint with duplicate indices: numpy assignment keeps the LAST occurrence
        _, first_rev = np.unique(idx[::-1], return_index=True)
        keep = np.sort(idx.size - 1 - first_rev)
        self._dedup = keep if keep.size != idx.size else None
        self._idx_dev = {}
        super().__init__(shape=(self.outer * idx.size * self.inner, size), dtype=np.dtype(dtype), is_symmetric=False)

    def _idx(self, dev):
        if dev not in self._idx_dev:
            full = torch.from_numpy(self.sampling_indices).to(dev)
            if self._dedup is None:
                self._idx_dev[dev] = (full, None, None)
            else:
                keep = torch.from_numpy(self._dedup.astype(np.int64)).to(dev)
                self._idx_dev[dev] = (full, keep, torch.from_numpy(self.sampling_indices[self._dedup]).to(dev))
        return self._idx_dev[dev]

    def _apply(self, x):
        full, _, _ = self._idx(x.device)
        out = torch.empty(self.shape[0], dtype=x.dtype, device=x.device)
        return K.take(out, x, full, self.outer, self.n, self.inner)

    def _adjoint(self, y):
        full, keep, kept_idx = self._idx(y.device)
        out = torch.empty(self.shape[1], dtype=y.dtype, device=y.device)
        if keep is None:
            return K.put(out, y, full, self.outer, self.n, self.inner)
        m = full.numel()
        ysel = K.take(torch.empty(self.outer * keep.numel() * self.inner, dtype=y.dtype, device=y.device), y, keep,
                      self.outer, m, self.inner)
        return K.put(out, ysel, kept_idx, self.outer, self.n, self.inner)


def SubSampling(size: int, sampling_indices, shape: Optional[tuple] = None, axis: int = 0, dtype: str = 'float64',
                inplace: bool = True) -> LinearOperator:
    return SubSamplingOperator(size, sampling_indices, shape, axis, dtype, inplace)


class Masking(LinearOperator):
    r"""Boolean-mask sampling ``x[mask]`` (reference ``sampling.py:125-196``)."""

    def __init__(self, size: int, sampling_bool, dtype: type = np.float64):
        sb = sampling_bool.detach().cpu().numpy() if isinstance(sampling_bool, torch.Tensor) else np.asarray(sampling_bool)
        self.sampling_bool = sb.reshape(-1).astype(bool)
        self.input_size = size
        self.nb_of_samples = int(np.count_nonzero(self.sampling_bool))
        if self.sampling_bool.size != size:
            raise ValueError('Invalid size of boolean sampling array.')
        self._idx_dev = {}
        super(Masking, self).__init__(shape=(self.nb_of_samples, self.input_size), dtype=dtype)

    def _idx(self, dev):
        if dev not in self._idx_dev:
            mask = torch.from_numpy(self.sampling_bool.view(np.uint8)).to(dev)
            idx = K.mask_to_indices(mask)
            if idx.numel() != self.nb_of_samples:
                raise RuntimeError('device mask compaction disagrees with the host popcount')
            self._idx_dev[dev] = idx
        return self._idx_dev[dev]

    def _apply(self, x):
        out = torch.empty(self.nb_of_samples, dtype=x.dtype, device=x.device)
        return K.take(out, x, self._idx(x.device), 1, self.input_size, 1)

    def _adjoint(self, y):
        # zeros(input_size, dtype=self.dtype) then x[mask] = y   (reference sampling.py:193-196)
        out = torch.empty(self.input_size, dtype=y.dtype, device=y.device)
        return K.put(out, y, self._idx(y.device), 1, self.input_size, 1)


class DownSampling(Masking):
    r"""Regular down-sampling by integer factors (reference ``sampling.py:199-391``)."""

    def __init__(self, size: int, downsampling_factor: Union[int, tuple, list], shape: Optional[tuple] = None,
                 axis: Optional[int] = None, dtype: type = np.float64):
        if type(downsampling_factor) is int:
            if (shape is not None) and (axis is None):
                self.downsampling_factor = len(shape) * (downsampling_factor,)
            else:
                self.downsampling_factor = (downsampling_factor,)
        else:
            self.downsampling_factor = tuple(downsampling_factor)
        if shape is not None:
            if size != np.prod(shape):
                raise ValueError(f'Array size {size} is incompatible with array shape {shape}.')
            if (axis is not None) and (axis > len(shape) - 1):
                raise ValueError(f'Array size {size} is incompatible with array shape {shape}.')
        if (shape is None) and (len(self.downsampling_factor) > 1):
            raise ValueError('Please specify an array shape for multidimensional downsampling.')
        elif (shape is not None) and (axis is None) and (len(shape) != len(self.downsampling_factor)):
            raise ValueError(f'Inconsistent downsampling factors {downsampling_factor} for array of shape {shape}.')
        self.input_size = size
        self.input_shape = shape
        self.axis = axis
        self.downsampling_mask = self.compute_downsampling_mask()
        if self.input_shape is None:
            self.output_shape = None
        elif len(self.downsampling_factor) > 1:
            self.output_shape = tuple(-(-int(n) // int(f)) for n, f in zip(self.input_shape, self.downsampling_factor))
        else:
            out = list(self.input_shape)
            out[self.axis] = -(-int(self.input_shape[self.axis]) // int(self.downsampling_factor[0]))
            self.output_shape = tuple(out)
        super(DownSampling, self).__init__(size=self.input_size, sampling_bool=self.downsampling_mask, dtype=dtype)

    def compute_downsampling_mask(self) -> np.ndarray:
        """Boolean mask ``index % factor == 0`` per axis, flattened in C order."""
        if self.input_shape is None:
            return (np.arange(self.input_size) % self.downsampling_factor[0]) == 0
        nd = len(self.input_shape)
        if len(self.downsampling_factor) > 1:
            axes = [(np.arange(n) % f) == 0 for n, f in zip(self.input_shape, self.downsampling_factor)]
        else:
            ax = self.axis if self.axis >= 0 else nd + self.axis
            axes = [np.ones(n, dtype=bool) for n in self.input_shape]
            axes[ax] = (np.arange(self.input_shape[ax]) % self.downsampling_factor[0]) == 0
        mask = np.ones(self.input_shape, dtype=bool)
        for a, keep in enumerate(axes):
            mask &= keep.reshape([-1 if i == a else 1 for i in range(nd)])
        return mask.reshape(-1)


class Pooling(LinearOperator):
    r"""Sum / mean pooling over constant-size blocks tiling an array of up to three axes (reference
    ``sampling.py:394-536``, which calls ``skimage.measure.block_reduce`` with zero padding: ragged border blocks are
    completed with zeros and a mean divides by the FULL block size).  The adjoint is the reference's *unpooling*: the
    pooled value is repeated over its block and cropped to the input shape, without rescaling (``:529-536``)."""

    def __init__(self, shape: tuple, block_size: Union[tuple, list], pooling_func: str = 'mean',
                 dtype: type = np.float64):
        if len(shape) != len(block_size):
            raise ValueError(f'Inconsistent block size {block_size} for array of shape {shape}.')
        if pooling_func not in ['mean', 'sum', 'average']:
            raise ValueError(f'pooling_func must be one of: "mean" or "sum".')
        if pooling_func == 'average':
            # the reference stores pooling_func = None for 'average' and block_reduce then fails (sampling.py:503-505)
            raise NotImplementedError("pooling_func='average' does not work in the reference either; use 'mean'")
        if len(shape) > 3:
            raise NotImplementedError('Pooling is implemented for arrays of up to three axes')
        self.input_shape = tuple(int(d) for d in shape)
        self.block_size = tuple(int(b) for b in block_size)
        self.pooling_func = pooling_func
        self.output_shape = self.get_output_shape()
        self.output_size = int(np.prod(self.output_shape))
        self.input_size = int(np.prod(self.input_shape))
        pad = 3 - len(self.input_shape)
        self._dims = (1,) * pad + self.input_shape
        self._block = (1,) * pad + self.block_size
        self._scale = 1.0 / float(np.prod(self.block_size)) if pooling_func == 'mean' else 1.0
        super(Pooling, self).__init__(shape=(self.output_size, self.input_size), dtype=dtype)

    def get_output_shape(self) -> tuple:
        return tuple(-(-d // b) for d, b in zip(self.input_shape, self.block_size))

    def _apply(self, x):
        out = torch.empty(self.output_size, dtype=x.dtype, device=x.device)
        return K.pool_fwd(out, x, self._dims, self._block, self._scale)

    def _adjoint(self, y):
        out = torch.empty(self.input_size, dtype=y.dtype, device=y.device)
        return K.pool_adj(out, y, self._dims, self._block)


def _out_of_scope(name):
    def ctor(*args, **kwargs):
        raise NotImplementedError(f'{name} is outside the B200 hot path (SURVEY.md section 2)')
    ctor.__name__ = name
    return ctor


NNSampling = _out_of_scope('NNSampling')
GeneralisedVandermonde = _out_of_scope('GeneralisedVandermonde')
MappedDistanceMatrix = _out_of_scope('MappedDistanceMatrix')
