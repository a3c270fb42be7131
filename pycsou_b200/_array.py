"""API-edge array handling: numpy / lists / CPU tensors are copied to the device once; results mirror the input kind.

Inside the package everything is a contiguous 1-D CUDA tensor (PyTorch = device memory + streams only).
"""
from __future__ import annotations

from numbers import Number

import numpy as np
import torch

from ._kernels import require_cuda

_NP2T = {np.dtype('float32'): torch.float32, np.dtype('float64'): torch.float64,
         np.dtype('complex64'): torch.complex64, np.dtype('complex128'): torch.complex128}


class SlabRows:
    """This rank's rows of a vector that is defined on the whole (row-decomposed) image -- see ``opt/slab.py``.

    Stands in for the global vector when a functional is built (``SquaredL2Loss(dim=N, data=SlabRows(...))``): it has
    the global ``size`` for the constructors' shape checks and carries only the local rows.
    """

    def __init__(self, local, size: int):
        self.local, self.size = local, int(size)

    @property
    def shape(self):
        return (self.size,)

    def __neg__(self):
        return Negated(self)


class Negated:
    """``-a`` with the sign still pending: the losses shift by ``-data`` (reference ``func/loss.py:69,120``), and
    negating a large host array costs more than the upload -- the sign is applied on the device instead."""

    def __init__(self, a):
        self.a = a

    @property
    def shape(self):
        return tuple(self.a.shape)

    def __neg__(self):
        return self.a

    def __array__(self, dtype=None, copy=None):
        v = -np.asarray(self.a)
        return v if dtype is None else v.astype(dtype)


def torch_dtype(dtype) -> torch.dtype:
    """Map a numpy-style dtype spec ('float64', np.float32, type, torch dtype, None) to a torch dtype."""
    if dtype is None:
        return torch.float64
    if isinstance(dtype, torch.dtype):
        return dtype
    d = np.dtype(dtype)
    if d.kind in 'iub':
        return torch.float64
    try:
        return _NP2T[d]
    except KeyError:
        raise NotImplementedError(f'unsupported dtype {dtype}')


def np_dtype(tdtype: torch.dtype):
    return {v: k for k, v in _NP2T.items()}[tdtype]


def is_array(x) -> bool:
    return isinstance(x, (np.ndarray, torch.Tensor))


def device():
    require_cuda()
    return torch.device('cuda', torch.cuda.current_device())


def to_device(x, dtype: torch.dtype | None = None) -> torch.Tensor:
    """Contiguous flat CUDA tensor holding ``x`` (one H2D copy for host data, no copy for matching device data)."""
    if isinstance(x, SlabRows):
        raise NotImplementedError('row-decomposed data can only be used by a solver constructed with slab=SlabComm(...)')
    if isinstance(x, Negated):
        return -to_device(x.a, dtype)
    if isinstance(x, torch.Tensor):
        t = x
        if not t.is_cuda:
            t = t.to(device(), non_blocking=True)
    else:
        a = np.asarray(x)
        if a.dtype.kind in 'iub':
            a = a.astype(np.float64)
        t = torch.from_numpy(np.ascontiguousarray(a)).to(device(), non_blocking=True)
    if dtype is not None and t.dtype != dtype:
        if t.is_complex() and not dtype.is_complex:
            t = t.real            # numpy assignment into a real array drops the imaginary part
        t = t.to(dtype)
    t = t.reshape(-1)
    return t if t.is_contiguous() else t.contiguous()


def like_input(result, ref):
    """Return ``result`` (device tensor or python scalar) in the container kind of ``ref``."""
    if not isinstance(result, torch.Tensor):
        return result
    if isinstance(ref, torch.Tensor):
        return result if ref.is_cuda else result.cpu()
    return result.cpu().numpy()


def scalar(x) -> float:
    """Python float from a 0-/1-element device tensor (one D2H sync) or a number."""
    if isinstance(x, torch.Tensor):
        return float(x.item())
    return float(x)


def is_number(x) -> bool:
    return isinstance(x, Number) and not isinstance(x, bool)
