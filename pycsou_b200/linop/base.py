r"""Operator glue (mirror of the hot-path part of ``pycsou/linop/base.py``).

In scope: ``DiagonalOperator, IdentityOperator, NullOperator, HomothetyMap,
LinOpStack / LinOpVStack / LinOpHStack`` plus a minimal ``DenseLinearOperator`` (plain library GEMV through
``torch.mv`` -- explicit-matrix algebra is not part of the stencil path, it exists so that the reference's LASSO
example runs), ``PolynomialLinearOperator`` and a duck-typed ``PyLopLinearOperator``.  Sparse / dask /
Kronecker / Khatri-Rao / Block operators are out of scope and raise.
"""
from __future__ import annotations

from numbers import Number
from typing import Optional, Tuple, Union

import numpy as np
import torch

from .. import _kernels as K
from .._array import to_device, torch_dtype
from ..core.linop import LinearOperator
from ..core.map import DiffMapStack

__all__ = ['DenseLinearOperator', 'ExplicitLinearOperator', 'LinOpStack', 'LinOpVStack', 'LinOpHStack',
           'DiagonalOperator', 'IdentityOperator', 'NullOperator', 'HomothetyMap', 'PyLopLinearOperator',
           'SparseLinearOperator', 'DaskLinearOperator', 'BlockOperator', 'BlockDiagonalOperator',
           'PolynomialLinearOperator', 'KroneckerProduct', 'KroneckerSum', 'KhatriRaoProduct']


class DenseLinearOperator(LinearOperator):
    """Explicit dense matrix (reference ``linop/base.py:102-118``); GEMV by the vendor library, not a hot-path kernel."""

    def __init__(self, ndarray, is_symmetric: bool = False):
        a = ndarray.detach().cpu().numpy() if isinstance(ndarray, torch.Tensor) else np.asarray(ndarray)
        if a.ndim != 2:
            raise TypeError('Invalid input type.')
        super(DenseLinearOperator, self).__init__(shape=a.shape, dtype=a.dtype if a.dtype.kind in 'fc' else np.float64,
                                                  is_explicit=True, is_dense=True, is_symmetric=is_symmetric)
        self.mat = a
        self._mat_dev = None

    def _m(self, like):
        if self._mat_dev is None or self._mat_dev.dtype != like.dtype:
            self._mat_dev = torch.from_numpy(np.ascontiguousarray(self.mat)).to(like.device).to(like.dtype)
        return self._mat_dev

    def _apply(self, x):
        return torch.mv(self._m(x), x)

    def _adjoint(self, y):
        return torch.mv(self._m(y).conj().T, y)


ExplicitLinearOperator = DenseLinearOperator


class LinOpStack(LinearOperator, DiffMapStack):
    r"""Stack linear operators vertically (``axis=0``) or horizontally (``axis=1``) (reference ``base.py:159-292``).

    Vertical stack: outputs are concatenated, the adjoint splits its argument at ``cumsum(block_sizes)`` and sums
    the per-block adjoints (``base.py:272-284``).  Blocks run back-to-back on the current CUDA stream.
    """

    def __init__(self, *linops, axis: int, n_jobs: int = 1, joblib_backend: str = 'loky'):
        DiffMapStack.__init__(self, *linops, axis=axis, n_jobs=n_jobs, joblib_backend=joblib_backend)
        self.linops = self.maps
        self.is_explicit_list = [op.is_explicit for op in self.linops]
        self.is_dense_list = [op.is_dense for op in self.linops]
        self.is_sparse_list = [op.is_sparse for op in self.linops]
        self.is_dask_list = [op.is_dask for op in self.linops]
        self.is_symmetric_list = [op.is_symmetric for op in self.linops]
        dtypes = {np.dtype(op.dtype) for op in self.linops if op.dtype is not None}
        LinearOperator.__init__(self, shape=self.shape, dtype=dtypes.pop() if len(dtypes) == 1 else None,
                                is_explicit=bool(np.all(self.is_explicit_list)),
                                is_dense=bool(np.all(self.is_dense_list)),
                                is_sparse=bool(np.all(self.is_sparse_list)),
                                is_dask=bool(np.all(self.is_dask_list)),
                                is_symmetric=bool(np.all(self.is_symmetric_list)),
                                lipschitz_cst=self.lipschitz_cst)

    def _apply(self, x):
        if self.axis == 0:
            outs = [op._apply(x) for op in self.linops]
            dt = outs[0].dtype
            out = torch.empty(self.shape[0], dtype=dt, device=x.device)
            edges = [0] + [int(s) for s in self.sections]
            for i, o in enumerate(outs):
                out[edges[i]:edges[i + 1]].copy_(o)
            return out
        result = None
        for op, xi in zip(self.linops, self._split(x)):
            v = op._apply(xi)
            result = v if result is None else K.lincomb2(result, 1.0, result, 1.0, v)
        return result

    def _adjoint(self, y):
        if self.axis == 0:
            result = None
            for op, yi in zip(self.linops, self._split(y)):
                v = op._adjoint(yi)
                result = v if result is None else K.lincomb2(torch.empty_like(v), 1.0, result, 1.0, v)
            return result
        return torch.cat([op._adjoint(y).reshape(-1) for op in self.linops])


class LinOpVStack(LinOpStack):
    def __init__(self, *linops, n_jobs: int = 1, joblib_backend: str = 'loky'):
        super(LinOpVStack, self).__init__(*linops, axis=0, n_jobs=n_jobs, joblib_backend=joblib_backend)


class LinOpHStack(LinOpStack):
    def __init__(self, *linops, n_jobs: int = 1, joblib_backend: str = 'loky'):
        super(LinOpHStack, self).__init__(*linops, axis=1, n_jobs=n_jobs, joblib_backend=joblib_backend)


class DiagonalOperator(LinearOperator):
    r"""Element-wise scaling by ``diag`` (reference ``linop/base.py:551-579``)."""

    def __init__(self, diag):
        d = diag.detach().cpu().numpy() if isinstance(diag, torch.Tensor) else np.asarray(diag)
        self.diag = d.reshape(-1)
        dt = self.diag.dtype if self.diag.dtype.kind in 'fc' else np.dtype('float64')
        super(DiagonalOperator, self).__init__(shape=(self.diag.size, self.diag.size), dtype=dt,
                                               is_explicit=False, is_dense=False, is_sparse=False, is_dask=False,
                                               is_symmetric=bool(np.all(np.isreal(self.diag))))
        self.lipschitz_cst = self.diff_lipschitz_cst = np.max(d)      # sic: max, not max|.| (base.py:567)
        self._diag_dev = None

    def _d(self, like):
        if self._diag_dev is None or self._diag_dev.dtype != like.dtype or self._diag_dev.device != like.device:
            self._diag_dev = to_device(self.diag, like.dtype)
        return self._diag_dev

    def _apply(self, x):
        if self.diag.size == 1:
            return K.scale(torch.empty_like(x), float(np.real(self.diag[0])), x)
        if x.is_complex():
            return self._d(x) * x
        return K.mul(torch.empty_like(x), self._d(x), x)

    def _adjoint(self, y):
        if self.diag.size == 1:
            return K.scale(torch.empty_like(y), float(np.real(self.diag[0])), y)
        if y.is_complex():
            return self._d(y).conj() * y
        return K.mul(torch.empty_like(y), self._d(y), y)


class IdentityOperator(DiagonalOperator):
    def __init__(self, size: int, dtype: Optional[type] = None):
        self.diag = np.ones(shape=(1,), dtype=dtype)      # never materialised at full size on the device
        LinearOperator.__init__(self, shape=(size, size), dtype=np.dtype(dtype) if dtype is not None else np.dtype('float64'),
                                is_symmetric=True)
        self._diag_dev = None
        self.lipschitz_cst = self.diff_lipschitz_cst = 1

    def _apply(self, x):
        return x

    def _adjoint(self, y):
        return y


class NullOperator(LinearOperator):
    def __init__(self, shape: Tuple[int, int], dtype: Optional[type] = np.float64):
        super(NullOperator, self).__init__(shape=shape, dtype=dtype, is_explicit=False, is_dense=False,
                                           is_sparse=False, is_dask=False,
                                           is_symmetric=True if (shape[0] == shape[1]) else False)
        self.lipschitz_cst = self.diff_lipschitz_cst = 0

    def _apply(self, x):
        return torch.zeros(self.shape[0], dtype=x.dtype, device=x.device)

    def _adjoint(self, y):
        return torch.zeros(self.shape[1], dtype=y.dtype, device=y.device)

    def eigenvals(self, k: int, which='LM', **kwargs) -> np.ndarray:
        return np.zeros(shape=(k,), dtype=self.dtype)

    def singularvals(self, k: int, which='LM', **kwargs) -> np.ndarray:
        return np.zeros(shape=(k,), dtype=self.dtype)


class HomothetyMap(DiagonalOperator):
    r"""Scalar multiple of the identity, used for every ``alpha * f`` (reference ``linop/base.py:625-633``)."""

    def __init__(self, size: int, constant: Number):
        self.cst = constant
        super(HomothetyMap, self).__init__(diag=self.cst)
        self.dtype = None              # scale in whatever dtype flows through
        self.shape = (size, size)
        self.lipschitz_cst = self.diff_lipschitz_cst = constant

    def _apply(self, x):
        if isinstance(x, torch.Tensor):
            return K.scale(torch.empty_like(x), float(self.cst), x)
        return self.cst * x            # functional values are python floats

    _adjoint = _apply

    def jacobianT(self, arg: Optional[Number] = None) -> Number:
        return self.cst

    def _jacobianT(self, x=None):
        return self.cst


class PolynomialLinearOperator(LinearOperator):
    r""":math:`P(L)=\sum_k a_k L^k` by repeated application of a square operator (reference ``linop/base.py:636-712``):
    one application of ``L`` and one fused ``y += a_k z`` pass per degree.  As in the reference, the adjoint of an
    operator FLAGGED symmetric is the forward map itself (``:703-704``)."""

    def __init__(self, LinOp: LinearOperator, coeffs):
        self.coeffs = np.asarray(coeffs).astype(LinOp.dtype if LinOp.dtype is not None else 'float64')
        if np.iscomplexobj(self.coeffs):
            raise NotImplementedError('complex polynomial coefficients')
        if LinOp.shape[0] != LinOp.shape[1]:
            raise ValueError('Input linear operator must be square.')
        self.Linop = LinOp
        super(PolynomialLinearOperator, self).__init__(shape=LinOp.shape, dtype=LinOp.dtype,
                                                       is_explicit=LinOp.is_explicit, is_dense=LinOp.is_dense,
                                                       is_sparse=LinOp.is_sparse, is_dask=LinOp.is_dask,
                                                       is_symmetric=LinOp.is_symmetric)

    def _horner(self, x, step):
        z, y = x, K.scale(torch.empty_like(x), float(self.coeffs[0]), x)
        for c in self.coeffs[1:]:
            z = step(z)
            K.lincomb2(y, 1.0, y, float(c), z)
        return y

    def _apply(self, x):
        return self._horner(x, self.Linop._apply)

    def _adjoint(self, y):
        return self._horner(y, self.Linop._apply if self.is_symmetric else self.Linop._adjoint)


def _out_of_scope(name):
    def ctor(*args, **kwargs):
        raise NotImplementedError(f'{name} is outside the B200 hot path (explicit-matrix algebra, SURVEY.md section 2)')
    ctor.__name__ = name
    return ctor


class PyLopLinearOperator(LinearOperator):
    r"""Linear operator from a pylops-style object (reference ``linop/base.py:24-54``).

    The reference wraps a ``pylops.LinearOperator`` and forwards ``__call__`` / ``adjoint`` to its ``matvec`` /
    ``rmatvec`` on numpy arrays.  pylops' own operators are CPU objects (and pylops is not a dependency of this package),
    so here the adapter is duck-typed: ``PyLop`` is ANY object with ``shape``, ``matvec`` and ``rmatvec`` that maps 1-D
    **CUDA tensors** to CUDA tensors (``dtype`` and ``explicit`` are read when present).  That is the hook through which
    user-written device operators enter the solvers of ``pycsou_b200.opt.proxalgs``; an object that returns host arrays
    is refused (there is no CPU path).
    """

    def __init__(self, PyLop, is_symmetric: bool = False, is_dense: bool = False, is_sparse: bool = False,
                 lipschitz_cst: float = np.inf):
        for attr in ('shape', 'matvec', 'rmatvec'):
            if not hasattr(PyLop, attr):
                raise TypeError(f'PyLop must expose shape, matvec and rmatvec (missing {attr})')
        dt = np.dtype(getattr(PyLop, 'dtype', None) or np.float64)
        super(PyLopLinearOperator, self).__init__(shape=tuple(PyLop.shape), dtype=dt,
                                                  is_explicit=bool(getattr(PyLop, 'explicit', False)),
                                                  is_dense=is_dense, is_sparse=is_sparse, is_dask=False,
                                                  is_symmetric=is_symmetric, lipschitz_cst=lipschitz_cst)
        self.Op = PyLop

    @staticmethod
    def _dev(out, what):
        if not (isinstance(out, torch.Tensor) and out.is_cuda):
            raise TypeError(f'PyLopLinearOperator: {what} must return a CUDA tensor (pycsou_b200 has no CPU path)')
        return out.reshape(-1).contiguous()

    def _apply(self, x):
        return self._dev(self.Op.matvec(x), 'matvec')

    def _adjoint(self, y):
        return self._dev(self.Op.rmatvec(y), 'rmatvec')


SparseLinearOperator = _out_of_scope('SparseLinearOperator')
DaskLinearOperator = _out_of_scope('DaskLinearOperator')
BlockOperator = _out_of_scope('BlockOperator')
BlockDiagonalOperator = _out_of_scope('BlockDiagonalOperator')
KroneckerProduct = _out_of_scope('KroneckerProduct')
KroneckerSum = _out_of_scope('KroneckerSum')
KhatriRaoProduct = _out_of_scope('KhatriRaoProduct')
