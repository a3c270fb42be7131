r"""Abstract linear operators (mirror of ``pycsou/core/linop.py``, device-resident).

Kept from the reference (``core/linop.py:21-131, 279-321, 442-615``): ``__call__`` / ``matvec`` / ``adjoint`` /
``transpose`` / ``H`` / ``T`` / ``jacobianT`` / ``RangeGram`` / ``DomainGram``, attributes ``shape, dtype,
lipschitz_cst, diff_lipschitz_cst, is_symmetric, ...``, arithmetic dispatch, ``AdjointLinearOperator``,
``TransposeLinearOperator``, ``LinOpSum``, ``LinOpComp``, ``SymmetricLinearOperator``, ``UnitaryOperator``.

Replaced: ``compute_lipschitz_cst`` (reference: ARPACK ``svds``/``eigsh`` with a random start vector, not
repeatable) is a deterministic power iteration on :math:`K^\ast K` running on the device operator kernels with
fused norm reductions.  ``eigenvals`` / ``singularvals`` support ``k=1`` (largest magnitude) through the same
iteration.  Out of scope (explicit-matrix utilities): ``todense/tosparse/pinv/cond/...``
raise ``NotImplementedError``.
"""
from __future__ import annotations

from numbers import Number
from typing import Optional, Tuple

import numpy as np
import torch

from .. import _kernels as K
from .._array import is_array, like_input, to_device, torch_dtype
from .map import DifferentiableMap, DiffMapComp, DiffMapSum, Map, MapComp, MapSum


class LinearOperator(DifferentiableMap):
    r"""Base class for linear operators :math:`L:\mathbb{C}^N\to\mathbb{C}^M`."""

    def __init__(self, shape: Tuple[int, int], dtype: Optional[type] = None,
                 is_explicit: bool = False, is_dense: bool = False, is_sparse: bool = False, is_dask: bool = False,
                 is_symmetric: bool = False, lipschitz_cst: float = np.inf):
        DifferentiableMap.__init__(self, shape=shape, is_linear=True, lipschitz_cst=lipschitz_cst,
                                   diff_lipschitz_cst=lipschitz_cst)
        self.dtype = dtype
        self.is_explicit = is_explicit
        self.is_dense = is_dense
        self.is_sparse = is_sparse
        self.is_dask = is_dask
        self.is_symmetric = is_symmetric
        self.is_square = True if shape[0] == shape[1] else False

    # dtype used on the device: the operator's dtype when it has one (numpy semantics: outputs are allocated in
    # the operator dtype), otherwise the input's.
    def _tdtype(self, like: Optional[torch.Tensor] = None) -> torch.dtype:
        if self.dtype is not None:
            return torch_dtype(self.dtype)
        return like.dtype if like is not None else torch.float64

    def _in_dtype(self, arg):
        return torch_dtype(self.dtype) if self.dtype is not None else None

    def matvec(self, x):
        return self.__call__(x)

    def adjoint(self, y):
        if type(self)._adjoint is LinearOperator._adjoint:
            raise NotImplementedError('LinearOperator subclasses must implement adjoint or _adjoint')
        return like_input(self._adjoint(to_device(y, self._in_dtype(y))), y)

    def _adjoint(self, y: torch.Tensor):
        return self.adjoint(y)

    def transpose(self, y):
        yd = to_device(y, self._in_dtype(y))
        out = self._adjoint(yd.conj().resolve_conj() if yd.is_complex() else yd)
        return like_input(out.conj().resolve_conj() if out.is_complex() else out, y)

    def jacobianT(self, arg=None) -> 'LinearOperator':
        return self.get_adjointOp()

    def _jacobianT(self, x=None):
        return self.get_adjointOp()

    def get_adjointOp(self):
        return self if self.is_symmetric is True else AdjointLinearOperator(self)

    @property
    def H(self):
        return self.get_adjointOp()

    def get_transposeOp(self) -> 'TransposeLinearOperator':
        return TransposeLinearOperator(self)

    @property
    def T(self):
        return self.get_transposeOp()

    @property
    def RangeGram(self):
        return SymmetricLinearOperator(self * self.H)

    @property
    def DomainGram(self):
        return SymmetricLinearOperator(self.H * self)

    # -- spectral estimates ------------------------------------------------------------------------------------
    def _power_iteration(self, normal: bool, tol: float = 1e-10, maxiter: int = 400, seed: int = 0) -> float:
        """Largest |eigenvalue| of A (symmetric) or largest singular value (via ``A^H A``), on the device.

        The reference asks ARPACK (implicitly restarted Lanczos, ``tol=0``; ``core/linop.py:279-321``).  Here: a plain
        Lanczos recurrence on the device -- one operator application, one dot product, one fused ``w - alpha v - beta
        v_prev`` and one norm per step -- whose tridiagonal matrix is diagonalised on the host every few steps.  The
        largest Ritz value converges far faster than a power iteration (which stalls on the clustered top of a
        difference operator's spectrum) and is never above the true value.  ``self._spectral_converged`` records
        whether ``tol`` was reached within ``maxiter`` steps.
        """
        gen = torch.Generator(device='cpu').manual_seed(seed)
        dt = self._tdtype()
        rdt = torch.float64 if dt in (torch.float64, torch.complex128) else torch.float32
        tol = max(float(tol), 200 * float(torch.finfo(rdt).eps))
        n = self.shape[1]
        v = torch.randn(n, generator=gen, dtype=rdt).to(K_device()).to(dt)
        v = K.scale_by_norm(torch.empty_like(v), v, K.sumsq(v), 1.0, 2)
        v_prev, beta = None, 0.0
        alphas, betas = [], []
        theta, theta_prev = 0.0, None
        self._spectral_converged = False
        m = int(min(maxiter, n))
        for j in range(m):
            w = self._adjoint(self._apply(v)) if normal else self._apply(v)
            alpha = float(K.dot(K.real_view(w), K.real_view(v)).item())     # real: the operator is Hermitian
            alphas.append(alpha)
            if v_prev is None:
                w = K.lincomb2(torch.empty_like(w), 1.0, w, -alpha, v)
            else:
                w = K.lincomb3(torch.empty_like(w), 1.0, w, -alpha, v, -beta, v_prev)
            nrm2 = K.sumsq(w)
            beta = float(torch.sqrt(nrm2).item())
            last = j == m - 1
            if (j % 5) == 4 or last or beta == 0.0:
                ev = _tridiag_eigvals(alphas, betas)
                theta = float(max(abs(ev[0]), abs(ev[-1])))
                if theta == 0.0:
                    self._spectral_converged = True
                    return 0.0
                if beta <= 1e-14 * theta or (theta_prev is not None and abs(theta - theta_prev) <= tol * theta):
                    self._spectral_converged = True
                    break
                theta_prev = theta
            if last:
                break
            betas.append(beta)
            v_prev, v = v, K.scale_by_norm(torch.empty_like(w), w, nrm2, 1.0, 2)
        if m == n and not self._spectral_converged:          # the Krylov space is the whole space: exact up to rounding
            self._spectral_converged = True
        self._spectral_last_change = 0.0 if theta_prev is None else abs(theta - theta_prev) / theta
        return float(np.sqrt(theta)) if normal else theta

    def _lipschitz_upper_bound(self):
        """Analytic upper bound of the spectral norm, or None (overridden by the difference operators)."""
        return None

    def eigenvals(self, k: int, which='LM', **kwargs) -> np.ndarray:
        if not (self.is_symmetric or self.is_square):
            raise NotImplementedError(
                'The function eigenvals is only for square linear operator. For non square linear operators, use the method singularvals.')
        if k != 1 or which != 'LM' or not self.is_symmetric:
            raise NotImplementedError('only the largest-magnitude eigenvalue of symmetric operators (k=1) is supported')
        return np.array([self._power_iteration(False, **_pi_kwargs(kwargs))])

    def singularvals(self, k: int, which='LM', **kwargs) -> np.ndarray:
        if k != 1 or which != 'LM':
            raise NotImplementedError('only the largest singular value (k=1, which="LM") is supported')
        return np.array([self._power_iteration(True, **_pi_kwargs(kwargs))])

    def compute_lipschitz_cst(self, **kwargs):
        """Spectral norm by a Lanczos iteration on the device (``tol``/``maxiter`` kwargs as in the reference's ARPACK
        call, ``core/linop.py:279-321``).

        Step-size rules divide by this constant, so it must not UNDER-estimate the norm: when the iteration has not
        converged to ``tol`` (huge images: the top of a difference operator's spectrum is clustered) the analytic
        bound of the operator is used where one exists (``_lipschitz_upper_bound``), else the estimate is inflated by
        ten times its last observed change.
        """
        if self.is_symmetric is True:
            est = float(np.abs(self.eigenvals(k=1, **kwargs))[0])
        else:
            est = float(self.singularvals(k=1, **kwargs)[0])
        if not getattr(self, '_spectral_converged', True):
            bound = self._lipschitz_upper_bound()
            est = float(bound) if bound is not None else est * (1.0 + 10.0 * getattr(self, '_spectral_last_change', 0.0))
        self.lipschitz_cst = est
        self.diff_lipschitz_cst = self.lipschitz_cst

    # -- explicit-matrix utilities: out of scope for the hot path ---------------------------------------------
    def _out_of_scope(self, *a, **k):
        raise NotImplementedError('explicit-matrix / ARPACK utilities are outside the B200 hot path (SURVEY.md section 2)')

    todense = tosparse = tosciop = topylop = cond = pinv = _out_of_scope

    # -- arithmetic ------------------------------------------------------------------------------------------
    def __add__(self, other):
        if isinstance(other, LinearOperator):
            return LinOpSum(self, other)
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
        if isinstance(other, LinearOperator):
            return LinOpComp(self, other)
        if isinstance(other, DifferentiableMap):
            return DiffMapComp(self, other)
        if isinstance(other, Map):
            return MapComp(self, other)
        raise NotImplementedError

    def __rmul__(self, other):
        if isinstance(other, Number):
            other = self._wrap_scalar(other, self.shape[0])
        if isinstance(other, LinearOperator):
            return LinOpComp(other, self)
        if isinstance(other, DifferentiableMap):
            return DiffMapComp(other, self)
        if isinstance(other, Map):
            return MapComp(other, self)
        raise NotImplementedError


def K_device():
    from .._array import device
    return device()


def _tridiag_eigvals(alphas, betas):
    """Eigenvalues (ascending) of the Lanczos tridiagonal matrix: diagonal ``alphas``, off-diagonal ``betas``."""
    k = len(alphas)
    if k == 1:
        return np.array([alphas[0]])
    from scipy.linalg import eigvalsh_tridiagonal
    return eigvalsh_tridiagonal(np.asarray(alphas, dtype=np.float64), np.asarray(betas[:k - 1], dtype=np.float64))


def _pi_kwargs(kwargs):
    out = {}
    if 'tol' in kwargs and kwargs['tol']:
        out['tol'] = float(kwargs['tol'])
    if 'maxiter' in kwargs and kwargs['maxiter']:
        out['maxiter'] = int(kwargs['maxiter'])
    return out


def _flags(L: LinearOperator):
    return dict(is_explicit=L.is_explicit, is_dask=L.is_dask, is_dense=L.is_dense, is_sparse=L.is_sparse,
                is_symmetric=L.is_symmetric)


class AdjointLinearOperator(LinearOperator):
    def __init__(self, LinOp: LinearOperator):
        super(AdjointLinearOperator, self).__init__(shape=(LinOp.shape[1], LinOp.shape[0]), dtype=LinOp.dtype,
                                                    **_flags(LinOp))
        self.Linop = LinOp

    def _apply(self, y):
        return self.Linop._adjoint(y)

    def _adjoint(self, x):
        return self.Linop._apply(x)

    def compute_lipschitz_cst(self, **kwargs):
        if self.Linop.lipschitz_cst != np.inf:       # reference core/linop.py:502-506
            self.lipschitz_cst = self.Linop.lipschitz_cst
        else:
            LinearOperator.compute_lipschitz_cst(self, **kwargs)


class TransposeLinearOperator(LinearOperator):
    def __init__(self, LinOp: LinearOperator):
        super(TransposeLinearOperator, self).__init__(shape=(LinOp.shape[1], LinOp.shape[0]), dtype=LinOp.dtype,
                                                      **_flags(LinOp))
        self.Linop = LinOp

    def _apply(self, y):
        if y.is_complex():
            return self.Linop._adjoint(y.conj().resolve_conj()).conj().resolve_conj()
        return self.Linop._adjoint(y)

    def _adjoint(self, x):
        return self.Linop._apply(x)


def _pair_flags(A, B):
    return dict(is_explicit=A.is_explicit & B.is_explicit, is_dask=A.is_dask & B.is_dask,
                is_dense=A.is_dense & B.is_dense, is_sparse=A.is_sparse & B.is_sparse,
                is_symmetric=A.is_symmetric & B.is_symmetric)


class LinOpSum(LinearOperator, DiffMapSum):
    def __init__(self, LinOp1: LinearOperator, LinOp2: LinearOperator, dtype: Optional[type] = None):
        dtype = LinOp1.dtype if LinOp1.dtype is LinOp2.dtype else dtype
        DiffMapSum.__init__(self, map1=LinOp1, map2=LinOp2)
        LinearOperator.__init__(self, shape=self.shape, dtype=dtype, lipschitz_cst=self.lipschitz_cst,
                                **_pair_flags(LinOp1, LinOp2))
        self.LinOp1, self.LinOp2 = LinOp1, LinOp2

    def _adjoint(self, y):
        a, b = self.LinOp1._adjoint(y), self.LinOp2._adjoint(y)
        return K.lincomb2(torch.empty_like(a), 1.0, a, 1.0, b)


class LinOpComp(LinearOperator, DiffMapComp):
    def __init__(self, LinOp1: LinearOperator, LinOp2: LinearOperator, dtype: Optional[type] = None):
        dtype = LinOp1.dtype if LinOp1.dtype is LinOp2.dtype else dtype
        DiffMapComp.__init__(self, map1=LinOp1, map2=LinOp2)
        LinearOperator.__init__(self, shape=self.shape, dtype=dtype, lipschitz_cst=self.lipschitz_cst,
                                **_pair_flags(LinOp1, LinOp2))
        self.LinOp1, self.LinOp2 = LinOp1, LinOp2

    def _adjoint(self, y):
        return self.LinOp2._adjoint(self.LinOp1._adjoint(y))     # order reverses (core/linop.py:552-553)


class SymmetricLinearOperator(LinearOperator):
    def __init__(self, LinOp: LinearOperator):
        if LinOp.shape[0] != LinOp.shape[1]:
            raise TypeError('The input linear operator is not symmetric.')
        fl = _flags(LinOp)
        fl['is_symmetric'] = True
        super(SymmetricLinearOperator, self).__init__(shape=LinOp.shape, dtype=LinOp.dtype,
                                                      lipschitz_cst=LinOp.lipschitz_cst, **fl)
        self.LinOp = LinOp

    def _apply(self, x):
        return self.LinOp._apply(x)

    def _adjoint(self, y):
        return self._apply(y)


class UnitaryOperator(LinearOperator):
    def __init__(self, size: int, dtype: Optional[type] = None,
                 is_explicit: bool = False, is_dense: bool = False, is_sparse: bool = False, is_dask: bool = False,
                 is_symmetric: bool = False):
        super(UnitaryOperator, self).__init__(shape=(size, size), dtype=dtype, is_explicit=is_explicit,
                                              is_dense=is_dense, is_sparse=is_sparse, is_dask=is_dask,
                                              is_symmetric=is_symmetric)
        self.size = size
        self.lipschitz_cst = self.diff_lipschitz_cst = 1

    @property
    def RangeGram(self):
        from ..linop.base import IdentityOperator
        return IdentityOperator(size=self.size, dtype=self.dtype)

    @property
    def DomainGram(self):
        from ..linop.base import IdentityOperator
        return IdentityOperator(size=self.size, dtype=self.dtype)

    def eigenvals(self, k: int, which='LM', **kwargs) -> np.ndarray:
        return self.singularvals(k=k, which='LM', **kwargs)

    def singularvals(self, k: int, which='LM', **kwargs) -> np.ndarray:
        if k > min(self.shape[0], self.shape[1]):
            raise ValueError('The number of singular values must not exceed the smallest dimension size.')
        return np.ones(shape=(k,))

    def compute_lipschitz_cst(self, **kwargs):
        self.lipschitz_cst = self.diff_lipschitz_cst = 1

    def pinv(self, y, eps: Number = 0, **kwargs):
        return self.adjoint(y)

    @property
    def PinvOp(self):
        return self.H

    def cond(self, **kwargs):
        return 1
