r"""Penalty functionals on the device (mirror of the hot-path part of ``pycsou/func/penalty.py``).

Every penalty of the reference module is here: the hot-path ones and the leaves of 8(f)
rank 2 (root-find / Lambert-W proxes run on the device too).
"""
from __future__ import annotations

from numbers import Number
from typing import Optional, Union

import numpy as np
import torch

from .. import _kernels as K
from ..core.functional import DifferentiableFunctional, ProximableFunctional
from ..math.prox import proj_l1_ball, proj_l2_ball, proj_linfty_ball, proj_nonnegative_orthant, proj_segment, soft
from .base import IndicatorFunctional, LpNorm

__all__ = ['L2Norm', 'SquaredL2Norm', 'L2Ball', 'L1Norm', 'SquaredL1Norm', 'L1Ball', 'LInftyNorm', 'LInftyBall',
           'L21Norm', 'NonNegativeOrthant', 'Segment', 'RealLine', 'ImagLine', 'LogBarrier', 'ShannonEntropy',
           'QuadraticForm']


class L2Norm(LpNorm):
    r""":math:`\|x\|_2` (reference ``func/penalty.py:23-70``)."""

    def __init__(self, dim: int):
        super(L2Norm, self).__init__(dim=dim, proj_lq_ball=proj_l2_ball)

    def _apply(self, x):
        return float(torch.sqrt(K.sumsq(x)).item())

    def _prox(self, x, tau):
        # x - tau * proj_l2_ball(x / tau, 1)  ==  x * max(1 - tau/||x||, 0)   (one reduction + one scaling pass)
        return K.scale_by_norm(torch.empty_like(x), x, K.sumsq(x), tau, 0)


class SquaredL2Norm(DifferentiableFunctional):
    r""":math:`\|x\|_2^2`, gradient :math:`2x` (reference ``func/penalty.py:73-131``)."""

    def __init__(self, dim: int):
        super(SquaredL2Norm, self).__init__(dim=dim, data=None, is_linear=False, lipschitz_cst=np.inf,
                                            diff_lipschitz_cst=2)

    def _apply(self, x):
        return float(K.sumsq(x).item())

    def _jacobianT(self, x):
        return K.scale(torch.empty_like(x), 2.0, x)


def L2Ball(dim: int, radius: Number) -> IndicatorFunctional:
    """Indicator of the l2 ball (reference ``func/penalty.py:134-191``)."""
    return IndicatorFunctional(dim=dim, condition_func=lambda x: float(torch.sqrt(K.sumsq(x)).item()) <= radius,
                               projection_func=lambda x: proj_l2_ball.device(x, radius=radius),
                               kind='l2ball', params=(radius,))


class L1Norm(LpNorm):
    r""":math:`\|x\|_1`; prox = ``x - tau*clip(x/tau, -1, 1)`` (reference ``func/penalty.py:194-245``)."""

    def __init__(self, dim: int):
        super(L1Norm, self).__init__(dim=dim, proj_lq_ball=proj_linfty_ball)

    def _apply(self, x):
        return float(K.sumabs(x).item())

    def _prox(self, x, tau):
        return K.prox_l1(torch.empty_like(x), x, tau)

    def soft(self, x, tau: Number):
        return soft(x=x, tau=tau)


def LInftyBall(dim: int, radius: Number) -> IndicatorFunctional:
    """Indicator of the l-infinity ball (reference ``func/penalty.py:420-477``)."""
    def cond(x):
        mn, mx = K.minmax(x).tolist()
        return max(abs(mn), abs(mx)) <= radius
    return IndicatorFunctional(dim=dim, condition_func=cond,
                               projection_func=lambda x: proj_linfty_ball.device(x, radius=radius),
                               kind='segment', params=(-radius, radius))


class L21Norm(ProximableFunctional):
    r"""Mixed :math:`\ell_{2,1}` norm :math:`\sum_g \|x_g\|_2` (reference ``func/penalty.py:480-560``).

    ``groups`` is a label array of length ``dim``.  As in the reference, ``__new__`` degenerates to ``L1Norm`` when
    all labels are distinct (or ``groups`` is None) and to ``L2Norm`` when there is a single group
    (``penalty.py:525-530``).  The device kernels implement the *planes* layout ``groups == tile(arange(G), C)``
    (component ``c`` of group ``g`` at ``x[c*G + g]``, up to 4 components) -- the layout of a stacked ``Gradient``
    output, i.e. isotropic TV -- with one thread per group.  Any other label array (ragged groups, groups of any size,
    labels in any order) runs on the segmented kernels ``pcs_prox_l21_csr`` / ``pcs_reduce_sumgroupnorm_csr``: the labels
    are sorted once at construction into CSR form (members, offsets) and one warp works on each group.
    """

    def __new__(cls, dim: int, groups=None):
        if groups is None:
            return L1Norm(dim=dim)
        if isinstance(groups, PlanesGroups):
            if groups.n_comp == 1:
                return L1Norm(dim=dim)
            if groups.n_groups == 1:
                return L2Norm(dim=dim)
            return super(L21Norm, cls).__new__(cls)
        g = _labels(groups)
        n_unique = np.unique(g).size
        if n_unique == dim:
            return L1Norm(dim=dim)
        if n_unique == 1:
            return L2Norm(dim=dim)
        return super(L21Norm, cls).__new__(cls)

    def __init__(self, dim: int, groups):
        self.groups = groups
        self._csr = None                          # (members, offsets) on the device, general label arrays only
        if isinstance(groups, PlanesGroups) and groups.n_comp <= 4:
            # O(1) description of tile(arange(G), C): no label array is built
            if groups.size != dim:
                raise ValueError('groups must have one label per coordinate.')
            self.n_groups, self.n_comp = groups.n_groups, groups.n_comp
            self.groups_idxs = groups
            self._is_planes = True
            super(L21Norm, self).__init__(dim=dim, data=None, is_differentiable=False, is_linear=False)
            return
        g = _labels(groups)
        if g.size != dim:
            raise ValueError('groups must have one label per coordinate.')
        self.groups_idxs, inv, counts = np.unique(g, return_inverse=True, return_counts=True)
        inv = inv.reshape(-1)
        self.n_groups = int(self.groups_idxs.size)
        uniform = bool(np.all(counts == counts[0]))
        self.n_comp = int(counts[0]) if uniform else None
        # fast path: the planes layout of a stacked Gradient output (isotropic TV), up to 4 components per group
        self._is_planes = bool(uniform and self.n_comp <= 4 and
                               np.array_equal(inv, np.tile(np.arange(self.n_groups), self.n_comp)))
        if not self._is_planes:
            # any other label array (ragged groups, any group size): CSR form, members of a group in index order
            self._members_host = np.argsort(inv, kind='stable').astype(np.int64)
            self._offsets_host = np.concatenate(([0], np.cumsum(counts))).astype(np.int64)
        super(L21Norm, self).__init__(dim=dim, data=None, is_differentiable=False, is_linear=False)

    def _groups_csr(self, dev):
        if self._csr is None or self._csr[0].device != dev:
            self._csr = (torch.from_numpy(self._members_host).to(dev), torch.from_numpy(self._offsets_host).to(dev))
        return self._csr

    def _apply(self, x):
        if self._is_planes:
            return float(K.sumgroupnorm(x, self.n_groups, self.n_comp).item())
        return float(K.sumgroupnorm_csr(x, *self._groups_csr(x.device)).item())

    def _prox(self, x, tau):
        if self._is_planes:
            return K.prox_l21(torch.empty_like(x), x, tau, self.n_groups, self.n_comp)
        return K.prox_l21_csr(torch.empty_like(x), x, tau, *self._groups_csr(x.device))


class PlanesGroups:
    """O(1) stand-in for ``np.tile(np.arange(n_groups), n_comp)`` (per-pixel TV groups of a stacked ``Gradient``).

    ``np.asarray(PlanesGroups(...))`` materialises the label array, so the object can be handed to anything that expects
    the reference's ``groups`` argument; :class:`L21Norm` recognises it without building 2N labels (16 GB at 32768^2).
    """

    def __init__(self, n_groups: int, n_comp: int = 2):
        self.n_groups, self.n_comp = int(n_groups), int(n_comp)
        self.size = self.n_groups * self.n_comp

    def __len__(self):
        return self.size

    def __array__(self, dtype=None, copy=None):
        a = np.tile(np.arange(self.n_groups), self.n_comp)
        return a if dtype is None else a.astype(dtype)


def tv_groups(n_pixels: int, n_comp: int = 2) -> PlanesGroups:
    """Groups of isotropic TV on an image with ``n_pixels`` pixels: component c of pixel p has label p."""
    return PlanesGroups(n_pixels, n_comp)


def _labels(groups):
    if isinstance(groups, torch.Tensor):
        groups = groups.detach().cpu().numpy()
    return np.asarray(groups).reshape(-1)


def NonNegativeOrthant(dim: int) -> IndicatorFunctional:
    """Indicator of the non-negative orthant (reference ``func/penalty.py:563-612``)."""
    return IndicatorFunctional(dim=dim, condition_func=lambda x: float(K.minmax(x)[0].item()) >= 0,
                               projection_func=lambda x: proj_nonnegative_orthant.device(x), kind='nonneg')


def Segment(dim: int, a: Number = 0, b: Number = 1):
    """Indicator of the box ``[a, b]^dim`` (reference ``func/penalty.py:615-668``)."""
    def cond(x):
        mn, mx = K.minmax(x).tolist()
        return mn >= a and mx <= b
    return IndicatorFunctional(dim=dim, condition_func=cond,
                               projection_func=lambda x: proj_segment.device(x, a=a, b=b), kind='segment',
                               params=(a, b))


class SquaredL1Norm(ProximableFunctional):
    r""":math:`\|x\|_1^2` (reference ``func/penalty.py:248-317``).  The prox is ``soft(x, t)`` with ``t`` the root of
    ``sum max(|x| - t, 0) = t / (2 tau)``: the reference finds it by a descending sort and a cumulative sum
    (``prox_computation='sort'``, ``:310-316``) or by brentq on an equivalent equation (``'root'``, ``:300-309``); the
    device path finds the same root exactly by monotone Newton steps, each one a deterministic reduction
    (``K.threshold_newton``), whichever option is named.  Real arrays only."""

    def __init__(self, dim: int, prox_computation: str = 'sort'):
        self.prox_computation = prox_computation
        super(SquaredL1Norm, self).__init__(dim=dim, data=None, is_differentiable=False, is_linear=False)

    def _apply(self, x):
        return float(K.sumabs(x).item()) ** 2

    def _prox(self, x, tau):
        if self.prox_computation not in ('sort', 'root'):
            return None                      # the reference falls through and returns None (func/penalty.py:296-316)
        t = K.threshold_newton(x, 0.0, 1.0 / (2.0 * tau))
        if t <= 0.0:                         # x = 0 ('root' returns x, :308-309; 'sort' raises on the empty argmax)
            return x.clone()
        return K.prox_l1(torch.empty_like(x), x, t)


def L1Ball(dim: int, radius: Number) -> IndicatorFunctional:
    """Indicator of the l1 ball; prox = ``proj_l1_ball`` (reference ``func/penalty.py:319-377``)."""
    return IndicatorFunctional(dim=dim, condition_func=lambda x: float(K.sumabs(x).item()) <= radius,
                               projection_func=lambda x: proj_l1_ball.device(x, radius=radius),
                               kind='l1ball', params=(radius,))


class LInftyNorm(LpNorm):
    r""":math:`\max_i |x_i|`; prox ``x - tau * proj_l1_ball(x / tau, 1)`` (reference ``func/penalty.py:380-417``)."""

    def __init__(self, dim: int):
        super(LInftyNorm, self).__init__(dim=dim, proj_lq_ball=proj_l1_ball)

    def _apply(self, x):
        mn, mx = K.minmax(x).tolist()
        return max(abs(mn), abs(mx))


def RealLine(dim: int) -> IndicatorFunctional:
    """Indicator of the real line; prox = real part (reference ``func/penalty.py:671-717``)."""
    def cond(x):
        return (not x.is_complex()) or float(K.sumabs(x.imag.contiguous()).item()) == 0.0
    return IndicatorFunctional(dim=dim, condition_func=cond,
                               projection_func=lambda x: x.real.contiguous() if x.is_complex() else x,
                               kind='realline')


def ImagLine(dim: int) -> IndicatorFunctional:
    """Indicator of the imaginary line; prox returns the imaginary part as a real array, as the reference does
    (``func/penalty.py:720-767``: ``np.imag(x)``)."""
    def cond(x):
        re = x.real.contiguous() if x.is_complex() else x
        return float(K.sumabs(re).item()) == 0.0
    return IndicatorFunctional(dim=dim, condition_func=cond,
                               projection_func=lambda x: x.imag.contiguous() if x.is_complex() else torch.zeros_like(x),
                               kind='imagline')


class LogBarrier(ProximableFunctional):
    r""":math:`-\sum_i \log x_i` (``+inf`` unless ``x > 0``); prox ``(x + sqrt(x^2 + 4 tau)) / 2``
    (reference ``func/penalty.py:776-840``)."""

    def __init__(self, dim: int):
        super(LogBarrier, self).__init__(dim=dim, data=None, is_differentiable=False, is_linear=False)

    def _apply(self, x):
        return float(K.reduce_neglog(x).item())

    def _prox(self, x, tau):
        return K.prox_logbarrier(torch.empty_like(x), x, tau)


class ShannonEntropy(ProximableFunctional):
    r""":math:`\sum_i x_i \log x_i` (``0`` at 0, ``+inf`` below); prox ``tau W(exp(x/tau - 1)/tau)`` with the principal
    Lambert function (reference ``func/penalty.py:843-922``)."""

    def __init__(self, dim: int):
        super(ShannonEntropy, self).__init__(dim=dim, data=None, is_differentiable=False, is_linear=False)

    def _apply(self, x):
        return float(K.reduce_xlogx(x).item())

    def _prox(self, x, tau):
        return K.prox_shannon(torch.empty_like(x), x, tau)




class QuadraticForm(DifferentiableFunctional):
    r""":math:`x^T Q x` for a positive semi-definite operator ``Q`` (identity if ``None``), gradient :math:`2Qx`
    (reference ``func/penalty.py:925-990``).  ``Q`` is any operator of this package; real vectors."""

    def __init__(self, dim: int, linop=None):
        self.linop = linop
        beta = 2 if linop is None else 2 * linop.diff_lipschitz_cst
        super(QuadraticForm, self).__init__(dim=dim, data=None, is_linear=False, diff_lipschitz_cst=beta)

    def _apply(self, x):
        if self.linop is None:
            return float(K.sumsq(x).item())
        return float(K.dot(x, self.linop._apply(x)).item())

    def _jacobianT(self, x):
        q = x if self.linop is None else self.linop._apply(x)
        return K.scale(torch.empty_like(x), 2.0, q)
