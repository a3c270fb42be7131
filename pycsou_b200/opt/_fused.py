"""Fused-iteration planner: maps a solver's ``(F, G, H, K)`` onto the single-kernel iteration of ``csrc/pcs_pds.cu``.

``plan_pds(solver)`` inspects the expression trees exactly as the reference composes them
(``(1/2) * SquaredL2Loss(dim, y) * A`` -> ``DiffMapComp(DiffMapComp(HomothetyMap(.5), DiffMapShifted(SquaredL2Norm)), A)``,
``core/map.py:534-610``; ``lam * L21Norm`` -> ``ProxFuncPostComp``, ``core/functional.py:244-265``) and returns a plan whose
``step(x, z, diag_row)`` performs one whole ``PrimalDualSplitting.update_iterand`` + ``update_diagnostics``
(``opt/proxalgs.py:343-394``) in ONE kernel launch, or ``None`` when the problem is not a total-variation problem of
the supported shape (the solver then runs op by op on the stand-alone kernels).

Supported:  K = Gradient(2-D shape, kind='forward');  H = [lam *] L21Norm (TV groups) or [lam *] L1Norm (anisotropic);
G in {None, NonNegativeOrthant, Segment, [lam *] L1Norm};  F in {None, (1/2) SquaredL2Loss, (1/2) SquaredL2Loss * A}.
With ``A`` a separable ``Convolve2D`` (PSF up to 15 x 15) the data-term gradient is evaluated inside the kernel as
``(A^T A) x - A^T y`` where ``A^T A`` is the separable autocorrelation filter with exact corner corrections for the
zero "same" boundary (``composite_filter``) and ``A^T y`` is computed once per solve.  Any other differentiable F has
its gradient computed by the stand-alone kernels and handed to the fused primal/dual kernel.
"""
from __future__ import annotations

import ctypes as C
import os
import weakref
from numbers import Number

import numpy as np
import torch

from .. import _kernels as K
from .. import _lib
from .._array import to_device

MAX_HL = 14                      # PDS_MAXHL in csrc/pcs_pds.cu
DM_NONE, DM_IDENTITY, DM_CONV, DM_GIVEN = 0, 1, 2, 3
GM_NONE, GM_NONNEG, GM_SEGMENT, GM_L1 = 0, 1, 2, 3
HM_L21, HM_L1 = 0, 1


def composite_filter(h, off):
    """Autocorrelation form of ``T^T T`` for the 1-D zero-"same" convolution ``(T x)[i] = sum_k h[k] x[i + off - k]``.

    Returns ``(c, E_lo, E_hi)``: ``c`` (2 nh - 1 symmetric taps) such that away from the ends ``(T^T T x)[i] =
    sum_d c[d] x[i - (nh - 1) + d]``, and the corner blocks with ``T^T T = Toeplitz(c) - E_lo (+) E_hi`` where
    ``E_lo`` (``off`` x ``off``) acts on the first and ``E_hi`` (``nh-1-off`` squared) on the last coordinates.
    They exist because ``T x`` is truncated to the signal support before ``T^T`` is applied.
    """
    h = np.asarray(h, dtype=np.float64).reshape(-1)
    nh = h.size
    c = np.correlate(h, h, mode='full')                       # c[d + nh - 1] = sum_m h[m] h[m + d]
    n_lo, n_hi = int(off), int(nh - 1 - off)
    E_lo = np.zeros((n_lo, n_lo))
    for a in range(n_lo):
        for b in range(n_lo):                                  # terms with output index k < 0: m < off - a
            E_lo[a, b] = sum(h[m] * h[m + a - b] for m in range(0, off - a) if 0 <= m + a - b < nh)
    E_hi = np.zeros((n_hi, n_hi))
    for a in range(n_hi):                                      # a, b count from the first corrected coordinate n - n_hi
        for b in range(n_hi):                                  # terms with output index k > n - 1: m > off + n_hi - 1 - a
            E_hi[a, b] = sum(h[m] * h[m + a - b] for m in range(off + n_hi - a, nh) if 0 <= m + a - b < nh)
    return c, E_lo, E_hi


class _Match:
    """Result of pattern matching; raises ``_NoMatch`` when a piece is unsupported."""


class _NoMatch(Exception):
    pass


def _match_K(Kop, dim, any_kind=False):
    from ..linop.diff import GradientOperator
    if not isinstance(Kop, GradientOperator) or len(Kop.array_shape) != 2 or (Kop.kind != 'forward' and not any_kind):
        raise _NoMatch('K is not a 2-D forward-difference Gradient')
    H, W = Kop.array_shape
    if H * W != dim or H < 2 or W < 2:
        raise _NoMatch('shape mismatch')
    return H, W, Kop.steps


def _scaled(f):
    """Peel ``lam * f`` (ProxFuncPostComp with zero shift) -> (f, lam)."""
    from ..core.functional import ProxFuncPostComp
    lam = 1.0
    while isinstance(f, ProxFuncPostComp):
        if f.shift != 0 or not isinstance(f.scale, Number):
            raise _NoMatch('shifted functional')
        lam *= float(f.scale)
        f = f.prox_func
    return f, lam


def _match_H(Hf, N):
    from ..func.penalty import L1Norm, L21Norm
    f, lam = _scaled(Hf)
    if type(f) is L21Norm and f._is_planes and f.n_comp == 2 and f.n_groups == N:
        return HM_L21, lam
    if type(f) is L1Norm and f.dim == 2 * N:
        return HM_L1, lam
    raise _NoMatch('H is not a TV penalty')


def _match_G(Gf, tau):
    from ..func.base import IndicatorFunctional, NullProximableFunctional
    from ..func.penalty import L1Norm
    if isinstance(Gf, NullProximableFunctional):
        return GM_NONE, 0.0, 0.0
    if isinstance(Gf, IndicatorFunctional) and Gf.kind == 'nonneg':
        return GM_NONNEG, 0.0, 0.0
    if isinstance(Gf, IndicatorFunctional) and Gf.kind == 'segment':
        return GM_SEGMENT, float(Gf.params[0]), float(Gf.params[1])
    f, lam = _scaled(Gf)
    if type(f) is L1Norm:
        return GM_L1, tau * lam, 0.0
    raise _NoMatch('G has no fused prox')


def _half_sq_loss(F):
    """``(1/2) * SquaredL2Loss(dim, y)`` -> y (host or device array), else None."""
    from ..core.map import DiffMapComp, DiffMapShifted
    from ..func.penalty import SquaredL2Norm
    from ..linop.base import HomothetyMap
    if not (isinstance(F, DiffMapComp) and isinstance(F.map1, HomothetyMap) and F.map1.cst == 0.5):
        return None
    inner = F.map2
    if not (isinstance(inner, DiffMapShifted) and type(inner.map) is SquaredL2Norm):
        return None
    s = inner.shift
    if isinstance(s, Number):
        return None
    from .._array import Negated, SlabRows
    if isinstance(s, Negated):
        return s.a
    return -s if isinstance(s, (torch.Tensor, SlabRows)) else -np.asarray(s)


def _match_F(F, shape):
    """-> (data_mode, y or None, conv operator or None)."""
    from ..core.map import DiffMapComp
    from ..func.base import NullDifferentiableFunctional
    from ..linop.conv import Convolve2DOperator
    if isinstance(F, NullDifferentiableFunctional):
        return DM_NONE, None, None
    y = _half_sq_loss(F)
    if y is not None:
        return DM_IDENTITY, y, None
    if isinstance(F, DiffMapComp) and isinstance(F.map2, Convolve2DOperator):
        y = _half_sq_loss(F.map1)
        A = F.map2
        if (y is not None and A.is_separable and A.array_shape == tuple(shape)
                and max(A.filter.shape) - 1 <= MAX_HL and A.filter.shape[0] + 0 <= shape[0]
                and A.filter.shape[1] <= shape[1]):
            return DM_CONV, y, A
    return DM_GIVEN, None, None


class PdsTvPlan:
    """One-launch-per-iteration Condat-Vu TV solver state (device buffers, filter taps, workspace)."""

    def __init__(self, solver, H, W, steps, data_mode, y, A, g, h):
        self.solver = weakref.proxy(solver)      # no reference cycle: the solver (and its result buffers) die by refcount
        self.H, self.W, self.steps = H, W, steps
        self.data_mode = data_mode
        self.g_mode, self.g_a, self.g_b = g
        self.h_mode, self.lam = h
        dt = solver._dtype
        self.dtype = dt
        self.sfx = '_f32' if dt == torch.float32 else '_f64'
        dev = K_device()
        self.ntaps = 1
        self.crow = self.ccol = None
        self.edge = None
        self.nl = self.nr = self.nt = self.nb = 0
        self.b = None
        from .._array import SlabRows
        if isinstance(y, SlabRows):
            raise NotImplementedError('row-decomposed data needs a solver constructed with slab=SlabComm(...)')
        if data_mode == DM_IDENTITY:
            self.b = to_device(y, dt)
        elif data_mode == DM_CONV:
            ccol, Et, Eb = composite_filter(A.hcol, A.offset[0])
            crow, El, Er = composite_filter(A.hrow, A.offset[1])
            n = max(ccol.size, crow.size)

            def centred(c):
                out = np.zeros(n)
                p = (n - c.size) // 2
                out[p:p + c.size] = c
                return out
            self.ntaps = n
            self.crow = np.ascontiguousarray(centred(crow))
            self.ccol = np.ascontiguousarray(centred(ccol))
            pack = np.zeros((4, MAX_HL, MAX_HL))
            for k, E in enumerate((El, Er, Et, Eb)):
                pack[k, :E.shape[0], :E.shape[1]] = E
            self.nl, self.nr, self.nt, self.nb = El.shape[0], Er.shape[0], Et.shape[0], Eb.shape[0]
            self.edge = to_device(pack.reshape(-1), dt)
            self._A, self._y = A, to_device(y, dt)          # A^T y is formed by begin_solve(), once per solve
        nbytes = int(_lib.call_plain('pcs_pds_tv_workspace_bytes'))
        self.ws = torch.zeros(nbytes // 8, dtype=torch.float64, device=dev)
        self._xbuf, self._zbuf = [None, None], [None, None]
        self.strip_cols = int(os.environ.get('PCS_PDS_STRIP_COLS', '0'))
        self.nseg = int(os.environ.get('PCS_PDS_NSEG', '0'))
        self.use_tma = int(os.environ.get('PCS_PDS_TMA', '1'))
        if self.b is not None and self.b.numel() != H * W:
            raise _NoMatch('data size mismatch')
        if data_mode == DM_CONV and self._y.numel() != H * W:
            raise _NoMatch('data size mismatch')

    def begin_solve(self):
        """Per-solve precomputation: ``b = A^T y`` for the deconvolution problems (the kernel evaluates the data-term
        gradient as ``(A^T A) x - b``).  Runs at the start of EVERY solve -- part of the solve's cost, and a ``y`` that was
        modified in place between two solves is honoured."""
        if self.data_mode == DM_CONV:
            self.b = self._A._adjoint(self._y)

    def _out(self, pool, like):
        for k in (0, 1):
            if pool[k] is None:
                pool[k] = torch.empty_like(like)
            if pool[k].data_ptr() != like.data_ptr():
                return pool[k]
        raise RuntimeError('no free ping-pong buffer')

    def release(self, *tensors):
        """The final iterates are handed to the caller: forget them, so that the next solve cannot overwrite them."""
        for t in tensors:
            for pool in (self._xbuf, self._zbuf):
                for k in (0, 1):
                    if pool[k] is not None and t is not None and pool[k].data_ptr() == t.data_ptr():
                        pool[k] = None

    def step(self, x, z, drow):
        s = self.solver
        if self.data_mode == DM_GIVEN:
            b = s.F._gradient(x)
        else:
            if self.b is None and self.data_mode == DM_CONV:
                self.begin_solve()
            b = self.b
        xo, zo = self._out(self._xbuf, x), self._out(self._zbuf, z)
        crow = self.crow.ctypes.data_as(C.c_void_p) if self.crow is not None else None
        ccol = self.ccol.ctypes.data_as(C.c_void_p) if self.ccol is not None else None
        _lib.call('pcs_pds_tv_step' + self.sfx, xo.data_ptr(), zo.data_ptr(), x.data_ptr(), z.data_ptr(), K.ptr(b),
                  self.H, self.W, 0, self.H, 0, self.H, self.data_mode, self.ntaps, crow, ccol, K.ptr(self.edge),
                  self.nl, self.nr, self.nt, self.nb, self.g_mode, self.g_a, self.g_b, self.h_mode, self.lam,
                  float(s.tau), float(s.sigma), float(s.rho), self.steps[0], self.steps[1], self.strip_cols, self.nseg,
                  self.use_tma, drow.data_ptr(), self.ws.data_ptr(), K.stream())
        return xo, zo


def last_launch():
    """Kernel and launch geometry of this thread's last fused Condat-Vu update (``pcs_pds_tv_last_launch``)."""
    out = np.zeros(8, dtype=np.int64)
    _lib.check(_lib.call_plain('pcs_pds_tv_last_launch', out.ctypes.data_as(C.c_void_p)))
    keys = ('kernel', 'blocks', 'threads', 'smem', 'strips', 'strip_cols', 'n_fast', 'use_tma')
    d = dict(zip(keys, (int(v) for v in out)))
    d['kernel'] = {0: 'pds_tv_kernel', 1: 'pds_tv_ws_kernel', 2: 'pds_tv_persist_kernel',
                   3: 'pds_tv_resident_kernel', 4: 'pds_tv_ws3_kernel'}.get(d['kernel'], 'none')
    return d


class PdsGradPlan:
    """Two-launch-per-iteration Condat-Vu TV solver for ANY 2-D ``Gradient`` kind (``csrc/pcs_pds_grad.cu``): pycsou's
    default ``kind='centered'`` and ``'backward'`` reach one row / column behind as well as ahead, which the one-kernel
    update (forward differences) does not cover.  ``F`` in {None, 1/2 SquaredL2Loss, anything differentiable (its
    gradient is evaluated by the stand-alone kernels and handed to the primal kernel)}."""

    def __init__(self, solver, H, W, steps, kind, edge, data_mode, y, g, h):
        self.solver = weakref.proxy(solver)
        self.H, self.W, self.steps = H, W, steps
        self.kind, self.edge = K.KIND[kind], int(bool(edge))
        self.data_mode = data_mode
        self.g_mode, self.g_a, self.g_b = g
        self.h_mode, self.lam = h
        self.dtype = dt = solver._dtype
        self.sfx = '_f32' if dt == torch.float32 else '_f64'
        self.b = to_device(y, dt) if data_mode == DM_IDENTITY else None
        if self.b is not None and self.b.numel() != H * W:
            raise _NoMatch('data size mismatch')
        nbytes = int(_lib.call_plain('pcs_pds_tv_workspace_bytes'))
        self.ws = torch.zeros(nbytes // 8, dtype=torch.float64, device=K_device())
        self.u = torch.empty(H * W, dtype=dt, device=K_device())
        self._xbuf, self._zbuf = [None, None], [None, None]

    _out = PdsTvPlan._out
    release = PdsTvPlan.release

    def step(self, x, z, drow):
        s = self.solver
        b = s.F._gradient(x) if self.data_mode == DM_GIVEN else self.b
        xo, zo = self._out(self._xbuf, x), self._out(self._zbuf, z)
        st = K.stream()
        _lib.call('pcs_pds_grad_primal_step' + self.sfx, xo.data_ptr(), self.u.data_ptr(), x.data_ptr(), K.ptr(b),
                  z.data_ptr(), self.H, self.W, self.kind, self.edge, self.data_mode, self.g_mode, self.g_a, self.g_b,
                  float(s.tau), float(s.rho), self.steps[0], self.steps[1], drow.data_ptr(), self.ws.data_ptr(), st)
        _lib.call('pcs_pds_grad_dual_step' + self.sfx, zo.data_ptr(), self.u.data_ptr(), z.data_ptr(), self.H, self.W,
                  self.kind, self.edge, self.h_mode, self.lam, float(s.sigma), float(s.rho), self.steps[0], self.steps[1],
                  drow.data_ptr(), self.ws.data_ptr(), st)
        return xo, zo


def _run_many(self, x, z, n, drows):
    """``n`` consecutive updates enqueued by ONE library call (``pcs_pds_tv_run``); ``drows``: n x 4 diagnostics rows."""
    s = self.solver
    if self.b is None and self.data_mode == DM_CONV:
        self.begin_solve()
    xo, zo = self._out(self._xbuf, x), self._out(self._zbuf, z)
    crow = self.crow.ctypes.data_as(C.c_void_p) if self.crow is not None else None
    ccol = self.ccol.ctypes.data_as(C.c_void_p) if self.ccol is not None else None
    _lib.call('pcs_pds_tv_run' + self.sfx, x.data_ptr(), z.data_ptr(), xo.data_ptr(), zo.data_ptr(), K.ptr(self.b),
              self.H, self.W, self.data_mode, self.ntaps, crow, ccol, K.ptr(self.edge), self.nl, self.nr, self.nt, self.nb,
              self.g_mode, self.g_a, self.g_b, self.h_mode, self.lam, float(s.tau), float(s.sigma), float(s.rho),
              self.steps[0], self.steps[1], self.strip_cols, self.nseg, self.use_tma, int(n), drows.data_ptr(),
              self.ws.data_ptr(), K.stream())
    _lib.launch_count += int(n) - 1                      # one call, n kernel launches
    return (x, z) if n % 2 == 0 else (xo, zo)


PdsTvPlan.run_many = _run_many


def K_device():
    from .._array import device
    return device()


class CpsMriPlan:
    """Compressed-sensing MRI with a TV penalty (BASELINE.json configs[3]) on the kernels of ``csrc/pcs_cps.cu``.

    ``K = LinOpVStack(Masking * DFT2, Gradient(forward))``, ``H = ProxFuncHStack(L2BallLoss(m, y, eps), lam * L21Norm)``,
    ``F = G = None``, complex iterates; dual variable ``z = [u (m) ; w0 ; w1]``.  One update = three launches of this
    library (scatter, fused primal + TV-dual stencil, data-consistency ball) and the two DFTs (cuFFT through
    ``torch.fft``, where the reference calls pocketfft), instead of ~25 stand-alone kernels.
    """

    def __init__(self, solver, H, W, steps, mask_op, dft, y, eps, lam):
        self.solver = weakref.proxy(solver)      # no reference cycle: the solver (and its result buffers) die by refcount
        self.H, self.W, self.steps = H, W, steps
        self.N, self.m = H * W, int(mask_op.nb_of_samples)
        dt = solver._dtype
        self.dtype = dt
        self.sfx = '_c64' if dt == torch.complex64 else '_c128'
        dev = K_device()
        self.idx = mask_op._idx(dev)
        self.dft = dft
        self.y = to_device(y, dt)
        if self.y.numel() != self.m:
            raise _NoMatch('data size mismatch')
        self.eps, self.lam = float(eps), float(lam)
        self.ws = torch.zeros(int(_lib.call_plain('pcs_cps_mri_workspace_bytes')) // 8, dtype=torch.float64, device=dev)
        self.spec = torch.empty(self.N, dtype=dt, device=dev)      # Mask^T u: scattered once, then kept up to date
        self.v = torch.empty(self.N, dtype=dt, device=dev)         # 2 xt - x
        self.scale = 1.0 / np.sqrt(float(self.N))                  # the orthonormal scaling is folded into the kernels
        self._spec_of = None                                       # (data_ptr, version) of the z whose u is in spec
        self._xbuf, self._zbuf = [None, None], [None, None]

    _out = PdsTvPlan._out

    def release(self, *tensors):
        PdsTvPlan.release(self, *tensors)
        self._spec_of = None             # the caller owns the final z now: its address may come back with other contents

    def step(self, x, z, drow):
        s = self.solver
        m, N, shp = self.m, self.N, (self.H, self.W)
        tau, sigma, rho = float(s.tau), float(s.sigma), float(s.rho)
        xo, zo = self._out(self._xbuf, x), self._out(self._zbuf, z)
        st = K.stream()
        if self._spec_of != (z.data_ptr(), z._version):
            _lib.call('pcs_cps_mri_scatter' + self.sfx, self.spec.data_ptr(), z.data_ptr(), self.idx.data_ptr(), N, m, st)
        # unnormalised transforms (no separate scaling pass): ifft2 'forward' = plain inverse, fft2 'backward' = plain forward
        g = torch.fft.ifft2(self.spec.view(shp), norm='forward')      # (out= would add a copy pass: torch scales into it)
        _lib.call('pcs_cps_mri_tv_step' + self.sfx, xo.data_ptr(), zo[m:].data_ptr(), self.v.data_ptr(), x.data_ptr(),
                  z[m:].data_ptr(), g.data_ptr(), self.H, self.W, self.scale, self.lam, tau, sigma, rho,
                  self.steps[0], self.steps[1], drow.data_ptr(), self.ws.data_ptr(), st)
        V = torch.fft.fft2(self.v.view(shp), norm='backward')
        _lib.call('pcs_cps_mri_ball_step' + self.sfx, zo.data_ptr(), self.spec.data_ptr(), z.data_ptr(), V.data_ptr(),
                  self.idx.data_ptr(), self.y.data_ptr(), m, self.scale, self.eps, sigma, rho, drow.data_ptr(),
                  self.ws.data_ptr(), st)
        self._spec_of = (zo.data_ptr(), zo._version)
        return xo, zo


def plan_cps_mri(solver):
    """Fused plan for the compressed-sensing MRI problem shape (complex iterates), or None."""
    from ..core.functional import ProxFuncPreComp
    from ..core.linop import LinOpComp
    from ..func.base import IndicatorFunctional, NullDifferentiableFunctional, NullProximableFunctional, ProxFuncHStack
    from ..func.penalty import L21Norm
    from ..linop.base import LinOpStack
    from ..linop.diff import GradientOperator
    from ..linop.sampling import Masking
    from ..linop.transform import DFT2
    from .._array import Negated
    try:
        if not (isinstance(solver.F, NullDifferentiableFunctional) and isinstance(solver.G, NullProximableFunctional)):
            raise _NoMatch('F and G must be None')
        Kop, Hf = solver.K, solver.H
        if not (isinstance(Kop, LinOpStack) and Kop.axis == 0 and len(Kop.linops) == 2):
            raise _NoMatch('K is not a vertical stack of two operators')
        MF, Gr = Kop.linops
        if not (isinstance(MF, LinOpComp) and isinstance(MF.LinOp1, Masking) and isinstance(MF.LinOp2, DFT2)
                and MF.LinOp2.batch == 1):
            raise _NoMatch('first block is not Masking * DFT2')
        if not (isinstance(Gr, GradientOperator) and Gr.kind == 'forward' and len(Gr.array_shape) == 2
                and tuple(Gr.array_shape) == tuple(MF.LinOp2.array_shape)):
            raise _NoMatch('second block is not a forward-difference Gradient on the DFT grid')
        H, W = Gr.array_shape
        N = H * W
        if N != solver.dim or H < 2 or W < 2 or MF.LinOp1.input_size != N:
            raise _NoMatch('shape mismatch')
        if not (isinstance(Hf, ProxFuncHStack) and len(Hf.proxfuncs) == 2):
            raise _NoMatch('H is not a stack of two functionals')
        ball, tv = Hf.proxfuncs
        if not (isinstance(ball, ProxFuncPreComp) and ball.scale == 1 and isinstance(ball.prox_func, IndicatorFunctional)
                and ball.prox_func.kind == 'l2ball' and ball.dim == MF.LinOp1.nb_of_samples):
            raise _NoMatch('first functional is not an L2BallLoss on the samples')
        sh = ball.shift
        if isinstance(sh, Number):
            raise _NoMatch('scalar data')
        y = sh.a if isinstance(sh, Negated) else (-sh if isinstance(sh, torch.Tensor) else -np.asarray(sh))
        f, lam = _scaled(tv)
        if not (type(f) is L21Norm and f._is_planes and f.n_comp == 2 and f.n_groups == N):
            raise _NoMatch('second functional is not a TV penalty')
        if not (float(solver.tau) > 0 and float(solver.sigma) > 0 and lam > 0):
            raise _NoMatch('non-positive step')
        return CpsMriPlan(solver, H, W, Gr.steps, MF.LinOp1, MF.LinOp2, y, ball.prox_func.params[0], lam)
    except _NoMatch:
        return None


def plan_pds(solver):
    """Fused plan for a ``PrimalDualSplitting`` instance, or None (generic op-by-op path)."""
    if solver._H and solver._dtype in (torch.complex64, torch.complex128):
        return plan_cps_mri(solver)
    if not solver._H or solver._dtype not in (torch.float32, torch.float64):
        return None
    try:
        H, W, steps = _match_K(solver.K, solver.dim, any_kind=True)
        h = _match_H(solver.H, H * W)
        g = _match_G(solver.G, float(solver.tau))
        data_mode, y, A = _match_F(solver.F, (H, W))
        if not (float(solver.tau) > 0 and float(solver.sigma) > 0):
            raise _NoMatch('non-positive step')
        if solver.K.kind == 'forward':
            return PdsTvPlan(solver, H, W, steps, data_mode, y, A, g, h)
        # centred / backward differences: two kernels per update; a convolution data term is evaluated op by op
        if data_mode == DM_CONV:
            data_mode, y = DM_GIVEN, None
        if h[1] <= 0:
            raise _NoMatch('non-positive regularisation weight')
        return PdsGradPlan(solver, H, W, steps, solver.K.kind, solver.K.edge, data_mode, y, g, h)
    except _NoMatch:
        return None


def apgd_last_launch():
    """Kernel and launch geometry of this thread's last one-launch APGD solve (``pcs_apgd_last_launch``)."""
    out = np.zeros(4, dtype=np.int64)
    _lib.check(_lib.call_plain('pcs_apgd_last_launch', out.ctypes.data_as(C.c_void_p)))
    d = dict(zip(('kernel', 'blocks', 'threads', 'per_thread'), (int(v) for v in out)))
    d['kernel'] = {0: 'apgd_run_kernel', 1: 'apgd_cluster_kernel'}.get(d['kernel'], 'none')
    return d


class ApgdConv1dPlan:
    """Whole-solve plan for ``APGD`` with ``F = 1/2 ||h * x - y||^2`` on a flat vector (``csrc/pcs_apgd.cu``)."""

    ACC = {None: 0, 'BT': 1, 'CD': 2}

    def __init__(self, solver, A, y, g_mode, g_param):
        dt = solver._dtype
        self.solver, self.dtype = weakref.proxy(solver), dt
        self.sfx = '_f32' if dt == torch.float32 else '_f64'
        self.y = to_device(y, dt)
        self.h = to_device(A.filter, dt)
        self.nh, self.offset, self.n = int(A.filter.size), int(A.offset), int(A.n)
        self.g_mode, self.g_param = g_mode, float(g_param)
        self.n_iter = torch.zeros(2, dtype=torch.int64, device=K_device())

    def run(self, x, aux, t0, it0, diag_dev, single=False):
        """Updates ``it0, it0 + 1, ...`` under the reference's loop condition (``single``: exactly update ``it0``);
        ``diag_dev`` receives one row per update, starting at its first row.  Returns (x, aux, t, iter)."""
        s = self.solver
        x, aux = x.clone(), aux.clone()
        hi, lo, thr = (it0, it0, 0.0) if single else (s.max_iter, s.min_iter, s.accuracy_threshold)
        _lib.call('pcs_apgd_conv1d_run' + self.sfx, x.data_ptr(), aux.data_ptr(), self.y.data_ptr(), self.h.data_ptr(),
                  self.nh, self.offset, self.n, self.g_mode, self.g_param, float(s.tau), self.ACC[s.acceleration],
                  float(s.d), float(t0), int(it0), int(hi), int(lo), float(thr),
                  diag_dev.data_ptr(), self.n_iter.data_ptr(), K.stream())
        n_iter, t_bits = (int(v) for v in self.n_iter.cpu())
        t = float(np.int64(t_bits).view(np.float64))
        return x, aux, (t if n_iter > it0 else t0), n_iter


def plan_apgd(solver):
    """Whole-solve plan for an ``AcceleratedProximalGradientDescent`` instance, or None (generic op-by-op path)."""
    from ..core.map import DiffMapComp
    from ..func.base import IndicatorFunctional, NullProximableFunctional
    from ..func.penalty import L1Norm
    from ..linop.conv import Convolve1DOperator
    if solver._dtype not in (torch.float32, torch.float64) or solver.acceleration not in ApgdConv1dPlan.ACC:
        return None
    F = solver.F
    if not (isinstance(F, DiffMapComp) and isinstance(F.map2, Convolve1DOperator)):
        return None
    A, y = F.map2, _half_sq_loss(F.map1)
    from .._array import SlabRows
    if y is None or isinstance(y, SlabRows) or A.outer != 1 or A.inner != 1 or A.n != solver.dim:
        return None
    if A.n > int(_lib.call_plain('pcs_apgd_conv1d_run_max_n')) or A.filter.size > 256:
        return None
    if not float(solver.tau) > 0 or int(np.prod(np.shape(y))) != A.n:
        return None
    G = solver.G
    if isinstance(G, NullProximableFunctional):
        g = (0, 0.0)
    elif isinstance(G, IndicatorFunctional) and G.kind == 'nonneg':
        g = (2, 0.0)
    else:
        try:
            f, lam = _scaled(G)
        except _NoMatch:
            return None
        if type(f) is not L1Norm:
            return None
        g = (1, float(solver.tau) * lam)
    return ApgdConv1dPlan(solver, A, y, *g)
