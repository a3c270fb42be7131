"""Thin typed wrappers: torch CUDA tensors -> libpcsb200 C-ABI calls.

PyTorch is used only for device memory and streams.  Every function launches hand-written CUDA kernels on
``torch.cuda.current_stream()``; tensors must be contiguous 1-D CUDA tensors of dtype float32/float64
(complex tensors are passed as interleaved real views where an entry point supports it).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib

_SUFFIX = {torch.float32: '_f32', torch.float64: '_f64'}
_REAL_OF = {torch.complex64: torch.float32, torch.complex128: torch.float64}
_ws_cache = {}


def require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError('pycsou_b200 needs a CUDA device (B200); there is no CPU fallback')


def stream():
    return torch.cuda.current_stream().cuda_stream


def sfx(t: torch.Tensor) -> str:
    try:
        return _SUFFIX[_REAL_OF.get(t.dtype, t.dtype)]
    except KeyError:
        raise NotImplementedError(f'dtype {t.dtype} is not supported (float32/float64/complex64/complex128 only)')


def real_view(t: torch.Tensor) -> torch.Tensor:
    """Interleaved (re, im) real view of a complex tensor; identity for real tensors."""
    return torch.view_as_real(t).reshape(-1) if t.is_complex() else t


def ptr(t):
    return 0 if t is None else t.data_ptr()


def reduce_ws(device) -> torch.Tensor:
    """Zero-initialised reduction workspace (partials + ticket), one per (device, stream)."""
    key = (device.index if device.index is not None else torch.cuda.current_device(), stream())
    ws = _ws_cache.get(key)
    if ws is None:
        nbytes = int(_lib.call_plain('pcs_reduce_workspace_bytes'))
        ws = torch.zeros(nbytes // 8, dtype=torch.float64, device=device)
        _ws_cache[key] = ws
    return ws


def _chk(*ts):
    for t in ts:
        if t is not None and not (t.is_cuda and t.is_contiguous()):
            raise ValueError('expected contiguous CUDA tensors')


# ------------------------------------------------------------------------------------------- elementwise
def lincomb2(out, a, x, b, y):
    """out = a*x + b*y (real coefficients; complex tensors handled through their real view)."""
    _chk(out, x, y)
    o, xr, yr = real_view(out), real_view(x), real_view(y)
    _lib.call('pcs_lincomb2' + sfx(x), ptr(o), float(a), ptr(xr), float(b), ptr(yr), xr.numel(), stream())
    return out


def lincomb3(out, a, x, b, y, c, z):
    _chk(out, x, y, z)
    o, xr, yr, zr = real_view(out), real_view(x), real_view(y), real_view(z)
    _lib.call('pcs_lincomb3' + sfx(x), ptr(o), float(a), ptr(xr), float(b), ptr(yr), float(c), ptr(zr), xr.numel(),
              stream())
    return out


def scale(out, a, x):
    _chk(out, x)
    o, xr = real_view(out), real_view(x)
    _lib.call('pcs_scale' + sfx(x), ptr(o), float(a), ptr(xr), xr.numel(), stream())
    return out


def mul(out, x, y):
    _chk(out, x, y)
    if x.is_complex() or y.is_complex():
        raise NotImplementedError('complex element-wise product')
    _lib.call('pcs_mul' + sfx(x), ptr(out), ptr(x), ptr(y), x.numel(), stream())
    return out


def prox_l1(out, x, tau):
    _chk(out, x)
    if x.is_complex():
        raise NotImplementedError('L1Norm.prox is real-only in the reference (ordered comparisons, math/prox.py:254)')
    _lib.call('pcs_prox_l1' + sfx(x), ptr(out), ptr(x), float(tau), x.numel(), stream())
    return out


def _real_only(x, what):
    if x.is_complex():
        raise NotImplementedError(f'{what} is defined for real vectors')


def prox_logbarrier(out, x, tau):
    _chk(out, x); _real_only(x, 'LogBarrier.prox')
    _lib.call('pcs_prox_logbarrier' + sfx(x), ptr(out), ptr(x), float(tau), x.numel(), stream())
    return out


def prox_shannon(out, x, tau):
    _chk(out, x); _real_only(x, 'ShannonEntropy.prox')
    _lib.call('pcs_prox_shannon' + sfx(x), ptr(out), ptr(x), float(tau), x.numel(), stream())
    return out


def prox_kl(out, x, data, tau):
    _chk(out, x, data); _real_only(x, 'KLDivergence.prox')
    _lib.call('pcs_prox_kl' + sfx(x), ptr(out), ptr(x), ptr(data), float(tau), x.numel(), stream())
    return out


def reduce_neglog(x):
    _chk(x); _real_only(x, 'LogBarrier')
    out = _red_out(x)
    _lib.call('pcs_reduce_neglog' + sfx(x), ptr(x), x.numel(), ptr(out), ptr(reduce_ws(x.device)), stream())
    return out


def reduce_xlogx(x):
    _chk(x); _real_only(x, 'ShannonEntropy')
    out = _red_out(x)
    _lib.call('pcs_reduce_xlogx' + sfx(x), ptr(x), x.numel(), ptr(out), ptr(reduce_ws(x.device)), stream())
    return out


def reduce_kl(x, data):
    _chk(x, data); _real_only(x, 'KLDivergence')
    out = _red_out(x)
    _lib.call('pcs_reduce_kl' + sfx(x), ptr(x), ptr(data), x.numel(), ptr(out), ptr(reduce_ws(x.device)), stream())
    return out


def proj_nonneg(out, x):
    _chk(out, x)
    _lib.call('pcs_proj_nonneg' + sfx(x), ptr(out), ptr(x), x.numel(), stream())
    return out


def proj_segment(out, x, a, b):
    _chk(out, x)
    _lib.call('pcs_proj_segment' + sfx(x), ptr(out), ptr(x), float(a), float(b), x.numel(), stream())
    return out


def prox_l21(out, x, tau, n_groups, n_comp):
    _chk(out, x)
    _lib.call('pcs_prox_l21' + sfx(x), ptr(real_view(out)), ptr(real_view(x)), float(tau), int(n_groups), int(n_comp),
              int(x.is_complex()), stream())
    return out


# ------------------------------------------------------------------------------------------- reductions
def _red_out(x, k=1):
    return torch.empty(k, dtype=torch.float64, device=x.device)


def sumsq(x, out=None):
    """||x||^2 as a 1-element float64 device tensor (no host sync)."""
    _chk(x)
    out = _red_out(x) if out is None else out
    xr = real_view(x)
    _lib.call('pcs_reduce_sumsq' + sfx(x), ptr(xr), xr.numel(), ptr(out), ptr(reduce_ws(x.device)), stream())
    return out


def sumabs(x, out=None):
    _chk(x)
    if x.is_complex():
        raise NotImplementedError('complex L1 norm')
    out = _red_out(x) if out is None else out
    _lib.call('pcs_reduce_sumabs' + sfx(x), ptr(x), x.numel(), ptr(out), ptr(reduce_ws(x.device)), stream())
    return out


def diffsumsq(x, y, out=None):
    _chk(x, y)
    out = _red_out(x) if out is None else out
    xr, yr = real_view(x), real_view(y)
    _lib.call('pcs_reduce_diffsumsq' + sfx(x), ptr(xr), ptr(yr), xr.numel(), ptr(out), ptr(reduce_ws(x.device)),
              stream())
    return out


def dot(x, y, out=None):
    _chk(x, y)
    if x.is_complex():
        raise NotImplementedError('complex dot product')
    out = _red_out(x) if out is None else out
    _lib.call('pcs_reduce_dot' + sfx(x), ptr(x), ptr(y), x.numel(), ptr(out), ptr(reduce_ws(x.device)), stream())
    return out


def minmax(x):
    _chk(x)
    out = _red_out(x, 2)
    _lib.call('pcs_reduce_minmax' + sfx(x), ptr(x), x.numel(), ptr(out), ptr(reduce_ws(x.device)), stream())
    return out


def threshold_newton(x, radius, c0, batch=4, max_steps=4096):
    """Exact root ``t`` of ``sum max(|x| - t, 0) = radius + c0 t`` by monotone Newton steps on the device (the l1-ball
    and squared-l1 thresholds; see ``pcs_threshold_newton_step``).  ``batch`` steps are enqueued between host reads;
    the iteration has converged when the count of entries above the threshold stops changing."""
    _chk(x); _real_only(x, 'threshold_newton')
    state = torch.zeros(4, dtype=torch.float64, device=x.device)
    name, ws, st = 'pcs_threshold_newton_step' + sfx(x), ptr(reduce_ws(x.device)), stream()
    src, dst, done = -1, 0, 0
    while done < max_steps:
        for _ in range(batch):
            _lib.call(name, ptr(x), x.numel(), float(radius), float(c0), ptr(state), src, dst, ws, st)
            src, dst = dst, 1 - dst
        done += batch
        vals = state.tolist()
        (s_new, c_new), (s_old, c_old) = (vals[2 * src], vals[2 * src + 1]), (vals[2 * dst], vals[2 * dst + 1])
        t_new = max((s_new - radius) / (c_new + c0), 0.0) if c_new + c0 > 0 else 0.0
        t_old = max((s_old - radius) / (c_old + c0), 0.0) if c_old + c0 > 0 else 0.0
        # same active set, or an entry within rounding of the root flipping in and out of it
        if c_new == c_old or abs(t_new - t_old) <= 4e-16 * abs(t_new):
            return t_new
    raise RuntimeError('threshold_newton did not converge')


def sumgroupnorm(x, n_groups, n_comp):
    _chk(x)
    out = _red_out(x)
    _lib.call('pcs_reduce_sumgroupnorm' + sfx(x), ptr(real_view(x)), int(n_groups), int(n_comp), int(x.is_complex()),
              ptr(out), ptr(reduce_ws(x.device)), stream())
    return out


def prox_l21_csr(out, x, tau, members, offsets):
    """L21Norm.prox for arbitrary (ragged) groups given in CSR form (int64 device tensors)."""
    _chk(out, x, members, offsets)
    _lib.call('pcs_prox_l21_csr' + sfx(x), ptr(real_view(out)), ptr(real_view(x)), float(tau), ptr(members), ptr(offsets),
              int(offsets.numel()) - 1, int(x.is_complex()), stream())
    return out


def sumgroupnorm_csr(x, members, offsets):
    _chk(x, members, offsets)
    out = _red_out(x)
    _lib.call('pcs_reduce_sumgroupnorm_csr' + sfx(x), ptr(real_view(x)), ptr(members), ptr(offsets),
              int(offsets.numel()) - 1, int(x.is_complex()), ptr(out), ptr(reduce_ws(x.device)), stream())
    return out


def scale_by_norm(out, x, sumsq_dev, param, mode):
    """mode 0: L2Norm.prox factor max(1 - param/||x||, 0); mode 1: l2-ball projection factor."""
    _chk(out, x)
    xr = real_view(x)
    _lib.call('pcs_scale_by_norm' + sfx(x), ptr(real_view(out)), ptr(xr), ptr(sumsq_dev), float(param), int(mode),
              xr.numel(), stream())
    return out


# ------------------------------------------------------------------------------------------- stencils
KIND = {'forward': 0, 'centered': 1, 'backward': 2}


def firstderiv(out, x, outer, n, inner, kind, edge, step, adjoint=False, accumulate=False):
    _chk(out, x)
    e = 2 if x.is_complex() else 1
    name = 'pcs_firstderiv_adj' if adjoint else 'pcs_firstderiv_fwd'
    _lib.call(name + sfx(x), ptr(real_view(out)), ptr(real_view(x)), int(outer), int(n), int(inner) * e,
              KIND[kind], int(bool(edge)), float(step), int(accumulate), stream())
    return out


def cumsum(out, x, outer, n, inner, step, adjoint=False):
    """Running sum along one axis times ``step`` (forward) / the same from the far end (adjoint)."""
    _chk(out, x)
    e = 2 if x.is_complex() else 1
    name = 'pcs_cumsum_adj' if adjoint else 'pcs_cumsum_fwd'
    _lib.call(name + sfx(x), ptr(real_view(out)), ptr(real_view(x)), int(outer), int(n), int(inner) * e, float(step),
              stream())
    return out


def secondderiv(out, x, outer, n, inner, edge, step, adjoint=False, accumulate=False):
    _chk(out, x)
    e = 2 if x.is_complex() else 1
    name = 'pcs_secondderiv_adj' if adjoint else 'pcs_secondderiv_fwd'
    _lib.call(name + sfx(x), ptr(real_view(out)), ptr(real_view(x)), int(outer), int(n), int(inner) * e, 0,
              int(bool(edge)), float(step), int(accumulate), stream())
    return out


def grad2d(out, x, H, W, kind, edge, step0, step1, adjoint=False):
    _chk(out, x)
    if x.is_complex():
        raise NotImplementedError
    name = 'pcs_grad2d_adj' if adjoint else 'pcs_grad2d_fwd'
    _lib.call(name + sfx(x), ptr(out), ptr(x), int(H), int(W), KIND[kind], int(bool(edge)), float(step0), float(step1),
              stream())
    return out


def laplacian2d(out, x, H, W, w0, w1, step0, step1, edge, adjoint=False):
    _chk(out, x)
    if x.is_complex():
        raise NotImplementedError
    name = 'pcs_laplacian2d_adj' if adjoint else 'pcs_laplacian2d_fwd'
    _lib.call(name + sfx(x), ptr(out), ptr(x), int(H), int(W), float(w0), float(w1), float(step0), float(step1),
              int(bool(edge)), stream())
    return out


# ------------------------------------------------------------------------------------------- convolution
def conv1d(out, x, h, offset, outer, n, inner, adjoint=False):
    _chk(out, x, h)
    if x.is_complex():
        raise NotImplementedError
    name = 'pcs_conv1d_adj' if adjoint else 'pcs_conv1d_fwd'
    _lib.call(name + sfx(x), ptr(out), ptr(x), ptr(h), h.numel(), int(offset), int(outer), int(n), int(inner),
              stream())
    return out


def conv2d_direct(out, x, h2d, off0, off1, H, W, adjoint=False):
    _chk(out, x, h2d)
    name = 'pcs_conv2d_direct_adj' if adjoint else 'pcs_conv2d_direct_fwd'
    _lib.call(name + sfx(x), ptr(out), ptr(x), ptr(h2d), h2d.shape[0], h2d.shape[1], int(off0), int(off1), int(H),
              int(W), stream())
    return out


def spectral_mul(a, b, conj_b=False):
    """a <- a * b or a * conj(b) in place (complex tensors of equal size; csrc/pcs_conv.cu)."""
    if not (a.is_complex() and b.is_complex() and a.dtype == b.dtype and a.numel() == b.numel()):
        raise ValueError('spectral_mul needs two complex tensors of the same dtype and size')
    if not (a.is_contiguous() and b.is_contiguous()):
        raise ValueError('spectral_mul needs contiguous tensors')
    _lib.call('pcs_spectral_mul' + sfx(a), ptr(a), ptr(b), int(bool(conj_b)), a.numel(), stream())
    return a


def crop_shift(out, full, H, W, P0, P1, s0, s1, scale, sub=None):
    """out[i, j] = scale * full[(i + s0) % P0, (j + s1) % P1] - sub[i, j]  (csrc/pcs_conv.cu)."""
    _chk(out, full, sub)
    _lib.call('pcs_crop_shift' + sfx(out), ptr(out), ptr(full), ptr(sub), int(H), int(W), int(P0), int(P1), int(s0),
              int(s1), float(scale), stream())
    return out


def conv2d_sep(out, x, hcol, hrow, off0, off1, H, W, adjoint=False, sub=None):
    _chk(out, x, hcol, hrow, sub)
    name = 'pcs_conv2d_sep_adj' if adjoint else 'pcs_conv2d_sep_fwd'
    _lib.call(name + sfx(x), ptr(out), ptr(x), ptr(hcol), ptr(hrow), hcol.numel(), hrow.numel(), int(off0), int(off1),
              ptr(sub), int(H), int(W), stream())
    return out


# ------------------------------------------------------------------------------------------- sampling
def take(out, x, idx, outer, n, inner):
    _chk(out, x, idx)
    _lib.call('pcs_take', ptr(out), ptr(x), ptr(idx), int(outer), int(n), idx.numel(), int(inner), x.element_size(),
              stream())
    return out


def put(out, y, idx, outer, n, inner):
    _chk(out, y, idx)
    _lib.call('pcs_put', ptr(out), ptr(y), ptr(idx), int(outer), int(n), idx.numel(), int(inner), y.element_size(),
              stream())
    return out


def pool_fwd(out, x, dims, block, scale):
    """Block sum / mean of a C-order array of up to three axes (``dims``, ``block`` padded to length 3 with ones)."""
    _chk(out, x); _real_only(x, 'Pooling')
    _lib.call('pcs_pool_fwd' + sfx(x), ptr(out), ptr(x), *map(int, dims), *map(int, block), float(scale), stream())
    return out


def pool_adj(out, y, dims, block):
    _chk(out, y); _real_only(y, 'Pooling')
    _lib.call('pcs_pool_adj' + sfx(y), ptr(out), ptr(y), *map(int, dims), *map(int, block), stream())
    return out


def mask_to_indices(mask_u8: torch.Tensor) -> torch.Tensor:
    """Sorted positions of the set entries of a uint8 device mask (construction-time; one host sync)."""
    _chk(mask_u8)
    n = mask_u8.numel()
    ws = torch.empty(int(_lib.call_plain('pcs_mask_workspace_bytes', n)) // 8, dtype=torch.int64, device=mask_u8.device)
    total = torch.empty(1, dtype=torch.int64, device=mask_u8.device)
    _lib.call('pcs_mask_count', ptr(mask_u8), n, ptr(ws), ptr(total), stream())
    m = int(total.item())
    idx = torch.empty(m, dtype=torch.int64, device=mask_u8.device)
    if m == 0:
        return idx
    _lib.call('pcs_mask_indices', ptr(mask_u8), n, ptr(ws), ptr(idx), stream())
    return idx
