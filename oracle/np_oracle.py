"""Stand-alone numpy restatement of the pycsou proximal-solver hot path (the parity oracle).

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``): imported by ``tests/``,
``__graft_entry__.smoke()`` and the CPU-baseline legs of ``bench.py`` -- never by ``pycsou_b200``.

Every function cites the reference ``file:line`` (relative to /root/reference) it restates.  The
module depends on numpy/scipy only, so it travels to the GPU box where the reference tree does not
exist.  Pinning: ``tests/test_oracle_doctests.py`` (the reference's own doctest literals) and
``tests/test_oracle_golden.py`` (fixtures produced by the reference classes themselves through
``oracle.ref_shim``; generator ``tests/golden/gen_golden.py``).

Deliberate deviations from the reference, each verified equal on the reference's feasible range:
* ``L21.prox``/``__call__`` are vectorised (the reference loops over groups in Python, Theta(N^2)).
* No ``deepcopy`` of iterands and no pandas ``DataFrame``: diagnostics are returned as a dict of lists
  with the reference's column names.
"""
from __future__ import annotations

import numpy as np

from oracle import pylops_standin as pl

INF = np.inf


# --------------------------------------------------------------------------------------------------
# math/prox.py
# --------------------------------------------------------------------------------------------------
def sign(x):
    """math/prox.py:61-64 -- NB returns conj(x)/|x| (conjugates complex input)."""
    x = np.asarray(x)
    y = np.asarray(0 * x)
    nz = np.abs(x) != 0
    y[nz] = np.conj(x[nz]) / np.abs(x[nz])
    return y


def soft(x, tau):
    """math/prox.py:114."""
    return np.clip(np.abs(x) - tau, a_min=0, a_max=None) * sign(x)


def proj_l2_ball(x, radius):
    """math/prox.py:207-210."""
    nrm = np.linalg.norm(x)
    return x if nrm <= radius else radius * x / nrm


def proj_l1_ball(x, radius):
    """math/prox.py:157-164: the argument itself when feasible, else soft-thresholding at the brentq root of
    mu -> sum max(|x| - mu, 0) - radius on [0, max|x|]."""
    from scipy import optimize as sciop
    if np.sum(np.abs(x)) <= radius:
        return x
    mu_star = sciop.brentq(lambda mu: np.sum(np.clip(np.abs(x) - mu, a_min=0, a_max=None)) - radius,
                           a=0, b=np.max(np.abs(x)))
    return soft(x, mu_star)


def proj_linfty_ball(x, radius):
    """math/prox.py:253-256 (the reference clips in place; the oracle works on a copy)."""
    y = np.array(x, copy=True)
    y[y > radius] = radius
    y[y < -radius] = -radius
    return y


def proj_nonnegative_orthant(x):
    """math/prox.py:295-297."""
    y = np.array(np.real(x), copy=True)
    y[y < 0] = 0
    return y


def proj_segment(x, a=0, b=1):
    """math/prox.py:340-343."""
    y = np.array(np.real(x), copy=True)
    y[y < a] = a
    y[y > b] = b
    return y


# --------------------------------------------------------------------------------------------------
# core/functional.py + func/base.py + func/penalty.py + func/loss.py
# --------------------------------------------------------------------------------------------------
class Prox:
    """ProximableFunctional surface (core/functional.py:157-207)."""
    dim: int

    def __call__(self, x):
        raise NotImplementedError

    def prox(self, x, tau):
        raise NotImplementedError

    def fenchel_prox(self, z, sigma):
        """core/functional.py:207 (Moreau identity)."""
        return z - sigma * self.prox(z / sigma, 1 / sigma)

    def __rmul__(self, lam):
        """core/functional.py:244-246 (positive scalars only)."""
        if not (np.isscalar(lam) and lam > 0):
            raise NotImplementedError
        return PostComp(self, lam)


class PostComp(Prox):
    """ProxFuncPostComp, core/functional.py:253-265."""

    def __init__(self, f, scale):
        self.f, self.scale, self.dim = f, scale, f.dim

    def __call__(self, x):
        return self.scale * self.f(x)

    def prox(self, x, tau):
        return self.f.prox(x, tau * self.scale)


class PreCompShift(Prox):
    """ProxFuncPreComp with scale=1, shift=-data: ``ProximableLoss`` (func/loss.py:69; functional.py:298-299)."""

    def __init__(self, f, data):
        self.f, self.shift, self.dim = f, -np.asarray(data), f.dim

    def __call__(self, x):
        return self.f(1 * x + self.shift)

    def prox(self, x, tau):
        return (self.f.prox(1 * x + self.shift, tau * (1 ** 2)) - self.shift) / 1


class LpNorm(Prox):
    """func/base.py:239-240: prox = x - tau * proj_lq_ball(x / tau, radius=1)."""
    _proj = None

    def __init__(self, dim):
        self.dim = dim

    def prox(self, x, tau):
        return x - tau * type(self)._proj(x / tau, radius=1)


class L1Norm(LpNorm):
    """func/penalty.py:194-245."""
    _proj = staticmethod(proj_linfty_ball)

    def __call__(self, x):
        return np.sum(np.abs(x))


class L2Norm(LpNorm):
    """func/penalty.py:23-70."""
    _proj = staticmethod(proj_l2_ball)

    def __call__(self, x):
        return np.linalg.norm(x)


class L21Norm(Prox):
    """func/penalty.py:480-560, vectorised.  ``groups`` is an arbitrary label array of length dim."""

    def __init__(self, dim, groups):
        self.dim = dim
        self.groups = np.asarray(groups).reshape(-1)
        self.labels, self.inv = np.unique(self.groups, return_inverse=True)

    def _norms(self, x):
        return np.sqrt(np.bincount(self.inv, weights=np.abs(x) ** 2, minlength=self.labels.size))

    def __call__(self, x):
        return np.sum(self._norms(x))

    def prox(self, x, tau):
        with np.errstate(divide='ignore'):
            scale = np.clip(1 - tau / self._norms(x), a_min=0, a_max=None)  # penalty.py:554
        return scale[self.inv] * x


class TVL21Norm(Prox):
    """L21 for the TV layout ``groups = tile(arange(N), C)`` (one group per pixel): fast path of L21Norm."""

    def __init__(self, dim, n_groups):
        self.dim, self.n = dim, n_groups

    def __call__(self, x):
        return np.sum(np.linalg.norm(np.reshape(x, (-1, self.n)), axis=0))

    def prox(self, x, tau):
        X = np.reshape(x, (-1, self.n))
        with np.errstate(divide='ignore'):
            scale = np.clip(1 - tau / np.linalg.norm(X, axis=0), a_min=0, a_max=None)
        return (scale[None, :] * X).reshape(-1)


class Indicator(Prox):
    """IndicatorFunctional, func/base.py:140-168."""

    def __init__(self, dim, cond, proj):
        self.dim, self.cond, self.proj = dim, cond, proj

    def __call__(self, x):
        return 0 if self.cond(x) else INF

    def prox(self, x, tau):
        return self.proj(x)


def NonNegativeOrthant(dim):
    """func/penalty.py:610-612."""
    return Indicator(dim, lambda x: np.all(x >= 0), proj_nonnegative_orthant)


def Segment(dim, a=0, b=1):
    """func/penalty.py:666-668."""
    return Indicator(dim, lambda x: np.all((x >= a) & (x <= b)), lambda x: proj_segment(x, a=a, b=b))


def L2Ball(dim, radius):
    """func/penalty.py:189-191."""
    return Indicator(dim, lambda x: np.linalg.norm(x) <= radius, lambda x: proj_l2_ball(x, radius=radius))


class NullProx(Prox):
    """NullProximableFunctional, func/base.py:194-212."""

    def __init__(self, dim):
        self.dim = dim

    def __call__(self, x):
        return 0

    def prox(self, x, tau):
        return x


class ProxHStack(Prox):
    """ProxFuncHStack, func/base.py:80-89 (separable prox over concatenated blocks)."""

    def __init__(self, *fs):
        self.fs = fs
        self.sections = np.cumsum([f.dim for f in fs])
        self.dim = int(self.sections[-1])

    def __call__(self, x):
        return sum(f(xi) for f, xi in zip(self.fs, np.split(x, self.sections)))

    def prox(self, x, tau):
        return np.concatenate([f.prox(xi, tau) for f, xi in zip(self.fs, np.split(x, self.sections))], axis=0)


class HalfSquaredL2Loss:
    """``F = (1/2) * SquaredL2Loss(dim, data=y) * G`` (doc/general/examples.rst:37-38).

    gradient follows the reference's chain rule literally (core/map.py:609-610,
    571-572; func/penalty.py:129-131; linop/base.py:632-633):  G^T( (2 * (G x + (-y))) * 0.5 ).
    ``G=None`` means no operator (pure denoising).  beta = diff_lipschitz_cst = 0.5 * 2 * ||G||^2
    (core/map.py:601-605).
    """

    def __init__(self, data, G=None, G_lipschitz=1.0):
        self.y, self.G = np.asarray(data), G
        self.diff_lipschitz_cst = 0.5 * 2 * (G_lipschitz if G is not None else 1.0) * \
            (G_lipschitz if G is not None else 1.0)
        self.dim = G.shape[1] if G is not None else self.y.size

    def __call__(self, x):
        r = (self.G.matvec(x) if self.G is not None else x) + (-self.y)
        return 0.5 * np.linalg.norm(r) ** 2

    def gradient(self, x):
        r = (self.G.matvec(x) if self.G is not None else x) + (-self.y)
        g = (2 * r) * 0.5
        return self.G.rmatvec(g) if self.G is not None else g


class NullDiff:
    """NullDifferentiableFunctional, func/base.py:171-191."""

    def __init__(self, dim):
        self.dim, self.diff_lipschitz_cst = dim, 0

    def __call__(self, x):
        return 0

    def gradient(self, x):
        return np.zeros(shape=self.dim)


# --------------------------------------------------------------------------------------------------
# linop/sampling.py (own code of the reference, not pylops)
# --------------------------------------------------------------------------------------------------
class Masking:
    """linop/sampling.py:166-196."""

    def __init__(self, size, sampling_bool, dtype=np.float64):
        self.sampling_bool = np.asarray(sampling_bool).reshape(-1).astype(bool)
        if self.sampling_bool.size != size:
            raise ValueError('Invalid size of boolean sampling array.')
        self.input_size, self.nb_of_samples = size, int(self.sampling_bool.sum())
        self.shape, self.dtype = (self.nb_of_samples, size), dtype

    def matvec(self, x):
        return x[self.sampling_bool]

    def rmatvec(self, y):
        x = np.zeros(shape=self.input_size, dtype=self.dtype)
        x[self.sampling_bool] = y
        return x


def downsampling_mask(size, factor, shape=None, axis=None):
    """linop/sampling.py:323-329 (factor normalisation) and 362-391 (compute_downsampling_mask)."""
    if type(factor) is int:
        factor = len(shape) * (factor,) if (shape is not None and axis is None) else (factor,)
    else:
        factor = tuple(factor)
    if shape is None:
        return (np.arange(size) % factor) == 0
    if len(factor) > 1:
        mask = True
        for ax in range(len(shape)):
            idx = (np.arange(shape[ax]) % factor[ax]).reshape((shape[ax],) + (len(shape) - 1) * (1,))
            mask = mask * (np.swapaxes(idx, 0, ax) == 0)
        return np.asarray(mask).reshape(-1)
    mask = np.swapaxes(np.zeros(shape=shape, dtype=bool), 0, axis)
    mask[(np.arange(shape[axis]) % factor) == 0, ...] = True
    return np.swapaxes(mask, 0, axis).reshape(-1)


class LinOpVStackOracle:
    """LinOpVStack, linop/base.py:159-314: concatenated outputs; adjoint splits at cumsum(block sizes) and sums."""

    def __init__(self, *ops):
        self.ops = ops
        self.sections = np.cumsum([op.shape[0] for op in ops])
        self.shape = (int(self.sections[-1]), ops[0].shape[1])

    def matvec(self, x):
        return np.concatenate([np.asarray(op.matvec(x)).flatten() for op in self.ops], axis=0)

    def rmatvec(self, y):
        res = 0
        for op, yi in zip(self.ops, np.split(y, self.sections)):
            res = res + op.rmatvec(yi)
        return res


class FFT2Ortho:
    """Unitary 2-D DFT (the reference ships none: linop/transform.py is empty; UnitaryOperator API core/linop.py:574-615)."""

    def __init__(self, shape):
        self.img_shape = tuple(shape)
        n = int(np.prod(shape))
        self.shape = (n, n)

    def matvec(self, x):
        return np.fft.fft2(np.reshape(x, self.img_shape), norm='ortho').reshape(-1)

    def rmatvec(self, y):
        return np.fft.ifft2(np.reshape(y, self.img_shape), norm='ortho').reshape(-1)


class Comp:
    """LinOpComp, core/linop.py:540-553 (adjoint reverses the order)."""

    def __init__(self, A, B):
        self.A, self.B, self.shape = A, B, (A.shape[0], B.shape[1])

    def matvec(self, x):
        return self.A.matvec(self.B.matvec(x))

    def rmatvec(self, y):
        return self.B.rmatvec(self.A.rmatvec(y))


# --------------------------------------------------------------------------------------------------
# opt/proxalgs.py + core/solver.py
# --------------------------------------------------------------------------------------------------
def pds_step_sizes(beta, K_lipschitz, has_H=True):
    """opt/proxalgs.py:280-301 (set_step_sizes) and 312-316 (set_momentum_term)."""
    if beta > 0:
        if not has_H:
            tau, sigma = 2 / beta, 0
        else:
            L = K_lipschitz
            tau = sigma = (1 / L ** 2) * ((-beta / 4) + np.sqrt((beta ** 2 / 16) + L ** 2))
        rho = 0.9
    else:
        if not has_H:
            tau, sigma = 1, 0
        else:
            tau = sigma = 1 / K_lipschitz
        rho = 1
    return tau, sigma, rho


def _rel_improvement(old, new):
    """opt/proxalgs.py:372-377."""
    n = np.linalg.norm(old)
    return INF if n == 0 else np.linalg.norm(old - new) / n


def pds_update(x, z, F, G, H, K, tau, sigma, rho):
    """One Condat-Vu update, opt/proxalgs.py:343-355.  ``H is None`` <=> ``_H False`` (z untouched)."""
    Kadj_z = K.rmatvec(z) if H is not None else np.zeros(shape=x.size)  # NullOperator.adjoint linop/base.py:615
    x_temp = G.prox(x - tau * F.gradient(x) - tau * Kadj_z, tau=tau)
    if H is not None:
        u = 2 * x_temp - x
        z_temp = H.fenchel_prox(z + sigma * K.matvec(u), sigma=sigma)
        z = rho * z_temp + (1 - rho) * z
    x = rho * x_temp + (1 - rho) * x
    return x, z


def pds_iterate(dim, F=None, G=None, H=None, K=None, tau=None, sigma=None, rho=None, x0=None, z0=None,
                max_iter=500, min_iter=10, accuracy_threshold=1e-3, K_lipschitz=None, beta=None,
                callback=None):
    """PrimalDualSplitting(...).iterate(): opt/proxalgs.py:126-244, 343-394 + core/solver.py:55-76.

    Returns (x, z, diagnostics, n_updates).  ``K`` is any object with matvec/rmatvec (``None`` with an
    ``H`` means identity, proxalgs.py:203-205).
    """
    F = NullDiff(dim) if F is None else F
    G = NullProx(dim) if G is None else G
    beta = F.diff_lipschitz_cst if beta is None else beta
    if H is not None and K is None:
        K, K_lipschitz = _Identity(H.dim), 1
    if tau is not None and sigma is not None:
        pass
    elif tau is not None:
        sigma = tau
    elif sigma is not None:
        tau = sigma
    else:
        if H is not None and K_lipschitz is None:
            raise ValueError('Please compute the Lipschitz constant of the linear operator K')
        tau, sigma, _ = pds_step_sizes(beta, K_lipschitz if H is not None else 0, has_H=H is not None)
    if rho is None:
        rho = 0.9 if beta > 0 else 1
    x = np.zeros(shape=(dim,), dtype=float) if x0 is None else np.asarray(x0)          # proxalgs.py:327
    z = (None if H is None else np.zeros(shape=(H.dim,), dtype=float)) if z0 is None else np.asarray(z0)
    cols = ['Iter', 'Relative Improvement (primal variable)'] + \
        (['Relative Improvement (dual variable)'] if H is not None else [])
    diag = {c: [] for c in cols}
    it = 0

    def metric():  # proxalgs.py:360-364
        return INF if it == 0 else diag[cols[1]][it - 1]

    while ((it <= max_iter) and (metric() > accuracy_threshold)) or (it <= min_iter):  # solver.py:65-66
        x_old, z_old = x, z
        x, z = pds_update(x, z, F, G, H, K, tau, sigma, rho)
        diag['Iter'].append(it)
        diag[cols[1]].append(_rel_improvement(x_old, x))
        if H is not None:
            diag[cols[2]].append(_rel_improvement(z_old, z))
        if callback is not None:
            callback(it, x, z)
        it += 1
    return x, z, diag, it


class _Identity:
    def __init__(self, n):
        self.shape = (n, n)

    def matvec(self, x):
        return 1.0 * x

    def rmatvec(self, y):
        return 1.0 * y


def apgd_iterate(dim, F=None, G=None, tau=None, acceleration='CD', beta=None, x0=None, max_iter=500, min_iter=10,
                 accuracy_threshold=1e-3, d=75., callback=None):
    """AcceleratedProximalGradientDescent(...).iterate(): opt/proxalgs.py:485-622 + core/solver.py:55-76.

    Returns (x, past_aux, past_t, diagnostics, n_updates).
    """
    F = NullDiff(dim) if F is None else F
    G = NullProx(dim) if G is None else G
    beta = F.diff_lipschitz_cst if beta is None else beta
    tau = 1 / beta if tau is None else tau                               # proxalgs.py:573
    x = np.zeros(shape=(dim,), dtype=float) if x0 is None else np.asarray(x0)
    x_old, t_old = 0 * x, 1                                              # proxalgs.py:557
    diag = {'Iter': [], 'Relative Improvement': []}
    it = 0

    def metric():
        return INF if it == 0 else diag['Relative Improvement'][it - 1]

    while ((it <= max_iter) and (metric() > accuracy_threshold)) or (it <= min_iter):
        x_prev = x
        x_temp = G.prox(x - tau * F.gradient(x), tau=tau)                # proxalgs.py:591
        if acceleration == 'BT':
            t = (1 + np.sqrt(1 + 4 * t_old ** 2)) / 2
        elif acceleration == 'CD':
            t = (it + d) / d
        else:
            t = t_old = 1
        a = (t_old - 1) / t
        x = x_temp + a * (x_temp - x_old)                                # proxalgs.py:599
        x_old, t_old = x_temp, t
        diag['Iter'].append(it)
        diag['Relative Improvement'].append(_rel_improvement(x_prev, x))
        if callback is not None:
            callback(it, x, x_temp)
        it += 1
    return x, x_old, t_old, diag, it


# --------------------------------------------------------------------------------------------------
# operator factories with the reference's defaults (linop/diff.py, linop/conv.py, linop/sampling.py)
# --------------------------------------------------------------------------------------------------
def FirstDerivative(size, shape=None, axis=0, step=1.0, edge=True, dtype='float64', kind='forward'):
    """linop/diff.py:24-130 (defaults: kind='forward', edge=True)."""
    return pl.FirstDerivative(N=size, dims=shape, dir=axis, sampling=step, edge=edge, dtype=dtype, kind=kind)


def SecondDerivative(size, shape=None, axis=0, step=1.0, edge=True, dtype='float64'):
    """linop/diff.py:133-219."""
    return pl.SecondDerivative(N=size, dims=shape, dir=axis, sampling=step, edge=edge, dtype=dtype)


def Gradient(shape, step=1., edge=True, dtype='float64', kind='centered'):
    """linop/diff.py:777-882 (defaults: kind='centered', edge=True)."""
    return pl.Gradient(dims=shape, sampling=step, edge=edge, dtype=dtype, kind=kind)


def Laplacian(shape, weights=(1, 1), step=1., edge=True, dtype='float64'):
    """linop/diff.py:885-957."""
    if np.isscalar(step):
        step = [step] * len(shape)
    return pl.Laplacian(dims=shape, weights=weights, sampling=step, edge=edge, dtype=dtype)


def _offset(n):
    """linop/conv.py:159-162 / 285-292."""
    return n // 2 - 1 if n % 2 == 0 else n // 2


def Convolve1D(size, filter, reshape_dims=None, axis=0, dtype='float64', method=None):
    """linop/conv.py:20-164."""
    filter = np.asarray(filter)
    return pl.Convolve1D(N=size, h=filter, dims=reshape_dims, dir=axis, dtype=dtype, method=method,
                         offset=_offset(filter.size))


def Convolve2D(size, filter, shape, dtype='float64', method='fft'):
    """linop/conv.py:167-295."""
    filter = np.asarray(filter)
    return pl.Convolve2D(N=size, h=filter, dims=shape, nodir=None, dtype=dtype, method=method,
                         offset=(_offset(filter.shape[0]), _offset(filter.shape[1])))


def MovingAverage1D(window_size, shape, axis=0, dtype='float64'):
    """linop/conv.py:298-350 -> pylops.Smoothing1D(nsmooth, dims, dir): Convolve1D with ones(n)/n, n made odd, centred."""
    n = window_size + 1 if window_size % 2 == 0 else window_size
    shape = tuple(np.atleast_1d(shape))
    return pl.Convolve1D(N=int(np.prod(shape)), h=np.ones(n) / n, dims=shape if len(shape) > 1 else None, dir=axis,
                         dtype=dtype, method=None, offset=(n - 1) // 2)


def MovingAverage2D(window_shape, shape, dtype='float64'):
    """linop/conv.py:351-418 -> pylops.Smoothing2D(nsmooth, dims): Convolve2D with ones/(n0 n1), sizes made odd."""
    n0, n1 = (w + 1 if w % 2 == 0 else w for w in window_shape)
    return pl.Convolve2D(N=int(np.prod(shape)), h=np.ones((n0, n1)) / (n0 * n1), dims=shape, nodir=None, dtype=dtype,
                         method='fft', offset=((n0 - 1) // 2, (n1 - 1) // 2))


def SubSampling(size, sampling_indices, shape=None, axis=0, dtype='float64', inplace=True):
    """linop/sampling.py:25-122."""
    return pl.Restriction(M=size, iava=sampling_indices, dims=shape, dir=axis, dtype=dtype, inplace=inplace)


def DownSampling(size, downsampling_factor, shape=None, axis=None, dtype=np.float64):
    """linop/sampling.py:199-391."""
    return Masking(size, downsampling_mask(size, downsampling_factor, shape, axis), dtype=dtype)


def power_iteration_lipschitz(op, n_iter=200, seed=0):
    """Largest singular value by power iteration on op^T op (replaces ARPACK svds, core/linop.py:279-321)."""
    rng = np.random.default_rng(seed)
    v = rng.standard_normal(op.shape[1])
    v /= np.linalg.norm(v)
    s = 0.0
    for _ in range(n_iter):
        w = op.rmatvec(op.matvec(v))
        s = np.linalg.norm(w)
        if s == 0:
            return 0.0
        v = w / s
    return float(np.sqrt(s))


class LInftyNorm(LpNorm):
    """func/penalty.py:380-417."""
    _proj = staticmethod(proj_l1_ball)

    def __call__(self, x):
        return np.max(np.abs(x))


def L1Ball(dim, radius):
    """func/penalty.py:375-377."""
    return Indicator(dim, lambda x: np.sum(np.abs(x)) <= radius, lambda x: proj_l1_ball(x, radius=radius))


class SquaredL1Norm(Prox):
    """func/penalty.py:248-317 (both prox computations)."""

    def __init__(self, dim, prox_computation='sort'):
        self.dim, self.prox_computation = dim, prox_computation

    def __call__(self, x):
        return np.sum(np.abs(x)) ** 2

    def prox(self, x, tau):
        if self.prox_computation == 'root':                                  # :300-309
            from scipy import optimize as sciop
            if not np.linalg.norm(x) > 0:
                return x
            lam = lambda mu: np.clip(np.abs(x) * np.sqrt(tau / mu) - 2 * tau, a_min=0, a_max=None)
            mu_star = sciop.brentq(lambda mu: np.sum(lam(mu)) - 1, a=1e-12, b=np.max(np.abs(x) ** 2) / (4 * tau))
            return lam(mu_star) * x / (lam(mu_star) + 2 * tau)
        y = np.sort(np.abs(x))[::-1]                                         # :310-316
        csum = np.cumsum(y)
        factor = 2 * tau / (1 + (np.arange(y.size) + 1) * 2 * tau)
        k = np.max(np.nonzero(y - factor * csum > 0))
        return soft(x, factor[k] * csum[k])


def RealLine(dim):
    """func/penalty.py:715-717."""
    return Indicator(dim, lambda x: bool(np.all(np.isreal(x))), lambda x: np.real(x))


def ImagLine(dim):
    """func/penalty.py:765-767 (the prox is np.imag(x): a REAL array)."""
    return Indicator(dim, lambda x: bool(np.all(np.real(x) == 0)), lambda x: np.imag(x))


def ConsistencyLoss(dim, data):
    """func/loss.py:584-586."""
    return Indicator(dim, lambda x: bool(np.allclose(x, data)), lambda x: data)


class LogBarrier(Prox):
    """func/penalty.py:776-840."""

    def __init__(self, dim):
        self.dim = dim

    def __call__(self, x):
        x = np.asarray(x, dtype=float)
        y = 0 * x - np.inf
        y[x > 0] = np.log(x[x > 0])
        return -y.sum()

    def prox(self, x, tau):
        return (x + np.sqrt(x ** 2 + 4 * tau)) / 2


class ShannonEntropy(Prox):
    """func/penalty.py:843-922."""

    def __init__(self, dim):
        self.dim = dim

    def __call__(self, x):
        x = np.asarray(x, dtype=float)
        y = 0 * x + np.inf
        y[x == 0] = 0
        y[x > 0] = x[x > 0] * np.log(x[x > 0])
        return y.sum()

    def prox(self, x, tau):
        from scipy.special import lambertw
        return np.real(tau * lambertw(np.exp(-1 + (x / tau)) / tau, k=0))


class KLDivergence(Prox):
    """func/loss.py:597-682."""

    def __init__(self, dim, data):
        self.dim, self.data = dim, np.asarray(data, dtype=float)

    def __call__(self, x):
        x = np.asarray(x, dtype=float)
        z = 0 * x + np.inf
        m = (x > 0) * (self.data > 0)
        z[m] = self.data[m] * np.log(self.data[m] / x[m])
        z[(x == 0) * (self.data >= 0)] = 0
        return np.sum(z - self.data + x)

    def prox(self, x, tau):
        return (x - tau + np.sqrt((x - tau) ** 2 + 4 * tau * self.data)) / 2


class QuadraticForm:
    """func/penalty.py:925-990: x^T Q x, gradient 2 Q x (Q = identity when linop is None)."""

    def __init__(self, dim, linop=None):
        self.dim, self.linop = dim, linop
        self.diff_lipschitz_cst = 2 if linop is None else 2 * linop.diff_lipschitz_cst

    def __call__(self, x):
        return np.dot(x.conj(), x if self.linop is None else self.linop.matvec(x))

    def gradient(self, x):
        return 2 * x if self.linop is None else 2 * self.linop.matvec(x)


class FirstDirectionalDerivative:
    """linop/diff.py:380-486 -> pylops.FirstDirectionalDerivative(dims, v, sampling, edge, dtype, kind)
    = Sum(axis 0) * Diagonal(v) * Gradient: y = sum_i v_i * d_i x; adjoint = Gradient^T (v_i * y stacked)."""

    def __init__(self, shape, directions, step=1., edge=True, dtype='float64', kind='centered'):
        self.dims = tuple(shape)
        self.N = int(np.prod(shape))
        nd = len(self.dims)
        v = np.asarray(directions, dtype=float)
        self.v = np.repeat(v, self.N).reshape(nd, self.N) if v.ndim == 1 else v.reshape(nd, self.N)
        self.G = pl.Gradient(dims=shape, sampling=step, edge=edge, dtype=dtype, kind=kind)
        self.shape = (self.N, self.N)

    def matvec(self, x):
        return (self.v * self.G.matvec(x).reshape(len(self.dims), self.N)).sum(axis=0)

    def rmatvec(self, y):
        return self.G.rmatvec((self.v * y[None, :]).reshape(-1))


class SecondDirectionalDerivative:
    """linop/diff.py:489-606: KillEdge * pylops.SecondDirectionalDerivative, the latter being -D^T D with D the centered
    first directional derivative; KillEdge zeroes the two outermost samples of every axis."""

    def __init__(self, shape, directions, step=1., edge=True, dtype='float64'):
        self.D = FirstDirectionalDerivative(shape, directions, step, edge, dtype, 'centered')
        kill = np.ones(shape)
        for axis in range(len(shape)):
            kill = np.swapaxes(kill, axis, 0)
            kill[-2:] = 0
            kill[:2] = 0
            kill = np.swapaxes(kill, 0, axis)
        self.kill = kill.reshape(-1)
        self.shape = self.D.shape

    def matvec(self, x):
        return self.kill * (-self.D.rmatvec(self.D.matvec(x)))

    def rmatvec(self, z):
        return -self.D.rmatvec(self.D.matvec(self.kill * z))


class Integration1D:
    """linop/diff.py:1071-1137 -> pylops.CausalIntegration(N, dims, dir, sampling, halfcurrent=False):
    matvec = sampling * cumsum along the axis; rmatvec = sampling * reversed cumsum (flip, cumsum, flip)."""

    def __init__(self, size, shape=None, axis=0, step=1., dtype='float64'):
        self.dims = (size,) if shape is None else tuple(shape)
        self.axis, self.step, self.shape = axis, step, (size, size)

    def matvec(self, x):
        return (self.step * np.cumsum(x.reshape(self.dims), axis=self.axis)).reshape(-1)

    def rmatvec(self, y):
        Y = np.flip(y.reshape(self.dims), axis=self.axis)
        return (self.step * np.flip(np.cumsum(Y, axis=self.axis), axis=self.axis)).reshape(-1)


# --------------------------------------------------------------------------------------------------
# linop/base.py:636-712 (PolynomialLinearOperator); linop/diff.py:222-377, 960-1068 (generalised derivatives)
# --------------------------------------------------------------------------------------------------
class PolynomialLinearOperator:
    """P(L) = sum_k a_k L^k by repeated application (linop/base.py:695-712); the adjoint of a SYMMETRIC-flagged L is
    the forward map itself (:703-704), else the same recursion on L's adjoint with conjugated coefficients."""

    def __init__(self, L, coeffs, is_symmetric=False):
        self.L, self.coeffs, self.is_symmetric, self.shape = L, np.asarray(coeffs, dtype=float), is_symmetric, L.shape

    def _horner(self, x, apply):
        z, y = x.astype(float), self.coeffs[0] * x
        for c in self.coeffs[1:]:
            z = apply(z)
            y = y + c * z
        return y

    def matvec(self, x):
        return self._horner(x, self.L.matvec)

    def rmatvec(self, y):
        return self.matvec(y) if self.is_symmetric else self._horner(y, self.L.rmatvec)


class _AffineOfOp:
    """a * Id + b * L (core/linop.py:442-485 sums and scalings)."""

    def __init__(self, a, b, L):
        self.a, self.b, self.L, self.shape = a, b, L, L.shape

    def matvec(self, x):
        return self.a * x + self.b * self.L.matvec(x)

    def rmatvec(self, y):
        return self.a * y + self.b * self.L.rmatvec(y)


class _Power:
    """L ** N = L * (L * ...) (core/map.py:319-327)."""

    def __init__(self, L, N):
        self.L, self.N, self.shape = L, int(N), L.shape

    def matvec(self, x):
        for _ in range(self.N):
            x = self.L.matvec(x)
        return x

    def rmatvec(self, y):
        for _ in range(self.N):
            y = self.L.rmatvec(y)
        return y


class _KillEdges:
    """DiagonalOperator(kill_edges) * Dgen (linop/diff.py:374-376, 1065-1067)."""

    def __init__(self, mask, L):
        self.m, self.L, self.shape = mask.reshape(-1), L, L.shape

    def matvec(self, x):
        return self.m * self.L.matvec(x)

    def rmatvec(self, y):
        return self.L.rmatvec(self.m * y)


def GeneralisedDerivative(size, shape=None, axis=0, step=1.0, edge=True, dtype='float64', kind_op='iterated',
                          kind_diff='centered', **kwargs):
    """linop/diff.py:334-377."""
    D = FirstDerivative(size, shape, axis, step, edge, dtype, kind_diff)
    D2 = SecondDerivative(size, shape, axis, step, edge, dtype)
    if kind_op == 'iterated':
        Dgen, order = _Power(D, kwargs['order']), kwargs['order']
    elif kind_op == 'sobolev':
        Dgen, order = _Power(_AffineOfOp(kwargs['constant'] ** 2, -1.0, D2), kwargs['order']), 2 * kwargs['order']
    elif kind_op == 'exponential':
        Dgen, order = _Power(_AffineOfOp(kwargs['constant'], 1.0, D), kwargs['order']), kwargs['order']
    elif kind_op == 'polynomial':
        Dgen, order = PolynomialLinearOperator(D, kwargs['coeffs'], is_symmetric=False), len(kwargs['coeffs']) - 1
    else:
        raise NotImplementedError
    kill = np.ones(size if shape is None else shape)
    kill = np.swapaxes(kill, axis, 0) if axis > 0 else kill
    if kind_diff in ('forward', 'centered'):
        kill[-order:] = 0
    if kind_diff in ('backward', 'centered'):
        kill[:order] = 0
    kill = np.swapaxes(kill, 0, axis) if axis > 0 else kill
    return _KillEdges(np.ascontiguousarray(kill), Dgen)


def GeneralisedLaplacian(shape, step=1., edge=True, dtype='float64', kind='iterated', **kwargs):
    """linop/diff.py:1038-1068 (the Laplacian is FLAGGED symmetric, :1039, whatever its edge rows)."""
    Delta = Laplacian(shape, step=step, edge=edge, dtype=dtype)
    if kind == 'iterated':
        Dgen, order = _Power(Delta, kwargs['order']), 2 * kwargs['order']
    elif kind == 'sobolev':
        Dgen, order = _Power(_AffineOfOp(kwargs['constant'] ** 2, -1.0, Delta), kwargs['order']), 2 * kwargs['order']
    elif kind == 'polynomial':
        Dgen, order = PolynomialLinearOperator(Delta, kwargs['coeffs'], is_symmetric=True), 2 * (len(kwargs['coeffs']) - 1)
    else:
        raise NotImplementedError
    kill = np.ones(shape)
    for ax in range(len(shape)):
        idx = [slice(None)] * len(shape)
        idx[ax] = slice(0, order)
        kill[tuple(idx)] = 0
        idx[ax] = slice(shape[ax] - order, None)
        kill[tuple(idx)] = 0
    return _KillEdges(kill, Dgen)


class Pooling:
    """linop/sampling.py:394-536 -> skimage.measure.block_reduce(image, block_size, func, cval=0): the array is padded
    with zeros up to a multiple of the block size, every block is summed or averaged over its FULL size; the adjoint
    repeats each pooled value over its block and crops (:529-536; not rescaled for 'mean').  skimage is not in this
    image, so this restatement is pinned by the reference's doctest literals (:405-424) only."""

    def __init__(self, shape, block_size, pooling_func='mean'):
        if len(shape) != len(block_size):
            raise ValueError(f'Inconsistent block size {block_size} for array of shape {shape}.')
        if pooling_func not in ['mean', 'sum', 'average']:
            raise ValueError('pooling_func must be one of: "mean" or "sum".')
        self.input_shape, self.block_size, self.func = tuple(shape), tuple(block_size), pooling_func
        self.output_shape = tuple(-(-d // b) for d, b in zip(shape, block_size))
        self.shape = (int(np.prod(self.output_shape)), int(np.prod(shape)))

    def matvec(self, x):
        pad = [(0, o * b - d) for d, b, o in zip(self.input_shape, self.block_size, self.output_shape)]
        X = np.pad(np.asarray(x).reshape(self.input_shape), pad, mode='constant', constant_values=0)
        split = [v for o, b in zip(self.output_shape, self.block_size) for v in (o, b)]
        blocks = X.reshape(split)
        axes = tuple(range(1, 2 * len(self.block_size), 2))
        return (np.mean(blocks, axis=axes) if self.func == 'mean' else np.sum(blocks, axis=axes)).reshape(-1)

    def rmatvec(self, y):
        X = np.copy(np.asarray(y).reshape(self.output_shape))
        for ax in range(len(self.input_shape)):
            X = np.repeat(X, self.block_size[ax], axis=ax)
            idx = [slice(None)] * X.ndim
            idx[ax] = slice(0, self.input_shape[ax])
            X = X[tuple(idx)]
        return X.reshape(-1)
