"""The golden cases of tests/golden/gen_golden.py rebuilt on the numpy oracle (same inputs, same parameters)."""
import numpy as np

from oracle import np_oracle as O
from pycsou_b200.util import synth

SQRT8 = np.sqrt(8.)


def c1(acc='CD', k=500, dtype=np.float64):
    truth, h, y = synth.spikes_1d(4096, 32, 0)
    N = y.size
    Gop = O.Convolve1D(size=N, filter=h)
    lam = 0.1 * np.max(np.abs(Gop.rmatvec(y)))
    F = O.HalfSquaredL2Loss(y, Gop, G_lipschitz=1.0)
    return O.apgd_iterate(N, F=F, G=lam * O.L1Norm(N), tau=1.0, acceleration=acc, beta=1.0, max_iter=k - 1,
                          min_iter=k - 1, accuracy_threshold=0, d=75.)


def c2(k, G=None, kind='forward', H='l21', shape=(48, 64), lam=0.1, seed=0, stop_rule=False, groups=None):
    _, y = synth.tv_denoise_problem(shape, 0.1, seed)
    N = y.size
    K = O.Gradient(shape, kind=kind)
    F = O.HalfSquaredL2Loss(y)
    if H == 'l21':
        Hf = lam * (O.TVL21Norm(2 * N, N) if groups is None else O.L21Norm(2 * N, groups))
    else:
        Hf = lam * O.L1Norm(2 * N)
    kw = {} if stop_rule else dict(max_iter=k - 1, min_iter=k - 1, accuracy_threshold=0)
    return O.pds_iterate(N, F=F, G=G(N) if G is not None else None, H=Hf, K=K, K_lipschitz=SQRT8, **kw)


def c3(k, tag='sep', shape=(48, 64), lam=0.02):
    psf = synth.gaussian_psf_2d(15, 2.5) if tag == 'sep' else synth.rotated_gaussian_psf_2d(15)
    _, h, y = synth.tv_deconv_problem(shape, psf, 0.02, 0)
    N = y.size
    A = O.Convolve2D(size=N, filter=h, shape=shape)
    F = O.HalfSquaredL2Loss(y, A, G_lipschitz=1.0)
    K = O.Gradient(shape, kind='forward')
    return O.pds_iterate(N, F=F, G=O.NonNegativeOrthant(N), H=lam * O.TVL21Norm(2 * N, N), K=K, K_lipschitz=SQRT8,
                         beta=1.0, max_iter=k - 1, min_iter=k - 1, accuracy_threshold=0)


def c3_objective(x, tag='sep', shape=(48, 64), lam=0.02):
    psf = synth.gaussian_psf_2d(15, 2.5) if tag == 'sep' else synth.rotated_gaussian_psf_2d(15)
    _, h, y = synth.tv_deconv_problem(shape, psf, 0.02, 0)
    N = y.size
    A = O.Convolve2D(size=N, filter=h, shape=shape)
    K = O.Gradient(shape, kind='forward')
    return 0.5 * np.linalg.norm(A.matvec(x) - y) ** 2 + lam * O.TVL21Norm(2 * N, N)(K.matvec(x))


def examples_pbp():
    x = np.repeat([0, 2, 1, 3, 0, 2, 0], 10).astype(float)
    D = O.FirstDerivative(size=x.size, kind='forward')
    Gop = O.DownSampling(size=x.size, downsampling_factor=3)
    y = Gop.matvec(x)
    F = O.HalfSquaredL2Loss(y, Gop, G_lipschitz=1.0)
    return O.pds_iterate(x.size, F=F, G=0.01 * O.L1Norm(x.size), H=0.1 * O.L1Norm(x.size), K=D, K_lipschitz=2.0)


def misc(which, k, shape=(24, 32)):
    N = shape[0] * shape[1]
    _, y = synth.tv_denoise_problem(shape, 0.1, 1)
    kw = dict(max_iter=k - 1, min_iter=k - 1, accuracy_threshold=0)
    if which == 'cps':
        K = O.Gradient(shape, kind='forward')
        return O.pds_iterate(N, G=O.PreCompShift(O.L1Norm(N), y), H=0.3 * O.TVL21Norm(2 * N, N), K=K,
                             K_lipschitz=SQRT8, rho=1, **kw)
    if which == 'drs':
        return O.pds_iterate(N, G=O.PreCompShift(O.L1Norm(N), y), H=O.Segment(N, 0.2, 0.6), K=None, tau=0.7,
                             sigma=1 / 0.7, rho=1, **kw)
    if which == 'fbs':
        return O.pds_iterate(N, F=O.HalfSquaredL2Loss(y), G=0.2 * O.L1Norm(N), rho=1, **kw)
    raise ValueError(which)


def c4(k, mask, y, shape=(32, 48), lam=0.05, eps=None):
    N = shape[0] * shape[1]
    M = O.Masking(N, mask, dtype=np.complex128)
    m = M.shape[0]
    K = O.LinOpVStackOracle(O.Comp(M, O.FFT2Ortho(shape)), O.Gradient(shape, kind='forward', dtype='complex128'))
    H = O.ProxHStack(O.PreCompShift(O.L2Ball(m, eps), y), lam * O.TVL21Norm(2 * N, N))
    return O.pds_iterate(N, G=None, H=H, K=K, tau=1 / 3, sigma=1 / 3, rho=1, x0=np.zeros(N, dtype=np.complex128),
                         z0=np.zeros(m + 2 * N, dtype=np.complex128), max_iter=k - 1, min_iter=k - 1,
                         accuracy_threshold=0)
