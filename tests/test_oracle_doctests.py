"""The reference's own doctest literals for the hot path, transcribed and checked against the oracle.

Each test cites the reference docstring lines it pins.
"""
import numpy as np
import scipy.signal as sps

from oracle import np_oracle as O


def hann(n):  # scipy.signal.hann was removed; the doctests use the symmetric window
    return sps.windows.hann(n)


def test_sign_soft_literals():
    # math/prox.py:51-58, 102-107
    x = np.array([-3., -1., 0., 2.])
    assert np.array_equal(O.sign(x), np.array([-1., -1., 0., 1.]))
    assert np.allclose(O.soft(np.arange(-3, 4).astype(float), 1), [-2, -1, 0, 0, 0, 1, 2])
    assert np.isclose(O.soft(np.array(3 + 1j), 0.1), 2.905131670194949 - 0.9683772233983162j)


def test_projection_literals():
    # math/prox.py:193-197, 239-243, 277-279, 322-324
    x = np.arange(10).astype(float)
    assert np.allclose(O.proj_l2_ball(x, 2), 2 * x / np.linalg.norm(x))
    assert np.array_equal(O.proj_linfty_ball(np.arange(-3, 4).astype(float), 2), [-2, -2, -1, 0, 1, 2, 2])
    assert np.array_equal(O.proj_nonnegative_orthant(np.array([-1., 1.])), [0, 1])
    assert np.allclose(O.proj_segment(np.linspace(-3, 3, 5), a=-2, b=1), [-2., -1.5, 0., 1., 1.])
    assert np.array_equal(O.proj_segment(np.arange(10.) - 3, 1, 4), [1, 1, 1, 1, 1, 2, 3, 4, 4, 4])


def test_penalty_literals():
    # func/penalty.py:37-46, 209-218, 494-509, 597-604, 653-660
    x = np.arange(10).astype(float)
    assert np.isclose(O.L2Norm(10)(x), 16.881943016134134)
    assert np.isclose((3 * O.L2Norm(10))(x), 50.645829048402405)
    assert np.isclose(O.L1Norm(10)(x), 45) and np.isclose((3 * O.L1Norm(10))(x), 135)
    tau = 1.2
    assert np.allclose(O.L1Norm(10).prox(x, tau), O.soft(x, tau))
    assert np.allclose(O.L2Norm(10).prox(x, tau), np.clip(1 - tau / np.linalg.norm(x), 0, None) * x)
    groups = np.concatenate((np.ones(5), 2 * np.ones(5)))
    l21 = O.L21Norm(10, groups)
    assert np.isclose(l21(x), 21.44594499772297)
    ref = np.concatenate((O.L2Norm(5).prox(x[:5], tau), O.L2Norm(5).prox(x[5:], tau)))
    assert np.allclose(l21.prox(x, tau), ref)
    nn = O.NonNegativeOrthant(10)
    assert nn(np.arange(-5, 5)) == np.inf and nn(np.arange(10)) == 0
    assert np.array_equal(nn.prox(np.arange(-5., 5.), 1), np.clip(np.arange(-5., 5.), 0, None))
    seg = O.Segment(10, a=1, b=4)
    x1 = np.arange(10.) - 3
    assert seg(x1) == np.inf and seg(np.clip(x1, 1, 4)) == 0
    assert np.array_equal(seg.prox(x1, 1), [1, 1, 1, 1, 1, 2, 3, 4, 4, 4])


def test_fenchel_prox_moreau():
    # core/functional.py:124-132 (prox calculus), :207 (Moreau)
    rng = np.random.default_rng(0)
    z = rng.standard_normal(20)
    f = 0.3 * O.L1Norm(20)
    # conjugate of lam*L1 is the indicator of the lam-ball: fenchel prox = clip
    assert np.allclose(f.fenchel_prox(z, 0.7), np.clip(z, -0.3, 0.3))


def test_first_derivative_pins():
    # linop/diff.py:72-78 and core/map.py:418-423
    x = np.repeat([0., 2., 1., 3., 0., 2., 0.], 10)
    D = O.FirstDerivative(size=x.size)
    y = D.matvec(x)
    assert np.allclose(y, np.diff(x, append=0)) and np.count_nonzero(y) == 6
    assert abs(O.power_iteration_lipschitz(O.FirstDerivative(10, kind='forward'), 4000) - 1.9) < 0.1
    assert abs(O.power_iteration_lipschitz(O.FirstDerivative(10, kind='centered'), 4000) - 1.5) < 0.1


def test_gradient_is_stack_of_first_derivatives():
    # linop/diff.py:814-820, linop/base.py:213-233
    shape = (7, 9)
    rng = np.random.default_rng(1)
    x, z = rng.standard_normal(63), rng.standard_normal(126)
    for kind in ('forward', 'centered', 'backward'):
        G = O.Gradient(shape, kind=kind)
        D0 = O.FirstDerivative(63, shape=shape, axis=0, kind=kind)
        D1 = O.FirstDerivative(63, shape=shape, axis=1, kind=kind)
        assert np.allclose(G.matvec(x), np.concatenate((D0.matvec(x), D1.matvec(x))))
        assert np.allclose(G.rmatvec(z), D0.rmatvec(z[:63]) + D1.rmatvec(z[63:]))
        assert np.isclose(np.dot(G.matvec(x), z), np.dot(x, G.rmatvec(z)))


def test_laplacian_adjoint_dot():
    rng = np.random.default_rng(2)
    L = O.Laplacian((8, 6), weights=(1.5, 0.5), step=(1., 2.))
    x, y = rng.standard_normal(48), rng.standard_normal(48)
    assert np.isclose(np.dot(L.matvec(x), y), np.dot(x, L.rmatvec(y)))


def test_convolve1d_pin():
    # linop/conv.py:67-73
    sig = np.repeat([0., 1., 0.], 10)
    filt = hann(5)
    filt[filt.size // 2:] = 0
    C = O.Convolve1D(size=sig.size, filter=filt)
    assert np.allclose(C.matvec(sig), sps.convolve(sig, filt, mode='same', method='direct'))
    rng = np.random.default_rng(3)
    x, y = rng.standard_normal(30), rng.standard_normal(30)
    assert np.isclose(np.dot(C.matvec(x), y), np.dot(x, C.rmatvec(y)))
    # even-length filter: centred at nh//2-1 (conv.py:159-160) and still an exact adjoint pair
    Ce = O.Convolve1D(size=30, filter=np.array([1., 2., 3., 4.]))
    e = np.zeros(30)
    e[10] = 1
    assert np.allclose(Ce.matvec(e)[9:13], [1, 2, 3, 4])
    assert np.isclose(np.dot(Ce.matvec(x), y), np.dot(x, Ce.rmatvec(y)))


def test_convolve2d_pin():
    # linop/conv.py:209-217
    sig = np.zeros((100, 100))
    sig[48:53, 48:53] = 1
    f = hann(25)
    f[f.size // 2:] = 0
    filt = f[None, :] * f[:, None]
    C = O.Convolve2D(size=sig.size, filter=filt, shape=sig.shape)
    assert np.allclose(C.matvec(sig.ravel()).reshape(sig.shape), sps.convolve(sig, filt, mode='same', method='fft'))
    rng = np.random.default_rng(4)
    x, y = rng.standard_normal(10000), rng.standard_normal(10000)
    assert np.isclose(np.dot(C.matvec(x), y), np.dot(x, C.rmatvec(y)))


def test_sampling_literals():
    # linop/sampling.py:72-89 (SubSampling), 141-154 (Masking), 223-253 (DownSampling)
    x = np.arange(9).reshape(3, 3)
    S = O.SubSampling(size=9, sampling_indices=[0, 2])
    assert np.array_equal(S.matvec(x.reshape(-1)), [0, 2])
    assert np.array_equal(S.rmatvec(S.matvec(x.reshape(-1))).reshape(3, 3), [[0, 0, 2], [0, 0, 0], [0, 0, 0]])
    S = O.SubSampling(size=9, sampling_indices=[0, 2], shape=(3, 3), axis=1)
    assert np.array_equal(S.matvec(x.reshape(-1)).reshape(3, 2), [[0, 2], [3, 5], [6, 8]])
    assert np.array_equal(S.rmatvec(S.matvec(x.reshape(-1))).reshape(3, 3), [[0, 0, 2], [3, 0, 5], [6, 0, 8]])
    sb = np.zeros((3, 3), dtype=bool)
    sb[[1, 2], [0, 2]] = True
    M = O.Masking(size=9, sampling_bool=sb)
    assert np.array_equal(M.matvec(x.reshape(-1)), [3, 8])
    assert np.array_equal(M.rmatvec(M.matvec(x.reshape(-1))).reshape(3, 3), [[0, 0, 0], [3, 0, 0], [0, 0, 8]])
    x = np.arange(10)
    D = O.DownSampling(size=10, downsampling_factor=3)
    assert np.array_equal(D.matvec(x), [0, 3, 6, 9])
    assert np.array_equal(D.rmatvec(D.matvec(x)), [0, 0, 0, 3, 0, 0, 6, 0, 0, 9])
    x = np.arange(20).reshape(4, 5)
    D = O.DownSampling(size=20, downsampling_factor=3, shape=(4, 5), axis=-1)
    assert np.array_equal(D.matvec(x.ravel()).reshape(4, 2), [[0, 3], [5, 8], [10, 13], [15, 18]])
    D = O.DownSampling(size=20, downsampling_factor=(2, 3), shape=(4, 5))
    assert np.array_equal(D.matvec(x.ravel()).reshape(2, 2), [[0, 3], [10, 13]])


def test_squared_l2_loss_gradient():
    # func/loss.py:193-207: gradient 2 G^T(Gx - y) for SquaredL2Loss*G; half of it for (1/2)*...
    rng = np.random.default_rng(5)
    Gop = O.FirstDerivative(12, kind='centered')
    y, x = rng.standard_normal(12), rng.standard_normal(12)
    F = O.HalfSquaredL2Loss(y, Gop, G_lipschitz=1.5)
    assert np.allclose(F.gradient(x), Gop.rmatvec(Gop.matvec(x) - y))
    assert np.isclose(F.diff_lipschitz_cst, 1.5 ** 2)


def test_entropy_and_barrier_functionals():
    # func/penalty.py:793-798 (LogBarrier), :867-872 (ShannonEntropy); func/loss.py:632-638 (KLDivergence)
    x1 = np.arange(10)
    f = O.LogBarrier(10)
    assert f(x1) == np.inf and f(x1 + 2) == -17.502307845873887
    assert np.array_equal(np.round(f.prox(x1, tau=1)), [1, 2, 2, 3, 4, 5, 6, 7, 8, 9])
    g = O.ShannonEntropy(10)
    x2 = np.zeros(10)
    x2[0] = 10
    assert g(x1) == 79.05697962199447 and g(x2) == 23.02585092994046
    assert np.array_equal(np.round(g.prox(x1, tau=2)), [0, 0, 1, 1, 1, 2, 2, 3, 3, 4])
    k = O.KLDivergence(10, np.arange(10))
    assert k(2 * np.arange(10)) == 13.80837687480246
    assert np.array_equal(np.round(k.prox(2 * np.arange(10), tau=1)), [0, 2, 4, 6, 8, 10, 12, 14, 16, 18])


def test_l1_ball_linfty_and_squared_l1():
    # math/prox.py:143-147 (proj_l1_ball); func/penalty.py:262-269 (SquaredL1Norm), :355-362 (L1Ball), :393-398
    # (LInftyNorm), :704-709 (RealLine), :755-758 (ImagLine); func/loss.py:571-577 (ConsistencyLoss)
    x = np.linspace(-1, 1, 5)
    assert np.allclose(O.proj_l1_ball(x, radius=2), [-0.75, -0.25, 0., 0.25, 0.75], atol=1e-11, rtol=0)
    assert np.isclose(np.linalg.norm(O.proj_l1_ball(x, radius=2), ord=1), 2.0, atol=1e-11, rtol=0)
    x1 = np.arange(10)
    assert O.SquaredL1Norm(10, 'sort')(x1) == 2025
    assert np.allclose(O.SquaredL1Norm(10, 'sort').prox(x1, tau=1), O.SquaredL1Norm(10, 'root').prox(x1, tau=1))
    x2 = x1 / np.linalg.norm(x1, ord=1)
    ball = O.L1Ball(10, radius=10)
    assert (ball(x1), ball(x2)) == (np.inf, 0)
    assert np.isclose(np.linalg.norm(ball.prox(x1, tau=1), ord=1), 10.0, atol=1e-10, rtol=0)
    assert np.allclose(ball.prox(x2, tau=1), x2)
    assert O.LInftyNorm(10)(x1) == 9
    xc = np.arange(10) + 1j
    assert (O.RealLine(10)(xc), O.RealLine(10)(np.real(xc))) == (np.inf, 0)
    assert np.allclose(O.RealLine(10).prox(xc, tau=1), np.real(xc))
    assert (O.ImagLine(10)(xc), O.ImagLine(10)(1j * np.imag(xc))) == (np.inf, 0)
    assert np.allclose(O.ImagLine(10).prox(xc, tau=1), np.imag(xc))
    loss = O.ConsistencyLoss(10, x1)
    assert (loss(2 * x1), loss(x1)) == (np.inf, 0) and np.allclose(loss.prox(2 * x1, tau=1), x1)


def test_pooling_doctest():
    # linop/sampling.py:405-424
    x = np.arange(24).reshape(4, 6)
    P = O.Pooling(x.shape, (2, 3), 'mean')
    assert P.output_shape == (2, 2) and np.array_equal(P.matvec(x.flatten()).reshape(2, 2), [[4., 7.], [16., 19.]])
    assert np.array_equal(P.rmatvec(P.matvec(x.flatten())).reshape(4, 6),
                          np.repeat(np.repeat([[4., 7.], [16., 19.]], 2, axis=0), 3, axis=1))
    P = O.Pooling(x.shape, (2, 3), 'sum')
    assert np.array_equal(P.matvec(x.flatten()).reshape(2, 2), [[24, 42], [96, 114]])
    assert np.array_equal(P.rmatvec(P.matvec(x.flatten())).reshape(4, 6),
                          np.repeat(np.repeat([[24, 42], [96, 114]], 2, axis=0), 3, axis=1))
    P = O.Pooling((5, 7), (2, 3), 'mean')          # ragged: zero padding counts in the mean
    X = np.arange(35.).reshape(5, 7)
    assert P.output_shape == (3, 3) and P.matvec(X.flatten())[-1] == X[4, 6] / 6
    assert P.rmatvec(np.arange(9.)).shape == (35,)
