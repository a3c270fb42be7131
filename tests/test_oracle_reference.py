"""Oracle vs the live, unmodified reference (build container only: needs /root/reference)."""
import warnings

import numpy as np
import pytest

from oracle import np_oracle as O

pytestmark = pytest.mark.needs_reference
warnings.filterwarnings('ignore')


@pytest.fixture(scope='module')
def ref():
    from oracle.ref_shim import load_reference
    return load_reference()


def test_functionals_match_reference(ref):
    from pycsou.func import penalty as P, loss as L
    rng = np.random.default_rng(0)
    x = rng.standard_normal(40)
    for tau in (0.05, 0.7, 3.0):
        assert np.array_equal(P.L1Norm(40).prox(x.copy(), tau), O.L1Norm(40).prox(x, tau))
        assert np.array_equal(P.L2Norm(40).prox(x.copy(), tau), O.L2Norm(40).prox(x, tau))
        assert np.array_equal((0.3 * P.L1Norm(40)).fenchel_prox(x.copy(), tau), (0.3 * O.L1Norm(40)).fenchel_prox(x, tau))
        g = np.tile(np.arange(20), 2)
        assert np.allclose(P.L21Norm(40, g).prox(x.copy(), tau), O.TVL21Norm(40, 20).prox(x, tau), rtol=1e-15, atol=0)
        assert np.allclose(P.L21Norm(40, g).prox(x.copy(), tau), O.L21Norm(40, g).prox(x, tau), rtol=1e-15, atol=0)
        y = rng.standard_normal(40)
        assert np.array_equal(L.L1Loss(40, y).prox(x.copy(), tau), O.PreCompShift(O.L1Norm(40), y).prox(x, tau))
        assert np.array_equal(L.L2BallLoss(40, y, radius=2.).prox(x.copy(), tau),
                              O.PreCompShift(O.L2Ball(40, 2.), y).prox(x, tau))
    assert np.array_equal(P.NonNegativeOrthant(40).prox(x.copy(), 1), O.NonNegativeOrthant(40).prox(x, 1))
    assert np.array_equal(P.Segment(40, -0.5, 0.2).prox(x.copy(), 1), O.Segment(40, -0.5, 0.2).prox(x, 1))


def test_loss_gradient_chain_matches_reference(ref):
    from pycsou.func.loss import SquaredL2Loss
    from pycsou.linop.conv import Convolve2D
    rng = np.random.default_rng(1)
    shape = (20, 17)
    h = rng.random((5, 7))
    y, x = rng.standard_normal(340), rng.standard_normal(340)
    A = Convolve2D(size=340, filter=h, shape=shape)
    A.lipschitz_cst = A.diff_lipschitz_cst = 3.0
    F = (1 / 2) * SquaredL2Loss(dim=340, data=y) * A
    Fo = O.HalfSquaredL2Loss(y, O.Convolve2D(340, h, shape), G_lipschitz=3.0)
    assert np.array_equal(F.gradient(x), Fo.gradient(x))
    assert F.diff_lipschitz_cst == Fo.diff_lipschitz_cst == 9.0


def test_step_size_rule_matches_reference(ref):
    from pycsou.func.loss import SquaredL2Loss
    from pycsou.func.penalty import L1Norm
    from pycsou.linop.diff import FirstDerivative
    from pycsou.opt.proxalgs import PDS
    D = FirstDerivative(size=30)
    D.lipschitz_cst = D.diff_lipschitz_cst = 1.7
    for beta_scale in (1.0, 0.25):
        F = beta_scale * SquaredL2Loss(dim=30, data=np.ones(30))
        s = PDS(dim=30, F=F, H=L1Norm(30), K=D, verbose=None)
        tau, sigma, rho = O.pds_step_sizes(2 * beta_scale, 1.7)
        assert (s.tau, s.sigma, s.rho) == (tau, sigma, rho)
    s = PDS(dim=30, F=None, H=L1Norm(30), K=D, verbose=None)
    assert (s.tau, s.sigma, s.rho) == O.pds_step_sizes(0, 1.7)
    s = PDS(dim=30, F=SquaredL2Loss(dim=30, data=np.ones(30)), verbose=None)
    assert (s.tau, s.sigma, s.rho) == O.pds_step_sizes(2, 0, has_H=False)


def test_entropy_and_barrier_match_reference(ref):
    from pycsou.func import penalty as P, loss as L
    rng = np.random.default_rng(2)
    x, d = 4 * rng.standard_normal(50), 3 * rng.random(50)
    xp = np.abs(x) + 0.1
    for tau in (0.05, 0.7, 3.0):
        assert np.array_equal(P.LogBarrier(50).prox(x.copy(), tau), O.LogBarrier(50).prox(x, tau))
        assert np.array_equal(P.ShannonEntropy(50).prox(x.copy(), tau), O.ShannonEntropy(50).prox(x, tau))
        assert np.array_equal(L.KLDivergence(50, d).prox(x.copy(), tau), O.KLDivergence(50, d).prox(x, tau))
    assert P.LogBarrier(50)(xp) == O.LogBarrier(50)(xp) and P.LogBarrier(50)(x) == O.LogBarrier(50)(x) == np.inf
    assert P.ShannonEntropy(50)(xp) == O.ShannonEntropy(50)(xp)
    assert L.KLDivergence(50, d)(xp) == O.KLDivergence(50, d)(xp)


def test_l1_ball_linfty_and_squared_l1_match_reference(ref):
    from pycsou.func import penalty as P, loss as L
    from pycsou.math import prox as M
    rng = np.random.default_rng(8)
    x = 3 * rng.standard_normal(200)
    for radius in (0.5, 7.0, 1e4):
        assert np.array_equal(M.proj_l1_ball(x.copy(), radius), O.proj_l1_ball(x, radius))
        assert np.array_equal(P.L1Ball(200, radius).prox(x.copy(), 1.0), O.L1Ball(200, radius).prox(x, 1.0))
        assert P.L1Ball(200, radius)(x) == O.L1Ball(200, radius)(x)
    for tau in (0.01, 0.4, 5.0):
        assert np.array_equal(P.LInftyNorm(200).prox(x.copy(), tau), O.LInftyNorm(200).prox(x, tau))
        for how in ('sort', 'root'):
            assert np.array_equal(P.SquaredL1Norm(200, how).prox(x.copy(), tau), O.SquaredL1Norm(200, how).prox(x, tau))
    assert P.LInftyNorm(200)(x) == O.LInftyNorm(200)(x) and P.SquaredL1Norm(200)(x) == O.SquaredL1Norm(200)(x)
    xc = x[:100] + 1j * x[100:]
    assert np.array_equal(P.RealLine(100).prox(xc, 1.0), O.RealLine(100).prox(xc, 1.0))
    assert np.array_equal(P.ImagLine(100).prox(xc, 1.0), O.ImagLine(100).prox(xc, 1.0))
    assert P.RealLine(100)(xc) == O.RealLine(100)(xc) and P.ImagLine(100)(1j * x[:100]) == O.ImagLine(100)(1j * x[:100])
    assert L.ConsistencyLoss(200, x)(x + 1) == O.ConsistencyLoss(200, x)(x + 1) == np.inf


def test_generalised_derivatives_match_reference(ref):
    """GeneralisedDerivative / GeneralisedLaplacian / PolynomialLinearOperator run live on the pylops stand-in."""
    from pycsou.linop.diff import GeneralisedDerivative as GD, GeneralisedLaplacian as GL
    rng = np.random.default_rng(9)
    x, y = rng.standard_normal(60), rng.standard_normal(60)
    cases = [dict(kind_op='iterated', order=3), dict(kind_op='sobolev', order=2, constant=0.3),
             dict(kind_op='exponential', order=3, constant=-0.2),
             dict(kind_op='polynomial', coeffs=0.5 * np.array([1e-2, 0, -0.2, 0, 1]))]
    for kw in cases:
        for kind_diff in ('forward', 'backward', 'centered'):
            A, Ao = GD(size=60, kind_diff=kind_diff, **kw), O.GeneralisedDerivative(60, kind_diff=kind_diff, **kw)
            assert np.allclose(A(x), Ao.matvec(x), rtol=1e-13, atol=1e-13)
            assert np.allclose(A.adjoint(y), Ao.rmatvec(y), rtol=1e-13, atol=1e-13)
        A = GD(size=60, shape=(6, 10), axis=1, step=0.5, **kw)
        Ao = O.GeneralisedDerivative(60, shape=(6, 10), axis=1, step=0.5, **kw)
        assert np.allclose(A(x), Ao.matvec(x), rtol=1e-13, atol=1e-13)
        assert np.allclose(A.adjoint(y), Ao.rmatvec(y), rtol=1e-13, atol=1e-13)
    x, y = rng.standard_normal(20 * 22), rng.standard_normal(20 * 22)
    for kw in (dict(kind='iterated', order=2), dict(kind='sobolev', order=2, constant=0.4),
               dict(kind='polynomial', coeffs=[0.5, -1.0, 0.25])):
        A, Ao = GL(shape=(20, 22), **kw), O.GeneralisedLaplacian((20, 22), **kw)
        assert np.allclose(A(x), Ao.matvec(x), rtol=1e-13, atol=1e-13)
        assert np.allclose(A.adjoint(y), Ao.rmatvec(y), rtol=1e-13, atol=1e-13)
