"""CPU tests of the host-side mirror of the reference API: construction, shapes, arithmetic dispatch, Lipschitz
propagation, step-size rules and constructor errors (no kernel is launched)."""
import numpy as np
import pytest

from oracle import np_oracle as O
from pycsou_b200.core import DifferentiableMap, LinearOperator, ProximableFunctional
from pycsou_b200.core.functional import ProxFuncPostComp, ProxFuncPreComp
from pycsou_b200.core.linop import AdjointLinearOperator, LinOpComp, LinOpSum
from pycsou_b200.core.map import DiffMapComp, DiffMapShifted, DiffMapSum, MapSum
from pycsou_b200.func.base import IndicatorFunctional, NullProximableFunctional
from pycsou_b200.func.loss import L1Loss, SquaredL2Loss
from pycsou_b200.func.penalty import L1Norm, L2Norm, L21Norm, NonNegativeOrthant, Segment, SquaredL2Norm
from pycsou_b200.linop.base import DiagonalOperator, HomothetyMap, IdentityOperator, LinOpVStack, NullOperator
from pycsou_b200.linop.conv import Convolve1D, Convolve2D
from pycsou_b200.linop.diff import FirstDerivative, Gradient, Laplacian
from pycsou_b200.linop.sampling import DownSampling, Masking, SubSampling
from pycsou_b200.opt.proxalgs import APGD, CPS, DRS, FBS, PDS, PrimalDualSplitting
from pycsou_b200.util import synth


def test_shapes_and_defaults():
    G = Gradient((6, 7))
    assert G.shape == (84, 42) and G.kind == 'centered' and G.edge is True          # diff.py:777-778 defaults
    D = FirstDerivative(10)
    assert D.shape == (10, 10) and D.kind == 'forward'                             # diff.py:24-25 defaults
    assert Laplacian((5, 6)).shape == (30, 30)
    assert Convolve1D(20, np.ones(4)).offset == 1 and Convolve1D(20, np.ones(5)).offset == 2    # conv.py:159-162
    C = Convolve2D(30, np.ones((4, 3)), (5, 6))
    assert C.offset == (1, 1) and C.is_separable
    assert not Convolve2D(30, synth.rotated_gaussian_psf_2d(5), (5, 6)).is_separable
    M = Masking(6, [1, 0, 0, 1, 1, 0])
    assert M.shape == (3, 6) and M.nb_of_samples == 3 and M.input_size == 6
    Ds = DownSampling(20, (2, 3), shape=(4, 5))
    assert Ds.output_shape == (2, 2) and Ds.shape == (4, 20)
    assert np.array_equal(Ds.downsampling_mask, O.downsampling_mask(20, (2, 3), (4, 5)))
    assert DownSampling(20, 3, shape=(4, 5), axis=-1).output_shape == (4, 2)
    assert SubSampling(9, [0, 2], shape=(3, 3), axis=1).shape == (6, 9)
    with pytest.raises(ValueError):
        Masking(5, [1, 0])
    with pytest.raises(NotImplementedError):
        FirstDerivative(5, kind='sideways')


def test_arithmetic_dispatch_and_lipschitz_rules():
    G = Gradient((4, 5), kind='forward')
    G.lipschitz_cst = G.diff_lipschitz_cst = 2.0
    y = np.arange(40.)
    loss = SquaredL2Loss(dim=40, data=y)
    assert isinstance(loss, DiffMapShifted) and loss.diff_lipschitz_cst == 2
    half = (1 / 2) * loss
    assert isinstance(half, DiffMapComp) and isinstance(half.map1, HomothetyMap) and half.diff_lipschitz_cst == 1.0
    F = half * G
    assert isinstance(F, DiffMapComp) and F.shape == (1, 20) and F.diff_lipschitz_cst == 4.0   # beta = ||G||^2
    assert isinstance(G.H, AdjointLinearOperator) and G.H.shape == (20, 40) and G.H.lipschitz_cst == np.inf
    assert isinstance(G.H * G, LinOpComp) and (G.H * G).shape == (20, 20)
    assert isinstance(G + G, LinOpSum) and (G + G).lipschitz_cst == 4.0
    f = 3 * L1Norm(10)
    assert isinstance(f, ProxFuncPostComp) and f.scale == 3
    assert isinstance(L1Norm(10) * 2, ProxFuncPreComp)
    with pytest.raises(NotImplementedError):
        (-1) * L1Norm(10)                                   # functional.py:245: positive scalars only
    assert isinstance(L1Loss(10, np.ones(10)), ProxFuncPreComp)
    assert isinstance(SquaredL2Norm(10) + SquaredL2Norm(10), DiffMapSum)
    assert isinstance(L1Norm(10) + SquaredL2Norm(10), MapSum)
    assert isinstance(L21Norm(10, np.arange(10)), L1Norm) and isinstance(L21Norm(10, np.zeros(10)), L2Norm)
    assert isinstance(L21Norm(10, None), L1Norm)
    l21 = L21Norm(10, np.tile(np.arange(5), 2))
    assert isinstance(l21, L21Norm) and (l21.n_groups, l21.n_comp, l21._is_planes) == (5, 2, True)
    assert not L21Norm(10, np.repeat(np.arange(5), 2))._is_planes
    rag = L21Norm(10, [0, 0, 0, 1, 1, 1, 1, 2, 2, 2])       # ragged groups: CSR form (members sorted by group, offsets)
    assert not rag._is_planes and rag.n_groups == 3 and rag.n_comp is None
    assert rag._offsets_host.tolist() == [0, 3, 7, 10] and rag._members_host.tolist() == list(range(10))
    perm = L21Norm(6, [5., 2., 5., 2., 9., 9.])             # labels in any order: groups sorted by label value
    assert perm._members_host.tolist() == [1, 3, 0, 2, 4, 5] and perm._offsets_host.tolist() == [0, 2, 4, 6]
    V = LinOpVStack(G, IdentityOperator(20))
    assert V.shape == (60, 20) and np.isclose(V.lipschitz_cst, np.sqrt(5.))
    assert DiagonalOperator(np.array([-3., 2.])).lipschitz_cst == 2.0      # sic: max, not max|.| (base.py:567)
    assert NullOperator((3, 4)).lipschitz_cst == 0 and IdentityOperator(4).lipschitz_cst == 1
    ind = NonNegativeOrthant(7)
    assert isinstance(ind, IndicatorFunctional) and ind.kind == 'nonneg' and Segment(7, 1, 2).params == (1, 2)


def test_pds_constructor_matches_reference_rules():
    K = Gradient((4, 5), kind='forward')
    y = np.ones(20)
    F = (1 / 2) * SquaredL2Loss(dim=20, data=y)
    H = 0.1 * L21Norm(40, np.tile(np.arange(20), 2))
    with pytest.raises(ValueError, match='compute the Lipschitz constant'):
        PDS(dim=20, F=F, H=H, K=K)                         # proxalgs.py:285-290
    K.lipschitz_cst = K.diff_lipschitz_cst = np.sqrt(8.)
    s = PDS(dim=20, F=F, H=H, K=K)
    assert (s.tau, s.sigma, s.rho) == O.pds_step_sizes(1.0, np.sqrt(8.)) and s.beta == 1.0
    assert s.x0.dtype == np.float64 and s.z0.shape == (40,) and s.verbose == 1 and (s.max_iter, s.min_iter) == (500, 10)
    s = PDS(dim=20, F=None, H=H, K=K)
    assert (s.tau, s.sigma, s.rho) == O.pds_step_sizes(0, np.sqrt(8.)) and s.beta == 0
    s = PDS(dim=20, F=F)
    assert (s.tau, s.sigma, s.rho) == (2.0, 0, 0.9) and s.z0 is None and s._H is False
    assert isinstance(s.K, NullOperator) and isinstance(s.H, NullProximableFunctional)
    s = PDS(dim=20, F=F, H=H, K=K, tau=0.3)
    assert s.tau == s.sigma == 0.3
    s = PDS(dim=20, H=0.1 * L1Norm(20))
    assert isinstance(s.K, IdentityOperator) and s.tau == 1.0
    with pytest.raises(ValueError, match='proper dimension'):
        PDS(dim=21, F=F)
    with pytest.raises(ValueError, match='proper dimension'):
        PDS(dim=20, G=L1Norm(19))
    with pytest.raises(ValueError, match='inconsistent'):
        PDS(dim=20, H=0.1 * L1Norm(39), K=K)
    with pytest.raises(TypeError):
        PDS(dim=20, F='nope')
    with pytest.raises(TypeError):
        PDS(dim=20, G=SquaredL2Norm(20))
    with pytest.raises(TypeError):
        PDS(dim=20, H=L1Norm(20), K=3)
    A = Convolve2D(20, np.ones((3, 3)), (4, 5))
    with pytest.raises(ValueError, match='Lipschitz-continuous'):
        PDS(dim=20, F=F * A)                               # diff_lipschitz_cst = inf and no beta
    assert PDS(dim=20, F=F * A, beta=2.0).beta == 2.0
    assert PDS is PrimalDualSplitting


def test_other_solvers_constructor():
    y = np.ones(20)
    F = (1 / 2) * SquaredL2Loss(dim=20, data=y)
    a = APGD(dim=20, F=F, G=0.1 * L1Norm(20))
    assert a.tau == 1.0 and a.acceleration == 'CD' and a.d == 75. and a.init_iterand['past_t'] == 1
    with pytest.raises(ZeroDivisionError):
        APGD(dim=20, G=L1Norm(20))                         # 1/beta with F=None, as the reference
    assert [round(a._next_t(i, 1)[0], 12) for i in (0, 75)] == [1.0, 2.0]
    K = Gradient((4, 5), kind='forward')
    K.lipschitz_cst = K.diff_lipschitz_cst = np.sqrt(8.)
    c = CPS(dim=20, G=L1Loss(20, y), H=L1Norm(40), K=K)
    assert c.rho == 1 and c.beta == 0 and np.isclose(c.tau, 1 / np.sqrt(8.))
    d = DRS(dim=20, G=L1Loss(20, y), H=Segment(20), tau=0.5)
    assert d.sigma == 2.0 and d.rho == 1 and isinstance(d.K, IdentityOperator)
    f = FBS(dim=20, F=F, G=L1Norm(20))
    assert f.tau == 2.0 and f.rho == 1 and f._H is False


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        Gradient((4, 5))(np.ones(20))
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        PDS(dim=20, F=(1 / 2) * SquaredL2Loss(dim=20, data=np.ones(20)), verbose=None).iterate()
