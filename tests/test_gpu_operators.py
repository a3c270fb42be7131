"""GPU parity: every stand-alone operator / prox kernel (through the C ABI) vs the numpy oracle.

fp64: relative L2 <= 1e-13 (same formulas, different summation order for convolutions);
fp32: <= 2e-6.  Index / mask work: bit-exact.
"""
import numpy as np
import pytest
import torch

from oracle import np_oracle as O
from tests.helpers import rel_l2

pytestmark = pytest.mark.gpu

DT = [(np.float64, 1e-13), (np.float32, 3e-6)]


def rng(seed=0):
    return np.random.default_rng(seed)


@pytest.mark.parametrize('dtype,tol', DT)
@pytest.mark.parametrize('kind', ['forward', 'centered', 'backward'])
@pytest.mark.parametrize('edge', [True, False])
@pytest.mark.parametrize('shape,axis', [(None, 0), ((7, 9), 0), ((7, 9), 1), ((4, 5, 6), 1), ((4, 5, 6), -1), ((2, 3), 0),
                                        ((40, 520), 0), ((40, 520), 1), ((3, 20, 132), 1), ((2050,), 0)])
def test_first_derivative(dtype, tol, kind, edge, shape, axis):
    from pycsou_b200.linop.diff import FirstDerivative
    n = 37 if shape is None else int(np.prod(shape))
    x, z = rng(1).standard_normal(n).astype(dtype), rng(2).standard_normal(n).astype(dtype)
    D = FirstDerivative(n, shape=shape, axis=axis, step=0.7, edge=edge, dtype=np.dtype(dtype).name, kind=kind)
    Do = O.FirstDerivative(n, shape=shape, axis=axis, step=0.7, edge=edge, kind=kind)
    assert rel_l2(D(x), Do.matvec(x.astype(np.float64))) < tol
    assert rel_l2(D.adjoint(z), Do.rmatvec(z.astype(np.float64))) < tol
    assert isinstance(D(x), np.ndarray) and D(x).dtype == dtype


@pytest.mark.parametrize('dtype,tol', DT)
@pytest.mark.parametrize('edge', [True, False])
@pytest.mark.parametrize('shape,axis', [(None, 0), ((7, 9), 0), ((7, 9), 1), ((3, 4), 1)])
def test_second_derivative(dtype, tol, edge, shape, axis):
    from pycsou_b200.linop.diff import SecondDerivative
    n = 23 if shape is None else int(np.prod(shape))
    x, z = rng(1).standard_normal(n).astype(dtype), rng(2).standard_normal(n).astype(dtype)
    D = SecondDerivative(n, shape=shape, axis=axis, step=1.3, edge=edge, dtype=np.dtype(dtype).name)
    Do = O.SecondDerivative(n, shape=shape, axis=axis, step=1.3, edge=edge)
    assert rel_l2(D(x), Do.matvec(x.astype(np.float64))) < tol
    assert rel_l2(D.adjoint(z), Do.rmatvec(z.astype(np.float64))) < tol


@pytest.mark.parametrize('dtype,tol', DT)
@pytest.mark.parametrize('kind', ['forward', 'centered', 'backward'])
@pytest.mark.parametrize('shape', [(48, 64), (5, 3), (2, 2), (1, 9), (9, 1), (130, 257), (4, 5, 6)])
def test_gradient(dtype, tol, kind, shape):
    from pycsou_b200.linop.diff import Gradient
    if kind == 'centered' and min(shape) < 2:
        pytest.skip('pylops centered+edge needs 2 samples')
    n, nd = int(np.prod(shape)), len(shape)
    x, z = rng(1).standard_normal(n).astype(dtype), rng(2).standard_normal(nd * n).astype(dtype)
    step = tuple(0.5 + 0.25 * i for i in range(nd))
    G = Gradient(shape, step=step, kind=kind, dtype=np.dtype(dtype).name)
    Go = O.Gradient(shape, step=list(step), kind=kind)
    assert G.shape == (nd * n, n)
    assert rel_l2(G(x), Go.matvec(x.astype(np.float64))) < tol
    assert rel_l2(G.adjoint(z), Go.rmatvec(z.astype(np.float64))) < tol
    # adjoint dot test on the device result itself
    assert abs(np.dot(G(x).astype(np.float64), z) - np.dot(x, G.adjoint(z).astype(np.float64))) < 1e3 * tol * n


@pytest.mark.parametrize('dtype,tol', DT)
@pytest.mark.parametrize('edge', [True, False])
@pytest.mark.parametrize('shape', [(19, 33), (70, 1028), (6, 6), (5, 40), (3, 3), (130, 516)])
def test_laplacian(dtype, tol, edge, shape):
    """(19, 33): scalar tiles; (70, 1028), (130, 516): vector tiles, several tiles per row; (6, 6): smallest image on the
    tiled path with edge rows; (5, 40), (3, 3): scalar fallback kernel."""
    from pycsou_b200.linop.diff import Laplacian
    n = shape[0] * shape[1]
    x, z = rng(1).standard_normal(n).astype(dtype), rng(2).standard_normal(n).astype(dtype)
    L = Laplacian(shape, weights=(1.5, 0.5), step=(1., 2.), edge=edge, dtype=np.dtype(dtype).name)
    Lo = O.Laplacian(shape, weights=(1.5, 0.5), step=(1., 2.), edge=edge)
    assert rel_l2(L(x), Lo.matvec(x.astype(np.float64))) < tol
    assert rel_l2(L.adjoint(z), Lo.rmatvec(z.astype(np.float64))) < tol


@pytest.mark.parametrize('dtype,tol', DT)
@pytest.mark.parametrize('nh', [1, 4, 5, 65])
def test_convolve1d(dtype, tol, nh):
    from pycsou_b200.linop.conv import Convolve1D
    n = 300
    h = rng(3).random(nh)
    x = rng(1).standard_normal(n).astype(dtype)
    C, Co = Convolve1D(n, h, dtype=np.dtype(dtype).name), O.Convolve1D(n, h)
    assert rel_l2(C(x), Co.matvec(x.astype(np.float64))) < tol
    assert rel_l2(C.adjoint(x), Co.rmatvec(x.astype(np.float64))) < tol
    # along an axis of a 2-D array
    shape = (12, 25)
    for axis in (0, 1):
        C = Convolve1D(n, h[:min(nh, 9)], reshape_dims=shape, axis=axis, dtype=np.dtype(dtype).name)
        Co = O.Convolve1D(n, h[:min(nh, 9)], reshape_dims=shape, axis=axis)
        assert rel_l2(C(x), Co.matvec(x.astype(np.float64))) < 10 * tol
        assert rel_l2(C.adjoint(x), Co.rmatvec(x.astype(np.float64))) < 10 * tol


@pytest.mark.parametrize('dtype,tol', DT)
@pytest.mark.parametrize('hshape,sep', [((15, 15), True), ((15, 15), False), ((3, 7), False), ((4, 6), True),
                                        ((1, 1), True), ((25, 25), False)])
@pytest.mark.parametrize('shape', [(48, 64), (70, 33), (5, 4)])
def test_convolve2d(dtype, tol, hshape, sep, shape):
    from pycsou_b200.linop.conv import Convolve2D
    r = rng(4)
    h = np.outer(r.random(hshape[0]), r.random(hshape[1])) if sep else r.random(hshape)
    n = shape[0] * shape[1]
    x = r.standard_normal(n).astype(dtype)
    C, Co = Convolve2D(n, h, shape, dtype=np.dtype(dtype).name), O.Convolve2D(n, h, shape)
    assert C.is_separable == (sep or h.size == 1 or min(hshape) == 1)
    assert rel_l2(C(x), Co.matvec(x.astype(np.float64))) < 20 * tol
    assert rel_l2(C.adjoint(x), Co.rmatvec(x.astype(np.float64))) < 20 * tol


def test_sampling_bit_exact():
    from pycsou_b200.linop.sampling import DownSampling, Masking, SubSampling
    r = rng(5)
    for dtype in (np.float32, np.float64, np.complex64, np.complex128):
        n = 5000
        x = (r.standard_normal(n) + (1j * r.standard_normal(n) if np.dtype(dtype).kind == 'c' else 0)).astype(dtype)
        mask = r.random(n) < 0.3
        M, Mo = Masking(n, mask, dtype=dtype), O.Masking(n, mask, dtype=dtype)
        y = M(x)
        assert y.dtype == dtype and np.array_equal(y, Mo.matvec(x))
        assert np.array_equal(M.adjoint(y), Mo.rmatvec(y))
        assert M.nb_of_samples == Mo.nb_of_samples and M.shape == Mo.shape
    x = r.standard_normal(20 * 30)
    for kw in (dict(downsampling_factor=3), dict(downsampling_factor=(2, 3), shape=(20, 30)),
               dict(downsampling_factor=4, shape=(20, 30), axis=-1), dict(downsampling_factor=4, shape=(20, 30), axis=0)):
        D, Do = DownSampling(600, **kw), O.DownSampling(600, **kw)
        assert np.array_equal(D.downsampling_mask, Do.sampling_bool)
        assert np.array_equal(D(x), Do.matvec(x)) and np.array_equal(D.adjoint(D(x)), Do.rmatvec(Do.matvec(x)))
    idx = np.array([5, 0, 17, 5, 3, 17])          # duplicates: the last occurrence wins in the adjoint
    S, So = SubSampling(600, idx), O.SubSampling(600, idx)
    y = r.standard_normal(6)
    assert np.array_equal(S(x), So.matvec(x)) and np.array_equal(S.adjoint(y), So.rmatvec(y))
    for axis in (0, 1):
        ia = [0, 2, 7]
        S, So = SubSampling(600, ia, shape=(20, 30), axis=axis), O.SubSampling(600, ia, shape=(20, 30), axis=axis)
        y = S(x)
        assert np.array_equal(y, So.matvec(x)) and np.array_equal(S.adjoint(y), So.rmatvec(y))
    # empty mask / empty input
    M = Masking(10, np.zeros(10, bool))
    assert M(np.arange(10.)).size == 0 and np.array_equal(M.adjoint(np.zeros(0)), np.zeros(10))


@pytest.mark.parametrize('dtype,tol', DT)
def test_prox_and_projections(dtype, tol):
    from pycsou_b200.func import penalty as P, loss as L
    r = rng(6)
    n = 4001
    x = r.standard_normal(n).astype(dtype)
    x64 = x.astype(np.float64)
    for tau in (0.05, 0.7, 3.0):
        assert rel_l2(P.L1Norm(n).prox(x, tau), O.L1Norm(n).prox(x64, tau)) < tol
        assert rel_l2(P.L2Norm(n).prox(x, tau * 20), O.L2Norm(n).prox(x64, tau * 20)) < tol
        assert rel_l2((0.3 * P.L1Norm(n)).fenchel_prox(x, tau), (0.3 * O.L1Norm(n)).fenchel_prox(x64, tau)) < tol
        y = r.standard_normal(n).astype(dtype)
        assert rel_l2(L.L1Loss(n, y).prox(x, tau), O.PreCompShift(O.L1Norm(n), y.astype(np.float64)).prox(x64, tau)) < 10 * tol
        assert rel_l2(L.L2BallLoss(n, y, radius=2.).prox(x, tau),
                      O.PreCompShift(O.L2Ball(n, 2.), y.astype(np.float64)).prox(x64, tau)) < 10 * tol
    m = 2000
    z = r.standard_normal(2 * m).astype(dtype)
    z[:5] = 0
    z[m:m + 5] = 0          # zero groups: 1 - tau/0 -> -inf -> 0
    g = np.tile(np.arange(m), 2)
    f, fo = P.L21Norm(2 * m, g), O.TVL21Norm(2 * m, m)
    assert isinstance(f, P.L21Norm)
    assert rel_l2(f.prox(z, 0.4), fo.prox(z.astype(np.float64), 0.4)) < tol
    assert abs(f(z) - fo(z.astype(np.float64))) < 1e3 * tol * fo(z.astype(np.float64))
    assert rel_l2((0.1 * f).fenchel_prox(z, 0.5), (0.1 * fo).fenchel_prox(z.astype(np.float64), 0.5)) < 10 * tol
    # arbitrary (non-planes) equal-size groups
    gi = np.repeat(np.arange(m), 2)
    f2, f2o = P.L21Norm(2 * m, gi), O.L21Norm(2 * m, gi)
    assert rel_l2(f2.prox(z, 0.4), f2o.prox(z.astype(np.float64), 0.4)) < tol
    assert np.array_equal(P.NonNegativeOrthant(n).prox(x, 1), O.NonNegativeOrthant(n).prox(x, 1))
    assert np.array_equal(P.Segment(n, -0.5, 0.2).prox(x, 1), O.Segment(n, -0.5, 0.2).prox(x, 1))
    assert P.NonNegativeOrthant(n)(x) == np.inf and P.NonNegativeOrthant(n)(np.abs(x)) == 0
    assert P.Segment(n, -9, 9)(x) == 0 and P.Segment(n, 0, 1)(x) == np.inf
    assert abs(P.L1Norm(n)(x) - np.sum(np.abs(x64))) < 1e3 * tol * n
    assert abs(P.SquaredL2Norm(n)(x) - np.sum(x64 ** 2)) < 1e3 * tol * n


def test_l21norm_reference_doctest_pin():
    """The reference's own doctest for L21Norm (/root/reference/pycsou/func/penalty.py:494-509): two groups of five."""
    from pycsou_b200.func import penalty as P
    x = np.arange(10, dtype=np.float64)
    groups = np.concatenate((np.ones(5), 2 * np.ones(5)))
    f = P.L21Norm(dim=x.size, groups=groups)
    assert type(f) is P.L21Norm
    assert type(P.L21Norm(dim=x.size, groups=np.arange(10))) is P.L1Norm          # penalty.py:497-500
    assert type(P.L21Norm(dim=x.size, groups=np.ones(10))) is P.L2Norm            # penalty.py:501-504
    assert abs(f(x) - 21.44594499772297) < 1e-13                                  # penalty.py:505-506
    tau = 0.5
    # penalty.py:507-509: prox == per-group L2Norm prox
    l2 = P.L2Norm(dim=5)
    expect = np.concatenate((l2.prox(x[0:5], tau=tau), l2.prox(x[5:10], tau=tau)))
    assert np.allclose(f.prox(x, tau=tau), expect, rtol=1e-14, atol=0)
    assert rel_l2(f.prox(x, tau), O.L21Norm(10, groups).prox(x, tau)) < 1e-15


@pytest.mark.parametrize('dtype,tol', DT)
def test_l21norm_general_groups(dtype, tol):
    """Arbitrary label arrays (penalty.py:532-560): ragged groups, groups longer than a warp, labels in any order,
    a zero group (1 - tau/0 -> -inf -> 0), complex input."""
    from pycsou_b200.func import penalty as P
    r = rng(11)
    n = 5000
    labels = r.integers(0, 37, size=n).astype(np.float64) * 1.5 - 3        # ragged, unordered, non-integer labels
    labels[:40] = 99.0                                                     # one group that is entirely zero
    x = r.standard_normal(n).astype(dtype)
    x[:40] = 0
    f, fo = P.L21Norm(n, labels), O.L21Norm(n, labels)
    assert type(f) is P.L21Norm and not f._is_planes
    x64 = x.astype(np.float64)
    for tau in (0.3, 5.0, 40.0):
        assert rel_l2(f.prox(x, tau), fo.prox(x64, tau)) < tol
    assert abs(f(x) - fo(x64)) < 10 * tol * fo(x64)
    assert rel_l2((0.2 * f).fenchel_prox(x, 0.7), (0.2 * fo).fenchel_prox(x64, 0.7)) < 10 * tol
    # many small ragged groups (sizes 1..7) and five-component planes (more than the planes kernel takes)
    sizes = r.integers(1, 8, size=900)
    lab2 = r.permutation(np.repeat(np.arange(sizes.size), sizes))
    x2 = r.standard_normal(lab2.size).astype(dtype)
    assert rel_l2(P.L21Norm(lab2.size, lab2).prox(x2, 0.8), O.L21Norm(lab2.size, lab2).prox(x2.astype(np.float64), 0.8)) < tol
    lab3 = np.tile(np.arange(300), 5)
    x3 = r.standard_normal(lab3.size).astype(dtype)
    assert rel_l2(P.L21Norm(lab3.size, lab3).prox(x3, 1.1), O.L21Norm(lab3.size, lab3).prox(x3.astype(np.float64), 1.1)) < tol
    # complex input: |x|^2 over real and imaginary parts (np.linalg.norm of a complex group)
    ctype = np.complex64 if dtype == 'float32' else np.complex128
    xc = (r.standard_normal(n) + 1j * r.standard_normal(n)).astype(ctype)
    assert rel_l2(f.prox(xc, 2.0), fo.prox(xc.astype(np.complex128), 2.0)) < tol
    assert abs(f(xc) - fo(xc.astype(np.complex128))) < 10 * tol * fo(xc.astype(np.complex128))


@pytest.mark.parametrize('dtype,tol', DT)
def test_loss_gradient_and_composition(dtype, tol):
    from pycsou_b200.func.loss import SquaredL2Loss
    from pycsou_b200.linop.conv import Convolve2D
    r = rng(7)
    shape = (40, 37)
    n = shape[0] * shape[1]
    h = r.random((5, 7))
    y, x = r.standard_normal(n).astype(dtype), r.standard_normal(n).astype(dtype)
    A = Convolve2D(n, h, shape, dtype=np.dtype(dtype).name)
    A.lipschitz_cst = A.diff_lipschitz_cst = 3.0
    F = (1 / 2) * SquaredL2Loss(dim=n, data=y) * A
    Fo = O.HalfSquaredL2Loss(y.astype(np.float64), O.Convolve2D(n, h, shape), G_lipschitz=3.0)
    assert F.diff_lipschitz_cst == Fo.diff_lipschitz_cst == 9.0
    assert rel_l2(F.gradient(x), Fo.gradient(x.astype(np.float64))) < 20 * tol
    assert abs(F(x) - Fo(x.astype(np.float64))) < 1e3 * tol * Fo(x.astype(np.float64))


def test_lipschitz_power_iteration():
    """compute_lipschitz_cst (reference core/linop.py:279-321, ARPACK) = the largest singular value: compared with a
    dense SVD of the oracle operator to 1e-6 relative (the reference's ARPACK call is not repeatable), and
    never below it by more than that -- the Condat-Vu step rule divides by it."""
    from pycsou_b200.linop.conv import Convolve2D
    from pycsou_b200.linop.diff import FirstDerivative, Gradient

    def dense_norm(op, n):
        M = np.stack([op.matvec(e) for e in np.eye(n)], axis=1)
        return np.linalg.svd(M, compute_uv=False)[0]

    for kind in ('forward', 'centered'):
        D = FirstDerivative(10, kind=kind)
        D.compute_lipschitz_cst()
        true = dense_norm(O.FirstDerivative(10, kind=kind), 10)
        assert abs(D.lipschitz_cst - true) <= 1e-9 * true            # core/map.py:418-423: ~2.0 / ~1.5 for n = 10
    shape = (24, 20)
    G = Gradient(shape, kind='forward')
    G.compute_lipschitz_cst()
    true = dense_norm(O.Gradient(shape, kind='forward'), shape[0] * shape[1])
    assert abs(G.lipschitz_cst - true) <= 1e-6 * true and G._spectral_converged
    h = rng(3).random((5, 7))
    A = Convolve2D(shape[0] * shape[1], h, shape)
    A.compute_lipschitz_cst()
    true = dense_norm(O.Convolve2D(shape[0] * shape[1], h, shape), shape[0] * shape[1])
    assert abs(A.lipschitz_cst - true) <= 1e-6 * true
    # a large image: the top of the spectrum is clustered, a short run does not converge -> the analytic bound (>= ||K||,
    # within 1e-6 of it at this size) is used instead of an under-estimate
    G = Gradient((512, 512), kind='forward')
    G.compute_lipschitz_cst(maxiter=40)
    assert not G._spectral_converged and G.lipschitz_cst == np.sqrt(8.)
    G.compute_lipschitz_cst(tol=1e-9, maxiter=2000)
    true = np.sqrt(8.) * np.cos(np.pi / 1024)          # 2-D forward difference with a zero last row / column
    assert abs(G.lipschitz_cst - true) <= 1e-6 * true or G.lipschitz_cst == np.sqrt(8.)


def test_torch_in_torch_out_and_errors():
    from pycsou_b200.linop.diff import Gradient
    G = Gradient((8, 8), kind='forward')
    x = torch.arange(64, dtype=torch.float64, device='cuda')
    y = G(x)
    assert isinstance(y, torch.Tensor) and y.is_cuda and y.shape == (128,)
    assert isinstance(G * x, torch.Tensor)
    with pytest.raises(ValueError):
        from pycsou_b200 import _kernels as K
        K.grad2d(x, x, 8, 8, 'forward', True, 1, 1)     # aliased stencil output


@pytest.mark.parametrize('dtype,tol', [('float64', 1e-12), ('float32', 2e-5)])
@pytest.mark.parametrize('shape,kshape', [((70, 90), (21, 17)), ((64, 64), (32, 30)), ((40, 130), (9, 7)), ((33, 35), (33, 35))])
def test_convolve2d_fft_path_matches_oracle(shape, kshape, dtype, tol, monkeypatch):
    """Large non-separable PSFs take the FFT path (cuFFT + fused spectral multiply / crop kernels): forward, adjoint and
    the fused residual against the scipy-based oracle; odd and even filter sizes; a filter as large as the image."""
    import torch
    from oracle import np_oracle as O
    from pycsou_b200.linop.conv import Convolve2D
    monkeypatch.setenv('PCS_CONV2D_PATH', 'fft' if kshape[0] * kshape[1] <= 225 else '')
    rng = np.random.default_rng(3)
    N = shape[0] * shape[1]
    h = rng.random(kshape) + 0.05
    h /= h.sum()
    A = Convolve2D(size=N, filter=h, shape=shape, dtype=dtype)
    assert A.use_fft and not A.is_separable
    Ao = O.Convolve2D(N, h, shape)
    x, y = rng.standard_normal(N).astype(dtype), rng.standard_normal(N).astype(dtype)
    assert rel_l2(A(x), Ao.matvec(x.astype(np.float64))) < tol
    assert rel_l2(A.adjoint(y), Ao.rmatvec(y.astype(np.float64))) < tol
    xd, yd = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    r = A._residual(xd, yd).cpu().numpy()
    assert rel_l2(r, Ao.matvec(x.astype(np.float64)) - y) < tol
    # adjointness <A x, y> = <x, A^T y>
    lhs, rhs = float(np.dot(A(x).astype(np.float64), y)), float(np.dot(x, A.adjoint(y).astype(np.float64)))
    assert abs(lhs - rhs) <= (1e-10 if dtype == 'float64' else 1e-3) * max(1.0, abs(lhs))


def test_convolve2d_fft_and_direct_paths_agree(monkeypatch):
    from pycsou_b200.linop.conv import Convolve2D
    rng = np.random.default_rng(4)
    shape, N = (50, 60), 3000
    h = rng.random((11, 13))
    x = rng.standard_normal(N)
    monkeypatch.setenv('PCS_CONV2D_PATH', 'direct')
    Ad = Convolve2D(size=N, filter=h, shape=shape)
    monkeypatch.setenv('PCS_CONV2D_PATH', 'fft')
    Af = Convolve2D(size=N, filter=h, shape=shape)
    assert Af.use_fft and not Ad.use_fft
    assert rel_l2(Af(x), Ad(x)) < 1e-13 and rel_l2(Af.adjoint(x), Ad.adjoint(x)) < 1e-13


def test_moving_average_operators():
    """MovingAverage1D / MovingAverage2D (reference conv.py:298-418) against the oracle; even windows become odd."""
    from pycsou_b200.linop.conv import MovingAverage1D, MovingAverage2D
    r = rng(9)
    for win, shape, axis in ((5, (300,), 0), (4, (40, 30), 0), (7, (40, 30), 1), (3, (6, 20, 10), 1)):
        n = int(np.prod(shape))
        x = r.standard_normal(n)
        A, Ao = MovingAverage1D(win, shape, axis=axis), O.MovingAverage1D(win, shape, axis=axis)
        assert A.shape == (n, n)
        assert rel_l2(A(x), Ao.matvec(x)) < 1e-13 and rel_l2(A.adjoint(x), Ao.rmatvec(x)) < 1e-13
    for win, shape in (((5, 3), (40, 50)), ((4, 6), (33, 47))):
        n = shape[0] * shape[1]
        x = r.standard_normal(n)
        A, Ao = MovingAverage2D(win, shape), O.MovingAverage2D(win, shape)
        assert A.is_separable
        assert rel_l2(A(x), Ao.matvec(x)) < 1e-13 and rel_l2(A.adjoint(x), Ao.rmatvec(x)) < 1e-13
    # constant signal: the interior is reproduced exactly (the window sums to one)
    y = MovingAverage1D(5, (50,))(np.ones(50))
    assert np.allclose(y[2:-2], 1.0) and y[0] < 1.0


@pytest.mark.parametrize('dtype,tol', [('float64', 1e-12), ('float32', 2e-6)])
def test_entropy_and_barrier_functionals(dtype, tol):
    """LogBarrier / ShannonEntropy / KLDivergence (reference func/penalty.py:776-922, func/loss.py:597-682): the
    reference's doctest literals, values incl. the +inf cases, proxes against the oracle (scipy's Lambert W)."""
    from pycsou_b200.func.loss import KLDivergence
    from pycsou_b200.func.penalty import LogBarrier, ShannonEntropy
    x1 = np.arange(10).astype(dtype)
    f, fo = LogBarrier(10), O.LogBarrier(10)
    assert f(x1) == np.inf and np.isclose(f(x1 + 2), -17.502307845873887, rtol=tol * 10)
    assert np.array_equal(np.round(f.prox(x1, tau=1)), [1, 2, 2, 3, 4, 5, 6, 7, 8, 9])
    g, go = ShannonEntropy(10), O.ShannonEntropy(10)
    x2 = np.zeros(10, dtype=dtype)
    x2[0] = 10
    assert np.isclose(g(x1), 79.05697962199447, rtol=tol * 10) and np.isclose(g(x2), 23.02585092994046, rtol=tol * 10)
    assert g(x1 - 1) == np.inf
    assert np.array_equal(np.round(g.prox(x1, tau=2)), [0, 0, 1, 1, 1, 2, 2, 3, 3, 4])
    y = np.arange(10).astype(dtype)
    k, ko = KLDivergence(10, y), O.KLDivergence(10, y)
    assert np.isclose(k(2 * y), 13.80837687480246, rtol=tol * 10)
    assert k(y - 1) == np.inf
    assert np.array_equal(np.round(k.prox(2 * y, tau=1)), [0, 2, 4, 6, 8, 10, 12, 14, 16, 18])
    r = rng(12)
    n = 20000
    x = (r.standard_normal(n) * 5).astype(dtype)
    d = r.random(n).astype(dtype) * 3
    for tau in (0.05, 1.0, 7.5):
        assert rel_l2(f.__class__(n).prox(x, tau), O.LogBarrier(n).prox(x.astype(np.float64), tau)) < tol
        assert rel_l2(KLDivergence(n, d).prox(x, tau), O.KLDivergence(n, d.astype(np.float64)).prox(x.astype(np.float64), tau)) < tol
        assert rel_l2(ShannonEntropy(n).prox(x, tau), O.ShannonEntropy(n).prox(x.astype(np.float64), tau)) < 20 * tol
    xp = np.abs(x) + 0.1
    assert np.isclose(LogBarrier(n)(xp), O.LogBarrier(n)(xp.astype(np.float64)), rtol=10 * tol)
    assert np.isclose(ShannonEntropy(n)(xp), O.ShannonEntropy(n)(xp.astype(np.float64)), rtol=10 * tol)
    assert np.isclose(KLDivergence(n, d)(xp), O.KLDivergence(n, d.astype(np.float64))(xp.astype(np.float64)), rtol=10 * tol)
    # Moreau identity through the base class (core/functional.py:207) and the solver-facing scaling 2 * f
    z = np.abs(x[:100]) + 0.5
    assert rel_l2((2 * LogBarrier(100)).fenchel_prox(z, 0.7), (2 * O.LogBarrier(100)).fenchel_prox(z.astype(np.float64), 0.7)) < 10 * tol


@pytest.mark.parametrize('dtype,tol', [('float64', 1e-10), ('float32', 2e-6)])
def test_l1_ball_linfty_and_squared_l1(dtype, tol):
    """proj_l1_ball, L1Ball, LInftyNorm, SquaredL1Norm ('sort' and 'root'), RealLine, ImagLine, ConsistencyLoss and
    the *Loss wrappers over them (reference math/prox.py:117-164, func/penalty.py:248-417, 671-767, func/loss.py:
    329-475, 536-586): the reference's doctest literals, then the oracle on random vectors.  The device finds the
    thresholds by exact Newton steps where the reference uses brentq (2e-12) or a sort: 1e-10 is the fp64 bar."""
    from pycsou_b200.func.loss import ConsistencyLoss, L1BallLoss, LInftyLoss, SquaredL1Loss
    from pycsou_b200.func.penalty import ImagLine, L1Ball, LInftyNorm, RealLine, SquaredL1Norm
    from pycsou_b200.math.prox import proj_l1_ball
    x = np.linspace(-1, 1, 5).astype(dtype)
    assert np.allclose(proj_l1_ball(x, radius=2), [-0.75, -0.25, 0., 0.25, 0.75], atol=10 * tol, rtol=0)
    x1 = np.arange(10).astype(dtype)
    assert SquaredL1Norm(10)(x1) == 2025 and LInftyNorm(10)(x1) == 9 and (3 * LInftyNorm(10))(x1) == 27
    assert np.allclose(SquaredL1Norm(10, 'sort').prox(x1, tau=1), SquaredL1Norm(10, 'root').prox(x1, tau=1))
    x2 = (x1 / np.linalg.norm(x1, ord=1)).astype(dtype)
    ball = L1Ball(10, radius=10)
    xf = (x2 * 0.999).astype(dtype)
    assert ball(x1) == np.inf and ball(xf) == 0
    assert np.isclose(np.linalg.norm(ball.prox(x1, tau=1), ord=1), 10.0, rtol=10 * tol)
    assert np.array_equal(ball.prox(xf, tau=1), xf)
    r = rng(21)
    n = 50001
    xs = (3 * r.standard_normal(n)).astype(dtype)
    xs[::7] = 0
    xs[5::11] = xs[3]                                # ties at and around the threshold
    x64 = xs.astype(np.float64)
    # the error is measured against ||x||: soft(x, t) = x - t cancels when few entries survive (radius 0.3 keeps
    # three of 50001), which fp32 cannot resolve to 2e-6 of the RESULT
    err = lambda a, b: float(np.linalg.norm(np.asarray(a, dtype=np.float64) - b) / np.linalg.norm(x64))
    for radius in (0.3, 40.0, 9000.0, 1e7):
        assert err(proj_l1_ball(xs, radius), O.proj_l1_ball(x64, radius)) < tol
        assert err(L1Ball(n, radius).prox(xs, 1.0), O.L1Ball(n, radius).prox(x64, 1.0)) < tol
    for tau in (1e-6, 0.02, 0.9, 30.0):
        assert err(LInftyNorm(n).prox(xs, tau), O.LInftyNorm(n).prox(x64, tau)) < tol
        for how in ('sort', 'root'):
            assert err(SquaredL1Norm(n, how).prox(xs, tau), O.SquaredL1Norm(n, 'sort').prox(x64, tau)) < tol
        assert err((2 * LInftyNorm(n)).fenchel_prox(xs, tau), (2 * O.LInftyNorm(n)).fenchel_prox(x64, tau)) < 10 * tol
    assert np.isclose(SquaredL1Norm(n)(xs), O.SquaredL1Norm(n)(x64), rtol=10 * tol)
    assert LInftyNorm(n)(xs) == O.LInftyNorm(n)(x64)
    assert np.array_equal(SquaredL1Norm(4).prox(np.zeros(4, dtype=dtype), 1.0), np.zeros(4))
    y = r.standard_normal(n).astype(dtype)
    y64 = y.astype(np.float64)
    shift = lambda f: O.PreCompShift(f, y64)
    assert rel_l2(L1BallLoss(n, y, radius=25.0).prox(xs, 1.0), shift(O.L1Ball(n, 25.0)).prox(x64, 1.0)) < tol
    assert rel_l2(LInftyLoss(n, y).prox(xs, 0.4), shift(O.LInftyNorm(n)).prox(x64, 0.4)) < tol
    assert rel_l2(SquaredL1Loss(n, y).prox(xs, 0.4), shift(O.SquaredL1Norm(n)).prox(x64, 0.4)) < tol
    cdt = 'complex128' if dtype == 'float64' else 'complex64'
    xc = (np.arange(10) + 1j).astype(cdt)
    assert (RealLine(10)(xc), RealLine(10)(np.real(xc).astype(dtype))) == (np.inf, 0)
    assert np.array_equal(RealLine(10).prox(xc, tau=1), np.real(xc))
    assert (ImagLine(10)(xc), ImagLine(10)((1j * np.imag(xc)).astype(cdt))) == (np.inf, 0)
    assert np.array_equal(ImagLine(10).prox(xc, tau=1), np.imag(xc))
    loss = ConsistencyLoss(10, x1)
    assert (loss(2 * x1), loss(x1)) == (np.inf, 0) and np.array_equal(loss.prox(2 * x1, tau=1), x1)


def test_generalised_derivatives_and_polynomial_operator():
    """GeneralisedDerivative / GeneralisedLaplacian / PolynomialLinearOperator (reference linop/diff.py:222-377,
    960-1068; linop/base.py:636-712) against the oracle (itself checked against the live reference), forward and
    adjoint, plus the reference's doctests (diff.py:283-288: order-1 iterated == FirstDerivative; base.py:664-669)."""
    from pycsou_b200.linop.base import DenseLinearOperator, PolynomialLinearOperator
    from pycsou_b200.linop.diff import FirstDerivative, GeneralisedDerivative as GD, GeneralisedLaplacian as GL
    xs = np.linspace(-5, 5, 100)
    z = np.sinc(xs)
    assert np.allclose(GD(size=100, kind_op='iterated', order=1, kind_diff='forward') * z,
                       FirstDerivative(size=100, kind='forward') * z)
    L = DenseLinearOperator(np.arange(64).reshape(8, 8).astype(float))
    PL = PolynomialLinearOperator(LinOp=L, coeffs=[1 / 2, 2, 1])
    x8 = np.arange(8).astype(float)
    assert np.allclose(PL(x8), x8 / 2 + 2 * L(x8) + (L ** 2)(x8))
    r = rng(22)
    x, y = r.standard_normal(60), r.standard_normal(60)
    cases = [dict(kind_op='iterated', order=3), dict(kind_op='sobolev', order=2, constant=0.3),
             dict(kind_op='exponential', order=3, constant=-0.2),
             dict(kind_op='polynomial', coeffs=0.5 * np.array([1e-2, 0, -0.2, 0, 1]))]
    for kw in cases:
        for kind_diff in ('forward', 'backward', 'centered'):
            A, Ao = GD(size=60, kind_diff=kind_diff, **kw), O.GeneralisedDerivative(60, kind_diff=kind_diff, **kw)
            assert rel_l2(A(x), Ao.matvec(x)) < 1e-12 and rel_l2(A.adjoint(y), Ao.rmatvec(y)) < 1e-12
        A = GD(size=60, shape=(6, 10), axis=1, step=0.5, **kw)
        Ao = O.GeneralisedDerivative(60, shape=(6, 10), axis=1, step=0.5, **kw)
        assert rel_l2(A(x), Ao.matvec(x)) < 1e-12 and rel_l2(A.adjoint(y), Ao.rmatvec(y)) < 1e-12
    shape = (70, 83)
    n = shape[0] * shape[1]
    x, y = r.standard_normal(n), r.standard_normal(n)
    for kw in (dict(kind='iterated', order=2), dict(kind='sobolev', order=2, constant=0.4),
               dict(kind='polynomial', coeffs=[0.5, -1.0, 0.25])):
        A, Ao = GL(shape=shape, **kw), O.GeneralisedLaplacian(shape, **kw)
        assert rel_l2(A(x), Ao.matvec(x)) < 1e-12 and rel_l2(A.adjoint(y), Ao.rmatvec(y)) < 1e-12
    A32 = GL(shape=shape, dtype='float32', kind='iterated', order=2)
    assert rel_l2(A32(x.astype(np.float32)), O.GeneralisedLaplacian(shape, kind='iterated', order=2).matvec(x)) < 5e-6


@pytest.mark.parametrize('dtype,tol', [('float64', 1e-13), ('float32', 1e-6)])
def test_pooling(dtype, tol):
    """Pooling (reference linop/sampling.py:394-536): the doctest literals, then ragged 1-D / 2-D / 3-D shapes against
    the oracle restatement of block_reduce (zero padding counts in the mean) and of the repeat-and-crop adjoint."""
    from pycsou_b200.linop.sampling import Pooling
    x = np.arange(24).reshape(4, 6).astype(dtype)
    P = Pooling(shape=x.shape, block_size=(2, 3), pooling_func='mean', dtype=dtype)
    assert P.output_shape == (2, 2) and P.shape == (4, 24)
    assert np.array_equal((P * x.flatten()).reshape(P.output_shape), [[4., 7.], [16., 19.]])
    assert np.array_equal(P.adjoint(P * x.flatten()).reshape(x.shape),
                          np.repeat(np.repeat([[4., 7.], [16., 19.]], 2, axis=0), 3, axis=1))
    P = Pooling(shape=x.shape, block_size=(2, 3), pooling_func='sum', dtype=dtype)
    assert np.array_equal((P * x.flatten()).reshape(P.output_shape), [[24, 42], [96, 114]])
    r = rng(23)
    for shape, block in [((1000,), (7,)), ((257, 300), (10, 20)), ((130, 97), (1, 5)), ((12, 33, 41), (5, 4, 3)),
                         ((64, 64), (64, 64)), ((5, 7), (9, 9))]:
        for func in ('mean', 'sum'):
            P, Po = Pooling(shape, block, func, dtype=dtype), O.Pooling(shape, block, func)
            assert P.output_shape == Po.output_shape and P.shape == Po.shape
            xv = r.standard_normal(P.shape[1]).astype(dtype)
            yv = r.standard_normal(P.shape[0]).astype(dtype)
            assert rel_l2(P(xv), Po.matvec(xv.astype(np.float64))) < tol
            assert np.array_equal(P.adjoint(yv), Po.rmatvec(yv))
    with pytest.raises(ValueError):
        Pooling((4, 6), (2,))
    with pytest.raises(ValueError):
        Pooling((4, 6), (2, 3), 'max')


def test_leaves_against_reference_fixture():
    """The device leaves of SURVEY 8(f) against tests/golden/leaves_rank23.npz (outputs of the reference's own classes,
    tests/golden/gen_golden.py): fp64, 1e-10 relative L2 measured against ||x|| for the threshold proxes."""
    from tests import oracle_cases as C
    from tests.helpers import load_golden
    from pycsou_b200.func.loss import KLDivergence, L1BallLoss, LInftyLoss
    from pycsou_b200.func.penalty import LInftyNorm, LogBarrier, ShannonEntropy, SquaredL1Norm
    from pycsou_b200.linop.diff import GeneralisedDerivative, GeneralisedLaplacian
    from pycsou_b200.math.prox import proj_l1_ball
    meta, g = load_golden('leaves_rank23')
    x, d, y = C.leaf_inputs(meta['n'], meta['seed'])
    n = x.size
    err = lambda a, b: float(np.linalg.norm(np.asarray(a) - b) / np.linalg.norm(x))
    for i, tau in enumerate(C.LEAF_TAUS):
        assert err(LogBarrier(n).prox(x, tau), g[f'logbarrier_{i}']) < 1e-12
        assert err(ShannonEntropy(n).prox(x, tau), g[f'shannon_{i}']) < 1e-12
        assert err(KLDivergence(n, d).prox(x, tau), g[f'kl_{i}']) < 1e-12
        assert err(LInftyNorm(n).prox(x, tau), g[f'linfty_{i}']) < 1e-10
        assert err(SquaredL1Norm(n, 'sort').prox(x, tau), g[f'sql1_sort_{i}']) < 1e-10
        assert err(SquaredL1Norm(n, 'root').prox(x, tau), g[f'sql1_root_{i}']) < 1e-10
        assert err(LInftyLoss(n, y).prox(x, tau), g[f'linfty_loss_{i}']) < 1e-10
    for i, r in enumerate(C.LEAF_RADII):
        assert err(proj_l1_ball(x, r), g[f'l1ball_{i}']) < 1e-10
        assert err(L1BallLoss(n, y, radius=r).prox(x, 1.0), g[f'l1ball_loss_{i}']) < 1e-10
    xp, v = np.abs(x) + 0.1, meta['values']
    assert np.isclose(LogBarrier(n)(xp), v['logbarrier'], rtol=1e-12) and np.isclose(ShannonEntropy(n)(xp), v['shannon'], rtol=1e-12)
    assert np.isclose(KLDivergence(n, d)(xp), v['kl'], rtol=1e-12) and LInftyNorm(n)(x) == v['linfty']
    assert np.isclose(SquaredL1Norm(n)(x), v['sql1'], rtol=1e-12)
    for i, kw in enumerate(C.GD_CASES):
        for kd in ('forward', 'backward', 'centered'):
            A = GeneralisedDerivative(size=n, shape=(20, 20), axis=1, step=0.5, kind_diff=kd, **kw)
            assert rel_l2(A(x), g[f'gd_{i}_{kd}_fwd']) < 1e-12 and rel_l2(A.adjoint(y), g[f'gd_{i}_{kd}_adj']) < 1e-12
    for i, kw in enumerate(C.GL_CASES):
        A = GeneralisedLaplacian(shape=(20, 20), **kw)
        assert rel_l2(A(x), g[f'gl_{i}_fwd']) < 1e-12 and rel_l2(A.adjoint(y), g[f'gl_{i}_adj']) < 1e-12


def test_quadratic_form():
    """QuadraticForm (reference func/penalty.py:925-990): value x^T Q x and gradient 2 Q x with Q = L^T L built from
    this package's operators, and as the smooth term of a forward-backward solve."""
    from pycsou_b200.func.penalty import QuadraticForm
    from pycsou_b200.linop.diff import Laplacian
    r = rng(13)
    shape, n = (24, 30), 720
    Lp = Laplacian(shape=shape)
    Lp.lipschitz_cst = Lp.diff_lipschitz_cst = 8.0
    Q = Lp.H * Lp
    Q.lipschitz_cst = Q.diff_lipschitz_cst = 64.0
    x = r.standard_normal(n)
    Lo = O.Laplacian(shape)
    Qx = Lo.rmatvec(Lo.matvec(x))
    F = QuadraticForm(n, linop=Q)
    assert np.isclose(F(x), np.dot(x, Qx), rtol=1e-12) and rel_l2(F.gradient(x), 2 * Qx) < 1e-13
    assert F.diff_lipschitz_cst == 128.0
    F0 = QuadraticForm(n)
    assert np.isclose(F0(x), np.dot(x, x), rtol=1e-13) and rel_l2(F0.gradient(x), 2 * x) < 1e-15 and F0.diff_lipschitz_cst == 2


@pytest.mark.parametrize('shape', [(30, 44), (6, 10, 12)])
def test_directional_derivatives(shape):
    """First/SecondDirectionalDerivative, DirectionalGradient, DirectionalLaplacian (reference diff.py:380-774) against
    the oracle restatement of pylops' Sum * Diagonal(v) * Gradient, constant and per-pixel direction fields."""
    from pycsou_b200.linop.diff import (DirectionalGradient, DirectionalLaplacian, FirstDirectionalDerivative,
                                         SecondDirectionalDerivative)
    r = rng(21)
    nd, n = len(shape), int(np.prod(shape))
    x, y = r.standard_normal(n), r.standard_normal(n)
    v_const = r.standard_normal(nd)
    v_const /= np.linalg.norm(v_const)
    v_field = r.standard_normal((nd,) + shape)
    for v in (v_const, v_field):
        for kind in ('centered', 'forward'):
            D, Do = FirstDirectionalDerivative(shape, v, step=0.5, kind=kind), O.FirstDirectionalDerivative(shape, v, step=0.5, kind=kind)
            assert rel_l2(D(x), Do.matvec(x)) < 1e-13 and rel_l2(D.adjoint(y), Do.rmatvec(y)) < 1e-13
        D2, D2o = SecondDirectionalDerivative(shape, v), O.SecondDirectionalDerivative(shape, v)
        assert rel_l2(D2(x), D2o.matvec(x)) < 1e-12 and rel_l2(D2.adjoint(y), D2o.rmatvec(y)) < 1e-12
        assert abs(np.dot(D2(x), y) - np.dot(x, D2.adjoint(y))) < 1e-9 * n
    D1, D1b = FirstDirectionalDerivative(shape, v_const), FirstDirectionalDerivative(shape, v_field)
    DG = DirectionalGradient([D1, D1b])
    assert DG.shape == (2 * n, n)
    assert rel_l2(DG(x), np.concatenate([D1(x), D1b(x)])) < 1e-15
    S1, S2 = SecondDirectionalDerivative(shape, v_const), SecondDirectionalDerivative(shape, v_field)
    DL = DirectionalLaplacian([S1, S2], weights=[3., 0.5])
    assert rel_l2(DL(x), S1(x) + 0.5 * S2(x)) < 1e-13           # sic: the first weight is not applied (diff.py:770-773)


@pytest.mark.parametrize('dtype,tol', [('float64', 1e-12), ('float32', 2e-5)])
def test_integration1d(dtype, tol):
    """Integration1D (reference diff.py:1071-1137): every axis position, short and long contiguous lines, adjointness."""
    from pycsou_b200.linop.diff import Integration1D
    r = rng(31)
    for shape, axis in (((24,), 0), ((5000,), 0), ((40, 30), 0), ((40, 30), 1), ((3, 4100), 1), ((6, 7, 9), 1), ((6, 7, 9), -1)):
        n = int(np.prod(shape))
        x, y = r.standard_normal(n).astype(dtype), r.standard_normal(n).astype(dtype)
        A = Integration1D(n, shape if len(shape) > 1 else None, axis=axis, step=0.25, dtype=dtype)
        Ao = O.Integration1D(n, shape if len(shape) > 1 else None, axis=axis, step=0.25)
        assert rel_l2(A(x), Ao.matvec(x.astype(np.float64))) < tol
        assert rel_l2(A.adjoint(y), Ao.rmatvec(y.astype(np.float64))) < tol
        lhs, rhs = np.dot(A(x).astype(np.float64), y), np.dot(x, A.adjoint(y).astype(np.float64))
        assert abs(lhs - rhs) < (1e-9 if dtype == 'float64' else 1e-2) * max(1.0, abs(lhs))
    sig = np.array([0, 0, 0, 1, 0, 0, 0, 0, 0, 2, 0, 0, 0, 0, -1, 0, 0, 0, 0, 2, 0, 0, 0, 0], dtype=dtype)
    assert np.array_equal(Integration1D(sig.size, dtype=dtype)(sig), np.cumsum(sig))       # the docstring's example signal
