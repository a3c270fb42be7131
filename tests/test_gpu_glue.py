"""GPU parity of the composition glue of SURVEY.md 8(f)-1 / rows a9, a18, a20: ``LinOpHStack``, ``DiffFuncHStack``,
``MapStack``, ``ProxFuncAffineSum``, ``ProxFuncPreCompUnitOp``, vector ``DiagonalOperator``, ``PyLopLinearOperator``.

Each case is checked against the closed-form numpy expression of the reference's definition (cited) and, when the
reference package is present (``/root/reference`` here, the ``oracle/_ref`` archive on the GPU box), against the
reference's OWN classes fed with the same operators restated on numpy.
"""
import numpy as np
import pytest
import torch

from oracle import np_oracle as O
from oracle.ref_shim import reference_available
from tests.helpers import rel_l2

pytestmark = pytest.mark.gpu
TOL = 1e-13


def rng(seed=0):
    return np.random.default_rng(seed)


def test_diagonal_operator_vector_and_complex():
    """linop/base.py:551-579: y = diag * x, adjoint = conj(diag) * y, lipschitz_cst = np.max(diag) (sic)."""
    from pycsou_b200.linop.base import DiagonalOperator
    r = rng(1)
    n = 3001
    d, x = r.standard_normal(n), r.standard_normal(n)
    D = DiagonalOperator(d)
    assert D.shape == (n, n) and D.is_symmetric and D.lipschitz_cst == np.max(d)
    assert rel_l2(D(x), d * x) < TOL and rel_l2(D.adjoint(x), d * x) < TOL
    assert rel_l2((D * D)(x), d * d * x) < TOL and rel_l2(D.H(x), d * x) < TOL
    dc = r.standard_normal(n) + 1j * r.standard_normal(n)
    xc = r.standard_normal(n) + 1j * r.standard_normal(n)
    Dc = DiagonalOperator(dc)
    assert not Dc.is_symmetric
    assert rel_l2(Dc(xc), dc * xc) < TOL and rel_l2(Dc.adjoint(xc), dc.conj() * xc) < TOL
    x32 = x.astype(np.float32)
    assert rel_l2(DiagonalOperator(d.astype(np.float32))(x32), d.astype(np.float32) * x32) < 1e-6
    if reference_available():
        from oracle.ref_shim import load_reference
        load_reference()
        from pycsou.linop.base import DiagonalOperator as RefDiag
        assert np.array_equal(RefDiag(d)(x), np.asarray(D(x))) and RefDiag(d).lipschitz_cst == D.lipschitz_cst


def test_linop_hstack_vstack_and_stacked_functionals():
    """linop/base.py:159-336 (LinOpStack/HStack/VStack), func/base.py:21-121 (ProxFuncHStack, DiffFuncHStack),
    core/map.py:613-1001 (MapStack/DiffMapStack)."""
    from pycsou_b200.core.map import DiffMapHStack, MapHStack, MapVStack
    from pycsou_b200.func.base import DiffFuncHStack, ProxFuncHStack
    from pycsou_b200.func.loss import SquaredL2Loss
    from pycsou_b200.func.penalty import L1Norm, L2Norm, SquaredL2Norm
    from pycsou_b200.linop.base import DiagonalOperator, LinOpHStack, LinOpVStack
    from pycsou_b200.linop.conv import Convolve2D
    from pycsou_b200.linop.diff import FirstDerivative
    r = rng(2)
    shape = (24, 30)
    n = shape[0] * shape[1]
    h, d = r.random((3, 5)), r.standard_normal(n)
    A, Dg, Dv = Convolve2D(n, h, shape), DiagonalOperator(d), FirstDerivative(n, shape=shape, axis=1, kind='centered')
    Ao, Dvo = O.Convolve2D(n, h, shape), O.FirstDerivative(n, shape=shape, axis=1, kind='centered')
    for op in (A, Dg, Dv):
        op.lipschitz_cst = op.diff_lipschitz_cst = 2.0
    x3, y1 = r.standard_normal(3 * n), r.standard_normal(n)
    Hs = LinOpHStack(A, Dg, Dv)                                  # R^{3n} -> R^n : A x1 + D x2 + Dv x3
    assert Hs.shape == (n, 3 * n)
    want = Ao.matvec(x3[:n]) + d * x3[n:2 * n] + Dvo.matvec(x3[2 * n:])
    assert rel_l2(Hs(x3), want) < TOL
    assert rel_l2(Hs.adjoint(y1), np.concatenate([Ao.rmatvec(y1), d * y1, Dvo.rmatvec(y1)])) < TOL
    Vs = LinOpVStack(A, Dg, Dv)                                  # R^n -> R^{3n}
    assert Vs.shape == (3 * n, n)
    assert rel_l2(Vs(y1), np.concatenate([Ao.matvec(y1), d * y1, Dvo.matvec(y1)])) < TOL
    assert rel_l2(Vs.adjoint(x3), Ao.rmatvec(x3[:n]) + d * x3[n:2 * n] + Dvo.rmatvec(x3[2 * n:])) < TOL
    # <Hs x, y> == <x, Hs^H y>
    assert abs(np.dot(np.asarray(Hs(x3)), y1) - np.dot(x3, np.asarray(Hs.adjoint(y1)))) < 1e-9 * np.linalg.norm(x3) * np.linalg.norm(y1)
    # stacked functionals: value = sum of block values, prox / gradient block-wise (func/base.py:46-53, 96-103)
    Ph = ProxFuncHStack(L1Norm(n), 0.5 * L2Norm(2 * n))
    assert Ph.dim == 3 * n
    assert abs(Ph(x3) - (np.abs(x3[:n]).sum() + 0.5 * np.linalg.norm(x3[n:]))) < 1e-10
    want = np.concatenate([O.L1Norm(n).prox(x3[:n], 0.3), (0.5 * O.L2Norm(2 * n)).prox(x3[n:], 0.3)])
    assert rel_l2(Ph.prox(x3, 0.3), want) < TOL
    y2 = r.standard_normal(2 * n)
    Dh = DiffFuncHStack(SquaredL2Norm(n), (1 / 2) * SquaredL2Loss(2 * n, y2))
    assert abs(Dh(x3) - (np.sum(x3[:n] ** 2) + 0.5 * np.sum((x3[n:] - y2) ** 2))) < 1e-9 * np.sum(x3 ** 2)
    assert rel_l2(Dh.gradient(x3), np.concatenate([2 * x3[:n], x3[n:] - y2])) < TOL
    # plain map stacks (core/map.py:613-760): vertical = concatenated outputs, horizontal = summed outputs
    Mv = MapVStack(A, Dg)
    assert rel_l2(Mv(y1), np.concatenate([Ao.matvec(y1), d * y1])) < TOL
    Mh = MapHStack(A, Dg)
    assert rel_l2(Mh(x3[:2 * n]), Ao.matvec(x3[:n]) + d * x3[n:2 * n]) < TOL
    Dh2 = DiffMapHStack(A, Dg)
    assert rel_l2(Dh2(x3[:2 * n]), Ao.matvec(x3[:n]) + d * x3[n:2 * n]) < TOL


def test_prox_func_affine_sum_and_precomp_unitary():
    """core/functional.py:268-283 (prox_{f + <a,.> + c}(x) = prox_f(x - tau a)) and :302-313 (prox_{f o U} = U^H prox_f U)."""
    from pycsou_b200.core.functional import ProxFuncAffineSum, ProxFuncPreCompUnitOp
    from pycsou_b200.func.base import ExplicitLinearFunctional
    from pycsou_b200.func.penalty import L1Norm, L2Norm
    from pycsou_b200.linop.transform import DFT2
    r = rng(3)
    n = 2000
    a, x = r.standard_normal(n), r.standard_normal(n)
    lin = ExplicitLinearFunctional(a)
    f = L1Norm(n) + lin                                          # ProximableFunctional.__add__ (functional.py:215-222)
    assert isinstance(f, ProxFuncAffineSum)
    assert abs(f(x) - (np.abs(x).sum() + a @ x)) < 1e-10
    assert rel_l2(f.prox(x, 0.4), O.L1Norm(n).prox(x - 0.4 * a, 0.4)) < TOL
    g = ProxFuncAffineSum(0.7 * L2Norm(n), lin, intercept=2.5)
    assert abs(g(x) - (0.7 * np.linalg.norm(x) + a @ x + 2.5)) < 1e-10
    assert rel_l2(g.prox(x, 1.3), (0.7 * O.L2Norm(n)).prox(x - 1.3 * a, 1.3)) < TOL
    # f o U with U the unitary 2-D DFT
    shape = (32, 48)
    N = shape[0] * shape[1]
    U = DFT2(shape, dtype='complex128')
    xc = r.standard_normal(N) + 1j * r.standard_normal(N)
    from pycsou_b200.func.penalty import L21Norm
    groups = np.arange(N) // 4                                   # (L1Norm.prox is real-only, as in the reference)
    fu = L21Norm(N, groups) * U                                  # functional.py:229-234 -> ProxFuncPreCompUnitOp
    assert isinstance(fu, ProxFuncPreCompUnitOp)
    Ux = np.fft.fft2(xc.reshape(shape), norm='ortho').reshape(-1)
    fo = O.L21Norm(N, groups)
    assert abs(fu(xc) - fo(Ux)) < 1e-9 * fo(Ux)
    want = np.fft.ifft2(fo.prox(Ux, 0.6).reshape(shape), norm='ortho').reshape(-1)
    assert rel_l2(fu.prox(xc, 0.6), want) < 1e-12
    # (the reference's own ProxFuncAffineSum.prox goes through pylops' todense(): not runnable without pylops, so the
    # closed form of functional.py:281-283 above is the pin)


class _DeviceForwardDifference:
    """A user-written device operator with the pylops interface: forward difference along a flat vector."""

    def __init__(self, n):
        self.shape, self.dtype, self.explicit = (n, n), np.float64, False

    def matvec(self, x):
        y = torch.zeros_like(x)
        y[:-1] = x[1:] - x[:-1]
        return y

    def rmatvec(self, z):
        y = torch.zeros_like(z)
        y[:-1] -= z[:-1]
        y[1:] += z[:-1]
        return y


def test_pylop_linear_operator_adapter_in_a_solver():
    """linop/base.py:24-54: an operator given by matvec / rmatvec enters the solvers like a native one (generic op-by-op
    path); same iterates as the native FirstDerivative and as the numpy oracle."""
    from pycsou_b200.func.loss import SquaredL2Loss
    from pycsou_b200.func.penalty import L1Norm
    from pycsou_b200.linop.base import PyLopLinearOperator
    from pycsou_b200.linop.diff import FirstDerivative
    from pycsou_b200.opt.proxalgs import PDS
    n, k = 500, 25
    y = np.repeat([0., 2, 1, 3, 0], n // 5) + 0.05 * rng(4).standard_normal(n)
    P = PyLopLinearOperator(_DeviceForwardDifference(n), lipschitz_cst=2.0)
    Dn = FirstDerivative(n, kind='forward')
    Dn.lipschitz_cst = Dn.diff_lipschitz_cst = 2.0
    xr = rng(5).standard_normal(n)
    assert P.shape == (n, n) and rel_l2(P(xr), Dn(xr)) < TOL and rel_l2(P.adjoint(xr), Dn.adjoint(xr)) < TOL
    assert rel_l2(P.H(xr), Dn.adjoint(xr)) < TOL and rel_l2((2 * P)(xr), 2 * np.asarray(Dn(xr))) < TOL
    kw = dict(dim=n, F=(1 / 2) * SquaredL2Loss(dim=n, data=y), H=0.3 * L1Norm(dim=n), max_iter=k - 1, min_iter=k - 1,
              accuracy_threshold=0, verbose=None)
    ea, _, _ = PDS(K=P, **kw).iterate()
    eb, _, _ = PDS(K=Dn, **kw).iterate()
    assert rel_l2(ea['primal_variable'], eb['primal_variable']) < TOL and rel_l2(ea['dual_variable'], eb['dual_variable']) < TOL
    xo, zo, _, _ = O.pds_iterate(n, F=O.HalfSquaredL2Loss(y), H=0.3 * O.L1Norm(n), K=O.FirstDerivative(n, kind='forward'),
                                 K_lipschitz=2.0, max_iter=k - 1, min_iter=k - 1, accuracy_threshold=0)
    assert rel_l2(ea['primal_variable'], xo) < 1e-12 and rel_l2(ea['dual_variable'], zo) < 1e-12

    class Host:
        shape = (n, n)

        def matvec(self, x):
            return x.cpu().numpy()
        rmatvec = matvec
    with pytest.raises(TypeError):
        PyLopLinearOperator(Host())(xr)                          # no CPU path: host results are refused
    with pytest.raises(TypeError):
        PyLopLinearOperator(object())
