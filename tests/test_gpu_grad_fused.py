"""GPU parity of the two-kernel Condat-Vu TV update for ANY 2-D Gradient kind (csrc/pcs_pds_grad.cu, PdsGradPlan):
pycsou's default ``kind='centered'`` (with and without pylops' edge rows) and ``'backward'``, against the numpy oracle
(reference update ``/root/reference/pycsou/opt/proxalgs.py:343-355`` on the pylops stand-in operators) and against this
package's op-by-op path.  Widths that are / are not 16-byte multiples (vector and scalar tiles), non-unit steps, every G
and H the plan accepts, every data mode (none / identity / gradient supplied by a convolution data term).
"""
import numpy as np
import pytest

from oracle import np_oracle as O
from tests.helpers import rel_l2

pytestmark = pytest.mark.gpu


def _build(shape, kind, edge, data, gname, hname, dtype, step=1., seed=0):
    from pycsou_b200.func.loss import SquaredL2Loss
    from pycsou_b200.func.penalty import L1Norm, L21Norm, NonNegativeOrthant, Segment
    from pycsou_b200.linop.conv import Convolve2D
    from pycsou_b200.linop.diff import Gradient
    rng = np.random.default_rng(seed)
    N = shape[0] * shape[1]
    y = rng.standard_normal(N)
    L = 3.0 / (min(step) if isinstance(step, tuple) else step)
    K = Gradient(shape=shape, kind=kind, edge=edge, step=step, dtype=dtype)
    K.lipschitz_cst = K.diff_lipschitz_cst = L
    Ko = O.Gradient(shape, step=step, edge=edge, kind=kind)
    beta = None
    if data == 'none':
        F = Fo = None
    elif data == 'identity':
        F, Fo = (1 / 2) * SquaredL2Loss(dim=N, data=y.astype(dtype)), O.HalfSquaredL2Loss(y)
    else:
        psf = rng.random((5, 7)) + 0.1
        psf /= psf.sum()
        F = (1 / 2) * SquaredL2Loss(dim=N, data=y.astype(dtype)) * Convolve2D(size=N, filter=psf, shape=shape, dtype=dtype)
        Fo = O.HalfSquaredL2Loss(y, O.Convolve2D(N, psf, shape), G_lipschitz=1.0)
        beta = 1.0
    G = {'none': None, 'nonneg': NonNegativeOrthant(dim=N), 'segment': Segment(dim=N, a=-0.3, b=0.4),
         'l1': 0.05 * L1Norm(dim=N)}[gname]
    Go = {'none': None, 'nonneg': O.NonNegativeOrthant(N), 'segment': O.Segment(N, -0.3, 0.4), 'l1': 0.05 * O.L1Norm(N)}[gname]
    H = 0.3 * L21Norm(dim=2 * N, groups=np.tile(np.arange(N), 2)) if hname == 'l21' else 0.3 * L1Norm(dim=2 * N)
    Ho = 0.3 * O.TVL21Norm(2 * N, N) if hname == 'l21' else 0.3 * O.L1Norm(2 * N)
    x0, z0 = rng.standard_normal(N), rng.standard_normal(2 * N)
    ours = dict(dim=N, F=F, G=G, H=H, K=K, beta=beta, x0=x0.astype(dtype), z0=z0.astype(dtype))
    orac = dict(F=Fo, G=Go, H=Ho, K=Ko, K_lipschitz=L, beta=beta, x0=x0, z0=z0)
    return ours, orac


CASES = [
    # shape, kind, edge, data, G, H, step
    ((40, 52), 'centered', True, 'identity', 'none', 'l21', 1.),
    ((33, 130), 'centered', False, 'identity', 'nonneg', 'l21', 1.),
    ((70, 301), 'centered', True, 'none', 'segment', 'l1', 1.),              # odd width: scalar tiles
    ((64, 64), 'centered', True, 'identity', 'l1', 'l21', (0.5, 2.0)),
    ((2, 2), 'centered', True, 'identity', 'none', 'l21', 1.),               # smallest image with edge rows
    ((3, 5), 'centered', False, 'identity', 'nonneg', 'l1', 1.),
    ((48, 136), 'backward', True, 'identity', 'nonneg', 'l21', 1.),
    ((17, 600), 'backward', False, 'none', 'none', 'l1', (2.0, 0.7)),        # more than one tile along the row
    ((100, 260), 'centered', True, 'conv', 'nonneg', 'l21', 1.),             # gradient of a convolution data term, handed over
    ((37, 129), 'backward', True, 'conv', 'segment', 'l21', 1.),
    ((260, 96), 'forward', True, 'identity', 'nonneg', 'l21', 1.),           # (forward goes to the one-kernel plan)
]


@pytest.mark.parametrize('dtype', ['float64', 'float32'])
@pytest.mark.parametrize('case', CASES, ids=[f'{c[0][0]}x{c[0][1]}-{c[1]}-edge{int(c[2])}-{c[3]}-{c[4]}-{c[5]}' for c in CASES])
def test_grad_plan_matches_oracle_and_op_by_op(case, dtype):
    from pycsou_b200.opt.proxalgs import PDS
    shape, kind, edge, data, gname, hname, step = case
    ours, orac = _build(shape, kind, edge, data, gname, hname, dtype, step)
    tol = 1e-11 if dtype == 'float64' else 3e-5
    for k in (1, 6):
        kw = dict(max_iter=k - 1, min_iter=k - 1, accuracy_threshold=0)
        s = PDS(verbose=None, **kw, **ours)
        est, _, diag = s.iterate()
        assert type(s._plan).__name__ == ('PdsTvPlan' if kind == 'forward' else 'PdsGradPlan')
        xo, zo, dref, _ = O.pds_iterate(ours['dim'], **kw, **orac)
        assert rel_l2(est['primal_variable'], xo) < tol and rel_l2(est['dual_variable'], zo) < tol
        g = PDS(verbose=None, fused=False, **kw, **ours)
        ref, _, dg = g.iterate()
        assert g._plan is None
        assert rel_l2(est['primal_variable'], ref['primal_variable']) < tol
        assert rel_l2(est['dual_variable'], ref['dual_variable']) < tol
        for col in list(diag.columns)[1:]:
            assert np.allclose(np.asarray(diag[col], float), np.asarray(dg[col], float), rtol=1e-6 if dtype == 'float64' else 2e-3)


def test_grad_plan_stop_rule_and_default_kind():
    """pycsou's Gradient defaults to kind='centered' (linop/diff.py:777): a user who keeps the default gets the fused plan,
    and the reference's stopping rule (core/solver.py:65-66) fires on the same update as in the oracle."""
    from pycsou_b200.func.loss import SquaredL2Loss
    from pycsou_b200.func.penalty import L21Norm
    from pycsou_b200.linop.diff import Gradient
    from pycsou_b200.opt.proxalgs import PDS
    from pycsou_b200.util import synth
    shape = (64, 96)
    N = shape[0] * shape[1]
    _, y = synth.tv_denoise_problem(shape, 0.1, 2)
    K = Gradient(shape=shape)                                   # default kind
    assert K.kind == 'centered'
    K.lipschitz_cst = K.diff_lipschitz_cst = 2.2
    s = PDS(dim=N, F=(1 / 2) * SquaredL2Loss(dim=N, data=y), H=0.1 * L21Norm(2 * N, np.tile(np.arange(N), 2)), K=K,
            max_iter=300, min_iter=5, accuracy_threshold=2e-3, verbose=None)
    est, _, diag = s.iterate()
    assert type(s._plan).__name__ == 'PdsGradPlan'
    xo, zo, _, n = O.pds_iterate(N, F=O.HalfSquaredL2Loss(y), H=0.1 * O.TVL21Norm(2 * N, N), K=O.Gradient(shape, kind='centered'),
                                 K_lipschitz=2.2, max_iter=300, min_iter=5, accuracy_threshold=2e-3)
    assert 6 < n < 300 and s.iter == n and len(diag) == n
    assert rel_l2(est['primal_variable'], xo) < 1e-10 and rel_l2(est['dual_variable'], zo) < 1e-10
