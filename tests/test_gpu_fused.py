"""GPU tests of the fused single-kernel Condat-Vu TV iteration (csrc/pcs_pds.cu) against the op-by-op path of the same
package (which is itself parity-checked against the oracle / golden fixtures in test_gpu_solvers.py) and against the
numpy oracle directly.  Shapes are chosen to exercise several column strips, several row segments, ragged last strips,
widths that are not 16-byte multiples (plain staging instead of TMA) and every image edge of the composite filter.
"""
import os

import numpy as np
import pytest

from tests.helpers import rel_l2

pytestmark = pytest.mark.gpu
SQRT8 = np.sqrt(8.)


def _problem(shape, data, gname, hname, dtype, step=1., psf_shape=(15, 15), seed=0):
    from pycsou_b200.func.loss import SquaredL2Loss
    from pycsou_b200.func.penalty import L1Norm, L21Norm, NonNegativeOrthant, Segment
    from pycsou_b200.linop.conv import Convolve2D
    from pycsou_b200.linop.diff import Gradient
    rng = np.random.default_rng(seed)
    N = shape[0] * shape[1]
    y = rng.standard_normal(N)
    K = Gradient(shape=shape, kind='forward', step=step, dtype=dtype)
    K.lipschitz_cst = K.diff_lipschitz_cst = SQRT8 / (min(step) if isinstance(step, tuple) else step)
    beta = None
    if data == 'none':
        F = None
    elif data == 'identity':
        F = (1 / 2) * SquaredL2Loss(dim=N, data=y.astype(dtype))
    else:
        kh, kw = psf_shape
        if data == 'sep':
            a, b = rng.random(kh) + 0.1, rng.random(kw) + 0.1
            psf = np.outer(a / a.sum(), b / b.sum())
        else:
            psf = rng.random((kh, kw)) + 0.1
            psf /= psf.sum()
        A = Convolve2D(size=N, filter=psf, shape=shape, dtype=dtype)
        F = (1 / 2) * SquaredL2Loss(dim=N, data=y.astype(dtype)) * A
        beta = 1.0
    G = {'none': None, 'nonneg': NonNegativeOrthant(dim=N), 'segment': Segment(dim=N, a=-0.3, b=0.4),
         'l1': 0.05 * L1Norm(dim=N)}[gname]
    H = 0.3 * L21Norm(dim=2 * N, groups=np.tile(np.arange(N), 2)) if hname == 'l21' else 0.3 * L1Norm(dim=2 * N)
    x0 = rng.standard_normal(N).astype(dtype)
    z0 = rng.standard_normal(2 * N).astype(dtype)      # non-zero everywhere, incl. the last row / column of each plane
    return dict(dim=N, F=F, G=G, H=H, K=K, beta=beta, x0=x0, z0=z0)


def _run(kw, k, fused):
    from pycsou_b200.opt.proxalgs import PDS
    s = PDS(max_iter=k - 1, min_iter=k - 1, accuracy_threshold=0, verbose=None, fused=fused, **kw)
    est, _, diag = s.iterate()
    return s, est, diag


CASES = [
    # shape, data, G, H, step, psf
    ((40, 50), 'identity', 'none', 'l21', 1., None),
    ((33, 130), 'identity', 'nonneg', 'l21', 1., None),
    ((70, 300), 'none', 'segment', 'l1', 1., None),
    ((64, 64), 'identity', 'l1', 'l21', (0.5, 2.0), None),
    ((40, 52), 'sep', 'nonneg', 'l21', 1., (15, 15)),
    ((100, 300), 'sep', 'nonneg', 'l21', 1., (15, 15)),
    ((257, 129), 'sep', 'none', 'l1', 1., (15, 15)),
    ((90, 131), 'sep', 'segment', 'l21', 1., (15, 15)),       # width not a 16-byte multiple -> plain staging
    ((64, 140), 'sep', 'nonneg', 'l21', 1., (5, 7)),
    ((64, 140), 'sep', 'nonneg', 'l21', 1., (4, 6)),          # even sizes: off-centre filter (conv.py:285-293)
    ((48, 136), 'sep', 'nonneg', 'l21', 1., (1, 3)),
    ((30, 40), 'sep', 'nonneg', 'l21', 1., (15, 15)),         # image barely larger than the filter support
    ((64, 140), 'nonsep', 'nonneg', 'l21', 1., (5, 5)),       # gradient by stand-alone kernels, fused primal/dual
    ((72, 256), 'sep', 'nonneg', 'l21', 1., (15, 15)),        # width = whole 128-column strips (no overlap column)
    ((48, 384), 'identity', 'nonneg', 'l21', 1., None),
    ((40, 264), 'sep', 'segment', 'l21', 1., (15, 15)),       # last strip narrower than the filter half-width
]


@pytest.mark.parametrize('dtype', ['float64', 'float32'])
@pytest.mark.parametrize('case', CASES, ids=[f'{c[0][0]}x{c[0][1]}-{c[1]}-{c[2]}-{c[3]}-psf{c[5][0]}x{c[5][1]}' if c[5] else f'{c[0][0]}x{c[0][1]}-{c[1]}-{c[2]}-{c[3]}' for c in CASES])
def test_fused_step_matches_op_by_op(case, dtype):
    shape, data, gname, hname, step, psf = case
    kw = _problem(shape, data, gname, hname, dtype, step=step, psf_shape=psf or (15, 15))
    tol = 1e-12 if dtype == 'float64' else 2e-5
    for k in (1, 4):
        s, est, diag = _run(kw, k, True)
        assert s._plan is not None, 'fused plan was not selected'
        _, ref, dref = _run(kw, k, False)
        assert rel_l2(est['primal_variable'], ref['primal_variable']) < tol
        assert rel_l2(est['dual_variable'], ref['dual_variable']) < tol
        for col in diag.columns[1:]:
            assert np.allclose(np.asarray(diag[col], float), np.asarray(dref[col], float), rtol=1e-6 if dtype == 'float64' else 1e-3)


@pytest.mark.parametrize('env', [dict(PCS_PDS_TMA='0'), dict(PCS_PDS_NSEG='1'), dict(PCS_PDS_NSEG='7'),
                                 dict(PCS_PDS_STRIP_COLS='37'), dict(PCS_PDS_STRIP_COLS='36'), dict(PCS_PDS_STRIP_COLS='100', PCS_PDS_NSEG='3')])
def test_fused_geometry_and_staging_are_bit_identical(env, monkeypatch):
    """Strip width, segment count and TMA vs plain staging must not change a single bit of the iterates."""
    kw = _problem((150, 260), 'sep', 'nonneg', 'l21', 'float64')
    _, base, _ = _run(kw, 3, True)
    for k_, v in env.items():
        monkeypatch.setenv(k_, v)
    _, est, _ = _run(kw, 3, True)
    assert np.array_equal(est['primal_variable'], base['primal_variable'])
    assert np.array_equal(est['dual_variable'], base['dual_variable'])


@pytest.mark.parametrize('shape', [(100, 300), (72, 256), (40, 264)])
def test_fused_single_role_kernel_fp32(shape, monkeypatch):
    """fp32 deconvolution normally runs the warp-specialised kernel; PCS_PDS_TMA=3 selects the single-role one."""
    kw = _problem(shape, 'sep', 'nonneg', 'l21', 'float32')
    _, ref, _ = _run(kw, 3, False)
    _, ws, _ = _run(kw, 3, True)
    monkeypatch.setenv('PCS_PDS_TMA', '3')
    _, est, _ = _run(kw, 3, True)
    assert rel_l2(est['primal_variable'], ref['primal_variable']) < 2e-5
    assert rel_l2(est['dual_variable'], ref['dual_variable']) < 2e-5
    assert rel_l2(est['primal_variable'], ws['primal_variable']) < 2e-6


def test_fused_matches_numpy_oracle_directly():
    from oracle import np_oracle as O
    from pycsou_b200.util import synth
    shape, k, lam = (80, 200), 12, 0.02
    _, psf, y = synth.tv_deconv_problem(shape, None, 0.02, 0)
    N = y.size
    from pycsou_b200.func.loss import SquaredL2Loss
    from pycsou_b200.func.penalty import L21Norm, NonNegativeOrthant
    from pycsou_b200.linop.conv import Convolve2D
    from pycsou_b200.linop.diff import Gradient
    from pycsou_b200.opt.proxalgs import PDS
    A = Convolve2D(size=N, filter=psf, shape=shape)
    K = Gradient(shape=shape, kind='forward')
    K.lipschitz_cst = K.diff_lipschitz_cst = SQRT8
    F = (1 / 2) * SquaredL2Loss(dim=N, data=y) * A
    s = PDS(dim=N, F=F, G=NonNegativeOrthant(dim=N), H=lam * L21Norm(dim=2 * N, groups=np.tile(np.arange(N), 2)), K=K,
            beta=1.0, max_iter=k - 1, min_iter=k - 1, accuracy_threshold=0, verbose=None)
    est, _, _ = s.iterate()
    assert type(s._plan).__name__ == 'PdsTvPlan' and s._plan.data_mode == 2
    Fo = O.HalfSquaredL2Loss(y, O.Convolve2D(N, psf, shape), G_lipschitz=1.0)
    xo, zo, _, _ = O.pds_iterate(N, F=Fo, G=O.NonNegativeOrthant(N), H=lam * O.TVL21Norm(2 * N, N),
                                 K=O.Gradient(shape, kind='forward'), K_lipschitz=SQRT8, beta=1.0, max_iter=k - 1,
                                 min_iter=k - 1, accuracy_threshold=0)
    assert rel_l2(est['primal_variable'], xo) < 1e-10 and rel_l2(est['dual_variable'], zo) < 1e-10
