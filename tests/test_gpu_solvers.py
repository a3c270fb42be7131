"""GPU parity of the solvers: pycsou_b200 (through the public, reference-shaped API) vs the golden fixtures produced by
the reference classes and vs the numpy oracle.  Written the way a reference user would write the script.

Tolerances (BASELINE.json north_star): fp64 iterates <= 1e-10 relative L2; fp32 objective <= 1e-4 relative.
Each case runs twice: fused kernels allowed, and forced op-by-op (``fused=False``).
"""
import numpy as np
import pytest

from pycsou_b200.util import synth
from tests import oracle_cases as C
from tests.helpers import load_golden, rel_l2

pytestmark = pytest.mark.gpu
TOL64 = 1e-10
SQRT8 = np.sqrt(8.)


def _imports():
    from pycsou_b200.func.loss import L1Loss, SquaredL2Loss
    from pycsou_b200.func.penalty import L1Norm, L21Norm, NonNegativeOrthant, Segment
    from pycsou_b200.linop.conv import Convolve1D, Convolve2D
    from pycsou_b200.linop.diff import FirstDerivative, Gradient
    from pycsou_b200.linop.sampling import DownSampling
    from pycsou_b200.opt.proxalgs import APGD, CPS, DRS, FBS, PDS
    return locals()


def tv_K(shape, kind='forward', dtype='float64'):
    from pycsou_b200.linop.diff import Gradient
    K = Gradient(shape=shape, kind=kind, edge=True, dtype=dtype)
    K.lipschitz_cst = K.diff_lipschitz_cst = SQRT8
    return K


@pytest.mark.parametrize('fused', [True, False])
@pytest.mark.parametrize('acc,ks', [('CD', (1, 2, 10, 100, 500)), ('BT', (1, 10, 100)), (None, (1, 10, 100))])
def test_c1_lasso_apgd(acc, ks, fused):
    m = _imports()
    meta, g = load_golden('c1_lasso_apgd')
    truth, h, y = synth.spikes_1d(4096, 32, 0)
    N = y.size
    Gop = m['Convolve1D'](size=N, filter=h)
    lam = 0.1 * np.max(np.abs(Gop.adjoint(y)))
    assert np.isclose(lam, meta['lam'], rtol=1e-12)
    for k in ks:
        F = (1 / 2) * m['SquaredL2Loss'](dim=N, data=y) * Gop
        G = lam * m['L1Norm'](dim=N)
        s = m['APGD'](dim=N, F=F, G=G, tau=1.0, acceleration=acc, beta=1.0, max_iter=k - 1, min_iter=k - 1,
                      accuracy_threshold=0, verbose=None, d=75., fused=fused)
        est, conv, diag = s.iterate()
        assert s.iter == k and conv is True
        assert rel_l2(est['iterand'], g[f'x_{acc}_k{k}']) < TOL64
    assert rel_l2(est['past_aux'], g[f'aux_{acc}_k{ks[-1]}']) < TOL64
    assert np.isclose(est['past_t'], float(g[f't_{acc}_k{ks[-1]}']), rtol=1e-15)
    ri = np.asarray(diag['Relative Improvement'], dtype=float)
    assert list(diag.columns) == ['Iter', 'Relative Improvement']
    assert ri[0] == np.inf and np.allclose(ri[1:], g[f'relimp_{acc}'][1:], rtol=1e-7)


@pytest.mark.parametrize('fused', [True, False])
def test_c2_tv_denoise_pds_fp64(fused):
    m = _imports()
    meta, g = load_golden('c2_tv_denoise_pds')
    shape = tuple(meta['shape'])
    _, y = synth.tv_denoise_problem(shape, 0.1, 0)
    N = y.size
    groups = np.tile(np.arange(N), 2)
    for k in (1, 2, 10, 50):
        F = (1 / 2) * m['SquaredL2Loss'](dim=N, data=y)
        s = m['PDS'](dim=N, F=F, H=meta['lam'] * m['L21Norm'](dim=2 * N, groups=groups), K=tv_K(shape), max_iter=k - 1,
                     min_iter=k - 1, accuracy_threshold=0, verbose=None, fused=fused)
        assert np.isclose(s.tau, meta['tau'], rtol=1e-15) and s.sigma == s.tau and s.rho == 0.9
        est, _, diag = s.iterate()
        assert s.iter == k
        assert rel_l2(est['primal_variable'], g[f'x_k{k}']) < TOL64
        assert rel_l2(est['dual_variable'], g[f'z_k{k}']) < TOL64
    # reference stop rule with the default max_iter/min_iter/threshold: same iteration count, same iterate
    F = (1 / 2) * m['SquaredL2Loss'](dim=N, data=y)
    s = m['PDS'](dim=N, F=F, H=meta['lam'] * m['L21Norm'](dim=2 * N, groups=groups), K=tv_K(shape), verbose=None,
                 fused=fused)
    est, _, diag = s.iterate()
    assert s.iter == meta['n_updates_stop'] and len(diag) == s.iter
    assert rel_l2(est['primal_variable'], g['x_stop']) < TOL64 and rel_l2(est['dual_variable'], g['z_stop']) < TOL64
    assert list(diag.columns) == ['Iter', 'Relative Improvement (primal variable)',
                                  'Relative Improvement (dual variable)']
    rp = np.asarray(diag['Relative Improvement (primal variable)'], dtype=float)
    rd = np.asarray(diag['Relative Improvement (dual variable)'], dtype=float)
    assert rp[0] == np.inf and rd[0] == np.inf
    assert np.allclose(rp[1:], g['relimp_primal'][1:], rtol=1e-7) and np.allclose(rd[1:], g['relimp_dual'][1:], rtol=1e-7)


@pytest.mark.parametrize('fused', [True, False])
@pytest.mark.parametrize('name', ['c2_tv_denoise_pds_nonneg', 'c2_tv_denoise_pds_segment', 'c2_aniso_tv_centered_pds'])
def test_c2_variants(name, fused):
    m = _imports()
    meta, g = load_golden(name)
    shape = tuple(meta['shape'])
    _, y = synth.tv_denoise_problem(shape, 0.1, 0)
    N = y.size
    for k in (1, 10, 50):
        F = (1 / 2) * m['SquaredL2Loss'](dim=N, data=y)
        if name.endswith('nonneg'):
            kw = dict(G=m['NonNegativeOrthant'](dim=N), H=0.1 * m['L21Norm'](2 * N, np.tile(np.arange(N), 2)), K=tv_K(shape))
        elif name.endswith('segment'):
            kw = dict(G=m['Segment'](dim=N, a=0.1, b=0.7), H=0.1 * m['L21Norm'](2 * N, np.tile(np.arange(N), 2)),
                      K=tv_K(shape))
        else:
            kw = dict(H=0.1 * m['L1Norm'](dim=2 * N), K=tv_K(shape, 'centered'))
        s = m['PDS'](dim=N, F=F, max_iter=k - 1, min_iter=k - 1, accuracy_threshold=0, verbose=None, fused=fused, **kw)
        est, _, _ = s.iterate()
        assert rel_l2(est['primal_variable'], g[f'x_k{k}']) < TOL64
        assert rel_l2(est['dual_variable'], g[f'z_k{k}']) < TOL64
        want = None if not fused else ('PdsGradPlan' if name.endswith('centered_pds') else 'PdsTvPlan')
        assert (type(s._plan).__name__ if s._plan is not None else None) == want     # centred differences are fused too


def _deconv_solver(m, shape, psf, y, lam, k, dtype, fused):
    N = y.size
    A = m['Convolve2D'](size=N, filter=psf, shape=shape, dtype=dtype)
    F = (1 / 2) * m['SquaredL2Loss'](dim=N, data=y.astype(dtype)) * A
    return m['PDS'](dim=N, F=F, G=m['NonNegativeOrthant'](dim=N),
                    H=lam * m['L21Norm'](2 * N, np.tile(np.arange(N), 2)), K=tv_K(shape, dtype=dtype), beta=1.0,
                    max_iter=k - 1, min_iter=k - 1, accuracy_threshold=0, verbose=None, dtype=dtype, fused=fused)


@pytest.mark.parametrize('fused', [True, False])
@pytest.mark.parametrize('tag', ['sep', 'nonsep'])
def test_c3_tv_deconv_pds_fp64(tag, fused):
    m = _imports()
    meta, g = load_golden(f'c3_tv_deconv_pds_{tag}')
    shape = tuple(meta['shape'])
    psf = synth.gaussian_psf_2d(15, 2.5) if tag == 'sep' else synth.rotated_gaussian_psf_2d(15)
    _, h, y = synth.tv_deconv_problem(shape, psf, 0.02, 0)
    for k in (1, 2, 10, 50):
        s = _deconv_solver(m, shape, h, y, meta['lam'], k, 'float64', fused)
        assert np.isclose(s.tau, meta['tau'], rtol=1e-15)
        est, _, _ = s.iterate()
        assert rel_l2(est['primal_variable'], g[f'x_k{k}']) < TOL64
        assert rel_l2(est['dual_variable'], g[f'z_k{k}']) < TOL64


@pytest.mark.parametrize('fused', [True, False])
@pytest.mark.parametrize('tag', ['sep', 'nonsep'])
def test_c3_tv_deconv_pds_fp32_objective(tag, fused):
    m = _imports()
    meta, g = load_golden(f'c3_tv_deconv_pds_{tag}')
    shape = tuple(meta['shape'])
    psf = synth.gaussian_psf_2d(15, 2.5) if tag == 'sep' else synth.rotated_gaussian_psf_2d(15)
    _, h, y = synth.tv_deconv_problem(shape, psf, 0.02, 0)
    s = _deconv_solver(m, shape, h, y, meta['lam'], 50, 'float32', fused)
    est, _, _ = s.iterate()
    assert est['primal_variable'].dtype == np.float32
    obj = C.c3_objective(est['primal_variable'].astype(np.float64), tag)
    assert abs(obj - meta['objective_k50']) <= 1e-4 * abs(meta['objective_k50'])


@pytest.mark.parametrize('fused', [True, False])
def test_examples_rst_pds(fused):
    """doc/general/examples.rst:19-47 verbatim (explicit Lipschitz constants as in the fixture)."""
    m = _imports()
    meta, g = load_golden('examples_pbp_pds')
    x = np.repeat([0, 2, 1, 3, 0, 2, 0], 10)
    D = m['FirstDerivative'](size=x.size, kind='forward')
    D.lipschitz_cst = D.diff_lipschitz_cst = 2.0
    Gop = m['DownSampling'](size=x.size, downsampling_factor=3)
    Gop.lipschitz_cst = Gop.diff_lipschitz_cst = 1.0
    y = Gop(x)
    l22_loss = (1 / 2) * m['SquaredL2Loss'](dim=Gop.shape[0], data=y)
    F = l22_loss * Gop
    H = 0.1 * m['L1Norm'](dim=D.shape[0])
    G = 0.01 * m['L1Norm'](dim=Gop.shape[1])
    pds = m['PDS'](dim=Gop.shape[1], F=F, G=G, H=H, K=D, verbose=None, fused=fused)
    est, converged, diag = pds.iterate()
    assert pds.iter == meta['n_updates'] and np.isclose(pds.tau, meta['tau'], rtol=1e-15)
    assert rel_l2(est['primal_variable'], g['x']) < TOL64 and rel_l2(est['dual_variable'], g['z']) < TOL64


@pytest.mark.parametrize('fused', [True, False])
def test_cps_drs_fbs(fused):
    m = _imports()
    meta, g = load_golden('misc_cps_drs_fbs')
    shape = tuple(meta['shape'])
    N = shape[0] * shape[1]
    _, y = synth.tv_denoise_problem(shape, 0.1, 1)
    for k in (1, 10, 40):
        kw = dict(max_iter=k - 1, min_iter=k - 1, accuracy_threshold=0, verbose=None, fused=fused)
        s = m['CPS'](dim=N, G=m['L1Loss'](dim=N, data=y), H=0.3 * m['L21Norm'](2 * N, np.tile(np.arange(N), 2)),
                     K=tv_K(shape), **kw)
        est, _, _ = s.iterate()
        assert rel_l2(est['primal_variable'], g[f'cps_x_k{k}']) < TOL64
        assert rel_l2(est['dual_variable'], g[f'cps_z_k{k}']) < TOL64
        s = m['DRS'](dim=N, G=m['L1Loss'](dim=N, data=y), H=m['Segment'](dim=N, a=0.2, b=0.6), tau=0.7, **kw)
        est, _, _ = s.iterate()
        assert rel_l2(est['primal_variable'], g[f'drs_x_k{k}']) < TOL64
        assert rel_l2(est['dual_variable'], g[f'drs_z_k{k}']) < TOL64
        s = m['FBS'](dim=N, F=(1 / 2) * m['SquaredL2Loss'](dim=N, data=y), G=0.2 * m['L1Norm'](dim=N), **kw)
        est, _, diag = s.iterate()
        assert est['dual_variable'] is None and list(diag.columns) == ['Iter', 'Relative Improvement (primal variable)']
        assert rel_l2(est['primal_variable'], g[f'fbs_x_k{k}']) < TOL64


def test_iterates_generator_and_torch_output():
    import torch
    m = _imports()
    shape = (24, 32)
    N = shape[0] * shape[1]
    _, y = synth.tv_denoise_problem(shape, 0.1, 1)
    F = (1 / 2) * m['SquaredL2Loss'](dim=N, data=y)
    mk = lambda **kw: m['PDS'](dim=N, F=F, H=0.1 * m['L21Norm'](2 * N, np.tile(np.arange(N), 2)), K=tv_K(shape),
                               verbose=None, **kw)
    s = mk()
    its = [it['primal_variable'] for it in s.iterates(5)]            # numpy in, numpy out (as the reference yields)
    assert all(isinstance(a, np.ndarray) for a in its)
    s2 = mk(max_iter=4, min_iter=4, accuracy_threshold=0)
    est, _, _ = s2.iterate()
    assert rel_l2(its[-1], est['primal_variable']) < 1e-14
    s3 = mk(max_iter=4, min_iter=4, accuracy_threshold=0, x0=torch.zeros(N, dtype=torch.float64, device='cuda'))
    est3, _, _ = s3.iterate()
    assert isinstance(est3['primal_variable'], torch.Tensor) and est3['primal_variable'].is_cuda


@pytest.mark.parametrize('output', ['numpy', 'torch'])
def test_iterates_yield_fresh_arrays_and_iterate_continues(output):
    """core/solver.py:88-103: every yielded iterand is a fresh array (collected WITHOUT copying, they must equal the
    op-by-op path's iterates one by one); core/solver.py:64-66: a second iterate() without reset() continues at
    self.iter -- a no-op once iter > max_iter, more updates after max_iter was raised -- and keeps earlier diagnostics."""
    m = _imports()
    shape = (40, 136)
    N = shape[0] * shape[1]
    _, psf, y = synth.tv_deconv_problem(shape, None, 0.02, 3)
    F = (1 / 2) * m['SquaredL2Loss'](dim=N, data=y) * m['Convolve2D'](size=N, filter=psf, shape=shape)
    mk = lambda **kw: m['PDS'](dim=N, F=F, G=m['NonNegativeOrthant'](dim=N),
                               H=0.1 * m['L21Norm'](2 * N, np.tile(np.arange(N), 2)), K=tv_K(shape), beta=1.0,
                               verbose=None, output=output, **kw)
    fused, generic = mk(), mk(fused=False)
    got = list(fused.iterates(6))                       # no .clone(): the plan must not recycle yielded buffers
    ref = list(generic.iterates(6))
    assert type(fused._plan).__name__ == 'PdsTvPlan' and generic._plan is None
    host = (lambda t: t.cpu().numpy()) if output == 'torch' else (lambda a: a)
    for a, b in zip(got, ref):
        assert isinstance(a['primal_variable'], np.ndarray if output == 'numpy' else __import__('torch').Tensor)
        assert rel_l2(host(a['primal_variable']), host(b['primal_variable'])) < 1e-13
        assert rel_l2(host(a['dual_variable']), host(b['dual_variable'])) < 1e-13
    assert len({a['primal_variable'].ctypes.data if output == 'numpy' else a['primal_variable'].data_ptr() for a in got}) == 6
    # continuation
    s = mk(max_iter=5, min_iter=5, accuracy_threshold=0)
    est, _, d1 = s.iterate()
    x6 = host(est['primal_variable']).copy()
    assert s.iter == 6 and rel_l2(x6, host(got[5]['primal_variable'])) < 1e-13
    est_b, _, d1b = s.iterate()                          # loop condition already false: nothing happens
    assert s.iter == 6 and np.array_equal(host(est_b['primal_variable']), x6) and len(d1b) == 6
    s.max_iter = s.min_iter = 9
    est2, _, d2 = s.iterate()                            # four more updates, from update 6 on
    assert s.iter == 10 and len(d2) == 10
    assert np.array_equal(host(est['primal_variable']), x6)       # the first result was not overwritten
    full = mk(max_iter=9, min_iter=9, accuracy_threshold=0)
    estf, _, df = full.iterate()
    assert rel_l2(host(est2['primal_variable']), host(estf['primal_variable'])) < 1e-13
    assert np.allclose(d2.to_numpy(dtype=float), df.to_numpy(dtype=float), rtol=1e-9)


def _c4_solver(shape, mask, y, lam, eps, k, dtype='complex128', output='numpy', fused=True):
    """BASELINE.json configs[3] exactly as tests/golden/gen_golden.py:c4_csmri builds it with the reference classes."""
    from pycsou_b200.func.base import ProxFuncHStack
    from pycsou_b200.func.loss import L2BallLoss
    from pycsou_b200.func.penalty import L21Norm
    from pycsou_b200.linop.base import LinOpVStack
    from pycsou_b200.linop.sampling import Masking
    from pycsou_b200.linop.transform import DFT2
    from pycsou_b200.opt.proxalgs import CPS
    N = shape[0] * shape[1]
    M = Masking(size=N, sampling_bool=mask, dtype=dtype)
    m = M.shape[0]
    MF = M * DFT2(shape, dtype=dtype)
    MF.lipschitz_cst = MF.diff_lipschitz_cst = 1.0
    Kop = LinOpVStack(MF, tv_K(shape, dtype=dtype))
    H = ProxFuncHStack(L2BallLoss(dim=m, data=y, radius=eps), lam * L21Norm(2 * N, np.tile(np.arange(N), 2)))
    return CPS(dim=N, G=None, H=H, K=Kop, tau=1 / 3, sigma=1 / 3, rho=1, x0=np.zeros(N, dtype=dtype),
               z0=np.zeros(m + 2 * N, dtype=dtype), max_iter=k - 1, min_iter=k - 1, accuracy_threshold=0,
               verbose=None, output=output, fused=fused), m


def test_c4_csmri_cps_complex128():
    """Compressed-sensing MRI: K = [Mask * DFT ; Gradient], H = [L2-ball data term ; TV], ChambollePock (complex)."""
    meta, g = load_golden('c4_csmri_cps')
    shape = tuple(meta['shape'])
    mask, y = g['mask'].astype(bool), g['y']
    for k in (1, 5, 20):
        s, m = _c4_solver(shape, mask, y, meta['lam'], meta['eps'], k)
        assert m == meta['m']                                   # sampling mask bit-exact: same sample count and order
        est, _, _ = s.iterate()
        assert type(s._plan).__name__ == 'CpsMriPlan'           # the fused kernels of csrc/pcs_cps.cu ran
        assert np.iscomplexobj(est['primal_variable'])
        assert rel_l2(est['primal_variable'], g[f'x_k{k}']) < TOL64
        assert rel_l2(est['dual_variable'], g[f'z_k{k}']) < TOL64


def test_c4_csmri_cps_complex64_vs_oracle():
    shape = (64, 96)
    N = shape[0] * shape[1]
    rng = np.random.default_rng(11)
    truth = synth.phantom(shape).reshape(-1) * np.exp(1j * 2 * np.pi * (np.arange(N) / N))
    mask = rng.random(N) < 0.3
    m = int(mask.sum())
    y = np.fft.fft2(truth.reshape(shape), norm='ortho').reshape(-1)[mask] + 0.01 * (
        rng.standard_normal(m) + 1j * rng.standard_normal(m))
    lam, eps, k = 0.05, 0.01 * np.sqrt(2 * m), 30
    s, _ = _c4_solver(shape, mask, y.astype(np.complex64), lam, eps, k, dtype='complex64')
    est, _, _ = s.iterate()
    xo, zo, _, _ = C.c4(k, mask, y, shape=shape, lam=lam, eps=eps)
    assert est['primal_variable'].dtype == np.complex64
    assert rel_l2(est['primal_variable'], xo) < 2e-4 and rel_l2(est['dual_variable'], zo) < 2e-4


@pytest.mark.parametrize('dtype,tol', [('complex128', 1e-12), ('complex64', 2e-5)])
@pytest.mark.parametrize('shape', [(70, 200), (33, 129), (2, 2), (64, 64)])
def test_c4_fused_matches_op_by_op(shape, dtype, tol):
    """Fused CPS-MRI iteration (csrc/pcs_cps.cu) against the op-by-op path (itself pinned by the golden fixture):
    ragged tiles, tiny images, a data ball that is active (small eps) and one that is not (huge eps)."""
    N = shape[0] * shape[1]
    rng = np.random.default_rng(5)
    truth = synth.phantom(shape).reshape(-1) * np.exp(1j * 2 * np.pi * (np.arange(N) / N))
    mask = rng.random(N) < 0.35
    mask[0] = True
    m = int(mask.sum())
    y = (np.fft.fft2(truth.reshape(shape), norm='ortho').reshape(-1)[mask]
         + 0.01 * (rng.standard_normal(m) + 1j * rng.standard_normal(m))).astype(dtype)
    for eps in (0.01 * np.sqrt(2 * m), 1e6):
        for k in (1, 6):
            sf, _ = _c4_solver(shape, mask, y, 0.05, eps, k, dtype=dtype)
            ef, _, df = sf.iterate()
            assert type(sf._plan).__name__ == 'CpsMriPlan'
            sg, _ = _c4_solver(shape, mask, y, 0.05, eps, k, dtype=dtype, fused=False)
            eg, _, dg = sg.iterate()
            assert sg._plan is None
            assert rel_l2(ef['primal_variable'], eg['primal_variable']) < tol
            assert rel_l2(ef['dual_variable'], eg['dual_variable']) < tol
            for col in df.columns[1:]:
                assert np.allclose(np.asarray(df[col], float), np.asarray(dg[col], float), rtol=1e-6 if dtype == 'complex128' else 1e-3)


def test_c4_fused_with_momentum_and_steps():
    """rho != 1 and non-unit sampling steps through PrimalDualSplitting directly."""
    from pycsou_b200.func.base import ProxFuncHStack
    from pycsou_b200.func.loss import L2BallLoss
    from pycsou_b200.func.penalty import L21Norm
    from pycsou_b200.linop.base import LinOpVStack
    from pycsou_b200.linop.diff import Gradient
    from pycsou_b200.linop.sampling import Masking
    from pycsou_b200.linop.transform import DFT2
    from pycsou_b200.opt.proxalgs import PDS
    shape, dtype = (40, 72), 'complex128'
    N = shape[0] * shape[1]
    rng = np.random.default_rng(8)
    mask = rng.random(N) < 0.4
    m = int(mask.sum())
    y = rng.standard_normal(m) + 1j * rng.standard_normal(m)
    x0 = rng.standard_normal(N) + 1j * rng.standard_normal(N)
    z0 = rng.standard_normal(m + 2 * N) + 1j * rng.standard_normal(m + 2 * N)
    out = {}
    for fused in (True, False):
        M = Masking(size=N, sampling_bool=mask, dtype=dtype)
        MF = M * DFT2(shape, dtype=dtype)
        MF.lipschitz_cst = MF.diff_lipschitz_cst = 1.0
        Gr = Gradient(shape=shape, kind='forward', step=(0.5, 2.0), dtype=dtype)
        Gr.lipschitz_cst = Gr.diff_lipschitz_cst = SQRT8 / 0.5
        Kop = LinOpVStack(MF, Gr)
        H = ProxFuncHStack(L2BallLoss(dim=m, data=y, radius=3.0), 0.2 * L21Norm(2 * N, np.tile(np.arange(N), 2)))
        s = PDS(dim=N, F=None, G=None, H=H, K=Kop, tau=0.1, sigma=0.2, rho=0.8, x0=x0, z0=z0, max_iter=4, min_iter=4,
                accuracy_threshold=0, verbose=None, fused=fused)
        est, _, _ = s.iterate()
        assert (type(s._plan).__name__ == 'CpsMriPlan') == fused
        out[fused] = est
    assert rel_l2(out[True]['primal_variable'], out[False]['primal_variable']) < 1e-12
    assert rel_l2(out[True]['dual_variable'], out[False]['dual_variable']) < 1e-12


def test_checkpoint_resume_is_bit_identical(tmp_path):
    """save_checkpoint / load_checkpoint: 8 + 12 updates through a checkpoint file == 20 updates in one go."""
    from pycsou_b200.func.loss import SquaredL2Loss
    from pycsou_b200.func.penalty import L21Norm, NonNegativeOrthant
    from pycsou_b200.linop.conv import Convolve2D
    from pycsou_b200.opt.proxalgs import PDS
    shape = (60, 140)
    _, psf, y = synth.tv_deconv_problem(shape, None, 0.02, 3)
    N = y.size

    def mk(k):
        F = (1 / 2) * SquaredL2Loss(dim=N, data=y) * Convolve2D(size=N, filter=psf, shape=shape)
        return PDS(dim=N, F=F, G=NonNegativeOrthant(dim=N), H=0.02 * L21Norm(dim=2 * N, groups=np.tile(np.arange(N), 2)),
                   K=tv_K(shape), beta=1.0, max_iter=k - 1, min_iter=k - 1, accuracy_threshold=0, verbose=None)
    full, _, _ = mk(20).iterate()
    a = mk(8)
    a.iterate()
    a.save_checkpoint(str(tmp_path / 'ck.pt'))
    b = mk(12)
    assert b.load_checkpoint(str(tmp_path / 'ck.pt')) == 8
    res, _, _ = b.iterate()
    assert np.array_equal(res['primal_variable'], full['primal_variable'])
    assert np.array_equal(res['dual_variable'], full['dual_variable'])
