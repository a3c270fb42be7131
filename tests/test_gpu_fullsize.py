"""GPU parity AT THE BENCHMARKED SIZES and on the benchmarked launch geometries (BASELINE.json configs 2-4).

The small-image tests never reach the code the headline numbers come from: at 4096^2 the fused Condat-Vu kernel runs a
ONE-WAVE grid of two CTAs per SM (33 strips x 9 segments), fp32 deconvolution runs the warp-specialised kernel with the
``%smid`` companion search and the uneven long/short segment split, and 1024^2 runs through the multi-update entry
point.  Here the same problems ``bench.py`` times are solved for a few updates on the numpy oracle (the reference's
algorithm, ``/root/reference/pycsou/opt/proxalgs.py:343-355`` driven by ``core/solver.py:64-76``) and on the GPU:

* fp64 iterates within 1e-10 relative L2 (north_star); fp32 objective within 1e-4 relative; masks bit-exact;
* the launch that ran is asserted to be the benchmarked one (``pcs_pds_tv_last_launch``).

``pcs_set_num_sms`` additionally drives the one-wave / uneven-split geometry on a small image, where it must not change
a single bit of the fp64 iterates.
"""
import numpy as np
import pytest

from oracle import np_oracle as O
from pycsou_b200.util import synth
from tests.helpers import rel_l2

pytestmark = pytest.mark.gpu
SQRT8 = np.sqrt(8.)
LAM3 = 0.02


def _mods():
    from pycsou_b200.func.loss import SquaredL2Loss
    from pycsou_b200.func.penalty import L21Norm, NonNegativeOrthant, tv_groups
    from pycsou_b200.linop.conv import Convolve2D
    from pycsou_b200.linop.diff import Gradient
    from pycsou_b200.opt._fused import last_launch
    from pycsou_b200.opt.proxalgs import PDS
    return locals()


def _tv_K(m, shape, dtype):
    K = m['Gradient'](shape=shape, kind='forward', dtype=dtype)
    K.lipschitz_cst = K.diff_lipschitz_cst = SQRT8
    return K


# ---------------------------------------------------------------------------------------------- C3 / north star: 4096^2
@pytest.fixture(scope='module')
def c3_oracle():
    """Five Condat-Vu updates of the 4096^2 TV-deconvolution of bench.py on the numpy oracle (fp64; ~3 s per update)."""
    size, k = 4096, 5
    shape, N = (size, size), size * size
    _, psf, y = synth.tv_deconv_problem(shape, synth.gaussian_psf_2d(15, 2.5), 0.02, seed=0)
    A = O.Convolve2D(N, psf, shape)
    Fo = O.HalfSquaredL2Loss(y, A, G_lipschitz=1.0)
    Ko = O.Gradient(shape, kind='forward')
    xo, zo, diag, n = O.pds_iterate(N, F=Fo, G=O.NonNegativeOrthant(N), H=LAM3 * O.TVL21Norm(2 * N, N), K=Ko,
                                    K_lipschitz=SQRT8, beta=1.0, max_iter=k - 1, min_iter=k - 1, accuracy_threshold=0)

    def objective(x):
        return 0.5 * np.linalg.norm(A.matvec(x) - y) ** 2 + LAM3 * O.TVL21Norm(2 * N, N)(Ko.matvec(x))
    return dict(shape=shape, N=N, k=k, psf=psf, y=y, x=xo, z=zo, diag=diag, objective=objective)


def _c3_solver(m, c, dtype, **kw):
    N, shape = c['N'], c['shape']
    A = m['Convolve2D'](size=N, filter=c['psf'], shape=shape, dtype=dtype)
    F = (1 / 2) * m['SquaredL2Loss'](dim=N, data=c['y'].astype(dtype)) * A
    H = LAM3 * m['L21Norm'](dim=2 * N, groups=m['tv_groups'](N))
    return m['PDS'](dim=N, F=F, G=m['NonNegativeOrthant'](dim=N), H=H, K=_tv_K(m, shape, dtype), beta=1.0,
                    max_iter=c['k'] - 1, min_iter=c['k'] - 1, accuracy_threshold=0, verbose=None, dtype=dtype, **kw)


def test_c3_north_star_fp64_4096(c3_oracle):
    m, c = _mods(), c3_oracle
    s = _c3_solver(m, c, 'float64')
    est, _, diag = s.iterate()
    L = m['last_launch']()
    # the benchmarked fp64 launch: single-role kernel, TMA staging, 33 strips of 127 columns, one wave of <= 296 CTAs
    assert type(s._plan).__name__ == 'PdsTvPlan' and s._plan.data_mode == 2
    assert L['kernel'] == 'pds_tv_kernel' and L['use_tma'] == 1 and L['strips'] == 33 and L['strip_cols'] == 126
    assert 148 < L['blocks'] <= 296
    ex, ez = rel_l2(est['primal_variable'], c['x']), rel_l2(est['dual_variable'], c['z'])
    print(f'4096^2 fp64, {c["k"]} updates: rel-L2 x {ex:.2e}, z {ez:.2e}, launch {L}')
    assert ex <= 1e-10 and ez <= 1e-10
    for col in ('Relative Improvement (primal variable)', 'Relative Improvement (dual variable)'):
        assert np.allclose(np.asarray(diag[col], float)[1:], np.asarray(c['diag'][col], float)[1:], rtol=1e-9)


def test_c3_fp32_4096_warp_specialised_uneven_split(c3_oracle):
    m, c = _mods(), c3_oracle
    s = _c3_solver(m, c, 'float32')
    est, _, _ = s.iterate()
    L = m['last_launch']()
    # the benchmarked fp32 launch: warp-specialised kernel, one wave, %smid companion search + uneven long/short split
    assert L['kernel'] == 'pds_tv_ws3_kernel' and L['use_tma'] == 1 and L['threads'] == 512
    assert L['n_fast'] >= 0 and 148 < L['blocks'] <= 296
    x = est['primal_variable'].astype(np.float64)
    fo, fg = c['objective'](c['x']), c['objective'](x)
    ex, ez = rel_l2(x, c['x']), rel_l2(est['dual_variable'], c['z'])
    print(f'4096^2 fp32, {c["k"]} updates: objective rel {abs(fg - fo) / fo:.2e}, rel-L2 x {ex:.2e}, z {ez:.2e}, launch {L}')
    assert abs(fg - fo) <= 1e-4 * fo                  # north_star: fp32 within 1e-4 relative error in the objective
    assert ex <= 2e-5 and ez <= 2e-5
    # the warp-specialised kernel against the single-role one at full size (same arithmetic, other schedule)
    import os
    os.environ['PCS_PDS_TMA'] = '3'
    try:
        s2 = _c3_solver(m, c, 'float32')
        est2, _, _ = s2.iterate()
        assert m['last_launch']()['kernel'] == 'pds_tv_kernel'
    finally:
        del os.environ['PCS_PDS_TMA']
    assert rel_l2(est2['primal_variable'], est['primal_variable']) <= 2e-6


# ---------------------------------------------------------------------------------------------- C2: 1024^2 denoising
@pytest.mark.parametrize('dtype', ['float64', 'float32'])
def test_c2_tv_denoise_1024(dtype):
    m = _mods()
    shape, k, lam = (1024, 1024), 20, 0.1
    N = shape[0] * shape[1]
    _, y = synth.tv_denoise_problem(shape, 0.1, 0)
    Ko = O.Gradient(shape, kind='forward')
    xo, zo, dref, _ = O.pds_iterate(N, F=O.HalfSquaredL2Loss(y), H=lam * O.TVL21Norm(2 * N, N), K=Ko, K_lipschitz=SQRT8,
                                    rho=0.9, max_iter=k - 1, min_iter=k - 1, accuracy_threshold=0)
    F = (1 / 2) * m['SquaredL2Loss'](dim=N, data=y.astype(dtype))
    s = m['PDS'](dim=N, F=F, H=lam * m['L21Norm'](dim=2 * N, groups=m['tv_groups'](N)), K=_tv_K(m, shape, dtype),
                 rho=0.9, max_iter=k - 1, min_iter=k - 1, accuracy_threshold=0, verbose=None, dtype=dtype)
    assert abs(s.tau - 0.3) < 0.05 and s.tau == s.sigma            # proxalgs.py:286-287 with beta = 1, L = sqrt(8)
    est, _, diag = s.iterate()
    L = m['last_launch']()
    # x and z of a 1024^2 image fit in the SMs' shared memory: the 20 updates run on chip (one slab of 7 rows per SM)
    assert L['kernel'] == 'pds_tv_resident_kernel' and L['blocks'] == 147 and L['threads'] == 512
    ex, ez = rel_l2(est['primal_variable'], xo), rel_l2(est['dual_variable'], zo)
    print(f'1024^2 {dtype}, {k} updates: rel-L2 x {ex:.2e}, z {ez:.2e}, launch {L}')
    if dtype == 'float64':
        assert ex <= 1e-10 and ez <= 1e-10
        col = 'Relative Improvement (primal variable)'
        assert np.allclose(np.asarray(diag[col], float)[1:], np.asarray(dref[col], float)[1:], rtol=1e-9)
    else:
        def obj(x):
            return 0.5 * np.linalg.norm(x - y) ** 2 + lam * O.TVL21Norm(2 * N, N)(Ko.matvec(x))
        fo, fg = obj(xo), obj(est['primal_variable'].astype(np.float64))
        assert abs(fg - fo) <= 1e-4 * fo and ex <= 2e-5 and ez <= 2e-5


# ---------------------------------------------------------------------------------------------- C4: 2048^2 complex64
def test_c4_csmri_2048_complex64():
    import bench
    import torch
    size, k = 2048, 5
    N = size * size
    mask, y, eps = bench.c4_problem(size, 0)
    xo, zo, _, _ = __import__('tests.oracle_cases', fromlist=['c4']).c4(k, mask, y, shape=(size, size), lam=bench.C4_LAM,
                                                                      eps=eps)
    s = bench.c4_solver(size, mask, y.astype(np.complex64), eps, k, 'numpy')
    est, _, _ = s.iterate()
    assert type(s._plan).__name__ == 'CpsMriPlan'
    # sampling indices are bit-exact: the device index list of the Masking operator == numpy's flatnonzero
    Mop = s.K.linops[0].LinOp1
    assert np.array_equal(Mop._idx(torch.device('cuda')).cpu().numpy(), np.flatnonzero(mask))
    ex, ez = rel_l2(est['primal_variable'], xo), rel_l2(est['dual_variable'], zo)
    print(f'2048^2 complex64 CS-MRI, {k} updates: rel-L2 x {ex:.2e}, z {ez:.2e}')
    assert est['primal_variable'].dtype == np.complex64
    assert ex <= 2e-5 and ez <= 2e-5


# ---------------------------------------------------------------------------------------------- forced geometries
@pytest.mark.parametrize('dtype', ['float64', 'float32'])
@pytest.mark.parametrize('nsm', [3, 4, 7])
def test_one_wave_uneven_split_geometry_on_a_small_image(dtype, nsm, monkeypatch):
    """With the SM count overridden (``pcs_set_num_sms``) the planner builds the ONE-WAVE grid of the 4096^2 runs
    (nsm < blocks <= 2 nsm; strips with one segment more than the others) on a 300 x 600 image, and -- for the
    warp-specialised fp32 kernel -- the uneven long/short split with the ``pds_assign_rows`` companion search.  Neither
    may change a single bit of the iterates relative to the default (many short segments) geometry."""
    from pycsou_b200 import _lib
    from tests.test_gpu_fused import _problem, _run
    m = _mods()
    kw = _problem((300, 600), 'sep', 'nonneg', 'l21', dtype)
    _, base, _ = _run(kw, 4, True)
    L0 = m['last_launch']()
    prev = _lib.call_plain('pcs_set_num_sms', nsm)
    try:
        _, est, _ = _run(kw, 4, True)
        L = m['last_launch']()
    finally:
        _lib.call_plain('pcs_set_num_sms', prev)
    assert L['kernel'] == L0['kernel'] and L['blocks'] != L0['blocks'] and L['blocks'] <= 2 * nsm
    if dtype == 'float32':
        assert L['kernel'] == 'pds_tv_ws3_kernel'
        if L['blocks'] > nsm:
            assert L['n_fast'] >= 0                       # uneven split + companion search were exercised
    assert np.array_equal(est['primal_variable'], base['primal_variable'])
    assert np.array_equal(est['dual_variable'], base['dual_variable'])


@pytest.mark.parametrize('dtype', ['float64', 'float32'])
@pytest.mark.parametrize('shape,data', [((1024, 1024), 'identity'), ((300, 520), 'none'), ((64, 136), 'identity')])
def test_persistent_multi_update_kernel_is_bit_identical(shape, data, dtype, monkeypatch):
    """pcs_pds_tv_run: all updates in one cooperative launch (grid barrier between updates) vs one launch per update --
    identical iterates and identical diagnostics rows, for odd and even update counts."""
    from tests.test_gpu_fused import _problem, _run
    m = _mods()
    kw = _problem(shape, data, 'nonneg', 'l21', dtype)
    monkeypatch.setenv('PCS_PDS_RESIDENT', '0')             # (the on-chip kernel would take these images: tested below)
    for k in (2, 7):
        monkeypatch.setenv('PCS_PDS_PERSIST', '0')
        _, ref, dref = _run(kw, k, True)
        assert m['last_launch']()['kernel'] == 'pds_tv_kernel'
        monkeypatch.setenv('PCS_PDS_PERSIST', '1')
        _, est, dest = _run(kw, k, True)
        assert m['last_launch']()['kernel'] == 'pds_tv_persist_kernel'
        assert np.array_equal(est['primal_variable'], ref['primal_variable'])
        assert np.array_equal(est['dual_variable'], ref['dual_variable'])
        assert np.array_equal(dest.to_numpy(dtype=float), dref.to_numpy(dtype=float))


@pytest.mark.parametrize('dtype', ['float64', 'float32'])
@pytest.mark.parametrize('shape,data,g,h', [((1024, 1024), 'identity', 'nonneg', 'l21'), ((300, 520), 'none', 'nonneg', 'l21'),
                                            ((64, 136), 'identity', 'segment', 'l1'), ((1000, 700), 'identity', 'none', 'l21'),
                                            ((5, 64), 'identity', 'nonneg', 'l21'), ((1036, 34), 'identity', 'l1', 'l21'),
                                            ((149, 1024), 'identity', 'nonneg', 'l1')])
def test_on_chip_resident_kernel_is_bit_identical(shape, data, g, h, dtype, monkeypatch):
    """pcs_pds_tv_run on images whose x and z fit in shared memory (pds_tv_resident_kernel: one slab of rows per SM kept on
    chip for all the updates, neighbours hand over boundary rows through global memory) vs one launch per update:
    identical iterates for 2, 3 and 8 updates; the diagnostics sums agree to rounding (other summation order).  Shapes:
    full slabs, a short last slab, one row per block, the tallest image that fits (fp64: 148 x 7 rows), two rows per slab."""
    from tests.test_gpu_fused import _problem, _run
    m = _mods()
    if dtype == 'float64' and shape[0] > 148 * 7:
        pytest.skip('taller than 148 slabs of 7 rows')
    kw = _problem(shape, data, g, h, dtype)
    for k in (2, 3, 8):
        monkeypatch.setenv('PCS_PDS_RESIDENT', '0')
        monkeypatch.setenv('PCS_PDS_PERSIST', '0')
        _, ref, dref = _run(kw, k, True)
        assert m['last_launch']()['kernel'] == 'pds_tv_kernel'
        monkeypatch.setenv('PCS_PDS_RESIDENT', '1')
        _, est, dest = _run(kw, k, True)
        assert m['last_launch']()['kernel'] == 'pds_tv_resident_kernel'
        assert np.array_equal(est['primal_variable'], ref['primal_variable'])
        assert np.array_equal(est['dual_variable'], ref['dual_variable'])
        a, b = dest.to_numpy(dtype=float), dref.to_numpy(dtype=float)
        assert np.allclose(a, b, rtol=1e-11 if dtype == 'float64' else 2e-5, atol=0, equal_nan=True)


def test_on_chip_resident_kernel_many_updates_in_chunks(monkeypatch):
    """More updates than one launch of the resident kernel has room for in its per-update partial sums (65536 / 147 = 445 at
    1024^2): the call is split into launches, each starting from the buffers the previous one left."""
    from tests.test_gpu_fused import _problem, _run
    m = _mods()
    kw = _problem((1024, 1024), 'identity', 'nonneg', 'l21', 'float32')
    monkeypatch.setenv('PCS_PDS_RESIDENT', '0')
    monkeypatch.setenv('PCS_PDS_PERSIST', '0')
    _, ref, dref = _run(kw, 451, True)
    monkeypatch.setenv('PCS_PDS_RESIDENT', '1')
    _, est, dest = _run(kw, 451, True)
    assert m['last_launch']()['kernel'] == 'pds_tv_resident_kernel'
    assert np.array_equal(est['primal_variable'], ref['primal_variable'])
    assert np.array_equal(est['dual_variable'], ref['dual_variable'])
    assert np.allclose(dest.to_numpy(dtype=float), dref.to_numpy(dtype=float), rtol=2e-5, atol=0, equal_nan=True)


@pytest.mark.parametrize('shape', [(4096, 4096), (300, 600), (72, 256), (40, 264)])
def test_three_role_fp32_kernel_is_bit_identical_to_the_two_role_one(shape, monkeypatch):
    """pds_tv_ws3_kernel (row-pass, column-pass and pointwise warps as three roles, ring with a spare box, mbarrier
    hand-over of ring boxes, setmaxnreg) against pds_tv_ws_kernel: identical iterates, full size and edge-strip shapes."""
    from tests.test_gpu_fused import _problem, _run
    m = _mods()
    kw = _problem(shape, 'sep', 'nonneg', 'l21', 'float32')
    k = 4 if shape[0] > 1000 else 9
    monkeypatch.setenv('PCS_PDS_WS3', '0')
    _, ref, _ = _run(kw, k, True)
    assert m['last_launch']()['kernel'] == 'pds_tv_ws_kernel'
    monkeypatch.setenv('PCS_PDS_WS3', '1')
    _, est, _ = _run(kw, k, True)
    assert m['last_launch']()['kernel'] == 'pds_tv_ws3_kernel'
    assert np.array_equal(est['primal_variable'], ref['primal_variable'])
    assert np.array_equal(est['dual_variable'], ref['dual_variable'])
