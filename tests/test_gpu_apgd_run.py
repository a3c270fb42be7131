"""GPU tests of the one-launch APGD solve (csrc/pcs_apgd.cu, planned by opt/_fused.py:plan_apgd) against the numpy oracle
and against the op-by-op path: iterates, the reference's stop test (number of updates), diagnostics, every acceleration
scheme, every supported G, filters with even length, sizes that are not multiples of the block, continuation."""
import time

import numpy as np
import pytest

from oracle import np_oracle as O
from pycsou_b200.util import synth
from tests.helpers import rel_l2

pytestmark = pytest.mark.gpu


def _build(n, nh, gname, acc, dtype='float64', seed=0, **kw):
    from pycsou_b200.func.loss import SquaredL2Loss
    from pycsou_b200.func.penalty import L1Norm, NonNegativeOrthant
    from pycsou_b200.linop.conv import Convolve1D
    from pycsou_b200.opt.proxalgs import APGD
    rng = np.random.default_rng(seed)
    truth = np.zeros(n)
    truth[rng.choice(n, max(n // 128, 3), replace=False)] = rng.standard_normal(max(n // 128, 3))
    h = synth.gaussian_psf_1d(nh, nh / 8.)
    A = Convolve1D(size=n, filter=h, dtype=dtype)
    y = np.asarray(A(truth)) + 0.01 * rng.standard_normal(n)
    lam = 0.1 * float(np.max(np.abs(np.asarray(A.adjoint(y)))))
    G = {'l1': lambda: lam * L1Norm(dim=n), 'nonneg': lambda: NonNegativeOrthant(dim=n), 'none': lambda: None}[gname]()
    F = (1 / 2) * SquaredL2Loss(dim=n, data=y) * A
    s = APGD(dim=n, F=F, G=G, tau=1.0, acceleration=acc, beta=1.0, verbose=None, dtype=dtype, **kw)
    Go = {'l1': lambda: lam * O.L1Norm(n), 'nonneg': lambda: O.NonNegativeOrthant(n), 'none': lambda: None}[gname]()
    Fo = O.HalfSquaredL2Loss(y, O.Convolve1D(n, h), G_lipschitz=1.0)
    return s, Fo, Go


@pytest.mark.parametrize('n,nh,gname,acc', [(4096, 65, 'l1', 'CD'), (4096, 65, 'l1', 'BT'), (1000, 33, 'nonneg', None),
                                            (5000, 64, 'l1', 'CD'), (257, 9, 'none', 'BT'), (8192, 129, 'l1', 'CD')])
def test_run_matches_oracle_and_counts_updates(n, nh, gname, acc):
    kw = dict(max_iter=300, min_iter=5, accuracy_threshold=3e-3)
    s, Fo, Go = _build(n, nh, gname, acc, **kw)
    est, conv, diag = s.iterate()
    assert type(s._plan).__name__ == 'ApgdConv1dPlan'
    xo, auxo, to, diago, no = O.apgd_iterate(n, F=Fo, G=Go, tau=1.0, acceleration=acc, beta=1.0, **kw)
    assert s.iter == no and 6 < no <= 301                       # the stop test inside the kernel is the reference's
    assert rel_l2(est['iterand'], xo) < 1e-10 and rel_l2(est['past_aux'], auxo) < 1e-10
    assert np.isclose(est['past_t'], to, rtol=1e-15)
    ri = np.asarray(diag['Relative Improvement'], dtype=float)
    assert len(ri) == no and ri[0] == np.inf
    assert np.allclose(ri[1:], np.asarray(diago['Relative Improvement'])[1:], rtol=1e-7)


def test_run_equals_op_by_op_and_is_one_launch():
    from pycsou_b200 import _lib
    kw = dict(max_iter=499, min_iter=499, accuracy_threshold=0)
    s, _, _ = _build(4096, 65, 'l1', 'CD', **kw)
    s.iterate()
    import torch
    torch.cuda.synchronize()
    c0, t0 = _lib.launch_count, time.perf_counter()
    est, _, _ = s.reset() or s.iterate()
    torch.cuda.synchronize()
    dt_fused, launches = time.perf_counter() - t0, _lib.launch_count - c0
    assert launches == 1 and s.iter == 500
    g, _, _ = _build(4096, 65, 'l1', 'CD', fused=False, **kw)
    g.iterate()
    g.reset()
    t0 = time.perf_counter()
    ref, _, _ = g.iterate()
    dt_generic = time.perf_counter() - t0
    assert rel_l2(est['iterand'], ref['iterand']) < 1e-11
    print(f'APGD 500 updates, N=4096 fp64: one launch {1e3 * dt_fused:.2f} ms, op by op {1e3 * dt_generic:.2f} ms')
    assert dt_fused < dt_generic


def test_run_fp32_continuation_and_verbose_path():
    kw = dict(max_iter=59, min_iter=59, accuracy_threshold=0)
    s, Fo, Go = _build(2048, 33, 'l1', 'CD', dtype='float32', **kw)
    est, _, _ = s.iterate()
    xo, *_ = O.apgd_iterate(2048, F=Fo, G=Go, tau=1.0, acceleration='CD', beta=1.0, **kw)
    assert est['iterand'].dtype == np.float32 and rel_l2(est['iterand'], xo) < 2e-4
    # stepping one update at a time (iterates generator -> single-update launches) gives the same iterates
    s2, _, _ = _build(2048, 33, 'l1', 'CD', **kw)
    last = None
    for last in s2.iterates(60):
        pass
    s3, _, _ = _build(2048, 33, 'l1', 'CD', **kw)
    full, _, _ = s3.iterate()
    assert isinstance(last['iterand'], np.ndarray) and rel_l2(last['iterand'], full['iterand']) < 1e-13


def _solve(cluster, *a, **kw):
    import os
    from pycsou_b200.opt._fused import apgd_last_launch
    old = os.environ.get('PCS_APGD_CLUSTER')
    os.environ['PCS_APGD_CLUSTER'] = '1' if cluster else '0'
    try:
        s, Fo, Go = _build(*a, **kw)
        est, _, diag = s.iterate()
        return s, est, np.asarray(diag['Relative Improvement'], dtype=float), apgd_last_launch(), (Fo, Go)
    finally:
        if old is None:
            del os.environ['PCS_APGD_CLUSTER']
        else:
            os.environ['PCS_APGD_CLUSTER'] = old


@pytest.mark.parametrize('n,nh,gname,acc', [(4096, 65, 'l1', 'CD'), (4096, 64, 'l1', 'BT'), (2048, 33, 'nonneg', None),
                                            (2049, 9, 'none', 'BT'), (8191, 255, 'l1', 'CD'), (3000, 1, 'l1', 'CD'),
                                            (6000, 200, 'nonneg', 'CD'), (8192, 129, 'l1', 'CD'), (512, 33, 'l1', 'CD'),
                                            (600, 255, 'l1', 'BT'), (1024, 64, 'nonneg', 'CD'), (1500, 129, 'l1', 'CD')])
@pytest.mark.parametrize('dtype', ['float64', 'float32'])
def test_cluster_kernel_equals_single_block_kernel(n, nh, gname, acc, dtype):
    """The thread-block-cluster kernel (n >= 512) against the single-block one: same number of updates under the live
    stop test (the cluster kernel starts an update before the sums of the previous one are in and drops it if they say
    stop), same diagnostics, iterates equal up to the order in which the taps are added."""
    kw = dict(max_iter=300, min_iter=5, accuracy_threshold=3e-3, dtype=dtype)
    s0, e0, r0, L0, _ = _solve(False, n, nh, gname, acc, **kw)
    s1, e1, r1, L1, _ = _solve(True, n, nh, gname, acc, **kw)
    assert L0['kernel'] == 'apgd_run_kernel' and L0['blocks'] == 1
    assert L1['kernel'] == 'apgd_cluster_kernel' and L1['threads'] == 160
    assert L1['blocks'] in (2, 4, 8, 16) and L1['blocks'] * 128 * L1['per_thread'] >= n          # the smallest cluster that holds n
    assert s0.iter == s1.iter and e0['past_t'] == e1['past_t'] and len(r0) == len(r1)
    tol = 1e-12 if dtype == 'float64' else 1e-4
    assert rel_l2(e1['iterand'], e0['iterand']) < tol and rel_l2(e1['past_aux'], e0['past_aux']) < tol
    assert np.allclose(r1[1:], r0[1:], rtol=1e-9 if dtype == 'float64' else 1e-2)


@pytest.mark.parametrize('kw', [dict(max_iter=40, min_iter=0, accuracy_threshold=0.0),        # runs into max_iter
                                dict(max_iter=300, min_iter=0, accuracy_threshold=5e-2),     # stops early: an update is dropped
                                dict(max_iter=3, min_iter=25, accuracy_threshold=1e9),       # min_iter forces updates past the test
                                dict(max_iter=300, min_iter=0, accuracy_threshold=np.inf),   # inf > inf is false: no update at all
                                dict(max_iter=0, min_iter=0, accuracy_threshold=0.0)])       # exactly one update
def test_cluster_kernel_loop_condition(kw):
    """core/solver.py:65-66 in the cluster kernel: every way the loop can end, against the oracle."""
    s, est, ri, L, (Fo, Go) = _solve(True, 4096, 65, 'l1', 'CD', **kw)
    xo, auxo, to, diago, no = O.apgd_iterate(4096, F=Fo, G=Go, tau=1.0, acceleration='CD', beta=1.0, **kw)
    assert s.iter == no
    if no > 0:
        assert L['kernel'] == 'apgd_cluster_kernel'
    assert rel_l2(est['iterand'], xo) < 1e-10 if np.linalg.norm(xo) else not np.any(est['iterand'])
    assert np.isclose(est['past_t'], to, rtol=1e-15)
    rio = np.asarray(diago['Relative Improvement'], dtype=float)
    assert len(ri) == len(rio) and np.allclose(ri[1:], rio[1:], rtol=1e-7)
