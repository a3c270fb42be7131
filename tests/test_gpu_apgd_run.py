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
