"""The benchmark problems built from the REFERENCE's own classes (``pycsou.opt.proxalgs`` solvers, ``pycsou.func``
functionals, ``pycsou.core`` composition rules), imported unmodified through ``oracle.ref_shim``.

TEST / BASELINE INFRASTRUCTURE ONLY (see ``oracle/__init__.py``): used by ``bench.py --impl reference``, by the
``cpu_baseline`` leg of ``bench.py`` and by the ``needs_reference`` tests.  The reference tree is found in
``/root/reference`` (build container) or in ``oracle/_ref`` / ``baseline/_ref`` (a copy made by ``oracle/make_ref.py``;
git-ignored, it travels to the GPU box with the snapshot).  What is NOT the reference's own code on this path, because
the reference does not contain it: the pylops operators (``oracle.pylops_standin``, un-vendored dependency) and the
per-pixel-group L21 (``ref_shim.vectorised_l21``; the reference's Python loop is Theta(N^2), 1.4e4 s per call at 1024^2).
"""
from __future__ import annotations

import time

import numpy as np


def c3_reference_solver(y, psf, shape, lam, iters):
    """``PrimalDualSplitting`` of the reference for BASELINE.json configs[2] (``doc/general/examples.rst:19-47`` pattern,
    ``opt/proxalgs.py:126-244``): exactly ``iters`` updates (``core/solver.py:65-66``)."""
    from oracle.ref_shim import load_reference, vectorised_l21
    load_reference()
    from pycsou.func.loss import SquaredL2Loss
    from pycsou.func.penalty import NonNegativeOrthant
    from pycsou.linop.conv import Convolve2D
    from pycsou.linop.diff import Gradient
    from pycsou.opt.proxalgs import PDS
    N = int(shape[0] * shape[1])
    A = Convolve2D(size=N, filter=psf, shape=tuple(shape))
    K = Gradient(shape=tuple(shape), kind='forward')
    K.lipschitz_cst = K.diff_lipschitz_cst = float(np.sqrt(8.))
    F = (1 / 2) * SquaredL2Loss(dim=N, data=y) * A
    return PDS(dim=N, F=F, G=NonNegativeOrthant(dim=N), H=lam * vectorised_l21(2 * N, N), K=K, beta=1.0,
               max_iter=iters - 1, min_iter=iters - 1, accuracy_threshold=0, verbose=None)


def c3_reference_sample(y, psf, shape, lam, iters):
    """Run ``iters`` updates on the reference solver; returns (seconds, x, z)."""
    s = c3_reference_solver(np.asarray(y, dtype=np.float64), psf, shape, lam, iters)
    t0 = time.perf_counter()
    est, _, _ = s.iterate()
    dt = time.perf_counter() - t0
    assert s.iter == iters
    return dt, np.asarray(est['primal_variable']), np.asarray(est['dual_variable'])
