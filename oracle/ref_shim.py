"""Import the UNMODIFIED reference ``pycsou`` in the build container.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Used by ``tests/golden/gen_golden.py`` and
by the ``needs_reference`` tests to pin ``oracle.np_oracle``, and by ``oracle/ref_problems.py`` (the CPU reference arm of
``bench.py``).  On the GPU box ``/root/reference`` does not exist; the archive ``oracle/_ref/pycsou_reference.zip``
(``oracle/make_ref.py``; git-ignored, shipped with the snapshot) is used there.

What the shim does:

1. restores the four numpy aliases the reference still uses and numpy 2 removed
   (``np.infty``, ``np.float``, ``np.alltrue``, ``np.asscalar``);
2. registers stand-in modules for absent third-party packages that the reference imports at module
   top level: ``pylops`` (+ ``pylops.signalprocessing``, ``pylops.optimization.leastsquares``) backed
   by ``oracle.pylops_standin``; empty ``dask`` / ``dask.array`` and ``skimage.measure`` stubs;
3. puts the reference tree on ``sys.path`` and returns the imported ``pycsou`` package.

It also offers ``VectorisedL21Norm``: the reference's ``L21Norm.prox`` loops over groups in Python
with a full boolean mask per group (func/penalty.py:551-560), which is Theta(N^2) for TV; the
vectorised subclass computes the same numbers for the ``groups = tile(arange(N), 2)`` layout.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np

_REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# the live tree (build container); the archive packed by oracle/make_ref.py (travels to the GPU box; imported through
# zipimport); a driver-made install
REFERENCE_ROOTS = ('/root/reference', os.path.join(_REPO, 'oracle', '_ref', 'pycsou_reference.zip'),
                   os.path.join(_REPO, 'baseline', '_ref'))


def _has_reference(root: str) -> bool:
    if os.path.isdir(root):
        return os.path.isdir(os.path.join(root, 'pycsou', 'opt'))
    if os.path.isfile(root) and root.endswith('.zip'):
        import zipfile
        try:
            with zipfile.ZipFile(root) as z:
                return 'pycsou/opt/proxalgs.py' in z.namelist()
        except zipfile.BadZipFile:
            return False
    return False


def reference_root():
    roots = REFERENCE_ROOTS
    if os.environ.get('PCS_REFERENCE_ROOT'):            # tests: force one root (e.g. the archive while the tree exists)
        roots = (os.environ['PCS_REFERENCE_ROOT'],)
    return next((r for r in roots if _has_reference(r)), None)


def reference_available() -> bool:
    return reference_root() is not None


def _restore_numpy_aliases():
    if not hasattr(np, 'infty'):
        np.infty = np.inf
    if not hasattr(np, 'float'):
        np.float = float
    if not hasattr(np, 'alltrue'):
        np.alltrue = np.all
    if not hasattr(np, 'asscalar'):
        np.asscalar = lambda a: np.asarray(a).item()


def _register_standins():
    from oracle import pylops_standin as ps

    if 'pylops' not in sys.modules:
        pylops = types.ModuleType('pylops')
        for name in ('LinearOperator', 'FirstDerivative', 'SecondDerivative', 'Gradient', 'Laplacian',
                     'Restriction', 'VStack'):
            setattr(pylops, name, getattr(ps, name))
        sp = types.ModuleType('pylops.signalprocessing')
        sp.Convolve1D, sp.Convolve2D = ps.Convolve1D, ps.Convolve2D
        opt = types.ModuleType('pylops.optimization')
        lsq = types.ModuleType('pylops.optimization.leastsquares')

        def _unavailable(*a, **k):
            raise NotImplementedError('pylops stand-in: not on the hot path')

        lsq.NormalEquationsInversion = _unavailable
        opt.leastsquares = lsq
        pylops.signalprocessing, pylops.optimization = sp, opt
        sys.modules.update({'pylops': pylops, 'pylops.signalprocessing': sp, 'pylops.optimization': opt,
                            'pylops.optimization.leastsquares': lsq})
    if 'dask' not in sys.modules:
        dask = types.ModuleType('dask')
        da = types.ModuleType('dask.array')
        core = types.ModuleType('dask.array.core')

        class Array:  # only used in isinstance checks / annotations
            pass

        core.Array = Array
        da.core, da.Array = core, Array
        dask.array = da
        sys.modules.update({'dask': dask, 'dask.array': da, 'dask.array.core': core})
    if 'skimage' not in sys.modules:
        sk = types.ModuleType('skimage')
        meas = types.ModuleType('skimage.measure')
        meas.block_reduce = lambda *a, **k: (_ for _ in ()).throw(NotImplementedError('skimage stub'))
        sk.measure = meas
        sys.modules.update({'skimage': sk, 'skimage.measure': meas})


_ref = None


def load_reference():
    """Return the reference ``pycsou`` package (imported once)."""
    global _ref
    if _ref is not None:
        return _ref
    root = reference_root()
    if root is None:
        raise ImportError('reference pycsou not present (expected in the build container only)')
    _restore_numpy_aliases()
    _register_standins()
    if root not in sys.path:
        sys.path.insert(0, root)
    import pycsou  # noqa: F401
    import pycsou.core, pycsou.func.penalty, pycsou.func.loss, pycsou.func.base  # noqa: F401,E401
    import pycsou.linop.base, pycsou.linop.diff, pycsou.linop.conv, pycsou.linop.sampling  # noqa: F401,E401
    import pycsou.math.prox, pycsou.opt.proxalgs  # noqa: F401,E401
    _ref = sys.modules['pycsou']
    return _ref


def vectorised_l21(dim: int, n_groups: int):
    """Reference-compatible L21 functional for ``groups = tile(arange(n_groups), dim // n_groups)``.

    Same numbers as the reference's ``L21Norm`` (func/penalty.py:548-560) -- checked in
    ``tests/test_oracle_reference.py`` at small sizes -- without its Theta(N^2) Python loop.
    """
    load_reference()
    from pycsou.core.functional import ProximableFunctional

    class VectorisedL21Norm(ProximableFunctional):
        def __init__(self):
            super().__init__(dim=dim, data=None, is_differentiable=False, is_linear=False)

        def __call__(self, x):
            return np.sum(np.linalg.norm(np.reshape(x, (dim // n_groups, n_groups)), axis=0))

        def prox(self, x, tau):
            X = np.reshape(x, (dim // n_groups, n_groups))
            with np.errstate(divide='ignore'):
                scale = np.clip(1 - tau / np.linalg.norm(X, axis=0), a_min=0, a_max=None)
            return (scale[None, :] * X).reshape(-1)

    return VectorisedL21Norm()
