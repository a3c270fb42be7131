r"""Driver loop of iterative algorithms (mirror of ``pycsou/core/solver.py``).

Same contract as the reference (``core/solver.py:25-134``): ``iterate()`` runs
``while (iter <= max_iter and stopping_metric() > accuracy_threshold) or iter <= min_iter`` -- inclusive bounds, so
at most ``max_iter + 1`` and at least ``min_iter + 1`` updates -- and returns ``(iterand, converged, diagnostics)``
with ``converged`` unconditionally ``True``; ``iterates(n)`` yields successive iterands without diagnostics;
``reset()`` zeroes ``iter`` / ``iterand`` only.

The reference deep-copies the iterand every iteration; here ``update_iterand`` returns fresh device buffers (or
ping-pongs), so ``old_iterand`` is kept by reference.
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import Any, Optional, Tuple

from .map import Map


class LazyDiagnostics:
    """The diagnostics table of a solve, materialised as the reference's ``pandas.DataFrame`` on first use
    (SURVEY.md 8(f)-4).  Building the frame costs ~0.2 ms of host time, during which the GPU would sit idle between two
    back-to-back solves; everything a ``DataFrame`` offers (``diag['col']``, ``.loc``, ``.iloc``, ``len()``, ``print``,
    ``.to_numpy()`` ...) is forwarded to the frame, ``diag.frame`` is the frame itself."""

    __slots__ = ('_build', '_frame')

    def __init__(self, build):
        object.__setattr__(self, '_build', build)
        object.__setattr__(self, '_frame', None)

    @property
    def frame(self):
        if self._frame is None:
            object.__setattr__(self, '_frame', self._build())
            object.__setattr__(self, '_build', None)
        return self._frame

    def __getattr__(self, name):
        return getattr(self.frame, name)

    def __setattr__(self, name, value):
        setattr(self.frame, name, value)

    def __getitem__(self, key):
        return self.frame[key]

    def __setitem__(self, key, value):
        self.frame[key] = value

    def __len__(self):
        return len(self.frame)

    def __iter__(self):
        return iter(self.frame)

    def __contains__(self, key):
        return key in self.frame

    def __array__(self, dtype=None, copy=None):
        import numpy as np
        return np.asarray(self.frame, dtype=dtype)

    def __eq__(self, other):
        return self.frame == (other.frame if isinstance(other, LazyDiagnostics) else other)

    __hash__ = None

    def __repr__(self):
        return repr(self.frame)

    def _repr_html_(self):
        return self.frame._repr_html_()


class GenericIterativeAlgorithm(ABC):
    def __init__(self, objective_functional: Map, init_iterand: Any, max_iter: int = 500, min_iter: int = 10,
                 accuracy_threshold: float = 1e-3, verbose: Optional[int] = None):
        self.objective_functional = objective_functional
        self.max_iter = max_iter
        self.min_iter = min_iter
        self.accuracy_threshold = accuracy_threshold
        self.verbose = verbose
        self.diagnostics = None
        self.iter = 0
        self.iterand = None
        self.init_iterand = init_iterand
        self.converged = False
        super(GenericIterativeAlgorithm, self).__init__()

    def iterate(self) -> Any:
        self.old_iterand = dict(self.init_iterand) if isinstance(self.init_iterand, dict) else self.init_iterand
        while ((self.iter <= self.max_iter) and (self.stopping_metric() > self.accuracy_threshold)) or (
                self.iter <= self.min_iter):
            self.iterand = self.update_iterand()
            self.update_diagnostics()
            if self.verbose is not None:
                if self.iter % self.verbose == 0:
                    self.print_diagnostics()
            self.old_iterand = dict(self.iterand) if isinstance(self.iterand, dict) else self.iterand
            self.iter += 1
        self.converged = True
        self.iterand = self.postprocess_iterand()
        return self.iterand, self.converged, self.diagnostics

    def postprocess_iterand(self) -> Any:
        return self.iterand

    def reset(self):
        self.iter = 0
        self.iterand = None

    def iterates(self, n: int) -> Tuple:
        self.reset()
        for _ in range(n):
            self.iterand = self.update_iterand()
            self.iter += 1
            yield self.iterand

    @abstractmethod
    def update_iterand(self) -> Any:
        pass

    @abstractmethod
    def print_diagnostics(self):
        pass

    @abstractmethod
    def stopping_metric(self):
        pass

    @abstractmethod
    def update_diagnostics(self):
        pass
