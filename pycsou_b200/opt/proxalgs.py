r"""Proximal algorithms (mirror of ``pycsou/opt/proxalgs.py``) running on the B200.

``PrimalDualSplitting`` (Condat-Vu, alias ``PDS``), ``AcceleratedProximalGradientDescent`` (``APGD``),
``ChambollePockSplitting`` (``CPS``), ``DouglasRachfordSplitting`` (``DRS``) and ``ForwardBackwardSplitting`` (``FBS``)
keep the reference's constructor signatures, defaults, validation errors, step-size rules
(``proxalgs.py:246-316``), update equations (``:343-355``, ``:586-601``), iterand-dict keys and diagnostics columns.

What differs by design
----------------------
* Iterates live in HBM as CUDA tensors.  ``iterate()`` pattern-matches ``(F, G, H, K)`` onto fused iteration kernels
  (``opt/_fused.py``: one or two launches per iteration, relative-improvement norms reduced inside the update kernel);
  anything it cannot match runs op-by-op on the stand-alone kernels.  There is no CPU path.
* The stop test needs the previous iteration's relative improvement on the host.  Instead of stalling the GPU every
  iteration the driver launches update ``k+1`` speculatively while the 32-byte metric of update ``k`` is in flight and
  discards it if the test says stop -- the iterate returned and the iteration count are exactly the reference's.
* Diagnostics are accumulated on the device and materialised as the reference's ``DataFrame`` at the end.
* Extra keyword-only options (defaults reproduce the reference): ``dtype`` (iterate precision; default float64 as the
  reference's ``np.zeros(dtype=np.float)``), ``output`` ('auto' | 'numpy' | 'torch'), ``fused`` (allow fused kernels).
"""
from __future__ import annotations

from numbers import Number
from typing import Any, Optional, Tuple

import os

import numpy as np
import torch

from .. import _kernels as K
from .._array import device, like_input, to_device, torch_dtype
from ..core.functional import ProximableFunctional
from ..core.linop import LinearOperator
from ..core.map import DifferentiableMap
from ..core.solver import GenericIterativeAlgorithm, LazyDiagnostics
from ..func.base import NullDifferentiableFunctional, NullProximableFunctional
from ..linop.base import IdentityOperator, NullOperator

__all__ = ['PrimalDualSplitting', 'PDS', 'AcceleratedProximalGradientDescent', 'APGD', 'ChambollePockSplitting', 'CPS',
           'DouglasRachfordSplitting', 'DRS', 'ForwardBackwardSplitting', 'FBS']


def _is_dev_tensor(a):
    return isinstance(a, torch.Tensor) and a.is_cuda


def _own_copy(a, dtype):
    """Device copy of ``a`` that the solver may overwrite (``to_device`` passes device tensors through)."""
    t = to_device(a, dtype)
    return t.clone() if isinstance(a, torch.Tensor) and t.data_ptr() == a.data_ptr() else t


def _to_host(t: torch.Tensor) -> np.ndarray:
    """Device tensor -> numpy array through a pinned buffer (a pageable read-back runs at a quarter of the PCIe rate).
    The pinned block comes from torch's caching host allocator and goes back to it when the array is dropped."""
    h = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
    h.copy_(t, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    return h.numpy()


class _DeviceLoop:
    """Shared driver: speculative stop test + device-side diagnostics (see module docstring)."""

    _n_metrics = 4          # [||d x||^2, ||x_old||^2, ||d z||^2, ||z_old||^2]

    def _alloc_diag(self, keep=0):
        """Diagnostics rows of this run; the first ``keep`` rows (updates of an earlier ``iterate()`` that this one
        continues, core/solver.py:64-66) are preserved."""
        rows = max(int(max(self.max_iter, self.min_iter)) + 2, int(keep) + 1)
        dev = getattr(self, '_diag_dev', None)
        if dev is not None and dev.shape[0] == rows and dev.device == device():
            dev[keep:].zero_()               # buffers (one of them pinned) are kept between solves of one solver
        else:
            new_dev = torch.zeros(rows, self._n_metrics, dtype=torch.float64, device=device())
            new_host = torch.zeros(rows, self._n_metrics, dtype=torch.float64).pin_memory()
            if keep and dev is not None:
                n = min(int(keep), dev.shape[0])
                new_dev[:n].copy_(dev[:n])
                new_host[:n].copy_(self._diag_host[:n])
            self._diag_dev, self._diag_host = new_dev, new_host
        self._diag_events = {}

    def _post_metric_copy(self, it):
        self._diag_host[it].copy_(self._diag_dev[it], non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self._diag_events[it] = ev

    def _metric_ready(self, it):
        ev = self._diag_events.pop(it, None)
        if ev is not None:
            ev.synchronize()

    @staticmethod
    def _rel(dss, ss):
        return np.inf if ss == 0 else float(np.sqrt(dss) / np.sqrt(ss))

    def _run_loop(self, state, it0=0):
        """Reference loop ``core/solver.py:64-76`` with a one-deep speculative update.  Returns the final state.

        ``it0``: the update count to continue from -- as in the reference, a second ``iterate()`` without ``reset()``
        carries on at ``self.iter`` (and does nothing if the loop condition already fails)."""
        self._alloc_diag(keep=it0)
        it = int(it0)
        spec = None                          # speculative (state, it) already launched for index `it`
        while True:
            # Updates up to min_iter run unconditionally (core/solver.py:65-66): a plan that can enqueue several updates
            # with one library call does so (no per-update Python / ctypes work between the kernels).
            if spec is None and self.verbose is None and it <= self.min_iter:
                n = int(self.min_iter) + 1 - it
                many = self._step_many(state, it, n) if n >= 2 else None
                if many is not None:
                    state = many
                    it += n
                    if self.min_iter < it <= self.max_iter:
                        self._post_metric_copy(it - 1)
                    continue
            if it <= self.min_iter:
                go = True
            elif it > self.max_iter:
                go = False
            elif it == 0:
                go = np.inf > self.accuracy_threshold
            else:
                if spec is None:             # overlap the metric read-back with the next update
                    spec = self._step(state, it)
                self._metric_ready(it - 1)
                go = self._primal_metric(it - 1) > self.accuracy_threshold
            if not go:
                break
            new_state = spec if spec is not None else self._step(state, it)
            spec = None
            show = self.verbose is not None and it % self.verbose == 0
            # the 32-byte metric of update `it` travels to the host only if the stop test of index it+1 (or the
            # verbose print) will read it; otherwise nothing sits between two update kernels on the stream and all
            # rows are fetched in one copy after the loop
            if show or self.min_iter < it + 1 <= self.max_iter:
                self._post_metric_copy(it)
            if show:
                self._metric_ready(it)
                self.iter = it
                print(self._diag_row(it))
            state = new_state
            it += 1
        self.iter = it
        self._diag_events.clear()
        self._reduce_diag(it)
        self._diag_host[:it].copy_(self._diag_dev[:it])      # synchronous: every update has finished after this
        return state

    def _reduce_diag(self, n):
        """Hook for distributed solvers: sum the per-rank partial norms of rows [0, n) (single GPU: nothing)."""

    def _step_many(self, state, it, n):
        """Hook: updates it .. it+n-1 enqueued by one call; returns the new state, or None (one call per update)."""
        return None

    def _primal_metric(self, it):
        d = self._diag_host[it]
        return self._rel(float(d[0]), float(d[1]))


class PrimalDualSplitting(_DeviceLoop, GenericIterativeAlgorithm):
    r"""Primal-dual splitting (Condat-Vu) for :math:`\min_x F(x) + G(x) + H(Kx)`; alias ``PDS``.

    Parameters are those of the reference (``opt/proxalgs.py:126-164``).  See the module docstring for the extra
    keyword-only options.
    """

    def __init__(self, dim: int, F: Optional[DifferentiableMap] = None, G: Optional[ProximableFunctional] = None,
                 H: Optional[ProximableFunctional] = None, K: Optional[LinearOperator] = None,
                 tau: Optional[float] = None, sigma: Optional[float] = None, rho: Optional[float] = None,
                 beta: Optional[float] = None, x0=None, z0=None,
                 max_iter: int = 500, min_iter: int = 10, accuracy_threshold: float = 1e-3, verbose: Optional[int] = 1,
                 *, dtype=None, output: str = 'auto', fused: bool = True, slab=None):
        self.dim = dim
        self._H = True
        self._slab = slab        # opt/slab.py SlabComm: one image row-decomposed over the GPUs of the process group
        if isinstance(F, DifferentiableMap):
            if F.shape[1] != dim:
                raise ValueError(f'F does not have the proper dimension: {F.shape[1]}!={dim}.')
            self.F = F
            if F.diff_lipschitz_cst < np.inf:
                self.beta = self.F.diff_lipschitz_cst if beta is None else beta
            elif (beta is not None) and isinstance(beta, Number):
                self.beta = beta
            else:
                raise ValueError('F must be a differentiable functional with Lipschitz-continuous gradient.')
        elif F is None:
            self.F = NullDifferentiableFunctional(dim=dim)
            self.beta = 0
        else:
            raise TypeError(f'F must be of type {DifferentiableMap}.')

        if isinstance(G, ProximableFunctional):
            if G.dim != dim:
                raise ValueError(f'G does not have the proper dimension: {G.dim}!={dim}.')
            self.G = G
        elif G is None:
            self.G = NullProximableFunctional(dim=dim)
        else:
            raise TypeError(f'G must be of type {ProximableFunctional}.')

        if isinstance(K, LinearOperator) and isinstance(H, ProximableFunctional):
            if (K.shape[1] != dim) or (K.shape[0] != H.dim):
                raise ValueError(
                    f'Operator K with shape {K.shape} is inconsistent with functional H with dimension {H.dim}.')

        if isinstance(H, ProximableFunctional):
            self.H = H
            if isinstance(K, LinearOperator):
                self.K = K
            elif K is None:
                self.K = IdentityOperator(size=H.dim)
                self.K.lipschitz_cst = self.K.diff_lipschitz_cst = 1
            else:
                raise TypeError(f'K must be of type {LinearOperator}.')
        elif H is None:
            self.H = NullProximableFunctional(dim=dim)
            self._H = False
            self.K = NullOperator(shape=(dim, dim))
            self.K.lipschitz_cst = self.K.diff_lipschitz_cst = 0
        else:
            raise TypeError(f'H must be of type {ProximableFunctional}.')

        if (tau is not None) and (sigma is not None):
            self.tau, self.sigma = tau, sigma
        elif (tau is not None) and (sigma is None):
            self.tau = self.sigma = tau
        elif (tau is None) and (sigma is not None):
            self.tau = self.sigma = sigma
        else:
            self.tau, self.sigma = self.set_step_sizes()

        self.rho = rho if rho is not None else self.set_momentum_term()

        self._user_x0, self._user_z0 = x0, z0
        self._dtype = self._pick_dtype(dtype, x0, z0)
        self._output = output
        self._allow_fused = fused
        self.x0 = x0 if x0 is not None else self.initialize_primal_variable()
        self.z0 = z0 if z0 is not None else self.initialize_dual_variable()

        objective_functional = (self.F + self.G) + (self.H * self.K)
        init_iterand = {'primal_variable': self.x0, 'dual_variable': self.z0}
        GenericIterativeAlgorithm.__init__(self, objective_functional=objective_functional, init_iterand=init_iterand,
                                           max_iter=max_iter, min_iter=min_iter,
                                           accuracy_threshold=accuracy_threshold, verbose=verbose)
        self._plan = None

    # -- reference helpers ---------------------------------------------------------------------------------------
    @staticmethod
    def _pick_dtype(dtype, x0, z0):
        if dtype is not None:
            return torch_dtype(dtype)
        for a in (x0, z0):
            if isinstance(a, torch.Tensor) and a.dtype in (torch.float32, torch.float64, torch.complex64,
                                                           torch.complex128):
                return a.dtype
            if isinstance(a, np.ndarray) and a.dtype.kind in 'fc':
                return torch_dtype(a.dtype)
        return torch.float64

    def set_step_sizes(self) -> Tuple[float, float]:
        r"""Balanced step sizes saturating the convergence inequality (reference ``proxalgs.py:280-301``)."""
        L = self.K.lipschitz_cst
        if self._H and not (L < np.inf):
            raise ValueError(
                'Please compute the Lipschitz constant of the linear operator K by calling its method "compute_lipschitz_cst()".')
        if self.beta > 0:
            if self._H is False:
                return 2 / self.beta, 0
            tau = (1 / L ** 2) * ((-self.beta / 4) + np.sqrt((self.beta ** 2 / 16) + L ** 2))
            return tau, tau
        if self._H is False:
            return 1, 0
        return 1 / L, 1 / L

    def set_momentum_term(self) -> float:
        return 0.9 if self.beta > 0 else 1

    def initialize_primal_variable(self):
        return np.zeros(shape=(self.dim,), dtype=float)

    def initialize_dual_variable(self):
        return None if self._H is False else np.zeros(shape=(self.H.dim,), dtype=float)

    # -- device state ----------------------------------------------------------------------------------------------
    def _init_state(self):
        dt = self._dtype
        if self._slab is not None:
            return self._plan.init_state(self._user_x0, self._user_z0)
        x = (torch.zeros(self.dim, dtype=dt, device=device()) if self._user_x0 is None
             else to_device(self._user_x0, dt).clone())
        if self._H is False:
            z = None
        elif self._user_z0 is None:
            z = torch.zeros(self.H.dim, dtype=dt, device=device())
        else:
            z = to_device(self._user_z0, dt).clone()
        return x, z

    def _generic_step(self, x, z, drow):
        """One update, op by op (reference ``proxalgs.py:343-355``); writes the four norms into ``drow``."""
        tau, sigma, rho = float(self.tau), float(self.sigma), float(self.rho)
        null_F = isinstance(self.F, NullDifferentiableFunctional)
        v = torch.empty_like(x)
        if self._H and not null_F:
            K.lincomb3(v, 1.0, x, -tau, self.F._gradient(x), -tau, self.K._adjoint(z))
        elif self._H:
            K.lincomb2(v, 1.0, x, -tau, self.K._adjoint(z))
        elif not null_F:
            K.lincomb2(v, 1.0, x, -tau, self.F._gradient(x))
        else:
            v.copy_(x)
        x_temp = self.G._prox(v, tau)
        z_new = z
        if self._H:
            u = K.lincomb2(torch.empty_like(x), 2.0, x_temp, -1.0, x)
            w = K.lincomb2(torch.empty_like(z), 1.0, z, sigma, self.K._apply(u))
            z_temp = self.H._fenchel_prox(w, sigma)
            z_new = K.lincomb2(torch.empty_like(z), rho, z_temp, 1.0 - rho, z)
            K.diffsumsq(z, z_new, out=drow[2:3])
            K.sumsq(z, out=drow[3:4])
        x_new = K.lincomb2(torch.empty_like(x), rho, x_temp, 1.0 - rho, x)
        K.diffsumsq(x, x_new, out=drow[0:1])
        K.sumsq(x, out=drow[1:2])
        return x_new, z_new

    def _step(self, state, it):
        x, z = state
        if self._plan is not None:
            return self._plan.step(x, z, self._diag_dev[it])
        return self._generic_step(x, z, self._diag_dev[it])

    def _step_many(self, state, it, n):
        plan = self._plan
        if (plan is None or self._slab is not None or not hasattr(plan, 'run_many') or getattr(plan, 'data_mode', 3) == 3
                or os.environ.get('PCS_PDS_BATCH', '1') == '0'):        # debug switch: one library call per update
            return None
        ev = getattr(self, '_kernel_events', None)         # bench.py: CUDA events around the update kernels of the solve
        if ev is None:
            return plan.run_many(state[0], state[1], n, self._diag_dev[it:it + n])
        import torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = plan.run_many(state[0], state[1], n, self._diag_dev[it:it + n])
        e1.record()
        ev.append((e0, e1, int(n)))
        return out

    def _make_plan(self):
        if self._slab is not None:
            from .slab import SlabPdsTvPlan, make_slab_plan
            if not isinstance(self._plan, SlabPdsTvPlan):     # the peer arena is collective and expensive: keep it
                self._plan = make_slab_plan(self, self._slab)
            return self._plan
        # a plan captures F's data, the filters and the prox modes; it is rebuilt only when one of them is replaced
        key = (id(self.F), id(self.G), id(self.H), id(self.K), float(self.tau), self._dtype, self._allow_fused)
        if getattr(self, '_plan_key', None) != key:
            self._plan = None
            if self._allow_fused:
                from ._fused import plan_pds
                self._plan = plan_pds(self)
            self._plan_key = key
        return self._plan

    # -- slab decomposition: the norms of the stop test are per-rank partial sums until all-reduced ----------------
    def _alloc_diag(self, keep=0):
        super()._alloc_diag(keep)
        self._diag_reduced = set(range(int(keep)))

    def _post_metric_copy(self, it):
        if self._slab is not None and it not in self._diag_reduced:
            self._slab.all_reduce_sum(self._diag_dev[it])
            self._diag_reduced.add(it)
        super()._post_metric_copy(it)

    def _metric_ready(self, it):
        super()._metric_ready(it)
        if self._slab is not None and self._plan is not None:
            self._plan.check()               # a neighbour that stopped answering is reported now, not after max_iter updates

    def _reduce_diag(self, n):
        if self._slab is None:
            return
        idx = [i for i in range(n) if i not in self._diag_reduced]
        if idx and idx == list(range(idx[0], n)):         # the usual case: a contiguous tail of rows, one call in place
            self._slab.all_reduce_sum(self._diag_dev[idx[0]:n])
            self._diag_reduced.update(idx)
        elif idx:
            sel = torch.tensor(idx, device=self._diag_dev.device)
            t = self._diag_dev.index_select(0, sel)
            self._slab.all_reduce_sum(t)
            self._diag_dev.index_copy_(0, sel, t)
            self._diag_reduced.update(idx)

    # -- public API (reference names) ------------------------------------------------------------------------------
    def _as_output(self, t):
        if t is None:
            return None
        mode = self._output
        if mode == 'auto':
            mode = 'torch' if (_is_dev_tensor(self._user_x0) or _is_dev_tensor(self._user_z0)) else 'numpy'
        return t if mode == 'torch' else _to_host(t)

    def iterate(self) -> Any:
        self._make_plan()
        fresh = self.iter == 0 or self.iterand is None
        if fresh and hasattr(self._plan, 'begin_solve'):
            self._plan.begin_solve()                  # per-solve precomputation (A^T y), inside the solve
        state = self._init_state() if fresh else self._state_from_iterand()
        x, z = self._run_loop(state, 0 if fresh else self.iter)
        self.converged = True
        if self._slab is not None:
            self._plan.check()
            x, z = self._plan.extract(x, z)          # this rank's rows
        elif self._plan is not None:
            self._plan.release(x, z)
        self._state = (x, z)
        self.iterand = {'primal_variable': self._as_output(x), 'dual_variable': self._as_output(z)}
        self._cont = None if self._slab is not None else (self.iterand, (x, z))
        self.old_iterand = self.iterand
        self.diagnostics = self._build_diagnostics(self.iter)
        return self.iterand, self.converged, self.diagnostics

    def _state_from_iterand(self, for_write=True):
        """Device state to continue from ``self.iterand``.  ``for_write``: the caller may overwrite the state buffers
        (the multi-update path ping-pongs into them), so tensors the user can see are copied first."""
        it = self.iterand
        cont = getattr(self, '_cont', None)
        if cont is not None and cont[0] is it:          # the iterand this solver produced last: its device state is kept
            x, z = cont[1]
            if for_write and self._slab is None and isinstance(it['primal_variable'], torch.Tensor):
                x, z = x.clone(), (None if z is None else z.clone())
            return x, z
        if self._slab is not None:
            return self._plan.init_state(it['primal_variable'], it['dual_variable'])
        # a caller-owned iterand: private copies (the multi-update path writes into its state buffers)
        x = _own_copy(it['primal_variable'], self._dtype)
        z = None if it['dual_variable'] is None else _own_copy(it['dual_variable'], self._dtype)
        return x, z

    def update_iterand(self) -> dict:
        """One update on the current iterand (used by ``iterates(n)``, core/solver.py:88-103).

        As in the reference every call returns FRESH arrays (numpy unless the solver was given device tensors / ``output=
        'torch'``): the buffers of a fused plan are handed over to the caller, never recycled under a yielded iterate."""
        if self._slab is not None:
            self._make_plan()
            if self.iter == 0 or self.iterand is None:
                x, z = self._plan.init_state(self._user_x0, self._user_z0)
            else:
                x, z = self._state_from_iterand(for_write=False)
            if not hasattr(self, '_scratch_row'):
                self._scratch_row = torch.zeros(self._n_metrics, dtype=torch.float64, device=device())
            x, z = self._plan.step(x, z, self._scratch_row)
            self._plan.check()
            xo, zo = self._plan.extract(x, z)             # fresh copies of this rank's rows
            out = {'primal_variable': self._as_output(xo), 'dual_variable': self._as_output(zo)}
            self._cont = (out, (x, z))
            return out
        if self._plan is None and self._allow_fused:
            self._make_plan()
        if self.iter == 0 or self.iterand is None:
            x, z = self._init_state()
        else:
            x, z = self._state_from_iterand(for_write=False)
        if not hasattr(self, '_scratch_row'):
            self._scratch_row = torch.zeros(self._n_metrics, dtype=torch.float64, device=device())
        if self._plan is not None:
            x, z = self._plan.step(x, z, self._scratch_row)
            self._plan.release(x, z)
        else:
            x, z = self._generic_step(x, z, self._scratch_row)
        out = {'primal_variable': self._as_output(x), 'dual_variable': self._as_output(z)}
        self._cont = (out, (x, z))
        return out

    # -- resumability: the state of a solve is (x, z, tau, sigma, rho) ------------------------
    def _ckpt_path(self, path: str) -> str:
        """Slab mode: one file per rank (``<path>.rank<r>of<world>``), each holding that rank's rows."""
        if self._slab is None:
            return path
        return f'{path}.rank{self._slab.rank}of{self._slab.world}'

    def save_checkpoint(self, path: str):
        """Write the current iterates and step sizes (``torch.save``).  Slab-decomposed solves (every rank calls it):
        each rank writes ITS rows of ``x`` and of both planes of ``z`` to ``<path>.rank<r>of<world>`` -- a 32768^2 solve
        never gathers its state on one rank."""
        if getattr(self, '_state', None) is None:
            raise RuntimeError('nothing to checkpoint: call iterate() first')
        x, z = self._state
        ck = {'primal_variable': x.cpu(), 'dual_variable': None if z is None else z.cpu(), 'updates': int(self.iter),
              'tau': float(self.tau), 'sigma': None if self.sigma is None else float(self.sigma),
              'rho': float(self.rho), 'dim': int(self.dim)}
        if self._slab is not None:
            lay = self._plan.lay
            ck.update(rows=(int(lay.r0), int(lay.r1)), world=int(self._slab.world), image=(int(lay.Hg), int(lay.W)))
        torch.save(ck, self._ckpt_path(path))
        if self._slab is not None:
            self._slab.barrier()

    def load_checkpoint(self, path: str) -> int:
        """Warm-start this solver from a checkpoint: the next ``iterate()`` continues from the saved iterates with the
        saved step sizes.  Returns the number of updates the checkpointed solve had done.  Slab mode: every rank loads
        its own file; the decomposition (world size, rows) must be the one the checkpoint was written with."""
        ck = torch.load(self._ckpt_path(path))
        if ck['dim'] != self.dim:
            raise ValueError('checkpoint of a problem of another size')
        if self._slab is not None:
            self._make_plan()
            lay = self._plan.lay
            if ck.get('world') != self._slab.world or tuple(ck.get('rows', ())) != (lay.r0, lay.r1):
                raise ValueError('checkpoint written with another slab decomposition')
        self._user_x0, self._user_z0 = ck['primal_variable'], ck['dual_variable']
        self.tau, self.rho = ck['tau'], ck['rho']
        if ck['sigma'] is not None:
            self.sigma = ck['sigma']
        self.reset()
        return ck['updates']

    def _diag_row(self, it):
        d = self._diag_host[it]
        row = {'Iter': it, 'Relative Improvement (primal variable)': self._rel(float(d[0]), float(d[1]))}
        if self._H:
            row['Relative Improvement (dual variable)'] = self._rel(float(d[2]), float(d[3]))
        return row

    def _build_diagnostics(self, n):
        d = self._diag_host[:n].numpy().copy()           # the pinned rows are reused by the next solve
        has_dual = self._H

        def build():
            from pandas import DataFrame
            with np.errstate(divide='ignore', invalid='ignore'):
                rel = np.where(d[:, 1::2] == 0, np.inf, np.sqrt(d[:, 0::2]) / np.sqrt(d[:, 1::2]))
            data = {'Iter': np.arange(n), 'Relative Improvement (primal variable)': rel[:, 0]}
            if has_dual:
                data['Relative Improvement (dual variable)'] = rel[:, 1]
            return DataFrame(data)
        return LazyDiagnostics(build)

    def print_diagnostics(self):
        print(dict(self.diagnostics.loc[self.iter]))

    def stopping_metric(self):
        if self.iter == 0:
            return np.inf
        return self.diagnostics.loc[self.iter - 1, 'Relative Improvement (primal variable)']

    def update_diagnostics(self):
        pass    # diagnostics are produced by the update kernels (see _generic_step / opt/_fused.py)


PDS = PrimalDualSplitting


class AcceleratedProximalGradientDescent(_DeviceLoop, GenericIterativeAlgorithm):
    r"""Accelerated proximal gradient descent (FISTA family) for :math:`\min_x F(x) + G(x)`; alias ``APGD``.

    Reference: ``opt/proxalgs.py:485-622``.  ``acceleration`` in {``'BT'``, ``'CD'``, ``None``}; ``d`` is Chambolle and
    Dossal's parameter.  Iterand keys: ``'iterand'``, ``'past_aux'``, ``'past_t'``.
    """

    def __init__(self, dim: int, F: Optional[DifferentiableMap] = None, G: Optional[ProximableFunctional] = None,
                 tau: Optional[float] = None, acceleration: Optional[str] = 'CD', beta: Optional[float] = None,
                 x0=None, max_iter: int = 500, min_iter: int = 10,
                 accuracy_threshold: float = 1e-3, verbose: Optional[int] = 1, d: float = 75.,
                 *, dtype=None, output: str = 'auto', fused: bool = True):
        self.dim = dim
        self.acceleration = acceleration
        self.d = d
        if isinstance(F, DifferentiableMap):
            if F.shape[1] != dim:
                raise ValueError(f'F does not have the proper dimension: {F.shape[1]}!={dim}.')
            self.F = F
            if F.diff_lipschitz_cst < np.inf:
                self.beta = self.F.diff_lipschitz_cst if beta is None else beta
            elif (beta is not None) and isinstance(beta, Number):
                self.beta = beta
            else:
                raise ValueError('F must be a differentiable functional with Lipschitz-continuous gradient.')
        elif F is None:
            self.F = NullDifferentiableFunctional(dim=dim)
            self.beta = 0
        else:
            raise TypeError(f'F must be of type {DifferentiableMap}.')

        if isinstance(G, ProximableFunctional):
            if G.dim != dim:
                raise ValueError(f'G does not have the proper dimension: {G.dim}!={dim}.')
            self.G = G
        elif G is None:
            self.G = NullProximableFunctional(dim=dim)
        else:
            raise TypeError(f'G must be of type {ProximableFunctional}.')

        self.tau = tau if tau is not None else self.set_step_size()
        self._user_x0 = x0
        self._dtype = PrimalDualSplitting._pick_dtype(dtype, x0, None)
        self._output = output
        self._allow_fused = fused
        self.x0 = x0 if x0 is not None else self.initialize_iterate()
        objective_functional = self.F + self.G
        init_iterand = {'iterand': self.x0, 'past_aux': 0 * self.x0, 'past_t': 1}
        GenericIterativeAlgorithm.__init__(self, objective_functional=objective_functional, init_iterand=init_iterand,
                                           max_iter=max_iter, min_iter=min_iter,
                                           accuracy_threshold=accuracy_threshold, verbose=verbose)
        self._plan = None

    def set_step_size(self) -> float:
        return 1 / self.beta        # ZeroDivisionError when F is None, as in the reference (proxalgs.py:573)

    def initialize_iterate(self):
        return np.zeros(shape=(self.dim,), dtype=float)

    def _next_t(self, it, t_old):
        """Momentum schedule (reference ``proxalgs.py:592-598``); returns (t, a)."""
        if self.acceleration == 'BT':
            t = (1 + np.sqrt(1 + 4 * t_old ** 2)) / 2
        elif self.acceleration == 'CD':
            t = (it + self.d) / self.d
        else:
            t = t_old = 1
        return t, (t_old - 1) / t

    def _init_state(self):
        dt = self._dtype
        x = (torch.zeros(self.dim, dtype=dt, device=device()) if self._user_x0 is None
             else to_device(self._user_x0, dt).clone())
        return x, torch.zeros_like(x), 1

    def _generic_step(self, x, x_old, t_old, it, drow):
        tau = float(self.tau)
        if isinstance(self.F, NullDifferentiableFunctional):
            v = x.clone()
        else:
            v = K.lincomb2(torch.empty_like(x), 1.0, x, -tau, self.F._gradient(x))
        x_temp = self.G._prox(v, tau)
        t, a = self._next_t(it, t_old)
        x_new = K.lincomb2(torch.empty_like(x), 1.0 + a, x_temp, -a, x_old)
        K.diffsumsq(x, x_new, out=drow[0:1])
        K.sumsq(x, out=drow[1:2])
        return x_new, x_temp, t

    def _step(self, state, it):
        x, x_old, t_old = state
        if self._plan is not None:
            x_new, x_temp, t, _ = self._plan.run(x, x_old, t_old, it, self._diag_dev[it:], single=True)
            return x_new, x_temp, t
        return self._generic_step(x, x_old, t_old, it, self._diag_dev[it])

    def _make_plan(self):
        key = (id(self.F), id(self.G), float(self.tau), self._dtype, self._allow_fused, self.acceleration)
        if getattr(self, '_plan_key', None) != key:
            self._plan = None
            if self._allow_fused:
                from ._fused import plan_apgd
                self._plan = plan_apgd(self)
            self._plan_key = key
        return self._plan

    def _as_output(self, t):
        mode = self._output
        if mode == 'auto':
            mode = 'torch' if _is_dev_tensor(self._user_x0) else 'numpy'
        return t if mode == 'torch' else _to_host(t)

    def iterate(self) -> Any:
        self._make_plan()
        fresh = self.iter == 0 or self.iterand is None
        if fresh:
            state = self._init_state()
        else:
            state = (to_device(self.iterand['iterand'], self._dtype), to_device(self.iterand['past_aux'], self._dtype),
                     self.iterand['past_t'])
        it0 = 0 if fresh else int(self.iter)         # a second iterate() without reset() continues (core/solver.py:64-66)
        if self._plan is not None and hasattr(self._plan, 'run') and self.verbose is None:
            # the whole loop -- updates, norms and the reference's stop test -- runs inside one kernel
            self._alloc_diag(keep=it0)
            x, aux, t, self.iter = self._plan.run(state[0], state[1], state[2], it0, self._diag_dev[it0:])
            self._diag_host[:self.iter].copy_(self._diag_dev[:self.iter])
        else:
            x, aux, t = self._run_loop(state, it0)
        self.converged = True
        self.iterand = {'iterand': self._as_output(x), 'past_aux': self._as_output(aux), 'past_t': t}
        self.old_iterand = self.iterand
        self.diagnostics = self._build_diagnostics(self.iter)
        return self.iterand, self.converged, self.diagnostics

    def update_iterand(self) -> Any:
        if self._plan is None and self._allow_fused:
            self._make_plan()
        cont = getattr(self, '_cont', None)
        if self.iter == 0 or self.iterand is None:
            x, x_old, t_old = self._init_state()
        elif cont is not None and cont[0] is self.iterand:      # the iterand this solver yielded last: device state kept
            x, x_old, t_old = cont[1]
        else:
            x, x_old, t_old = (to_device(self.iterand['iterand'], self._dtype),
                               to_device(self.iterand['past_aux'], self._dtype), self.iterand['past_t'])
        if not hasattr(self, '_scratch_row'):
            self._scratch_row = torch.zeros(self._n_metrics, dtype=torch.float64, device=device())
        self._diag_dev = self._scratch_row.reshape(1, -1).expand(1, -1)
        if self._plan is not None:
            x, aux, t, _ = self._plan.run(x, x_old, t_old, self.iter, self._scratch_row, single=True)
        else:
            x, aux, t = self._generic_step(x, x_old, t_old, self.iter, self._scratch_row)
        # fresh arrays every call, numpy unless the solver was given device tensors (core/solver.py:88-103)
        out = {'iterand': self._as_output(x), 'past_aux': self._as_output(aux), 'past_t': t}
        self._cont = (out, (x, aux, t))
        return out

    def _diag_row(self, it):
        d = self._diag_host[it]
        return {'Iter': it, 'Relative Improvement': self._rel(float(d[0]), float(d[1]))}

    def _build_diagnostics(self, n):
        d = self._diag_host[:n].numpy().copy()

        def build():
            from pandas import DataFrame
            with np.errstate(divide='ignore', invalid='ignore'):
                rel = np.where(d[:, 1] == 0, np.inf, np.sqrt(d[:, 0]) / np.sqrt(d[:, 1]))
            return DataFrame({'Iter': np.arange(n), 'Relative Improvement': rel})
        return LazyDiagnostics(build)

    def print_diagnostics(self):
        print(dict(self.diagnostics.loc[self.iter]))

    def stopping_metric(self):
        if self.iter == 0:
            return np.inf
        return self.diagnostics.loc[self.iter - 1, 'Relative Improvement']

    def update_diagnostics(self):
        pass


APGD = AcceleratedProximalGradientDescent


class ChambollePockSplitting(PrimalDualSplitting):
    r"""Chambolle-Pock: PDS with ``F=None`` and ``rho=1`` (reference ``proxalgs.py:628-713``); alias ``CPS``."""

    def __init__(self, dim: int, G: Optional[ProximableFunctional] = None,
                 H: Optional[ProximableFunctional] = None, K: Optional[LinearOperator] = None,
                 tau: Optional[float] = None, sigma: Optional[float] = None, rho: Optional[float] = 1,
                 x0=None, z0=None, max_iter: int = 500, min_iter: int = 10, accuracy_threshold: float = 1e-3,
                 verbose: Optional[int] = 1, **opts):
        super(ChambollePockSplitting, self).__init__(dim=dim, F=None, G=G, H=H, K=K, tau=tau, sigma=sigma, rho=rho,
                                                     x0=x0, z0=z0, max_iter=max_iter, min_iter=min_iter,
                                                     accuracy_threshold=accuracy_threshold, verbose=verbose, **opts)


CPS = ChambollePockSplitting


class DouglasRachfordSplitting(PrimalDualSplitting):
    r"""Douglas-Rachford: PDS with ``F=None, K=Id, sigma=1/tau, rho=1`` (reference ``proxalgs.py:719-778``); ``DRS``."""

    def __init__(self, dim: int, G: Optional[ProximableFunctional] = None,
                 H: Optional[ProximableFunctional] = None,
                 tau: float = 1., x0=None, z0=None,
                 max_iter: int = 500, min_iter: int = 10, accuracy_threshold: float = 1e-3, verbose: Optional[int] = 1,
                 **opts):
        super(DouglasRachfordSplitting, self).__init__(dim=dim, F=None, G=G, H=H, K=None, tau=tau, sigma=1 / tau, rho=1,
                                                       x0=x0, z0=z0, max_iter=max_iter, min_iter=min_iter,
                                                       accuracy_threshold=accuracy_threshold, verbose=verbose, **opts)


DRS = DouglasRachfordSplitting


class ForwardBackwardSplitting(PrimalDualSplitting):
    r"""Forward-backward: PDS with ``H=None, K=None`` (reference ``proxalgs.py:784-859``); alias ``FBS``."""

    def __init__(self, dim: int, F: Optional[DifferentiableMap] = None, G: Optional[ProximableFunctional] = None,
                 tau: Optional[float] = None, rho: Optional[float] = 1, x0=None,
                 max_iter: int = 500, min_iter: int = 10, accuracy_threshold: float = 1e-3, verbose: Optional[int] = 1,
                 **opts):
        super(ForwardBackwardSplitting, self).__init__(dim=dim, F=F, G=G, H=None, K=None, tau=tau, rho=rho,
                                                       x0=x0, max_iter=max_iter, min_iter=min_iter,
                                                       accuracy_threshold=accuracy_threshold, verbose=verbose, **opts)


FBS = ForwardBackwardSplitting
