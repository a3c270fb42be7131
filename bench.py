#!/usr/bin/env python
"""bench.py -- PDS iterations/sec and effective HBM GB/s on the 4096^2 TV-deconvolution workload (BASELINE.json).

Contract (one JSON line on stdout, printed by rank 0):

* a *step* is one solve: ``--iters`` (default 100) Condat-Vu updates of one synthetic ``--size``^2 TV-deconvolution
  problem (Convolve2D 15x15 PSF + Gradient/L21 + NonNegativeOrthant; BASELINE.json configs[2]), in **fp64** by default:
  the north-star line (``fp64-matching Condat-Vu TV-deconvolution at 4096^2``) and the arithmetic of the reference arm.
* ``value``: iterations/sec with the observation ``y`` already resident in HBM and the iterates staying there (CUDA
  events around the K timed steps, max over ranks); every solve forms ``A^T y`` anew.
* ``e2e``: the same metric through the public API with HOST buffers: every step copies ``y`` host->device, runs the
  solve and copies the primal and dual iterates back (wall clock, max over ranks).
* ``roofline``: algorithmic bytes per iteration (7*N*s) / measured time of one iteration, against the
  measured HBM copy bandwidth in MEASURED_PEAKS.json.
* ``parity``: the GPU iterates after ``--cpu-iters`` updates of the timed problem against the CPU reference arm's
  (fp64: relative L2 of x and z; fp32: relative error of the objective).  At N > 1 ``parity_vs_single_gpu``: the slab
  solve against the whole image solved on ONE GPU (fp64 must be identical on every pixel).
* ``secondary``: the other BASELINE.json configs, each with its own roofline fraction by SURVEY.md 8(d)'s byte counts:
  ``f32`` (configs[2] as written: fp32, with its own e2e), ``c1`` (1-D LASSO, 500 APGD updates), ``c2`` (TV denoising
  1024^2, L2-resident), ``c4`` (compressed-sensing MRI), ``c5_strong`` (ONE 32768^2 image over the N GPUs) and
  ``parity_mode`` (the stop test live on every update).
* ``cpu_baseline``: the reference's OWN solver classes (``oracle/_ref``) timed on this host (rank 0, N=1) on a bounded
  sample; the numpy oracle port when the reference package is not on the box.
* ``--impl reference``: the CPU reference arm (same workload, bounded sample per step).

Working set per iteration (x, y, z in/out: 7 arrays of N*s bytes = 470 MB fp32 / 940 MB fp64 at 4096^2) is larger
than the 126 MB L2, so no explicit L2 flush is needed between iterations (C1 / C2 are L2-resident and labelled so).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

LAMBDA = 0.02
NOISE = 0.02
# dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the fused kernel, from the committed `ncu --set full`
# captures named in NCU_SOURCE (bench.py cannot run under ncu itself)
NCU_DRAM_BYTES = {('f32', 4096, 'sep'): 445905152, ('f64', 4096, 'sep'): 957700864}
NCU_SOURCE = 'profiles/r02_ncu_pds_f32_ws3.txt, r02_ncu_pds_f64.txt (ncu --set full, one launch: dram read + write bytes)'


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument('--gpus', type=int, default=1)
    p.add_argument('--steps', type=int, default=5)
    p.add_argument('--warmup', type=int, default=3)
    p.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    p.add_argument('--c4-size', type=int, default=2048, help='side of each c4 (compressed-sensing MRI) image')
    p.add_argument('--c4-problems', type=int, default=8, help='independent c4 problems per GPU')
    p.add_argument('--workload', default='c3', choices=['c3', 'c4', 'c5'],
                   help="c3: BASELINE configs[2], one size^2 slab per GPU (weak scaling; the default and the metric's "
                        "config); c5: BASELINE configs[4], ONE 32768^2 image row-decomposed over the GPUs (strong scaling)")
    p.add_argument('--size', type=int, default=4096, help='image is size x size per GPU slab')
    p.add_argument('--c5-size', type=int, default=32768, help='side of the c5 image')
    p.add_argument('--iters', type=int, default=100, help='PDS updates per step (one solve)')
    p.add_argument('--dtype', default='f64', choices=['f32', 'f64'])
    p.add_argument('--psf', default='sep', choices=['sep', 'nonsep'])
    p.add_argument('--no-cpu-baseline', action='store_true')
    p.add_argument('--no-secondary', action='store_true', help='skip the secondary measurements (other configs / dtype)')
    p.add_argument('--cpu-size', type=int, default=0, help='CPU sample image size (default: --size)')
    p.add_argument('--cpu-iters', type=int, default=0, help='CPU reference updates per sample (default: 2; reference arm 1)')
    return p.parse_args()


# ------------------------------------------------------------------------------------------------ clocks sampler
class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons every 200 ms while the timed region runs."""
    FIELDS = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,'
              'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
              'clocks_event_reasons.sw_power_cap')
    NAMES = ('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap')

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', '-i', str(self.index), f'--query-gpu={self.FIELDS}', '--format=csv,noheader,nounits',
                 '-lms', '200'], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return self
        self.thread = threading.Thread(target=self._pump, daemon=True)
        self.thread.start()
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(',')])

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])), mx.append(float(r[1]))
                pw.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, flag in zip(self.NAMES, r[3:7]):
                if flag.lower().startswith('active'):
                    reasons.add(name)
        if not sm:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['no samples']}
        return {'sm_mhz': float(np.median(sm)), 'sm_max_mhz': float(max(mx)), 'power_w_max': float(max(pw)),
                'samples': len(sm), 'reasons': sorted(reasons)}


# ------------------------------------------------------------------------------------------------ helpers
def np_dt(tag):
    return np.float32 if tag == 'f32' else np.float64


def hbm_peak():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    try:
        with open(path) as f:
            return float(json.load(f)['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
    except Exception:
        return 6650.0, 'fallback (B200_PROFILING.md 6.65 TB/s)'


def _cpu_model():
    try:
        with open('/proc/cpuinfo') as f:
            for ln in f:
                if ln.startswith('model name'):
                    return f"{ln.split(':', 1)[1].strip()} x{os.cpu_count()}"
    except OSError:
        pass
    return f'unknown x{os.cpu_count()}'


class Dist:
    """Barrier / max-over-ranks helpers (no-ops on one GPU)."""

    def __init__(self, world):
        self.world = world

    def barrier(self):
        import torch
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()

    def max(self, v):
        if self.world == 1:
            return v
        import torch
        import torch.distributed as dist
        t = torch.tensor([v], dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def min(self, v):
        return -self.max(-v)


def time_solves(solver, steps, warmup, D):
    """``steps`` solves (reset + iterate) of a device-resident solver: (device ms max over ranks, C-ABI calls, last
    result, last diagnostics)."""
    import torch
    from pycsou_b200 import _lib
    est = diag = None
    for _ in range(warmup):
        solver.reset()
        solver.iterate()
    D.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n0 = _lib.launch_count
    solver._kernel_events = []            # (start, end, updates) around every run of update kernels inside the timed solves
    ev0.record()
    for _ in range(steps):
        solver.reset()
        est, _, diag = solver.iterate()
    ev1.record()
    D.barrier()
    kev, solver._kernel_events = solver._kernel_events, None
    k_ms, k_n = sum(a.elapsed_time(b) for a, b, _ in kev), sum(n for _, _, n in kev)
    time_solves.last_kernel = (D.max(k_ms), k_n)          # device ms inside the update kernels, updates they made
    return D.max(ev0.elapsed_time(ev1)), _lib.launch_count - n0, est, diag


def time_e2e(step_fn, steps, warmup, D):
    """Wall clock of ``steps`` calls of ``step_fn`` (host buffers in, host results out); max over ranks."""
    import torch
    out = None
    for _ in range(warmup):
        out = step_fn()                   # (the previous result stays alive during the next solve, as in the timed loop)
    D.barrier()
    each = []
    t0 = time.perf_counter()
    for _ in range(steps):
        t1 = time.perf_counter()
        out = step_fn()
        each.append(round(1e3 * (time.perf_counter() - t1), 2))
    torch.cuda.synchronize()
    total = D.max(time.perf_counter() - t0)
    D.barrier()
    return total, each, out


# ------------------------------------------------------------------------------------------------ C3 workload
def make_problem(size, psf_kind, rank=0):
    """Synthetic observation of the rank's size x size row block (seed = rank, so a block is the same for any GPU
    count)."""
    from pycsou_b200.util import synth
    psf = synth.gaussian_psf_2d(15, 2.5) if psf_kind == 'sep' else synth.rotated_gaussian_psf_2d(15)
    _, psf, y = synth.tv_deconv_problem((size, size), psf, NOISE, seed=rank, dtype=np.float64)
    return psf, y


def make_solver(y, psf, size, tdtype, iters, output, slab=None, rows=None, **kw):
    """The reference-facing construction (doc/general/examples.rst:19-47 pattern) for configs[2].

    With ``slab`` (N > 1 GPUs) the image is (world * size) x size, row-decomposed one size x size slab per GPU
    (configs[4] layout); ``y`` then holds this rank's rows.  ``rows``: image height when it is not ``size``."""
    from pycsou_b200.func.loss import SquaredL2Loss
    from pycsou_b200.func.penalty import L21Norm, NonNegativeOrthant, tv_groups
    from pycsou_b200.linop.conv import Convolve2D
    from pycsou_b200.linop.diff import Gradient
    from pycsou_b200.opt.proxalgs import PDS
    rows = rows if rows is not None else size * (slab.world if slab is not None else 1)
    N = rows * size
    A = Convolve2D(size=N, filter=psf, shape=(rows, size))
    K = Gradient(shape=(rows, size), kind='forward')
    K.lipschitz_cst = K.diff_lipschitz_cst = float(np.sqrt(8.))
    data = y if slab is None else slab.layout(rows, size).local(y)
    F = (1 / 2) * SquaredL2Loss(dim=N, data=data) * A
    H = LAMBDA * L21Norm(dim=2 * N, groups=tv_groups(N))
    opts = dict(max_iter=iters - 1, min_iter=iters - 1, accuracy_threshold=0)
    opts.update(kw)
    return PDS(dim=N, F=F, G=NonNegativeOrthant(dim=N), H=H, K=K, beta=1.0, verbose=None, dtype=tdtype, output=output,
               slab=slab, **opts)


def algorithmic_bytes_per_iter(size, itemsize):
    """read x, y, z(2N); write x, z(2N) = 7*N*s."""
    return 7 * size * size * itemsize


# ------------------------------------------------------------------------------------------------ CPU reference
def cpu_reference_sample(size, psf, y, iters, want_iterates=False):
    """``iters`` Condat-Vu updates of the CPU reference on the (size x size) problem, fp64 as the reference computes.

    The reference's OWN ``PrimalDualSplitting`` (``oracle/ref_problems.py``: reference solver, functionals and
    composition rules from ``oracle/_ref`` or ``/root/reference``; pylops stand-in operators; vectorised L21) when the
    reference package is on this machine, else the numpy oracle port.  Returns (seconds, threads, kind, x, z).
    The path is numpy ufuncs + scipy.signal fftconvolve: effectively single threaded."""
    from oracle.ref_shim import reference_available          # checker / CPU baseline only -- never on the product path
    if reference_available():
        from oracle import ref_problems as R
        dt, x, z = R.c3_reference_sample(y, psf, (size, size), LAMBDA, iters)
        return dt, 1, 'reference', x, z
    from oracle import np_oracle as O
    N = size * size
    Fo = O.HalfSquaredL2Loss(y, O.Convolve2D(N, psf, (size, size)), G_lipschitz=1.0)
    t0 = time.perf_counter()
    x, z, _, _ = O.pds_iterate(N, F=Fo, G=O.NonNegativeOrthant(N), H=LAMBDA * O.TVL21Norm(2 * N, N),
                               K=O.Gradient((size, size), kind='forward'), K_lipschitz=np.sqrt(8.), beta=1.0,
                               max_iter=iters - 1, min_iter=iters - 1, accuracy_threshold=0)
    return time.perf_counter() - t0, 1, 'port', x, z


def cpu_objective(size, psf, y, x):
    """Objective of configs[2] evaluated on the host in fp64 (for the fp32 criterion: 1e-4 relative)."""
    from oracle import np_oracle as O
    N = size * size
    A, Ko = O.Convolve2D(N, psf, (size, size)), O.Gradient((size, size), kind='forward')
    x = np.asarray(x, dtype=np.float64)
    return float(0.5 * np.linalg.norm(A.matvec(x) - y) ** 2 + LAMBDA * O.TVL21Norm(2 * N, N)(Ko.matvec(x)))


def reference_arm(args, rank, world):
    """CPU reference arm: rank 0 only; each step = a bounded sample (--cpu-iters updates) of the same workload."""
    if rank != 0:
        return
    size = args.cpu_size or args.size
    iters = args.cpu_iters or 1
    psf, y = make_problem(size, args.psf, 0)
    kind, cores = 'port', 1
    for _ in range(min(args.warmup, 1)):             # one warm-up update (fft plans, page faults); each costs seconds
        cpu_reference_sample(size, psf, y, 1)
    t = 0.0
    for _ in range(args.steps):
        dt, cores, kind, _, _ = cpu_reference_sample(size, psf, y, iters)
        t += dt
    ips = args.steps * iters / t
    what = ("the reference's own PrimalDualSplitting.iterate() (oracle/_ref; pylops stand-in operators, vectorised L21)"
            if kind == 'reference' else 'numpy oracle port of the reference path')
    sample = f'{iters} PDS update(s) of the {size}x{size} fp64 problem per step, {what}; 1 warm-up update'
    line = {
        'impl': 'reference', 'metric': 'pds_iterations_per_sec', 'value': ips, 'unit': 'iter/s', 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1e3 * t / args.steps, 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': {'workload': f'TV-deconvolution {size}x{size}, Convolve2D 15x15 {args.psf} PSF + Gradient/L21 + '
                               'NonNegativeOrthant, Condat-Vu PDS (BASELINE.json configs[2])',
                   'iters_per_step': iters, 'cpu': _cpu_model(), 'numpy': np.__version__},
        'cpu_baseline': {'value': ips, 'unit': 'iter/s', 'cores': cores, 'kind': kind, 'sample': sample},
        'e2e': {'value': ips, 'unit': 'iter/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ B200 arm (C3)
def c3_measure(args, dtype_tag, psf, y64, D, slab, rank, with_e2e=True):
    """Device-resident and end-to-end timing of configs[2] in one precision.  Returns a dict of raw numbers."""
    import torch
    size, iters, N = args.size, args.iters, args.size * args.size
    tdt = torch.float32 if dtype_tag == 'f32' else torch.float64
    itemsize = 4 if dtype_tag == 'f32' else 8
    world = D.world
    y_host = torch.from_numpy(y64.astype(np_dt(dtype_tag))).pin_memory()
    y_dev = y_host.to('cuda', non_blocking=True)
    torch.cuda.synchronize()

    solver = make_solver(y_dev, psf, size, tdt, iters, 'torch', slab)
    dev_ms, launches, est, _ = time_solves(solver, args.steps, args.warmup, D)
    kernel_ms, kernel_updates = time_solves.last_kernel
    plan = type(solver._plan).__name__ if getattr(solver, '_plan', None) is not None else 'generic (op-by-op)'
    from pycsou_b200.opt._fused import last_launch
    launch = last_launch()
    cs = est['primal_variable'].double().sum().reshape(1)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(cs)
    out = {'dev_ms': dev_ms, 'launches': launches, 'plan': plan, 'launch': launch, 'checksum': float(cs.item()),
           'itemsize': itemsize, 'kernel_ms': kernel_ms, 'kernel_updates': kernel_updates}
    del est, solver

    if with_e2e:
        def e2e_step():
            s = make_solver(y_host, psf, size, tdt, iters, 'numpy', slab)
            r, _, _ = s.iterate()
            return r
        e2e_s, each, res = time_e2e(e2e_step, args.steps, max(args.warmup, 3), D)
        assert res['primal_variable'].shape == (N,) and res['dual_variable'].shape == (2 * N,)
        out.update(e2e_s=e2e_s, e2e_each=each)
        # floor of the read-back: the same bytes copied device -> pinned host by every rank at the same time
        buf = torch.empty(3 * N, dtype=tdt, device='cuda')
        hbuf = torch.empty(3 * N, dtype=tdt, pin_memory=True)
        hbuf.copy_(buf)
        D.barrier()
        t0 = time.perf_counter()
        for _ in range(3):
            hbuf.copy_(buf, non_blocking=True)
            torch.cuda.synchronize()
        out['d2h_ms'] = D.max(1e3 * (time.perf_counter() - t0) / 3)
        del buf, hbuf, res
    return out


def gpu_iterates(psf, y64, size, dtype_tag, k, slab=None, y_local=None):
    """x, z (numpy) after exactly ``k`` updates through the public API."""
    import torch
    tdt = torch.float32 if dtype_tag == 'f32' else torch.float64
    y = torch.from_numpy((y64 if y_local is None else y_local).astype(np_dt(dtype_tag))).to('cuda')
    s = make_solver(y, psf, size, tdt, k, 'numpy', slab)
    est, _, _ = s.iterate()
    return est['primal_variable'], est['dual_variable']


def b200_arm(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise RuntimeError('bench.py needs a CUDA device (no CPU fallback on the product path)')
    torch.cuda.set_device(local_rank)
    from pycsou_b200 import _lib
    _lib.lib()                                                    # fails loudly if libpcsb200.so is missing
    D = Dist(world)
    size, iters, N = args.size, args.iters, args.size * args.size
    psf, y64 = make_problem(size, args.psf, rank)
    slab = None
    if world > 1:
        from pycsou_b200.opt.slab import SlabComm
        slab = SlabComm()

    sampler = ClockSampler(local_rank).start() if rank == 0 else None
    main = c3_measure(args, args.dtype, psf, y64, D, slab, rank)
    clocks = sampler.stop() if sampler is not None else None

    itemsize = main['itemsize']
    total_iters = world * args.steps * iters
    ips = total_iters / (main['dev_ms'] * 1e-3)
    ms_iter = main['dev_ms'] / (args.steps * iters)
    peak, peak_src = hbm_peak()
    alg = algorithmic_bytes_per_iter(size, itemsize)
    achieved_update = alg / (ms_iter * 1e-3) / 1e9          # whole solve / updates (A^T y, state set-up, host gaps included)
    # the dominant kernel alone: CUDA events around the runs of update kernels inside the SAME timed solves
    k_us = 1e3 * main['kernel_ms'] / main['kernel_updates'] if main.get('kernel_updates') else 1e3 * ms_iter
    achieved = alg / (k_us * 1e-6) / 1e9
    e2e_ips = total_iters / main['e2e_s']
    line = {
        'metric': 'pds_iterations_per_sec', 'value': ips, 'unit': 'iter/s', 'n_gpus': world, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': main['dev_ms'] / args.steps, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': args.dtype, 'data': 'synthetic',
        'config': {'workload': f'TV-deconvolution {size}x{size} per GPU, Convolve2D 15x15 {args.psf} PSF + '
                               'Gradient/L21 + NonNegativeOrthant, Condat-Vu PDS (BASELINE.json configs[2]; '
                               + ('fp64 = the north-star line and the reference arm\'s arithmetic)' if args.dtype == 'f64'
                                  else 'fp32 as configs[2] is written)'),
                   'iters_per_step': iters, 'ms_per_iteration': ms_iter, 'solver_path': main['plan'],
                   'kernel': main['launch'],
                   'parallelism': (f'one {world * size}x{size} image row-decomposed into {world} slabs of {size}x{size}, '
                                   'halo rows pushed to the neighbour GPUs over NVLink inside the update kernel, '
                                   'stop-test norms all-reduced; value counts SLAB-iterations per second (= N x image '
                                   'iterations/s: a weak-scaling aggregate, comparable with the CPU arm\'s one 4096^2 '
                                   'problem only if that is assumed linear in pixels)')
                   if world > 1 else 'single GPU',
                   'l2_policy': f'working set {alg / 1e6:.0f} MB per iteration > 126 MB L2 (no flush needed)',
                   'effective_gbs_per_gpu': achieved_update, 'checksum_x': main['checksum'],
                   'aty': 'A^T y is re-formed inside every timed solve'},
        'e2e': {'value': e2e_ips, 'unit': 'iter/s', 'h2d_bytes_per_step': N * itemsize,
                'd2h_bytes_per_step': 3 * N * itemsize, 'ms_per_step': 1e3 * main['e2e_s'] / args.steps,
                'ms_each_step_rank0': main['e2e_each'],
                'd2h_floor_ms': main['d2h_ms'],
                'note': 'd2h_floor_ms = the read-back bytes alone, copied to pinned host memory by all ranks at once'},
        'gpu_launches': main['launches'],
        'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak,
                     'traffic': NCU_DRAM_BYTES.get((args.dtype, size, args.psf)), 'traffic_source': NCU_SOURCE,
                     'peak_source': peak_src, 'algorithmic_bytes': alg,
                     'kernel_us': k_us, 'kernel_launches_timed': main.get('kernel_updates'),
                     'achieved_whole_solve': achieved_update, 'frac_whole_solve': achieved_update / peak,
                     'per': 'one launch of the fused update kernel = one PDS iteration: average duration by CUDA events '
                            'around the kernel runs inside the timed solves; *_whole_solve divides the solve time '
                            '(A^T y, state set-up, host gaps included) by the updates instead'},
        'clocks': clocks,
    }

    # ---- parity on the timed problem ----
    if world > 1:
        line['parity_vs_single_gpu'] = slab_vs_single_gpu(args, psf, y64, slab, rank, world, D)
    sec = {}
    if not args.no_secondary:
        other = 'f32' if args.dtype == 'f64' else 'f64'
        o = c3_measure(args, other, psf, y64, D, slab, rank)
        o_ms = o['dev_ms'] / (args.steps * iters)
        o_gbs = algorithmic_bytes_per_iter(size, o['itemsize']) / (o_ms * 1e-3) / 1e9
        o_kus = 1e3 * o['kernel_ms'] / o['kernel_updates'] if o.get('kernel_updates') else 1e3 * o_ms
        o_kgbs = algorithmic_bytes_per_iter(size, o['itemsize']) / (o_kus * 1e-6) / 1e9
        sec[other] = {'iter_per_s': world * 1e3 / o_ms, 'ms_per_iteration': o_ms, 'effective_gbs_per_gpu': o_gbs,
                      'roofline_frac': o_kgbs / peak, 'kernel_us': o_kus, 'roofline_frac_whole_solve': o_gbs / peak,
                      'kernel': o['launch'],
                      'e2e': {'value': total_iters / o['e2e_s'], 'unit': 'iter/s', 'h2d_bytes_per_step': N * o['itemsize'],
                              'd2h_bytes_per_step': 3 * N * o['itemsize'], 'ms_per_step': 1e3 * o['e2e_s'] / args.steps}}
        sec['parity_mode'] = parity_mode_measure(args, psf, y64, D, slab)
        if world == 1:
            # C1 and C2 are latency bound (state on chip): measured after two idle seconds, so that what they show is the
            # kernels and not how far the sustained runs above have pushed the board into its power cap
            time.sleep(2.0)
            sec['c1'] = c1_measure()
            sec['c2'] = c2_measure()
            sec['c1']['idle_before_s'] = sec['c2']['idle_before_s'] = 2.0
        sub = argparse.Namespace(**vars(args))
        sub.steps, sub.warmup, sub.no_cpu_baseline = 2, 1, True
        c4 = c4_arm(sub, rank, local_rank, world, emit=False)
        sec['c4'] = {k: c4[k] for k in ('value', 'unit', 'scaling', 'config', 'roofline', 'e2e')}
        sub.iters = 50
        c5 = c5_arm(sub, rank, local_rank, world, emit=False)
        sec['c5_strong'] = {k: c5[k] for k in ('value', 'unit', 'scaling', 'config', 'roofline', 'e2e')}
    if slab is not None:
        from pycsou_b200.opt.slab import release_arenas
        release_arenas()
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            csize = args.cpu_size or size
            k = args.cpu_iters or 2
            cpsf, cy = (psf, y64) if csize == size else make_problem(csize, args.psf, 0)
            cpu_reference_sample(csize, cpsf, cy, 1)                              # warm-up (fft plans, page faults)
            dt, cores, kind, xr, zr = cpu_reference_sample(csize, cpsf, cy, k)
            what = ("the reference's own PrimalDualSplitting.iterate() from oracle/_ref (pylops stand-in operators, "
                    'vectorised L21)') if kind == 'reference' else 'numpy oracle port'
            line['cpu_baseline'] = {'value': k / dt, 'unit': 'iter/s', 'cores': cores, 'kind': kind,
                                    'sample': f'{k} PDS updates of the {csize}x{csize} fp64 problem ({what}, '
                                              f'{_cpu_model()})'}
            # the same k updates on the GPU, both precisions, against the CPU arm's iterates
            x64, z64 = gpu_iterates(cpsf, cy, csize, 'f64', k)
            x32, _ = gpu_iterates(cpsf, cy, csize, 'f32', k)
            f_ref, f_32 = cpu_objective(csize, cpsf, cy, xr), cpu_objective(csize, cpsf, cy, x32)
            line['parity'] = {
                'updates': k, 'size': csize, 'against': f'cpu_baseline ({kind})',
                'x_rel_l2': float(np.linalg.norm(x64 - xr) / np.linalg.norm(xr)),
                'z_rel_l2': float(np.linalg.norm(z64 - zr) / np.linalg.norm(zr)),
                'obj_rel_f32': abs(f_32 - f_ref) / abs(f_ref),
                'x_rel_l2_f32': float(np.linalg.norm(x32.astype(np.float64) - xr) / np.linalg.norm(xr)),
                'bar': 'fp64 iterates <= 1e-10 relative L2; fp32 objective <= 1e-4 relative (north_star)'}
        if sec:
            line['secondary'] = sec
        print(json.dumps(line), flush=True)


def slab_vs_single_gpu(args, psf, y64, slab, rank, world, D, k=6):
    """Every rank also solves the WHOLE (world*size) x size image alone (single-GPU kernel path; the blocks of y are
    all-gathered over NCCL) and compares its rows of the slab solve with it.  fp64 must agree on every pixel."""
    import torch
    import torch.distributed as dist
    size = args.size
    out = {'updates': k}
    for tag in ('f64', 'f32'):
        tdt = torch.float64 if tag == 'f64' else torch.float32
        y_loc = torch.from_numpy(y64.astype(np_dt(tag))).to('cuda')
        y_all = torch.empty(world * y_loc.numel(), dtype=tdt, device='cuda')
        dist.all_gather_into_tensor(y_all, y_loc)
        s = make_solver(y_loc, psf, size, tdt, k, 'torch', slab)
        est, _, _ = s.iterate()
        w = make_solver(y_all, psf, size, tdt, k, 'torch', None, rows=world * size)
        ref, _, _ = w.iterate()
        lay = slab.layout(world * size, size)
        xr = ref['primal_variable'].view(world * size, size)[lay.r0:lay.r1].reshape(-1)
        zr = ref['dual_variable'].view(2, world * size, size)[:, lay.r0:lay.r1].reshape(-1)
        eq = torch.equal(est['primal_variable'], xr) and torch.equal(est['dual_variable'], zr)
        dx = float((est['primal_variable'] - xr).abs().max().item())
        dz = float((est['dual_variable'] - zr).abs().max().item())
        out[tag] = {'identical': bool(D.min(1.0 if eq else 0.0) > 0.5), 'x_maxabs': D.max(dx), 'z_maxabs': D.max(dz)}
        del s, w, est, ref, y_all
    out['what'] = (f'rows of the {world}-slab solve vs the whole {world * size}x{size} image solved by the single-GPU '
                   'kernel on every rank; max over ranks')
    return out


def parity_mode_measure(args, psf, y64, D, slab):
    """SURVEY.md 8(d) "parity mode": the reference's stop test live on EVERY update (min_iter = 0, a threshold that
    never fires), i.e. the 32-byte metric of update k-1 is read on the host before update k+1 is launched."""
    import torch
    tdt = torch.float32 if args.dtype == 'f32' else torch.float64
    y_dev = torch.from_numpy(y64.astype(np_dt(args.dtype))).to('cuda')
    s = make_solver(y_dev, psf, args.size, tdt, args.iters, 'torch', slab, min_iter=0, accuracy_threshold=1e-300)
    ms, _, _, _ = time_solves(s, 2, 1, D)
    ms_iter = ms / (2 * args.iters)
    return {'iter_per_s': D.world * 1e3 / ms_iter, 'ms_per_iteration': ms_iter, 'dtype': args.dtype,
            'what': 'stop test evaluated on the host after every update (speculative next update in flight)'}


# ------------------------------------------------------------------------------------------------ C1 / C2 (single GPU)
def c1_measure():
    """BASELINE.json configs[0]: 1-D LASSO, N = 4096 fp64, APGD('CD'), exactly 500 updates (one launch on a 16-block cluster)."""
    import torch
    from pycsou_b200.func.loss import SquaredL2Loss
    from pycsou_b200.func.penalty import L1Norm
    from pycsou_b200.linop.conv import Convolve1D
    from pycsou_b200.opt._fused import apgd_last_launch
    from pycsou_b200.opt.proxalgs import APGD
    from pycsou_b200.util import synth
    _, h, y = synth.spikes_1d(4096, 32, 0)
    N = y.size
    A = Convolve1D(size=N, filter=h)
    lam = 0.1 * float(np.max(np.abs(A.adjoint(y))))
    out = {}
    for tag, tdt in (('f64', torch.float64), ('f32', torch.float32)):
        F = (1 / 2) * SquaredL2Loss(dim=N, data=torch.from_numpy(y.astype(np_dt(tag))).cuda()) * \
            Convolve1D(size=N, filter=h, dtype=np_dt(tag).__name__)
        s = APGD(dim=N, F=F, G=lam * L1Norm(dim=N), tau=1.0, acceleration='CD', beta=1.0, max_iter=499, min_iter=499,
                 accuracy_threshold=0, verbose=None, d=75., dtype=tdt, output='torch')
        ms, _, _, _ = time_solves(s, 5, 2, Dist(1))
        ms_solve = ms / 5
        alg = 5 * N * (4 if tag == 'f32' else 8)
        out[tag] = {'solve_ms': ms_solve, 'iter_per_s': 500 / (ms_solve * 1e-3), 'us_per_update': 1e3 * ms_solve / 500,
                    'solver_path': type(s._plan).__name__ if s._plan is not None else 'generic',
                    'kernel': apgd_last_launch()}
    out['what'] = ('500 APGD updates, N=4096, ONE launch per solve on a thread-block cluster (16 SMs share the vector in '
                   'their shared memory, neighbours exchange edge values through distributed shared memory, stop test '
                   'live; DESIGN 3.7); latency bound (algorithmic HBM bytes are 5*N*s per SOLVE), no HBM roofline applies')
    return out


def c2_measure():
    """BASELINE.json configs[1]: TV denoising 1024^2, PDS, 200 updates, fp64 and fp32 (state resident on chip)."""
    import torch
    from pycsou_b200.func.loss import SquaredL2Loss
    from pycsou_b200.func.penalty import L21Norm, tv_groups
    from pycsou_b200.linop.diff import Gradient
    from pycsou_b200.opt._fused import last_launch
    from pycsou_b200.opt.proxalgs import PDS
    from pycsou_b200.util import synth
    shape, iters = (1024, 1024), 200
    N = shape[0] * shape[1]
    _, y = synth.tv_denoise_problem(shape, 0.1, 0)
    peak, _ = hbm_peak()
    out = {}
    for tag, tdt in (('f64', torch.float64), ('f32', torch.float32)):
        K = Gradient(shape=shape, kind='forward', dtype=np_dt(tag).__name__)
        K.lipschitz_cst = K.diff_lipschitz_cst = float(np.sqrt(8.))
        F = (1 / 2) * SquaredL2Loss(dim=N, data=torch.from_numpy(y.astype(np_dt(tag))).cuda())
        s = PDS(dim=N, F=F, H=0.1 * L21Norm(dim=2 * N, groups=tv_groups(N)), K=K, rho=0.9, max_iter=iters - 1,
                min_iter=iters - 1, accuracy_threshold=0, verbose=None, dtype=tdt, output='torch')
        ms, _, _, _ = time_solves(s, 5, 2, Dist(1))
        ms_iter = ms / (5 * iters)
        alg = 7 * N * (4 if tag == 'f32' else 8)
        gbs = alg / (ms_iter * 1e-3) / 1e9
        out[tag] = {'iter_per_s': 1e3 / ms_iter, 'us_per_update': 1e3 * ms_iter, 'effective_gbs': gbs,
                    'roofline_frac_L2_resident': gbs / peak, 'algorithmic_bytes': alg, 'kernel': last_launch()}
    out['what'] = ('TV denoising 1024^2, 200 Condat-Vu updates per solve.  x and z (25 / 13 MB) stay in the shared memory '
                   'of the 148 SMs for the whole solve (pds_tv_resident_kernel: a slab of 7 rows per SM, boundary rows '
                   'handed over through L2); the "fraction" divides the 7*N*s bytes an update would move through HBM by '
                   'the HBM figure -- no such traffic exists, the kernel is bound by the neighbour hand-over latency')
    return out


# ------------------------------------------------------------------------------------------------ C4 (independent problems)
C4_LAM, C4_FRAC, C4_NOISE = 0.05, 0.3, 0.01


def c4_problem(size, p):
    """Problem p of BASELINE.json configs[3]: phantom x smooth unit-modulus phase, mask =
    rng(p).random(N) < 0.3, y = Mask FFT2_ortho(truth) + 0.01 complex noise.  Host arrays (complex128)."""
    from pycsou_b200.util import synth
    N = size * size
    rng = np.random.default_rng(p)
    mask = rng.random(N) < C4_FRAC
    truth = synth.phantom((size, size)).reshape(-1) * np.exp(2j * np.pi * (np.arange(N) / N))
    m = int(mask.sum())
    y = np.fft.fft2(truth.reshape(size, size), norm='ortho').reshape(-1)[mask]
    y = y + C4_NOISE * (rng.standard_normal(m) + 1j * rng.standard_normal(m))
    return mask, y, C4_NOISE * np.sqrt(2 * m)


def c4_solver(size, mask, y, eps, iters, output, ctype='complex64'):
    from pycsou_b200.func.base import ProxFuncHStack
    from pycsou_b200.func.loss import L2BallLoss
    from pycsou_b200.func.penalty import L21Norm, tv_groups
    from pycsou_b200.linop.base import LinOpVStack
    from pycsou_b200.linop.diff import Gradient
    from pycsou_b200.linop.sampling import Masking
    from pycsou_b200.linop.transform import DFT2
    from pycsou_b200.opt.proxalgs import CPS
    N = size * size
    M = Masking(size=N, sampling_bool=mask, dtype=ctype)
    m = M.shape[0]
    MF = M * DFT2((size, size), dtype=ctype)
    MF.lipschitz_cst = MF.diff_lipschitz_cst = 1.0
    Gr = Gradient(shape=(size, size), kind='forward', dtype=ctype)
    Gr.lipschitz_cst = Gr.diff_lipschitz_cst = float(np.sqrt(8.))
    H = ProxFuncHStack(L2BallLoss(dim=m, data=y, radius=eps), C4_LAM * L21Norm(dim=2 * N, groups=tv_groups(N)))
    return CPS(dim=N, G=None, H=H, K=LinOpVStack(MF, Gr), tau=1 / 3, sigma=1 / 3, rho=1, max_iter=iters - 1,
               min_iter=iters - 1, accuracy_threshold=0, verbose=None, dtype=ctype, output=output)


def c4_cpu_sample(size, mask, y, eps, iters):
    """``iters`` Chambolle-Pock updates of one problem on the numpy oracle port (complex128, as the reference computes)."""
    from oracle import np_oracle as O            # checker / CPU baseline only -- never on the product path
    N = size * size
    M = O.Masking(N, mask, dtype=np.complex128)
    m = M.shape[0]
    Ko = O.LinOpVStackOracle(O.Comp(M, O.FFT2Ortho((size, size))), O.Gradient((size, size), kind='forward', dtype='complex128'))
    Ho = O.ProxHStack(O.PreCompShift(O.L2Ball(m, eps), y), C4_LAM * O.TVL21Norm(2 * N, N))
    t0 = time.perf_counter()
    O.pds_iterate(N, G=None, H=Ho, K=Ko, tau=1 / 3, sigma=1 / 3, rho=1, x0=np.zeros(N, dtype=np.complex128),
                  z0=np.zeros(m + 2 * N, dtype=np.complex128), max_iter=iters - 1, min_iter=iters - 1,
                  accuracy_threshold=0)
    return time.perf_counter() - t0, 1


def c4_arm(args, rank, local_rank, world, emit=True):
    """BASELINE.json configs[3]: independent compressed-sensing MRI problems, one subset per GPU, no data-path
    collective.  A step = --iters Chambolle-Pock updates of every problem of the rank."""
    import torch
    if not torch.cuda.is_available():
        raise RuntimeError('bench.py needs a CUDA device (no CPU fallback on the product path)')
    torch.cuda.set_device(local_rank)
    from pycsou_b200 import _lib
    _lib.lib()
    size, iters, P = args.c4_size, args.iters, args.c4_problems
    N = size * size
    D = Dist(world)

    probs = [c4_problem(size, rank * P + i) for i in range(P)]
    m_tot = sum(int(pr[0].sum()) for pr in probs)
    y_host = [torch.from_numpy(pr[1].astype(np.complex64)).pin_memory() for pr in probs]
    solvers = [c4_solver(size, pr[0], yh.to('cuda'), pr[2], iters, 'torch') for pr, yh in zip(probs, y_host)]
    for _ in range(args.warmup):
        for s in solvers:
            s.reset()
            s.iterate()
    sampler = ClockSampler(local_rank).start() if (rank == 0 and emit) else None
    D.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = _lib.launch_count
    ev0.record()
    for _ in range(args.steps):
        for s in solvers:
            s.reset()
            est, _, _ = s.iterate()
    ev1.record()
    D.barrier()
    launches = _lib.launch_count - launches0
    dev_ms = D.max(ev0.elapsed_time(ev1))
    plan = type(solvers[0]._plan).__name__ if solvers[0]._plan is not None else 'generic (op-by-op)'
    cs = est['primal_variable'].abs().double().sum().reshape(1)
    checksum = float(cs.item())
    del solvers, est

    # ---- end to end: host samples in, host image out, every problem of every step ----
    e2e_solvers = [c4_solver(size, pr[0], yh, pr[2], iters, 'numpy') for pr, yh in zip(probs, y_host)]

    def e2e_step():
        out = None
        for s, yh in zip(e2e_solvers, y_host):
            s._plan = None                                   # a new solve of new data: the data are uploaded again
            s._plan_key = None
            s.reset()
            out, _, _ = s.iterate()
        return out
    e2e_s, _, out = time_e2e(e2e_step, args.steps, 2, D)
    clocks = sampler.stop() if sampler is not None else None
    assert out['primal_variable'].shape == (N,)
    del e2e_solvers, out

    total = world * args.steps * P * iters                   # problem-iterations
    ips = total / (dev_ms * 1e-3)
    ms_pi = dev_ms / (args.steps * P * iters)
    peak, peak_src = hbm_peak()
    alg = (6 * N + 3 * m_tot / P) * 8                        # (6N + 3m) s per problem-iteration, FFT passes not counted
    achieved = alg / (ms_pi * 1e-3) / 1e9
    line = {
        'metric': 'cps_problem_iterations_per_sec', 'value': ips, 'unit': 'iter/s', 'n_gpus': world, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': dev_ms / args.steps, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'c64', 'data': 'synthetic',
        'config': {'workload': f'compressed-sensing MRI {size}x{size} complex64, 2-D DFT + random Masking (30 %) + TV, '
                               f'ChambollePock, {P} independent problems per GPU (BASELINE.json configs[3])',
                   'iters_per_step': iters, 'problems_per_gpu': P, 'ms_per_problem_iteration': ms_pi,
                   'solver_path': plan,
                   'parallelism': f'{world * P} independent problems, {P} per GPU, no data-path collective',
                   'l2_policy': f'working set {10 * N * 8 / 1e6:.0f} MB per problem > 126 MB L2 (no flush needed)',
                   'effective_gbs_per_gpu': achieved, 'checksum_abs_x': checksum},
        'e2e': {'value': total / e2e_s, 'unit': 'iter/s', 'h2d_bytes_per_step': int(m_tot * 8),
                'd2h_bytes_per_step': int(P * (3 * N * 8) + m_tot * 8), 'ms_per_step': 1e3 * e2e_s / args.steps},
        'gpu_launches': launches,
        'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak,
                     'traffic': None, 'peak_source': peak_src, 'algorithmic_bytes': alg,
                     'per': 'one problem-iteration: 3 kernels of this library + 2 cuFFT transforms whose passes are '
                            'NOT counted in the algorithmic bytes (SURVEY 8(d): the fraction is structurally < 50 %)'},
        'clocks': clocks,
    }
    if rank == 0 and emit:
        if world == 1 and not args.no_cpu_baseline:
            mask, y, eps = probs[0]
            k = args.cpu_iters or 2
            c4_cpu_sample(size, mask, y, eps, 1)
            dt, cores = c4_cpu_sample(size, mask, y, eps, k)
            line['cpu_baseline'] = {'value': k / dt, 'unit': 'iter/s', 'cores': cores, 'kind': 'port',
                                    'sample': f'{k} Chambolle-Pock updates of one {size}x{size} complex128 '
                                              f'problem (numpy oracle port, {_cpu_model()})'}
        print(json.dumps(line), flush=True)
    return line


# ------------------------------------------------------------------------------------------------ C5 (strong scaling)
def c5_rows_on_device(Hg, W, r0, r1, psf, tdt, torch):
    """Rows [r0, r1) of the synthetic Hg x W observation, built on the device: piecewise-constant phantom, blurred by
    the 15x15 PSF (this package's Convolve2D on the rows plus a 7-row margin), plus noise drawn per 1024-row block from
    a generator seeded with the block index -- the image is the same for every GPU count."""
    from pycsou_b200.linop.conv import Convolve2D
    R = 7
    lo, hi = max(r0 - R, 0), min(r1 + R, Hg)
    r = (torch.arange(lo, hi, device='cuda', dtype=torch.float64) / Hg)[:, None]
    c = (torch.arange(W, device='cuda', dtype=torch.float64) / W)[None, :]
    img = torch.zeros(hi - lo, W, dtype=tdt, device='cuda')
    img.masked_fill_((r >= 0.10) & (r < 0.45) & (c >= 0.15) & (c < 0.40), 0.8)
    img.masked_fill_((r >= 0.55) & (r < 0.90) & (c >= 0.10) & (c < 0.30), 0.4)
    img.masked_fill_((r >= 0.30) & (r < 0.70) & (c >= 0.55) & (c < 0.90), 0.6)
    img.masked_fill_(((r - 0.72) ** 2 + (c - 0.50) ** 2) < 0.12 ** 2, 1.0)
    A = Convolve2D(size=img.numel(), filter=psf, shape=tuple(img.shape))
    y = A(img.reshape(-1)).view(hi - lo, W)[r0 - lo:r0 - lo + (r1 - r0)].contiguous()
    del img
    blk = 1024
    for b in range(r0 // blk, (r1 - 1) // blk + 1):
        g = torch.Generator(device='cuda')
        g.manual_seed(1000 + b)
        nb = torch.randn(blk, W, generator=g, device='cuda', dtype=torch.float32)
        g0, g1 = max(r0, b * blk), min(r1, (b + 1) * blk)
        y[g0 - r0:g1 - r0] += (NOISE * nb[g0 - b * blk:g1 - b * blk]).to(tdt)
    return y.reshape(-1)


def c5_arm(args, rank, local_rank, world, emit=True):
    """BASELINE.json configs[4]: ONE c5_size^2 fp32 TV-deconvolution row-decomposed over the GPUs (strong scaling)."""
    import torch
    import torch.distributed as dist
    from pycsou_b200 import _lib
    from pycsou_b200.func.loss import SquaredL2Loss
    from pycsou_b200.func.penalty import L21Norm, NonNegativeOrthant, tv_groups
    from pycsou_b200.linop.conv import Convolve2D
    from pycsou_b200.linop.diff import Gradient
    from pycsou_b200.opt.proxalgs import PDS
    from pycsou_b200.opt.slab import SlabComm, SlabLayout, release_arenas
    from pycsou_b200.util import synth
    torch.cuda.set_device(local_rank)
    _lib.lib()
    dtag = args.dtype if emit else 'f32'                        # configs[4] is fp32; as a secondary line always fp32
    tdt = torch.float32 if dtag == 'f32' else torch.float64
    itemsize = 4 if dtag == 'f32' else 8
    Hg = W = args.c5_size
    N = Hg * W
    iters = args.iters
    D = Dist(world)
    psf = synth.gaussian_psf_2d(15, 2.5)
    slab = SlabComm() if world > 1 else None
    lay = slab.layout(Hg, W) if slab is not None else SlabLayout(Hg, W, 1, 0)
    y_dev = c5_rows_on_device(Hg, W, lay.r0, lay.r1, psf, tdt, torch)
    y_host = torch.empty(y_dev.shape, dtype=tdt, pin_memory=True)
    y_host.copy_(y_dev)
    torch.cuda.synchronize()

    def build(y, output):
        A = Convolve2D(size=N, filter=psf, shape=(Hg, W))
        K = Gradient(shape=(Hg, W), kind='forward')
        K.lipschitz_cst = K.diff_lipschitz_cst = float(np.sqrt(8.))
        F = (1 / 2) * SquaredL2Loss(dim=N, data=y if slab is None else lay.local(y)) * A
        H = LAMBDA * L21Norm(dim=2 * N, groups=tv_groups(N))
        return PDS(dim=N, F=F, G=NonNegativeOrthant(dim=N), H=H, K=K, beta=1.0, max_iter=iters - 1,
                   min_iter=iters - 1, accuracy_threshold=0, verbose=None, dtype=tdt, output=output, slab=slab)

    solver = build(y_dev, 'torch')
    sampler = ClockSampler(local_rank).start() if (rank == 0 and emit) else None
    dev_ms, launches, est, diag = time_solves(solver, args.steps, args.warmup, D)
    cs = est['primal_variable'].double().sum().reshape(1)
    if world > 1:
        dist.all_reduce(cs)
    rel_last = float(diag['Relative Improvement (primal variable)'].iloc[-1])
    plan = type(solver._plan).__name__
    del est, solver, y_dev

    # end to end: this rank's rows of y from pinned host memory every step; the step's result read back is the
    # diagnostics table (global, all-reduced) -- the iterates of a 32768^2 problem stay distributed in HBM
    def e2e_step():
        s = build(y_host, 'torch')
        _, _, d = s.iterate()
        return float(d['Relative Improvement (primal variable)'].iloc[-1])
    e2e_s, _, last = time_e2e(e2e_step, args.steps, 1, D)
    clocks = sampler.stop() if sampler is not None else None
    assert abs(last - rel_last) <= 1e-6 * abs(rel_last) + 1e-12
    if slab is not None and emit:
        release_arenas()

    total_iters = args.steps * iters
    ms_iter = dev_ms / total_iters
    peak, peak_src = hbm_peak()
    alg = 7 * N * itemsize
    achieved = alg / (ms_iter * 1e-3) / 1e9                      # whole job, all GPUs
    line = {
        'metric': 'pds_iterations_per_sec', 'value': total_iters / (dev_ms * 1e-3), 'unit': 'iter/s', 'n_gpus': world,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': dev_ms / args.steps, 'higher_is_better': True,
        'scaling': 'strong', 'vs_baseline': None, 'dtype': dtag, 'data': 'synthetic',
        'config': {'workload': f'TV-deconvolution {Hg}x{W} (ONE image), Convolve2D 15x15 sep PSF + Gradient/L21 + '
                               'NonNegativeOrthant, Condat-Vu PDS (BASELINE.json configs[4])',
                   'iters_per_step': iters, 'ms_per_iteration': ms_iter, 'solver_path': plan,
                   'parallelism': f'{world} row slabs of {lay.n_own}x{W}, halo rows pushed over NVLink inside the update '
                                  'kernel, stop-test norms all-reduced' if world > 1 else 'single GPU',
                   'l2_policy': f'working set {alg / world / 1e9:.1f} GB per GPU per iteration > 126 MB L2',
                   'effective_gbs_total': achieved, 'checksum_x': float(cs.item()), 'rel_improvement_last': rel_last},
        'e2e': {'value': total_iters / e2e_s, 'unit': 'iter/s', 'h2d_bytes_per_step': lay.n_own * W * itemsize,
                'd2h_bytes_per_step': iters * 32, 'ms_per_step': 1e3 * e2e_s / args.steps,
                'result_read_back': 'diagnostics table (iterates stay distributed in HBM)'},
        'gpu_launches': launches,
        'roofline': {'bound': 'hbm', 'achieved': achieved / world, 'peak': peak, 'unit': 'GB/s',
                     'frac': achieved / world / peak, 'traffic': None, 'peak_source': peak_src,
                     'algorithmic_bytes': alg // world, 'per': 'one PDS iteration, per GPU'},
        'clocks': clocks,
    }
    if rank == 0 and emit:
        print(json.dumps(line), flush=True)
    return line


def _json_only_stdout():
    """The contract is ONE JSON line on stdout.  Native libraries write there too (NCCL prints its version banner on
    stdout at N > 1): point file descriptor 1 at stderr for the whole run and keep the real stdout for `print` only."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real, 'w', buffering=1)


def main():
    args = parse_args()
    _json_only_stdout()
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    if args.impl == 'reference':
        reference_arm(args, rank, world)
        return
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=torch.device('cuda', local_rank))
    try:
        if args.workload == 'c5':
            if args.iters == 100:
                args.iters = 50
            c5_arm(args, rank, local_rank, world)
        elif args.workload == 'c4':
            c4_arm(args, rank, local_rank, world)
        else:
            b200_arm(args, rank, local_rank, world)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == '__main__':
    main()
