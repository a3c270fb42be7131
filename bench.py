#!/usr/bin/env python
"""bench.py -- PDS iterations/sec and effective HBM GB/s on the 4096^2 TV-deconvolution workload (BASELINE.json).

Contract (one JSON line on stdout, printed by rank 0):

* a *step* is one solve: ``--iters`` (default 100) Condat-Vu updates of one synthetic ``--size``^2 TV-deconvolution
  problem (Convolve2D 15x15 PSF + Gradient/L21 + NonNegativeOrthant; BASELINE.json configs[2]).
* ``value``: iterations/sec with the observation ``y`` already resident in HBM and the iterates staying there
  (CUDA events around the K timed steps, max over ranks).
* ``e2e``: the same metric through the public API with HOST buffers: every step copies ``y`` host->device, runs the
  solve and copies the primal and dual iterates back (wall clock, max over ranks).
* ``roofline``: algorithmic bytes per iteration (7*N*s) / measured time of one iteration, against the
  measured HBM copy bandwidth in MEASURED_PEAKS.json.
* ``cpu_baseline``: the numpy oracle port of the reference path timed on this host (rank 0, N=1) on a bounded sample.
* ``--impl reference``: the CPU reference arm (oracle port; the pure-Python reference cannot travel to the GPU box).

Working set per iteration (x, y, z in/out: 7 arrays of N*s bytes = 470 MB fp32 / 940 MB fp64 at 4096^2) is larger
than the 126 MB L2, so no explicit L2 flush is needed between iterations.
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
# dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the fused kernel from the committed `ncu --set full`
# captures (profiles/r01_ncu_pds_{f32,f64}_v5.txt); bench.py cannot run under ncu itself
NCU_DRAM_BYTES = {('f32', 4096, 'sep'): 441034240, ('f64', 4096, 'sep'): 962170368}
NCU_SOURCE = 'profiles/r01_ncu_pds_f32_v14_ws.txt, r01_ncu_pds_f64_v13.txt (ncu --set full, one launch)'


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
    p.add_argument('--dtype', default='f32', choices=['f32', 'f64'])
    p.add_argument('--psf', default='sep', choices=['sep', 'nonsep'])
    p.add_argument('--no-cpu-baseline', action='store_true')
    p.add_argument('--no-secondary', action='store_true', help='skip the fp64 / kernel-only secondary measurements')
    p.add_argument('--cpu-size', type=int, default=0, help='CPU sample image size (default: --size)')
    p.add_argument('--cpu-iters', type=int, default=2, help='CPU reference updates per sample')
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


# ------------------------------------------------------------------------------------------------ workload
def np_dt(tag):
    return np.float32 if tag == 'f32' else np.float64


def make_problem(size, psf_kind, rank=0):
    """Synthetic observation of the rank's size x size row block (seed = rank, so a block is the same for any GPU
    count)."""
    from pycsou_b200.util import synth
    psf = synth.gaussian_psf_2d(15, 2.5) if psf_kind == 'sep' else synth.rotated_gaussian_psf_2d(15)
    _, psf, y = synth.tv_deconv_problem((size, size), psf, NOISE, seed=rank, dtype=np.float64)
    return psf, y


def make_solver(y, psf, size, tdtype, iters, output, slab=None):
    """The reference-facing construction (doc/general/examples.rst:19-47 pattern) for configs[2].

    With ``slab`` (N > 1 GPUs) the image is (world * size) x size, row-decomposed one size x size slab per GPU
    (configs[4] layout); ``y`` then holds this rank's rows."""
    from pycsou_b200.func.loss import SquaredL2Loss
    from pycsou_b200.func.penalty import L21Norm, NonNegativeOrthant, tv_groups
    from pycsou_b200.linop.conv import Convolve2D
    from pycsou_b200.linop.diff import Gradient
    from pycsou_b200.opt.proxalgs import PDS
    rows = size * (slab.world if slab is not None else 1)
    N = rows * size
    A = Convolve2D(size=N, filter=psf, shape=(rows, size))
    K = Gradient(shape=(rows, size), kind='forward')
    K.lipschitz_cst = K.diff_lipschitz_cst = float(np.sqrt(8.))
    data = y if slab is None else slab.layout(rows, size).local(y)
    F = (1 / 2) * SquaredL2Loss(dim=N, data=data) * A
    H = LAMBDA * L21Norm(dim=2 * N, groups=tv_groups(N))
    return PDS(dim=N, F=F, G=NonNegativeOrthant(dim=N), H=H, K=K, beta=1.0, max_iter=iters - 1, min_iter=iters - 1,
               accuracy_threshold=0, verbose=None, dtype=tdtype, output=output, slab=slab)


def algorithmic_bytes_per_iter(size, itemsize):
    """read x, y, z(2N); write x, z(2N) = 7*N*s."""
    return 7 * size * size * itemsize


def hbm_peak():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    try:
        with open(path) as f:
            return float(json.load(f)['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
    except Exception:
        return 6650.0, 'fallback (B200_PROFILING.md 6.65 TB/s)'


# ------------------------------------------------------------------------------------------------ CPU reference
def cpu_reference_sample(size, psf, y, iters):
    """``iters`` Condat-Vu updates of the numpy oracle port (reference path restated; fp64 as the reference computes).

    Returns (seconds, threads).  The path is numpy ufuncs + scipy.signal fftconvolve: effectively single threaded.
    """
    from oracle import np_oracle as O            # checker / CPU baseline only -- never on the product path
    N = size * size
    Fo = O.HalfSquaredL2Loss(y, O.Convolve2D(N, psf, (size, size)), G_lipschitz=1.0)
    Ho = LAMBDA * O.TVL21Norm(2 * N, N)
    Ko = O.Gradient((size, size), kind='forward')
    t0 = time.perf_counter()
    O.pds_iterate(N, F=Fo, G=O.NonNegativeOrthant(N), H=Ho, K=Ko, K_lipschitz=np.sqrt(8.), beta=1.0,
                  max_iter=iters - 1, min_iter=iters - 1, accuracy_threshold=0)
    return time.perf_counter() - t0, 1


def reference_arm(args, rank, world):
    """CPU reference arm: rank 0 only; each step = a bounded sample (--cpu-iters updates) of the same workload."""
    if rank != 0:
        return
    size = args.cpu_size or args.size
    psf, y = make_problem(size, args.psf, 0)
    for _ in range(args.warmup):
        cpu_reference_sample(size, psf, y, 1)
    t = 0.0
    for _ in range(args.steps):
        dt, cores = cpu_reference_sample(size, psf, y, args.cpu_iters)
        t += dt
    ips = args.steps * args.cpu_iters / t
    sample = f'{args.cpu_iters} PDS updates of the {size}x{size} fp64 problem per step (warm-up steps: 1 update each)'
    line = {
        'impl': 'reference', 'metric': 'pds_iterations_per_sec', 'value': ips, 'unit': 'iter/s', 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1e3 * t / args.steps, 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': {'workload': f'TV-deconvolution {size}x{size}, Convolve2D 15x15 {args.psf} PSF + Gradient/L21 + '
                               'NonNegativeOrthant, Condat-Vu PDS (BASELINE.json configs[2])',
                   'iters_per_step': args.cpu_iters, 'cpu': _cpu_model(), 'numpy': np.__version__},
        'cpu_baseline': {'value': ips, 'unit': 'iter/s', 'cores': cores, 'kind': 'port', 'sample': sample},
        'e2e': {'value': ips, 'unit': 'iter/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line), flush=True)


def _cpu_model():
    try:
        with open('/proc/cpuinfo') as f:
            for ln in f:
                if ln.startswith('model name'):
                    return f"{ln.split(':', 1)[1].strip()} x{os.cpu_count()}"
    except OSError:
        pass
    return f'unknown x{os.cpu_count()}'


# ------------------------------------------------------------------------------------------------ B200 arm
def b200_arm(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise RuntimeError('bench.py needs a CUDA device (no CPU fallback on the product path)')
    torch.cuda.set_device(local_rank)
    from pycsou_b200 import _lib
    _lib.lib()                                                    # fails loudly if libpcsb200.so is missing
    tdt = torch.float32 if args.dtype == 'f32' else torch.float64
    size, iters, N = args.size, args.iters, args.size * args.size
    itemsize = 4 if args.dtype == 'f32' else 8

    psf, y64 = make_problem(size, args.psf, rank)
    y_host = torch.from_numpy(y64.astype(np_dt(args.dtype))).pin_memory()
    y_dev = y_host.to('cuda', non_blocking=True)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    slab = None
    if world > 1:
        from pycsou_b200.opt.slab import SlabComm
        slab = SlabComm()

    # ---- device-resident: y and the iterates stay in HBM ----
    solver = make_solver(y_dev, psf, size, tdt, iters, 'torch', slab)
    for _ in range(args.warmup):
        solver.reset()
        solver.iterate()
    sampler = ClockSampler(local_rank).start() if rank == 0 else None
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = _lib.launch_count
    ev0.record()
    for _ in range(args.steps):
        solver.reset()
        est, _, _ = solver.iterate()
    ev1.record()
    barrier()
    launches = _lib.launch_count - launches0
    dev_ms = max_over_ranks(ev0.elapsed_time(ev1))
    plan = type(solver._plan).__name__ if getattr(solver, '_plan', None) is not None else 'generic (op-by-op)'
    cs = est['primal_variable'].double().sum().reshape(1)
    if world > 1:
        dist.all_reduce(cs)
    checksum = float(cs.item())

    # ---- end to end: host y in, host x and z out, every step ----
    def e2e_step():
        s = make_solver(y_host, psf, size, tdt, iters, 'numpy', slab)
        out, _, _ = s.iterate()
        return out
    # warm-up with the buffer lifetimes of the timed loop: the previous result is still alive while the next solve runs,
    # so two sets of pinned read-back buffers exist before the timing starts (a fresh 201 MB cudaHostAlloc costs ~75 ms)
    out = None
    for _ in range(max(args.warmup, 3)):
        out = e2e_step()
    barrier()
    t0 = time.perf_counter()
    e2e_each = []
    for _ in range(args.steps):
        t1 = time.perf_counter()
        out = e2e_step()
        e2e_each.append(round(1e3 * (time.perf_counter() - t1), 2))
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    clocks = sampler.stop() if sampler is not None else None
    assert out['primal_variable'].shape == (N,) and out['dual_variable'].shape == (2 * N,)

    total_iters = world * args.steps * iters
    ips = total_iters / (dev_ms * 1e-3)
    ms_iter = dev_ms / (args.steps * iters)
    peak, peak_src = hbm_peak()
    alg = algorithmic_bytes_per_iter(size, itemsize)
    achieved = alg / (ms_iter * 1e-3) / 1e9
    line = {
        'metric': 'pds_iterations_per_sec', 'value': ips, 'unit': 'iter/s', 'n_gpus': world, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': dev_ms / args.steps, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': args.dtype, 'data': 'synthetic',
        'config': {'workload': f'TV-deconvolution {size}x{size} per GPU, Convolve2D 15x15 {args.psf} PSF + '
                               'Gradient/L21 + NonNegativeOrthant, Condat-Vu PDS (BASELINE.json configs[2])',
                   'iters_per_step': iters, 'ms_per_iteration': ms_iter, 'solver_path': plan,
                   'parallelism': (f'one {world * size}x{size} image row-decomposed into {world} slabs of {size}x{size}, '
                                   'halo rows pushed to the neighbour GPUs over NVLink inside the update kernel, '
                                   'stop-test norms all-reduced; value counts slab-iterations (N x image iterations/s)')
                   if world > 1 else 'single GPU',
                   'l2_policy': f'working set {alg / 1e6:.0f} MB per iteration > 126 MB L2 (no flush needed)',
                   'effective_gbs_per_gpu': achieved, 'checksum_x': checksum},
        'e2e': {'value': total_iters / e2e_s, 'unit': 'iter/s', 'h2d_bytes_per_step': N * itemsize,
                'd2h_bytes_per_step': 3 * N * itemsize, 'ms_per_step': 1e3 * e2e_s / args.steps,
                'ms_each_step_rank0': e2e_each},
        'gpu_launches': launches,
        'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak,
                     'traffic': NCU_DRAM_BYTES.get((args.dtype, size, args.psf)), 'traffic_source': NCU_SOURCE,
                     'peak_source': peak_src, 'algorithmic_bytes': alg,
                     'per': 'one PDS iteration (all kernels of the update)'},
        'clocks': clocks,
    }
    if slab is not None:
        from pycsou_b200.opt.slab import release_arenas
        release_arenas()
    if rank == 0:
        if world == 1 and not args.no_secondary:
            line['secondary'] = secondary_measurements(args, psf, y64, torch)
        if world == 1 and not args.no_cpu_baseline:
            csize = args.cpu_size or size
            cpsf, cy = (psf, y64) if csize == size else make_problem(csize, args.psf, 0)
            cpu_reference_sample(csize, cpsf, cy, 1)                              # warm-up (fft plans, page faults)
            dt, cores = cpu_reference_sample(csize, cpsf, cy, args.cpu_iters)
            line['cpu_baseline'] = {'value': args.cpu_iters / dt, 'unit': 'iter/s', 'cores': cores, 'kind': 'port',
                                    'sample': f'{args.cpu_iters} PDS updates of the {csize}x{csize} fp64 problem '
                                              f'(numpy oracle port, {_cpu_model()})'}
        print(json.dumps(line), flush=True)



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


def c4_arm(args, rank, local_rank, world):
    """BASELINE.json configs[3]: independent compressed-sensing MRI problems, one subset per GPU, no data-path
    collective.  A step = --iters Chambolle-Pock updates of every problem of the rank."""
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise RuntimeError('bench.py needs a CUDA device (no CPU fallback on the product path)')
    torch.cuda.set_device(local_rank)
    from pycsou_b200 import _lib
    _lib.lib()
    size, iters, P = args.c4_size, args.iters, args.c4_problems
    N = size * size

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    probs = [c4_problem(size, rank * P + i) for i in range(P)]
    m_tot = sum(int(pr[0].sum()) for pr in probs)
    y_host = [torch.from_numpy(pr[1].astype(np.complex64)).pin_memory() for pr in probs]
    solvers = [c4_solver(size, pr[0], yh.to('cuda'), pr[2], iters, 'torch') for pr, yh in zip(probs, y_host)]
    for _ in range(args.warmup):
        for s in solvers:
            s.reset()
            s.iterate()
    sampler = ClockSampler(local_rank).start() if rank == 0 else None
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = _lib.launch_count
    ev0.record()
    for _ in range(args.steps):
        for s in solvers:
            s.reset()
            est, _, _ = s.iterate()
    ev1.record()
    barrier()
    launches = _lib.launch_count - launches0
    dev_ms = max_over_ranks(ev0.elapsed_time(ev1))
    plan = type(solvers[0]._plan).__name__ if solvers[0]._plan is not None else 'generic (op-by-op)'
    cs = est['primal_variable'].abs().double().sum().reshape(1)
    checksum = float(cs.item())

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
    out = None
    for _ in range(3):                                       # (previous result alive during the next solve, as in the timed loop)
        out = e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out = e2e_step()
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    clocks = sampler.stop() if sampler is not None else None
    assert out['primal_variable'].shape == (N,)

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
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            mask, y, eps = probs[0]
            c4_cpu_sample(size, mask, y, eps, 1)
            dt, cores = c4_cpu_sample(size, mask, y, eps, args.cpu_iters)
            line['cpu_baseline'] = {'value': args.cpu_iters / dt, 'unit': 'iter/s', 'cores': cores, 'kind': 'port',
                                    'sample': f'{args.cpu_iters} Chambolle-Pock updates of one {size}x{size} complex128 '
                                              f'problem (numpy oracle port, {_cpu_model()})'}
        print(json.dumps(line), flush=True)


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


def c5_arm(args, rank, local_rank, world):
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
    tdt = torch.float32 if args.dtype == 'f32' else torch.float64
    itemsize = 4 if args.dtype == 'f32' else 8
    Hg = W = args.c5_size
    N = Hg * W
    iters = args.iters
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

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    solver = build(y_dev, 'torch')
    for _ in range(args.warmup):
        solver.reset()
        solver.iterate()
    sampler = ClockSampler(local_rank).start() if rank == 0 else None
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = _lib.launch_count
    ev0.record()
    for _ in range(args.steps):
        solver.reset()
        est, _, diag = solver.iterate()
    ev1.record()
    barrier()
    launches = _lib.launch_count - launches0
    dev_ms = max_over_ranks(ev0.elapsed_time(ev1))
    cs = est['primal_variable'].double().sum().reshape(1)
    if world > 1:
        dist.all_reduce(cs)
    rel_last = float(diag['Relative Improvement (primal variable)'].iloc[-1])
    plan = type(solver._plan).__name__
    del est, solver

    # end to end: this rank's rows of y from pinned host memory every step; the step's result read back is the
    # diagnostics table (global, all-reduced) -- the iterates of a 32768^2 problem stay distributed in HBM
    def e2e_step():
        s = build(y_host, 'torch')
        _, _, d = s.iterate()
        return float(d['Relative Improvement (primal variable)'].iloc[-1])
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        last = e2e_step()
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    clocks = sampler.stop() if sampler is not None else None
    assert abs(last - rel_last) <= 1e-6 * abs(rel_last) + 1e-12
    if slab is not None:
        release_arenas()

    total_iters = args.steps * iters
    ms_iter = dev_ms / total_iters
    peak, peak_src = hbm_peak()
    alg = 7 * N * itemsize
    achieved = alg / (ms_iter * 1e-3) / 1e9                      # whole job, all GPUs
    line = {
        'metric': 'pds_iterations_per_sec', 'value': total_iters / (dev_ms * 1e-3), 'unit': 'iter/s', 'n_gpus': world,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': dev_ms / args.steps, 'higher_is_better': True,
        'scaling': 'strong', 'vs_baseline': None, 'dtype': args.dtype, 'data': 'synthetic',
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
    if rank == 0:
        print(json.dumps(line), flush=True)


def secondary_measurements(args, psf, y64, torch):
    """Other precision of the same workload (device-resident iterations/sec); reported, not the headline."""
    out = {}
    other = 'f64' if args.dtype == 'f32' else 'f32'
    tdt = torch.float64 if other == 'f64' else torch.float32
    isz = 8 if other == 'f64' else 4
    y_dev = torch.from_numpy(y64.astype(np_dt(other))).to('cuda')
    solver = make_solver(y_dev, psf, args.size, tdt, args.iters, 'torch')
    for _ in range(2):
        solver.reset()
        solver.iterate()
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    reps = 3
    for _ in range(reps):
        solver.reset()
        solver.iterate()
    ev1.record()
    torch.cuda.synchronize()
    ms_iter = ev0.elapsed_time(ev1) / (reps * args.iters)
    gbs = algorithmic_bytes_per_iter(args.size, isz) / (ms_iter * 1e-3) / 1e9
    peak, _ = hbm_peak()
    out[other] = {'iter_per_s': 1e3 / ms_iter, 'ms_per_iteration': ms_iter, 'effective_gbs': gbs,
                  'roofline_frac': gbs / peak}
    return out


def main():
    args = parse_args()
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
