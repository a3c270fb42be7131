"""Generate tests/golden/*.npz by running the UNMODIFIED reference solver classes.

Run in the build container only (needs /root/reference):  ``python tests/golden/gen_golden.py``.

The reference (``pycsou.opt.proxalgs``, ``pycsou.func``, ``pycsou.core``, ``pycsou.linop.sampling``) is
imported as-is through ``oracle.ref_shim``; the operators it delegates to pylops use
``oracle.pylops_standin``; per-pixel-group L21 uses ``ref_shim.vectorised_l21`` (the reference's own
L21Norm is Theta(N^2), it is used as-is in the ``*_reflL21`` case at a tiny size to pin the vectorised one).

Each fixture stores the explicit step sizes used and the iterates after exactly k updates
(``max_iter = min_iter = k-1``, ``accuracy_threshold = 0``; core/solver.py:65-66).  Inputs are NOT
stored: they are regenerated from ``pycsou_b200.util.synth`` (deterministic) and guarded by a checksum.
"""
from __future__ import annotations

import json
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
warnings.filterwarnings('ignore')

from oracle.ref_shim import load_reference, vectorised_l21  # noqa: E402
from pycsou_b200.util import synth  # noqa: E402

load_reference()
from pycsou.core.linop import UnitaryOperator  # noqa: E402
from pycsou.func.base import ProxFuncHStack  # noqa: E402
from pycsou.func.loss import L1Loss, L2BallLoss, SquaredL2Loss  # noqa: E402
from pycsou.func.penalty import L1Norm, L21Norm, NonNegativeOrthant, Segment  # noqa: E402
from pycsou.linop.base import LinOpVStack  # noqa: E402
from pycsou.linop.conv import Convolve1D, Convolve2D  # noqa: E402
from pycsou.linop.diff import FirstDerivative, Gradient  # noqa: E402
from pycsou.linop.sampling import DownSampling, Masking  # noqa: E402
from pycsou.opt.proxalgs import APGD, CPS, DRS, FBS, PDS  # noqa: E402


checksum = synth.checksum


def save(name, meta, arrays):
    meta = dict(meta)
    path = os.path.join(HERE, name + '.npz')
    np.savez_compressed(path, meta=json.dumps(meta), **arrays)
    print(f'{name}: {os.path.getsize(path) / 1024:.1f} KiB  keys={sorted(arrays)[:6]}...')


def run_pds_at(ks, make):
    out = {}
    for k in ks:
        solver = make(k - 1)
        est, conv, diag = solver.iterate()
        assert solver.iter == k
        out[f'x_k{k}'] = np.asarray(est['primal_variable'])
        if est['dual_variable'] is not None:
            out[f'z_k{k}'] = np.asarray(est['dual_variable'])
    return out, solver


# ---------------------------------------------------------------------------------------------- C1
def c1_lasso():
    truth, h, y = synth.spikes_1d(4096, 32, 0)
    N = y.size
    Gop = Convolve1D(size=N, filter=h)
    lam = 0.1 * np.max(np.abs(Gop.adjoint(y)))
    arrays = {}
    for acc, ks in (('CD', (1, 2, 10, 100, 500)), ('BT', (1, 10, 100)), (None, (1, 10, 100))):
        for k in ks:
            F = (1 / 2) * SquaredL2Loss(dim=N, data=y) * Gop
            G = lam * L1Norm(dim=N)
            s = APGD(dim=N, F=F, G=G, tau=1.0, acceleration=acc, beta=1.0, max_iter=k - 1, min_iter=k - 1,
                     accuracy_threshold=0, verbose=None, d=75.)
            est, _, diag = s.iterate()
            assert s.iter == k
            arrays[f'x_{acc}_k{k}'] = est['iterand']
            if k == ks[-1]:
                arrays[f'aux_{acc}_k{k}'] = est['past_aux']
                arrays[f't_{acc}_k{k}'] = np.float64(est['past_t'])
                arrays[f'relimp_{acc}'] = np.asarray(diag['Relative Improvement'], dtype=np.float64)
    save('c1_lasso_apgd', dict(N=N, lam=lam, tau=1.0, d=75., y_checksum=checksum(y)), arrays)


# ---------------------------------------------------------------------------------------------- C2
def tv_common(shape, kind):
    N = shape[0] * shape[1]
    K = Gradient(shape=shape, kind=kind, edge=True)
    K.lipschitz_cst = K.diff_lipschitz_cst = np.sqrt(8.)
    return N, K


def c2_tv_denoise():
    shape = (48, 64)
    _, y = synth.tv_denoise_problem(shape, 0.1, 0)
    N, K = tv_common(shape, 'forward')
    lam = 0.1

    def make(m, G=None, H=None, rho=None):
        F = (1 / 2) * SquaredL2Loss(dim=N, data=y)
        H = lam * vectorised_l21(2 * N, N) if H is None else H
        return PDS(dim=N, F=F, G=G, H=H, K=K, rho=rho, max_iter=m, min_iter=m, accuracy_threshold=0, verbose=None)

    arrays, s = run_pds_at((1, 2, 10, 50), make)
    meta = dict(shape=shape, lam=lam, tau=float(s.tau), sigma=float(s.sigma), rho=float(s.rho), beta=float(s.beta),
                kind='forward', y_checksum=checksum(y))
    # reference stop rule with defaults (max_iter=500, min_iter=10, thr=1e-3)
    F = (1 / 2) * SquaredL2Loss(dim=N, data=y)
    s2 = PDS(dim=N, F=F, H=lam * vectorised_l21(2 * N, N), K=K, verbose=None)
    est, _, diag = s2.iterate()
    arrays['x_stop'], arrays['z_stop'] = est['primal_variable'], est['dual_variable']
    arrays['relimp_primal'] = np.asarray(diag['Relative Improvement (primal variable)'], dtype=np.float64)
    arrays['relimp_dual'] = np.asarray(diag['Relative Improvement (dual variable)'], dtype=np.float64)
    meta['n_updates_stop'] = int(s2.iter)
    save('c2_tv_denoise_pds', meta, arrays)

    # variants: positivity / box constraint on G, anisotropic TV (L1 on centred gradient)
    a2, s = run_pds_at((1, 10, 50), lambda m: make(m, G=NonNegativeOrthant(dim=N)))
    save('c2_tv_denoise_pds_nonneg', dict(meta, tau=float(s.tau), sigma=float(s.sigma), rho=float(s.rho)), a2)
    a3, s = run_pds_at((1, 10, 50), lambda m: make(m, G=Segment(dim=N, a=0.1, b=0.7)))
    save('c2_tv_denoise_pds_segment', dict(meta, a=0.1, b=0.7), a3)
    Nc, Kc = tv_common(shape, 'centered')

    def make_c(m):
        F = (1 / 2) * SquaredL2Loss(dim=N, data=y)
        return PDS(dim=N, F=F, G=None, H=lam * L1Norm(dim=2 * N), K=Kc, max_iter=m, min_iter=m,
                   accuracy_threshold=0, verbose=None)

    a4, s = run_pds_at((1, 10, 50), make_c)
    save('c2_aniso_tv_centered_pds', dict(meta, kind='centered', tau=float(s.tau), sigma=float(s.sigma)), a4)

    # the reference's own (Theta(N^2)) L21Norm inside the solver at a tiny size: pins the vectorised restatement
    shape_s = (12, 16)
    _, ys = synth.tv_denoise_problem(shape_s, 0.1, 0)
    Ns, Ks = tv_common(shape_s, 'forward')

    def make_s(m):
        F = (1 / 2) * SquaredL2Loss(dim=Ns, data=ys)
        H = lam * L21Norm(dim=2 * Ns, groups=np.tile(np.arange(Ns), 2))
        return PDS(dim=Ns, F=F, H=H, K=Ks, max_iter=m, min_iter=m, accuracy_threshold=0, verbose=None)

    a5, s = run_pds_at((1, 10, 30), make_s)
    save('c2_tv_denoise_pds_refL21', dict(shape=shape_s, lam=lam, tau=float(s.tau), sigma=float(s.sigma),
                                          rho=float(s.rho), kind='forward', y_checksum=checksum(ys)), a5)


# ---------------------------------------------------------------------------------------------- C3
def c3_tv_deconv():
    for tag, psf in (('sep', synth.gaussian_psf_2d(15, 2.5)), ('nonsep', synth.rotated_gaussian_psf_2d(15))):
        shape = (48, 64)
        _, h, y = synth.tv_deconv_problem(shape, psf, 0.02, 0)
        N, K = tv_common(shape, 'forward')
        A = Convolve2D(size=N, filter=h, shape=shape)
        lam = 0.02

        def make(m):
            F = (1 / 2) * SquaredL2Loss(dim=N, data=y) * A
            return PDS(dim=N, F=F, G=NonNegativeOrthant(dim=N), H=lam * vectorised_l21(2 * N, N), K=K, beta=1.0,
                       max_iter=m, min_iter=m, accuracy_threshold=0, verbose=None)

        arrays, s = run_pds_at((1, 2, 10, 50), make)
        # objective value after 50 updates (for the fp32 1e-4 objective criterion)
        x = arrays['x_k50']
        obj = 0.5 * np.linalg.norm(A(x) - y) ** 2 + lam * vectorised_l21(2 * N, N)(K(x))
        save(f'c3_tv_deconv_pds_{tag}', dict(shape=shape, lam=lam, tau=float(s.tau), sigma=float(s.sigma),
                                             rho=float(s.rho), beta=1.0, kind='forward', objective_k50=float(obj),
                                             y_checksum=checksum(y)), arrays)


# ------------------------------------------------------------------- examples.rst PDS + CPS/DRS/FBS
def misc_solvers():
    # doc/general/examples.rst:19-47 (1-D penalised basis pursuit), explicit Lipschitz constants
    x = np.repeat([0, 2, 1, 3, 0, 2, 0], 10).astype(float)
    D = FirstDerivative(size=x.size, kind='forward')
    D.lipschitz_cst = D.diff_lipschitz_cst = 2.0
    Gop = DownSampling(size=x.size, downsampling_factor=3)
    Gop.lipschitz_cst = Gop.diff_lipschitz_cst = 1.0
    y = Gop(x)
    F = (1 / 2) * SquaredL2Loss(dim=Gop.shape[0], data=y) * Gop
    pds = PDS(dim=Gop.shape[1], F=F, G=0.01 * L1Norm(dim=Gop.shape[1]), H=0.1 * L1Norm(dim=D.shape[0]), K=D,
              verbose=None)
    est, _, diag = pds.iterate()
    save('examples_pbp_pds', dict(n_updates=int(pds.iter), tau=float(pds.tau), sigma=float(pds.sigma),
                                  rho=float(pds.rho)),
         dict(x=est['primal_variable'], z=est['dual_variable'],
              relimp_primal=np.asarray(diag['Relative Improvement (primal variable)'], dtype=np.float64)))

    shape = (24, 32)
    N = shape[0] * shape[1]
    _, y = synth.tv_denoise_problem(shape, 0.1, 1)
    K = Gradient(shape=shape, kind='forward')
    K.lipschitz_cst = K.diff_lipschitz_cst = np.sqrt(8.)
    arrays = {}
    # CPS: G = L1 data term (ProximableLoss), H = TV
    for k in (1, 10, 40):
        s = CPS(dim=N, G=L1Loss(dim=N, data=y), H=0.3 * vectorised_l21(2 * N, N), K=K, max_iter=k - 1,
                min_iter=k - 1, accuracy_threshold=0, verbose=None)
        est, _, _ = s.iterate()
        arrays[f'cps_x_k{k}'], arrays[f'cps_z_k{k}'] = est['primal_variable'], est['dual_variable']
    meta = dict(shape=shape, cps_tau=float(s.tau), cps_sigma=float(s.sigma), cps_rho=float(s.rho))
    # DRS: G = L1 data term, H = Segment indicator, K = identity
    for k in (1, 10, 40):
        s = DRS(dim=N, G=L1Loss(dim=N, data=y), H=Segment(dim=N, a=0.2, b=0.6), tau=0.7, max_iter=k - 1,
                min_iter=k - 1, accuracy_threshold=0, verbose=None)
        est, _, _ = s.iterate()
        arrays[f'drs_x_k{k}'], arrays[f'drs_z_k{k}'] = est['primal_variable'], est['dual_variable']
    # FBS: F = 1/2||x-y||^2, G = lam L1
    for k in (1, 10, 40):
        s = FBS(dim=N, F=(1 / 2) * SquaredL2Loss(dim=N, data=y), G=0.2 * L1Norm(dim=N), max_iter=k - 1,
                min_iter=k - 1, accuracy_threshold=0, verbose=None)
        est, _, _ = s.iterate()
        arrays[f'fbs_x_k{k}'] = est['primal_variable']
    meta.update(fbs_tau=float(s.tau), fbs_rho=float(s.rho))
    save('misc_cps_drs_fbs', meta, arrays)


# ---------------------------------------------------------------------------------------------- C4
class _FFT2(UnitaryOperator):
    """Orthonormal 2-D DFT against the reference's UnitaryOperator API (the reference ships no DFT)."""

    def __init__(self, shape):
        self.img_shape = shape
        super().__init__(size=shape[0] * shape[1], dtype=np.complex128)

    def __call__(self, x):
        return np.fft.fft2(x.reshape(self.img_shape), norm='ortho').reshape(-1)

    def adjoint(self, y):
        return np.fft.ifft2(y.reshape(self.img_shape), norm='ortho').reshape(-1)


def c4_csmri():
    shape = (32, 48)
    N = shape[0] * shape[1]
    rng = np.random.default_rng(7)
    truth = synth.phantom(shape).reshape(-1) * np.exp(1j * 2 * np.pi * (np.arange(N) / N))
    mask = rng.random(N) < 0.3
    Fop = _FFT2(shape)
    M = Masking(size=N, sampling_bool=mask, dtype=np.complex128)
    m = M.shape[0]
    y = M(Fop(truth)) + 0.01 * (rng.standard_normal(m) + 1j * rng.standard_normal(m))
    Gr = Gradient(shape=shape, kind='forward', dtype='complex128')
    Gr.lipschitz_cst = Gr.diff_lipschitz_cst = np.sqrt(8.)
    MF = M * Fop
    MF.lipschitz_cst = MF.diff_lipschitz_cst = 1.0
    arrays = {}
    lam, eps = 0.05, 0.01 * np.sqrt(2 * m)
    for k in (1, 5, 20):
        K = LinOpVStack(MF, Gr)
        H = ProxFuncHStack(L2BallLoss(dim=m, data=y, radius=eps), lam * vectorised_l21(2 * N, N))
        s = CPS(dim=N, G=None, H=H, K=K, tau=1 / 3, sigma=1 / 3, rho=1, x0=np.zeros(N, dtype=np.complex128),
                z0=np.zeros(m + 2 * N, dtype=np.complex128), max_iter=k - 1, min_iter=k - 1, accuracy_threshold=0,
                verbose=None)
        est, _, _ = s.iterate()
        arrays[f'x_k{k}'], arrays[f'z_k{k}'] = est['primal_variable'], est['dual_variable']
    arrays['mask'] = mask
    arrays['y'] = y
    save('c4_csmri_cps', dict(shape=shape, lam=lam, eps=float(eps), tau=1 / 3, sigma=1 / 3, rho=1, m=int(m)), arrays)


from tests.oracle_cases import GD_CASES, GL_CASES, LEAF_RADII, LEAF_TAUS, leaf_inputs  # noqa: E402


def leaves():
    """Outputs of the reference's own proxes / operators for the SURVEY 8(f) rank-2 and rank-3 leaves."""
    from pycsou.func import loss as L, penalty as P
    from pycsou.linop.diff import GeneralisedDerivative, GeneralisedLaplacian
    from pycsou.math.prox import proj_l1_ball
    x, d, y = leaf_inputs()
    n = x.size
    arrays = {}
    for i, tau in enumerate(LEAF_TAUS):
        arrays[f'logbarrier_{i}'] = P.LogBarrier(n).prox(x.copy(), tau)
        arrays[f'shannon_{i}'] = P.ShannonEntropy(n).prox(x.copy(), tau)
        arrays[f'kl_{i}'] = L.KLDivergence(n, d).prox(x.copy(), tau)
        arrays[f'linfty_{i}'] = P.LInftyNorm(n).prox(x.copy(), tau)
        arrays[f'sql1_sort_{i}'] = P.SquaredL1Norm(n, 'sort').prox(x.copy(), tau)
        arrays[f'sql1_root_{i}'] = P.SquaredL1Norm(n, 'root').prox(x.copy(), tau)
        arrays[f'linfty_loss_{i}'] = L.LInftyLoss(n, y).prox(x.copy(), tau)
    for i, r in enumerate(LEAF_RADII):
        arrays[f'l1ball_{i}'] = np.array(proj_l1_ball(x.copy(), r))
        arrays[f'l1ball_loss_{i}'] = np.array(L.L1BallLoss(n, y, radius=r).prox(x.copy(), 1.0))
    xp = np.abs(x) + 0.1
    values = dict(logbarrier=float(P.LogBarrier(n)(xp)), shannon=float(P.ShannonEntropy(n)(xp)),
                  kl=float(L.KLDivergence(n, d)(xp)), linfty=float(P.LInftyNorm(n)(x)),
                  sql1=float(P.SquaredL1Norm(n)(x)))
    for i, kw in enumerate(GD_CASES):
        for kd in ('forward', 'backward', 'centered'):
            A = GeneralisedDerivative(size=n, shape=(20, 20), axis=1, step=0.5, kind_diff=kd, **kw)
            arrays[f'gd_{i}_{kd}_fwd'], arrays[f'gd_{i}_{kd}_adj'] = A(x), A.adjoint(y)
    for i, kw in enumerate(GL_CASES):
        A = GeneralisedLaplacian(shape=(20, 20), **kw)
        arrays[f'gl_{i}_fwd'], arrays[f'gl_{i}_adj'] = A(x), A.adjoint(y)
    save('leaves_rank23', dict(n=n, seed=31, x_checksum=checksum(x), d_checksum=checksum(d), y_checksum=checksum(y),
                               values=values), arrays)


if __name__ == '__main__':
    if sys.argv[1:] == ['leaves']:
        leaves()
        sys.exit(0)
    c1_lasso()
    c2_tv_denoise()
    c3_tv_deconv()
    misc_solvers()
    c4_csmri()
    leaves()
