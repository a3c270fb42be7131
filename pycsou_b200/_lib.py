"""ctypes binding of libpcsb200.so (C ABI declared in include/pcsb200.h).

The product path has NO fallback: if the library is missing or a call fails, an exception is raised.
Signatures are declared here once; ``call(name, *args)`` checks the return code and raises
``ValueError`` (bad argument) or ``RuntimeError`` (CUDA failure) with ``pcs_last_error()``.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('PCS_LIB_PATH') or os.path.join(_HERE, 'libpcsb200.so')   # override: A/B of builds

_lib = None
launch_count = 0   # number of C-ABI compute calls issued (bench.py reports it as gpu_launches evidence)

P, I64, I32, F64 = C.c_void_p, C.c_int64, C.c_int, C.c_double

# name -> argtypes (without dtype suffix); entries listed in REAL get _f32 and _f64 variants.
REAL = {
    'pcs_lincomb2': [P, F64, P, F64, P, I64, P],
    'pcs_lincomb3': [P, F64, P, F64, P, F64, P, I64, P],
    'pcs_scale': [P, F64, P, I64, P],
    'pcs_mul': [P, P, P, I64, P],
    'pcs_prox_l1': [P, P, F64, I64, P],
    'pcs_proj_nonneg': [P, P, I64, P],
    'pcs_prox_logbarrier': [P, P, F64, I64, P],
    'pcs_prox_shannon': [P, P, F64, I64, P],
    'pcs_prox_kl': [P, P, P, F64, I64, P],
    'pcs_reduce_neglog': [P, I64, P, P, P],
    'pcs_reduce_xlogx': [P, I64, P, P, P],
    'pcs_reduce_kl': [P, P, I64, P, P, P],
    'pcs_proj_segment': [P, P, F64, F64, I64, P],
    'pcs_prox_l21': [P, P, F64, I64, I32, I32, P],
    'pcs_prox_l21_csr': [P, P, F64, P, P, I64, I32, P],
    'pcs_reduce_sumgroupnorm_csr': [P, P, P, I64, I32, P, P, P],
    'pcs_reduce_sumsq': [P, I64, P, P, P],
    'pcs_reduce_sumabs': [P, I64, P, P, P],
    'pcs_reduce_diffsumsq': [P, P, I64, P, P, P],
    'pcs_reduce_dot': [P, P, I64, P, P, P],
    'pcs_reduce_minmax': [P, I64, P, P, P],
    'pcs_pool_fwd': [P, P, I64, I64, I64, I64, I64, I64, F64, P],
    'pcs_pool_adj': [P, P, I64, I64, I64, I64, I64, I64, P],
    'pcs_threshold_newton_step': [P, I64, F64, F64, P, I32, I32, P, P],
    'pcs_reduce_sumgroupnorm': [P, I64, I32, I32, P, P, P],
    'pcs_scale_by_norm': [P, P, P, F64, I32, I64, P],
    'pcs_prox_l2': [P, P, P, F64, I64, P],
    'pcs_proj_l2_ball': [P, P, P, F64, I64, P],
    'pcs_firstderiv_fwd': [P, P, I64, I64, I64, I32, I32, F64, I32, P],
    'pcs_firstderiv_adj': [P, P, I64, I64, I64, I32, I32, F64, I32, P],
    'pcs_secondderiv_fwd': [P, P, I64, I64, I64, I32, I32, F64, I32, P],
    'pcs_secondderiv_adj': [P, P, I64, I64, I64, I32, I32, F64, I32, P],
    'pcs_cumsum_fwd': [P, P, I64, I64, I64, F64, P],
    'pcs_cumsum_adj': [P, P, I64, I64, I64, F64, P],
    'pcs_grad2d_fwd': [P, P, I64, I64, I32, I32, F64, F64, P],
    'pcs_grad2d_adj': [P, P, I64, I64, I32, I32, F64, F64, P],
    'pcs_laplacian2d_fwd': [P, P, I64, I64, F64, F64, F64, F64, I32, P],
    'pcs_laplacian2d_adj': [P, P, I64, I64, F64, F64, F64, F64, I32, P],
    'pcs_conv1d_fwd': [P, P, P, I32, I32, I64, I64, I64, P],
    'pcs_conv1d_adj': [P, P, P, I32, I32, I64, I64, I64, P],
    'pcs_conv2d_direct_fwd': [P, P, P, I32, I32, I32, I32, I64, I64, P],
    'pcs_conv2d_direct_adj': [P, P, P, I32, I32, I32, I32, I64, I64, P],
    'pcs_conv2d_sep_fwd': [P, P, P, P, I32, I32, I32, I32, P, I64, I64, P],
    'pcs_conv2d_sep_adj': [P, P, P, P, I32, I32, I32, I32, P, I64, I64, P],
    'pcs_spectral_mul': [P, P, I32, I64, P],
    'pcs_crop_shift': [P, P, P, I64, I64, I64, I64, I64, I64, F64, P],
    # fused Condat-Vu TV iteration (csrc/pcs_pds.cu); argument order as in include/pcsb200.h
    'pcs_pds_tv_step': [P] * 5 + [I64] * 6 + [I32, I32, P, P, P] + [I32] * 4 + [I32, F64, F64, I32] + [F64] * 6 +
                       [I32] * 3 + [P, P, P],
    'pcs_pds_tv_run': [P] * 5 + [I64, I64] + [I32, I32, P, P, P] + [I32] * 4 + [I32, F64, F64, I32] + [F64] * 6 +
                      [I32] * 3 + [I64, P, P, P],
    # two-kernel Condat-Vu TV update for any 2-D Gradient kind (csrc/pcs_pds_grad.cu)
    'pcs_pds_grad_primal_step': [P] * 5 + [I64, I64, I32, I32, I32, I32] + [F64] * 6 + [P, P, P],
    'pcs_pds_grad_dual_step': [P] * 3 + [I64, I64, I32, I32, I32] + [F64] * 5 + [P, P, P],
    # a whole APGD solve (flat Convolve1D data term) in one launch (csrc/pcs_apgd.cu)
    'pcs_apgd_conv1d_run': [P, P, P, P, I32, I32, I64, I32, F64, F64, I32, F64, F64, I64, I64, I64, F64, P, P, P],
    # ... fused with the halo push / flag handshake of the slab decomposition (last argument: HOST int64[18])
    'pcs_pds_tv_step_slab': [P] * 5 + [I64] * 6 + [I32, I32, P, P, P] + [I32] * 4 + [I32, F64, F64, I32] + [F64] * 6 +
                            [I32] * 3 + [P, P, P, P],
}
PLAIN = {
    'pcs_take': ([P, P, P, I64, I64, I64, I64, I32, P], I32),
    'pcs_put': ([P, P, P, I64, I64, I64, I64, I32, P], I32),
    'pcs_mask_count': ([P, I64, P, P, P], I32),
    'pcs_mask_indices': ([P, I64, P, P, P], I32),
    'pcs_mask_workspace_bytes': ([I64], I64),
    'pcs_mask_static_max_n': ([], I64),
    'pcs_mask_to_indices_static': ([P, I64, P, P, I64, P], I32),
    'pcs_reduce_workspace_bytes': ([], I64),
    'pcs_pds_tv_workspace_bytes': ([], I64),
    'pcs_pds_tv_last_launch': ([P], I32),
    'pcs_apgd_last_launch': ([P], I32),
    # peer (NVLink) memory for the slab-decomposed solvers (csrc/pcs_peer.cu)
    'pcs_peer_handle_bytes': ([], I32),
    'pcs_peer_alloc': ([P, I64, P], I32),
    'pcs_peer_open': ([P, P], I32),
    'pcs_peer_close': ([P], I32),
    'pcs_peer_free': ([P], I32),
    'pcs_peer_can_access': ([I32, I32], I32),
    'pcs_apgd_conv1d_run_max_n': ([], I64),
    # fused ChambollePock iteration for compressed-sensing MRI (csrc/pcs_cps.cu); complex64 / complex128
    'pcs_cps_mri_workspace_bytes': ([], I64),
    'pcs_cps_mri_scatter_c64': ([P, P, P, I64, I64, P], I32),
    'pcs_cps_mri_scatter_c128': ([P, P, P, I64, I64, P], I32),
    'pcs_cps_mri_tv_step_c64': ([P] * 6 + [I64, I64] + [F64] * 7 + [P, P, P], I32),
    'pcs_cps_mri_tv_step_c128': ([P] * 6 + [I64, I64] + [F64] * 7 + [P, P, P], I32),
    'pcs_cps_mri_ball_step_c64': ([P] * 6 + [I64, F64, F64, F64, F64, P, P, P], I32),
    'pcs_cps_mri_ball_step_c128': ([P] * 6 + [I64, F64, F64, F64, F64, P, P, P], I32),
    'pcs_version': ([], I32),
    'pcs_num_sms': ([], I32),
    'pcs_set_num_sms': ([I32], I32),
}


def declared_symbols():
    """All C symbols this binding expects (tests check each is exported by the .so and declared in the header)."""
    names = list(PLAIN)
    for base in REAL:
        names += [base + '_f32', base + '_f64']
    return names + ['pcs_last_error']


def register_real(name, argtypes):
    REAL[name] = argtypes
    if _lib is not None:
        for suf in ('_f32', '_f64'):
            fn = getattr(_lib, name + suf)
            fn.argtypes, fn.restype = argtypes, I32


def register_plain(name, argtypes, restype=I32):
    PLAIN[name] = (argtypes, restype)
    if _lib is not None:
        fn = getattr(_lib, name)
        fn.argtypes, fn.restype = argtypes, restype


def lib():
    """Load the shared library (once).  Raises if it has not been built: there is no CPU fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f'{LIB_PATH} not found: build it with `python -m pycsou_b200._build` '
                               '(pycsou_b200 has no CPU fallback)')
        L = C.CDLL(LIB_PATH)
        for base, argtypes in REAL.items():
            for suf in ('_f32', '_f64'):
                fn = getattr(L, base + suf)
                fn.argtypes, fn.restype = argtypes, I32
        for name, (argtypes, restype) in PLAIN.items():
            fn = getattr(L, name)
            fn.argtypes, fn.restype = argtypes, restype
        L.pcs_last_error.argtypes, L.pcs_last_error.restype = [], C.c_char_p
        _lib = L
    return _lib


def call(name: str, *args):
    """Invoke a status-returning entry point; raise on failure."""
    global launch_count
    rc = getattr(lib(), name)(*args)
    launch_count += 1
    if rc != 0:
        msg = lib().pcs_last_error().decode(errors='replace')
        if rc == 1:
            raise ValueError(msg)
        raise RuntimeError(msg)


def check(rc: int):
    """Raise for a non-zero status returned by a ``call_plain`` entry point."""
    if rc != 0:
        msg = lib().pcs_last_error().decode(errors='replace')
        raise (ValueError if rc == 1 else RuntimeError)(msg)


def call_plain(name: str, *args):
    return getattr(lib(), name)(*args)
