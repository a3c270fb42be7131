"""Deterministic synthetic inputs for the BASELINE.json configs.

numpy only (no device code): shared by ``bench.py``, the parity tests and the golden-fixture generator so
that every arm sees byte-identical inputs.
"""
from __future__ import annotations

import numpy as np


def phantom(shape, dtype=np.float64) -> np.ndarray:
    """Piecewise-constant phantom in [0, 1]: three rectangles and one disk at fixed relative coordinates."""
    H, W = shape
    img = np.zeros(shape, dtype=np.float64)
    r, c = np.arange(H)[:, None] / H, np.arange(W)[None, :] / W
    img[(r >= 0.10) & (r < 0.45) & (c >= 0.15) & (c < 0.40)] = 0.8
    img[(r >= 0.55) & (r < 0.90) & (c >= 0.10) & (c < 0.30)] = 0.4
    img[(r >= 0.30) & (r < 0.70) & (c >= 0.55) & (c < 0.90)] = 0.6
    img[((r - 0.72) ** 2 + (c - 0.50) ** 2) < 0.12 ** 2] = 1.0
    return img.astype(dtype)


def gaussian_psf_1d(ntaps=65, sigma=8.0) -> np.ndarray:
    t = np.arange(ntaps) - ntaps // 2
    h = np.exp(-0.5 * (t / sigma) ** 2)
    return h / h.sum()


def gaussian_psf_2d(size=15, sigma=2.5) -> np.ndarray:
    """Sum-normalised isotropic Gaussian: exactly rank-1 (outer product of two 1-D Gaussians)."""
    g = gaussian_psf_1d(size, sigma)
    return np.outer(g, g)


def rotated_gaussian_psf_2d(size=15, sigmas=(3.5, 1.5), angle_deg=30.0) -> np.ndarray:
    """Sum-normalised anisotropic rotated Gaussian (NOT separable)."""
    t = np.arange(size) - size // 2
    yy, xx = np.meshgrid(t, t, indexing='ij')
    a = np.deg2rad(angle_deg)
    u, v = np.cos(a) * xx + np.sin(a) * yy, -np.sin(a) * xx + np.cos(a) * yy
    h = np.exp(-0.5 * ((u / sigmas[0]) ** 2 + (v / sigmas[1]) ** 2))
    return h / h.sum()


def conv2_same(x2d: np.ndarray, h: np.ndarray) -> np.ndarray:
    """Zero-padded 'same' 2-D convolution with an odd-sized filter (direct; used only to synthesise data)."""
    import scipy.signal as sps
    return sps.convolve(x2d, h, mode='same', method='fft')


def spikes_1d(n=4096, nspikes=32, seed=0):
    """Config 1: sparse spike train, Gaussian PSF, noisy blurred observation."""
    import scipy.signal as sps
    rng = np.random.default_rng(seed)
    truth = np.zeros(n)
    truth[rng.choice(n, nspikes, replace=False)] = rng.standard_normal(nspikes)
    h = gaussian_psf_1d(65, 8.0)
    y = sps.convolve(truth, h, mode='same', method='direct') + 0.01 * rng.standard_normal(n)
    return truth, h, y


def tv_denoise_problem(shape, noise=0.1, seed=0, dtype=np.float64):
    """Config 2: y = phantom + noise."""
    rng = np.random.default_rng(seed)
    truth = phantom(shape)
    y = truth + noise * rng.standard_normal(shape)
    return truth.reshape(-1).astype(dtype), y.reshape(-1).astype(dtype)


def tv_deconv_problem(shape, psf=None, noise=0.02, seed=0, dtype=np.float64):
    """Config 3 / 5: y = A phantom + noise with a 15x15 PSF (Gaussian sigma 2.5 unless given)."""
    rng = np.random.default_rng(seed)
    truth = phantom(shape)
    psf = gaussian_psf_2d(15, 2.5) if psf is None else psf
    y = conv2_same(truth, psf) + noise * rng.standard_normal(shape)
    return truth.reshape(-1).astype(dtype), psf.astype(dtype), y.reshape(-1).astype(dtype)


def tv_deconv_slab(shape, row0, row1, psf=None, noise=0.02, seed=0, dtype=np.float32, block_rows=1024):
    """Rows [row0,row1) of the config-5 observation, generated block-wise so every GPU count sees the same image.

    Noise for global row-block b (``block_rows`` rows) comes from ``default_rng(seed + b)``; the blur of the
    phantom is evaluated on the slab plus a PSF-radius halo of phantom rows.
    """
    H, W = shape
    psf = gaussian_psf_2d(15, 2.5) if psf is None else psf
    R = psf.shape[0] // 2
    lo, hi = max(row0 - R, 0), min(row1 + R, H)
    r = (np.arange(lo, hi)[:, None]) / H
    c = np.arange(W)[None, :] / W
    img = np.zeros((hi - lo, W))
    img[(r >= 0.10) & (r < 0.45) & (c >= 0.15) & (c < 0.40)] = 0.8
    img[(r >= 0.55) & (r < 0.90) & (c >= 0.10) & (c < 0.30)] = 0.4
    img[(r >= 0.30) & (r < 0.70) & (c >= 0.55) & (c < 0.90)] = 0.6
    img[((r - 0.72) ** 2 + (c - 0.50) ** 2) < 0.12 ** 2] = 1.0
    pad_top, pad_bot = R - (row0 - lo), R - (hi - row1)
    img = np.pad(img, ((pad_top, pad_bot), (0, 0)))
    blurred = conv2_same(img, psf)[R:R + (row1 - row0)]
    out = np.empty((row1 - row0, W), dtype=dtype)
    b0, b1 = row0 // block_rows, (row1 - 1) // block_rows
    for b in range(b0, b1 + 1):
        rng = np.random.default_rng(seed + b)
        nb = rng.standard_normal((block_rows, W))
        g0, g1 = max(row0, b * block_rows), min(row1, (b + 1) * block_rows)
        out[g0 - row0:g1 - row0] = blurred[g0 - row0:g1 - row0] + noise * nb[g0 - b * block_rows:g1 - b * block_rows]
    return out.reshape(-1)


def checksum(a) -> float:
    """Order-sensitive checksum guarding regenerated inputs against the ones the golden fixtures were made from."""
    v = np.ascontiguousarray(a).astype(np.complex128).view(np.float64)
    return float(np.sum(v * np.cos(np.arange(v.size) * 0.37)))
