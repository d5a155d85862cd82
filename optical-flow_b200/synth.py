"""Synthetic video generator for the BASELINE throughput configs.

Input DATA only (numpy): a smooth random texture translated by (1.5, 0.75) px per frame plus
noise; the "gesture" variant multiplies by a moving disc so that large regions are exactly
zero (interior det == 0, the case the validity mask exists for).  Not part of the flow path.
"""
from __future__ import annotations

import numpy as np


def _blur(img: np.ndarray, sigma: float) -> np.ndarray:
    rad = int(np.ceil(3 * sigma))
    x = np.arange(-rad, rad + 1, dtype=np.float64)
    k = np.exp(-0.5 * (x / sigma) ** 2)
    k /= k.sum()
    pad = np.pad(img, ((0, 0), (rad, rad)), mode="reflect")
    out = np.zeros_like(img)
    for i, kv in enumerate(k):
        out += kv * pad[:, i:i + img.shape[1]]
    pad = np.pad(out, ((rad, rad), (0, 0)), mode="reflect")
    out2 = np.zeros_like(img)
    for i, kv in enumerate(k):
        out2 += kv * pad[i:i + img.shape[0], :]
    return out2


def synth_frames(n_frames: int, height: int, width: int, seed: int = 1234,
                 gesture: bool = False, noise: float = 0.005) -> np.ndarray:
    """(n_frames, H, W) float64 in about [0, 1]."""
    rng = np.random.default_rng(seed)
    m = 64
    tex = _blur(rng.random((height + 2 * m, width + 2 * m)), 2.0)
    tex = (tex - tex.min()) / (tex.max() - tex.min())
    out = np.empty((n_frames, height, width), np.float64)
    yy, xx = np.mgrid[0:height, 0:width]
    for k in range(n_frames):
        dx = int(round(1.5 * k)) % m
        dy = int(round(0.75 * k)) % m
        f = tex[m + dy:m + dy + height, m + dx:m + dx + width].copy()
        f += np.random.default_rng(seed + 1 + k).normal(0.0, noise, f.shape)
        if gesture:
            cx = width * 0.3 + 6.0 * k
            cy = height * 0.5 + 2.0 * k
            rad = min(height, width) * 0.2
            f *= ((xx - cx) ** 2 + (yy - cy) ** 2 <= rad * rad)
        out[k] = f
    return out
