"""Seeded synthetic DEMs for benchmarks and tests: float32[n,n], 8 sinusoids
rescaled to [-0.25, 0.25], then 1 % of the texels multiplied by 4 as outliers that exercise the
1-99 % percentile clamp.  numpy only."""
from __future__ import annotations

import numpy as np


def synthetic_dem(seed: int, n: int) -> np.ndarray:
    rng = np.random.default_rng(seed)
    fx = rng.uniform(1, 32, 8)
    fy = rng.uniform(1, 32, 8)
    ph = rng.uniform(0, 2 * np.pi, 8)
    amp = 1.0 / (fx + fy)
    u = np.linspace(0.0, 1.0, n, dtype=np.float64)
    acc = np.zeros((n, n), np.float64)
    for k in range(8):
        a = 2 * np.pi * fx[k] * u           # along x (columns)
        b = 2 * np.pi * fy[k] * u + ph[k]   # along y (rows); sin(a+b) expanded -> O(n) trig calls
        acc += amp[k] * (np.outer(np.cos(b), np.sin(a)) + np.outer(np.sin(b), np.cos(a)))
    lo, hi = acc.min(), acc.max()
    dem = ((acc - lo) / (hi - lo) * 0.5 - 0.25).astype(np.float32)
    m = max(1, (n * n) // 100)
    idx = rng.choice(n * n, size=m, replace=False)
    dem.ravel()[idx] *= np.float32(4.0)
    return np.ascontiguousarray(dem)
