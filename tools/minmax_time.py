"""ms per call of the host-array percentile range (vf_dem_minmax, clamp = 1) and of a device image's (auto height range)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vulkan-forge_b200", "python"))
import numpy as np
from vulkan_forge import _vulkan_forge as ext
for n in (4096, 16384):
    a = np.random.default_rng(n).normal(size=(n, n)).astype(np.float32)
    ext.dem_minmax(a, True); ext.dem_minmax(a, True)
    t = []
    for _ in range(20):
        t0 = time.perf_counter(); r = ext.dem_minmax(a, True); t.append(time.perf_counter() - t0)
    print(f"min_max(clamp=true) {n}^2 host array: min {min(t) * 1e3:.3f} ms, median {sorted(t)[10] * 1e3:.3f} ms", r)
