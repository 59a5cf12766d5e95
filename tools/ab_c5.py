"""A/B helper: ms per frame of c5 (3840x2160, 16384^2 DEM, depth) and of c3; env knobs are read by the library."""
import os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vulkan-forge_b200", "python"))
import numpy as np
import vulkan_forge as vf
from vulkan_forge._synthetic import synthetic_dem
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
cam = ((3.0, 2.0, 3.0), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), 45.0, 0.1, 100.0)
sc = vf.Scene(3840, 2160, grid=n, colormap="magma")
sc.set_height_from_r32f(synthetic_dem(2, n))
sc.set_shading("northstar", depth=True); sc.set_sun(35.0, 120.0); sc.set_exposure(1.1); sc.set_spacing_exaggeration(1.0, 2.0); sc.set_auto_height_range()
sc.render_batch_device([cam] * 2)
t = [sc.render_batch_device([cam] * 4) / 4 for _ in range(5)]
print("c5-like G=%d ms/frame min %.4f med %.4f" % (n, min(t), statistics.median(t)))
