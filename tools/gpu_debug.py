"""Small-frame check of the terrain path against the oracle (debug aid for gpurun)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vulkan-forge_b200", "python")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import vf_oracle as o
import vulkan_forge as vf
from vulkan_forge._synthetic import synthetic_dem
W, H, G = 160, 120, int(sys.argv[1]) if len(sys.argv) > 1 else 96
dem = synthetic_dem(1, G)
lut = np.fromfile(os.path.join(ROOT, "tests", "golden", "lut_viridis.rgba"), np.uint8)
sc = vf.Scene(W, H, grid=G, colormap="viridis")
sc.set_height_from_r32f(dem)
for depth in (False, True):
    sc.set_shading("northstar", depth=depth)
    lo, hi = sc.set_auto_height_range()
    got = sc.render_rgba()
    ref = o.render_terrain(W, H, G, dem, sc.debug_uniforms_f32(), lut, mode=o.MODE_NORTHSTAR, depth=depth, h_min=float(lo), h_max=float(hi))
    print("G", G, "depth", depth, "equal", np.array_equal(got, ref), "ndiff", int((got != ref).any(-1).sum()), flush=True)
