"""Writes tests/golden/frames.npz: small terrain/triangle frames rendered by the CPU oracle
(the reference itself renders through wgpu, unavailable on this box -- see DESIGN.md), with the
exact inputs, so the GPU parity tests can also compare against committed bytes."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "vulkan-forge_b200", "python"))
import numpy as np
import vf_oracle as o
from vulkan_forge._synthetic import synthetic_dem

def lut(n): return np.fromfile(os.path.join(ROOT, "tests", "golden", f"lut_{n}.rgba"), np.uint8)

out = {}
out["triangle_48x32"] = o.render_triangle(48, 32)
dem = synthetic_dem(5, 48)
out["dem48"] = dem
cases = [("shipped_viridis", "viridis", o.MODE_SHIPPED, False, (3, 2, 3)),
         ("northstar_terrain", "terrain", o.MODE_NORTHSTAR, False, (2.5, 1.5, -3.0)),
         ("northstar_depth_magma", "magma", o.MODE_NORTHSTAR, True, (-3.0, 2.5, 1.0))]
for name, cmap, mode, depth, eye in cases:
    hmin, hmax = o.h_range(dem) if mode == o.MODE_NORTHSTAR else (-0.5, 0.5)
    u = o.look_at_uniforms(96, 64, eye, h_min=float(hmin), h_max=float(hmax),
                           sun=o.sun_dir(35.0, 120.0) if mode == o.MODE_NORTHSTAR else (0.5, 0.8, 0.6),
                           exposure=1.1 if mode == o.MODE_NORTHSTAR else 1.0)
    img = o.render_terrain(96, 64, 40, dem, u, lut(cmap), mode=mode, depth=depth, h_min=float(hmin), h_max=float(hmax))
    out[f"{name}_u44"] = u; out[f"{name}_rgba"] = img
    out[f"{name}_meta"] = np.array([mode, int(depth), float(hmin), float(hmax)], np.float64)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "frames.npz"), **out)
print({k: v.shape for k, v in out.items()})
