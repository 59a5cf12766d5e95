"""Tiny driver for ncu: renders a few frames of a BASELINE.json workload and exits.
usage: python tools/profile_frame.py [frames] [config]
  config: c3 (default; 1080p, DEM 4096^2, NORTHSTAR, no depth) | c3d (same with the depth test) |
          c3s (SHIPPED shading) | c4 (1024^2 orbit frame) | c5 (4K, DEM 16384^2, depth, magma, exaggeration 2)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vulkan-forge_b200", "python"))
import vulkan_forge as vf
from vulkan_forge._synthetic import synthetic_dem
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
cfg = sys.argv[2] if len(sys.argv) > 2 else "c3"
if cfg == "northstar":
    cfg = "c3"
W, H, G, seed, cmap = {"c5": (3840, 2160, 16384, 2, "magma"), "c4": (1024, 1024, 4096, 0, "terrain")}.get(cfg, (1920, 1080, 4096, 0, "terrain"))
sc = vf.Scene(W, H, grid=G, colormap=cmap)
sc.set_height_from_r32f(synthetic_dem(seed, G))
if cfg != "c3s":
    sc.set_shading("northstar", depth=cfg in ("c3d", "c5")); sc.set_sun(35.0, 120.0); sc.set_exposure(1.1)
    if cfg == "c5":
        sc.set_spacing_exaggeration(1.0, 2.0)
    sc.set_auto_height_range()
cams = vf.orbit_cameras(32) if cfg != "c5" else [((3.0, 2.0, 3.0), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), 45.0, 0.1, 100.0)] * 32
sc.render_batch_device(cams[:2])   # warm-up (module load, buffers)
ms = sc.render_batch_device(cams[:n])
print("config", cfg, "frames", n, "ms", round(ms, 3), "ms/frame", round(ms / n, 4))
sc.enable_timing(True)
for k in (0, 5, 11):
    sc.render_batch_device(cams[k:k + 1])
    m = sc.render_metrics(); print(k, {a: round(m[a], 4) for a in ("ms_geom", "ms_bin", "ms_raster", "ms_total", "gpu_ms")})
