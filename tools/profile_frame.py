"""Tiny driver for ncu: renders a few frames of the bench workload (c3) and exits.
usage: python tools/profile_frame.py [frames] [mode]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vulkan-forge_b200", "python"))
import vulkan_forge as vf
from vulkan_forge._synthetic import synthetic_dem
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
mode = sys.argv[2] if len(sys.argv) > 2 else "northstar"
sc = vf.Scene(1920, 1080, grid=4096, colormap="terrain")
sc.set_height_from_r32f(synthetic_dem(0, 4096))
if mode == "northstar":
    sc.set_shading("northstar", depth=False); sc.set_sun(35.0, 120.0); sc.set_exposure(1.1); sc.set_auto_height_range()
sc.render_batch_device(vf.orbit_cameras(32)[:2])   # warm-up (module load, buffers)
ms = sc.render_batch_device(vf.orbit_cameras(32)[:n])
print("frames", n, "ms", round(ms, 3), "ms/frame", round(ms / n, 4))
sc.enable_timing(True)
for k in (0, 5, 11):
    sc.render_batch_device(vf.orbit_cameras(32)[k:k + 1])
    print(k, {a: (round(v, 4) if isinstance(v, float) else v) for a, v in sc.render_metrics().items()})
