"""A/B helper: ms per frame of the bench workload (c3, 32 orbit frames per batch, device-resident),
min and median over several batches.  Tuning knobs are read from the environment by the library."""
import os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vulkan-forge_b200", "python"))
import vulkan_forge as vf
from vulkan_forge._synthetic import synthetic_dem
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 7
sc = vf.Scene(1920, 1080, grid=4096, colormap="terrain")
sc.set_height_from_r32f(synthetic_dem(0, 4096))
sc.set_shading("northstar", depth=False); sc.set_sun(35.0, 120.0); sc.set_exposure(1.1); sc.set_auto_height_range()
cams = vf.orbit_cameras(128)[:32]
sc.render_batch_device(cams[:4])
t = [sc.render_batch_device(cams) / len(cams) for _ in range(reps)]
print("ms/frame min %.4f med %.4f" % (min(t), statistics.median(t)))
