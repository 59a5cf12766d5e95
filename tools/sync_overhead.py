"""Where does the time of a synchronous render_batch_rgba call go?  Per-call wall clock vs the library's own total_ms."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vulkan-forge_b200", "python"))
import vulkan_forge as vf
from vulkan_forge._synthetic import synthetic_dem
sc = vf.Scene(1920, 1080, grid=4096, colormap="terrain")
sc.set_height_from_r32f(synthetic_dem(0, 4096))
sc.set_shading("northstar", depth=False); sc.set_auto_height_range()
cams = vf.orbit_cameras(128)[:32]
blocks = sc.camera_blocks(cams)
for label, arg in (("tuples", cams), ("blocks", blocks)):
    keep = None
    for k in range(6):
        t0 = time.perf_counter(); r = sc.render_batch_rgba(arg); t1 = time.perf_counter()
        m = sc.render_metrics()
        print(f"{label} call {k}: wall {1e3 * (t1 - t0):7.2f} ms  library total_ms {m['total_ms']:.2f} gpu_ms {m['gpu_ms']:.2f} readback_ms {m['readback_ms']:.2f}")
        t0 = time.perf_counter(); keep = r; del r; t1 = time.perf_counter()
