"""frames/s of Scene.render_batch_png_raw on the bench workload (c3), next to render_batch_rgba and device-only frames."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vulkan-forge_b200", "python"))
import vulkan_forge as vf
from vulkan_forge._synthetic import synthetic_dem
sc = vf.Scene(1920, 1080, grid=4096, colormap="terrain")
sc.set_height_from_r32f(synthetic_dem(0, 4096))
sc.set_shading("northstar", depth=False); sc.set_sun(35.0, 120.0); sc.set_exposure(1.1); sc.set_auto_height_range()
cams = vf.orbit_cameras(128)[:32]
for name, fn in (("device", lambda: sc.render_batch_device(cams)), ("rgba", lambda: sc.render_batch_rgba(cams)), ("png", lambda: sc.render_batch_png_raw(cams))):
    fn(); fn()
    t0 = time.perf_counter(); n = 5
    for _ in range(n): r = fn()
    dt = time.perf_counter() - t0
    extra = f" bytes/frame {int(r[1].mean())}" if name == "png" else ""
    m = sc.render_metrics()
    print(f"{name:7s} {n * len(cams) / dt:9.1f} frames/s{extra}  gpu_ms {m['gpu_ms']:.3f} readback_ms {m['readback_ms']:.3f} total_ms {m['total_ms']:.3f}")
