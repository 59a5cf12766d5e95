"""Per-step timing of bench.py's e2e_png leg (async DEM upload + synchronous render_batch_png_raw)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vulkan-forge_b200", "python"))
import torch
import vulkan_forge as vf
from vulkan_forge._synthetic import synthetic_dem
dem = torch.from_numpy(synthetic_dem(0, 4096)).pin_memory().numpy()
sc = vf.Scene(1920, 1080, grid=4096, colormap="terrain")
if "--torch-stream" in sys.argv:
    st = torch.cuda.Stream(); sc.set_stream(st.cuda_stream)
sc.set_height_from_r32f(dem)
sc.set_shading("northstar", depth=False); sc.set_sun(35.0, 120.0); sc.set_exposure(1.1); sc.set_auto_height_range()
blocks = sc.camera_blocks(vf.orbit_cameras(128)[:32])
for k in range(8):
    t0 = time.perf_counter(); sc.set_height_from_r32f_async(dem); t1 = time.perf_counter()
    buf, sizes = sc.render_batch_png_raw(blocks); t2 = time.perf_counter()
    m = sc.render_metrics()
    print(f"step {k}: upload call {1e3 * (t1 - t0):6.2f} ms, png call {1e3 * (t2 - t1):6.2f} ms | library total_ms {m['total_ms']:.2f} gpu_ms {m['gpu_ms']:.2f} upload_ms {m['upload_ms']:.2f} cap {buf.shape[1]}")
for k in range(4):
    t1 = time.perf_counter(); buf, sizes = sc.render_batch_png_raw(blocks); t2 = time.perf_counter()
    print(f"no upload {k}: png call {1e3 * (t2 - t1):6.2f} ms")
