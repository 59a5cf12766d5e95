"""Times every BASELINE.json config (c1..c5 + DEM stats + grid_generate) on one B200 next to the
CPU oracle on the box's host cores; prints one JSON object per line.  Run under gpurun:
  python tools/bench_configs.py > gpurun_out/configs.jsonl
GPU times: device-resident frames (render_batch_device, wall clock around a synchronous batch after
warm-up, L2 not flushed -- the headline protocol lives in bench.py).  CPU: oracle with all threads."""
import json, math, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vulkan-forge_b200", "python")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import vf_oracle as o
import vulkan_forge as vf
from vulkan_forge._synthetic import synthetic_dem

ext = vf._vulkan_forge
threads = len(os.sched_getaffinity(0))
quick = "--quick" in sys.argv


def lut(name):
    return np.fromfile(os.path.join(ROOT, "tests", "golden", f"lut_{name}.rgba"), np.uint8)


def emit(**kw):
    print(json.dumps(kw), flush=True)


def gpu_frames(sc, cams, reps=3):
    sc.render_batch_device(cams[:2])
    best = min(sc.render_batch_device(cams) for _ in range(reps))
    sc.enable_timing(True); sc.render_batch_device(cams[:1]); m = sc.render_metrics(); sc.enable_timing(False)
    return best / len(cams), m


def scene_case(tag, W, H, G, dem, cmap, mode, depth, cams, exag=1.0, cpu_frames=1):
    sc = ext.Scene(W, H, grid=G, colormap=cmap)
    if dem is not None:
        sc.set_height_from_r32f(dem)
    hmin, hmax = -0.5, 0.5
    if mode == "northstar":
        sc.set_shading("northstar", depth=depth); sc.set_sun(35.0, 120.0); sc.set_exposure(1.1)
        if exag != 1.0:
            sc.set_spacing_exaggeration(1.0, exag)
        hmin, hmax = sc.set_auto_height_range()
    ms, m = gpu_frames(sc, cams)
    dem_n = dem.size if dem is not None else 4
    algo = 4 * min(G * G, dem_n) + 4 * W * H + 1024 + 176
    sc.set_camera_look_at(*cams[0])
    u = sc.debug_uniforms_f32()
    cpu_ms = None
    if cpu_frames:
        d = dem if dem is not None else np.array([[0, .25], [.5, .75]], np.float32)
        t0 = time.perf_counter()
        for _ in range(cpu_frames):
            o.render_terrain(W, H, G, d, u, lut(cmap), mode=o.MODE_NORTHSTAR if mode == "northstar" else o.MODE_SHIPPED,
                             depth=depth, h_min=float(hmin), h_max=float(hmax), threads=threads)
        cpu_ms = (time.perf_counter() - t0) / cpu_frames * 1e3
    emit(config=tag, W=W, H=H, G=G, mode=mode, depth=depth, gpu_ms_per_frame=ms, gpu_fps=1e3 / ms, mpix_per_s=W * H / ms / 1e3,
         algorithmic_bytes=algo, hbm_gbs=algo / ms / 1e6, kernels_ms={k: m[k] for k in ("ms_geom", "ms_bin", "ms_raster", "ms_total")},
         big_prims=m["big_prims"], cpu_ms_per_frame=cpu_ms, cpu_threads=threads)


default_cam = ((3.0, 2.0, 3.0), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), 45.0, 0.1, 100.0)

# c1
t = [0.0] * 5
vf.render_triangle_rgba(256, 256)
for k in range(5):
    t0 = time.perf_counter(); vf.render_triangle_rgba(256, 256); t[k] = time.perf_counter() - t0
t0 = time.perf_counter(); o.render_triangle(256, 256); cpu = time.perf_counter() - t0
emit(config="c1 triangle 256x256", gpu_ms_per_call=min(t) * 1e3, cpu_ms_per_call=cpu * 1e3, note="host call incl. D2H of 256 KB")

# c2
dem128 = synthetic_dem(1, 128)
scene_case("c2 Scene 800x600 G=128 DEM128 shipped", 800, 600, 128, dem128, "viridis", "shipped", False, [default_cam] * 8)
scene_case("c2 TerrainSpike-equivalent 800x600 G=128 no DEM shipped", 800, 600, 128, None, "viridis", "shipped", False, [default_cam] * 8)

# c3
dem4096 = synthetic_dem(0, 4096)
cams = vf.orbit_cameras(32)[:8]
scene_case("c3 1080p DEM4096 G=4096 northstar", 1920, 1080, 4096, dem4096, "terrain", "northstar", False, cams, cpu_frames=2)
scene_case("c3 1080p DEM4096 G=4096 shipped", 1920, 1080, 4096, dem4096, "terrain", "shipped", False, cams, cpu_frames=2)
scene_case("c3' 1080p DEM4096 G=128 northstar", 1920, 1080, 128, dem4096, "terrain", "northstar", False, cams, cpu_frames=2)

# c4: 1024-frame orbit at 1024^2 (64-frame subsample unless --full)
n4 = 1024 if "--full" in sys.argv else 64
scene_case(f"c4 orbit 1024x1024 DEM4096 G=4096 ({n4} of 1024 frames)", 1024, 1024, 4096, dem4096, "terrain", "northstar", False,
           vf.orbit_cameras(1024)[::1024 // n4], cpu_frames=2)

# DEM stats + grid
for n, d in (("4096^2", dem4096),):
    ext.dem_minmax(d, True)
    t0 = time.perf_counter(); g = ext.dem_minmax(d, True); tg = time.perf_counter() - t0
    t0 = time.perf_counter(); r = o.min_max(d, True); tc = time.perf_counter() - t0
    emit(config=f"min_max(clamp=true) {n} host array", gpu_ms=tg * 1e3, cpu_ms=tc * 1e3, equal=bool(np.float32(g[0]) == r[0] and np.float32(g[1]) == r[1]),
         note="strided sample gathered on the host (256 KB cross the bus), order statistics selected on the device")
    t0 = time.perf_counter(); g = ext.dem_minmax(d, False); tg = time.perf_counter() - t0
    t0 = time.perf_counter(); r = o.min_max(d, False); tc = time.perf_counter() - t0
    emit(config=f"min_max(clamp=false) {n} host array", gpu_ms=tg * 1e3, cpu_ms=tc * 1e3, equal=bool(np.float32(g[0]) == r[0] and np.float32(g[1]) == r[1]),
         note="GPU time includes the 64 MiB H2D of a pageable host array, in 64 MiB chunks")
    rr = vf.Renderer(64, 64)
    rr.add_terrain(d, (1.0, 1.0), 1.0, "terrain")
    t0 = time.perf_counter(); rr.add_terrain(d, (1.0, 1.0), 2.0, "terrain"); ta = time.perf_counter() - t0
    t0 = time.perf_counter(); st = rr.terrain_stats(); ts = time.perf_counter() - t0
    emit(config=f"Renderer.add_terrain {n} (one H2D, scale + percentile range on the device) / terrain_stats()", add_terrain_ms=ta * 1e3, terrain_stats_ms=ts * 1e3)
    del rr
# PNG delivery of the c3 frame
sc = ext.Scene(1920, 1080, grid=4096, colormap="terrain")
sc.set_height_from_r32f(dem4096); sc.set_shading("northstar", depth=False); sc.set_sun(35.0, 120.0); sc.set_exposure(1.1); sc.set_auto_height_range()
for _ in range(4):
    buf, sizes = sc.render_batch_png_raw(cams)
t0 = time.perf_counter(); buf, sizes = sc.render_batch_png_raw(cams); tp = time.perf_counter() - t0
import zlib
emit(config="c3 frames as PNG files (render_batch_png_raw, 8 frames)", ms_per_frame=tp / len(cams) * 1e3, bytes_per_frame=int(sizes.mean()),
     raw_bytes_per_frame=1920 * 1080 * 4, zlib_level6_bytes_for_comparison=None)
del sc
vf.grid_generate(1024, 1024)
t0 = time.perf_counter(); a = vf.grid_generate(1024, 1024); tg = time.perf_counter() - t0
t0 = time.perf_counter(); b = o.make_grid(1024, 1024, 1.0, 1.0); tc = time.perf_counter() - t0
emit(config="grid_generate(1024,1024)", gpu_ms=tg * 1e3, cpu_ms=tc * 1e3, equal=all(np.array_equal(x, y) for x, y in zip(a, b)),
     note="GPU time includes 41.9 MB D2H into numpy arrays")

# c5
if not quick:
    del dem4096
    t0 = time.perf_counter(); dem16k = synthetic_dem(2, 16384); tgen = time.perf_counter() - t0
    scene_case("c5 4K DEM16384 G=16384 depth magma exag 2", 3840, 2160, 16384, dem16k, "magma", "northstar", True, [default_cam] * 4,
               exag=2.0, cpu_frames=1)
    emit(config="c5 synthetic DEM generation", cpu_s=tgen)
    ext.dem_minmax(dem16k, True)
    t0 = time.perf_counter(); g = ext.dem_minmax(dem16k, True); tg = time.perf_counter() - t0
    t0 = time.perf_counter(); r = o.min_max(dem16k, True); tc = time.perf_counter() - t0
    emit(config="min_max(clamp=true) 16384^2 host array", gpu_ms=tg * 1e3, cpu_ms=tc * 1e3, equal=bool(np.float32(g[0]) == r[0] and np.float32(g[1]) == r[1]))
