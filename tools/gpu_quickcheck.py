"""First-contact GPU check: CUDA path vs CPU oracle on a ladder of cases, with diff statistics
(not a test; tests/ holds the asserted versions).  Run under gpurun."""
import os, sys, time, json, traceback
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vulkan-forge_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import vf_oracle as o
import vulkan_forge as vf
from vulkan_forge._synthetic import synthetic_dem

def lut(name):
    return np.fromfile(os.path.join(ROOT, "tests", "golden", f"lut_{name}.rgba"), np.uint8)

def cmp_frames(tag, got, ref, gp=None, rp=None):
    d = np.abs(got.astype(int) - ref.astype(int)).max(-1)
    msg = f"{tag}: equal={np.array_equal(got, ref)} within1={(d<=1).mean():.6f} maxdiff={d.max()} ndiff={(d>0).sum()}"
    if gp is not None:
        msg += f" vis_mismatch={(gp != rp).sum()} covered={(rp>0).sum()}"
    print(msg, flush=True)

def step(name, fn):
    try:
        t0 = time.time(); fn(); print(f"[ok] {name} ({time.time()-t0:.2f}s)", flush=True)
    except Exception:
        print(f"[FAIL] {name}"); traceback.print_exc(); sys.stdout.flush()

def t_triangle():
    for (w, h) in [(16, 16), (64, 64), (256, 256), (32, 24), (37, 19), (512, 512)]:
        cmp_frames(f"triangle {w}x{h}", vf.render_triangle_rgba(w, h), o.render_triangle(w, h))

def t_grid():
    for (nx, nz, sp) in [(4, 3, (2.0, 1.0)), (3, 3, (2.0, 2.0)), (256, 256, (1.0, 1.0)), (1024, 1024, (0.5, 0.25))]:
        a = vf.grid_generate(nx, nz, sp); b = o.make_grid(nx, nz, *sp)
        print("grid", nx, nz, [np.array_equal(x, y) for x, y in zip(a, b)], flush=True)

def t_dem():
    from vulkan_forge import _vulkan_forge as ext
    for n in [16, 1000, 128*128, 65536, 65537, 10**6, 4096*4096]:
        rng = np.random.default_rng(n); a = rng.normal(size=n).astype(np.float32)
        for clamp in (False, True):
            g = ext.dem_minmax(a, clamp); r = o.min_max(a, clamp)
            print("minmax", n, clamp, g, tuple(float(x) for x in r), np.float32(g[0]) == r[0] and np.float32(g[1]) == r[1], flush=True)

def frame_case(tag, W, H, G, dem, cmap, mode, depth, cam=None, check_vis=True, threads=0):
    sc = vf.Scene(W, H, grid=G, colormap=cmap)
    if dem is not None: sc.set_height_from_r32f(dem)
    else: dem = np.array([[0, .25], [.5, .75]], np.float32)
    if cam: sc.set_camera_look_at(*cam)
    hmin, hmax = -0.5, 0.5
    if mode == "northstar":
        sc.set_shading("northstar", depth=depth); sc.set_sun(35.0, 120.0); sc.set_exposure(1.1)
        hmin, hmax = sc.set_auto_height_range()
    u = sc.debug_uniforms_f32()
    sc.enable_timing(True)
    got = sc.render_rgba(); m = sc.render_metrics()
    gp = sc.debug_visibility() if check_vis else None
    t0 = time.time()
    ref, rp, cnt = o.render_terrain(W, H, G, dem, u, lut(cmap), mode=o.MODE_NORTHSTAR if mode == "northstar" else o.MODE_SHIPPED,
                                    depth=depth, h_min=float(hmin), h_max=float(hmax), want_prim=True, want_counters=True, threads=threads)
    tcpu = time.time() - t0
    cmp_frames(tag, got, ref, gp, rp)
    print("   metrics", {k: (round(v, 4) if isinstance(v, float) else v) for k, v in m.items()}, "oracle_s", round(tcpu, 3), cnt, flush=True)
    return sc

def t_frames_small():
    frame_case("scene default dummy 320x240 G64", 320, 240, 64, None, "viridis", "shipped", False)
    d128 = synthetic_dem(1, 128)
    frame_case("c2 scene 800x600 G128 dem128", 800, 600, 128, d128, "viridis", "shipped", False)
    frame_case("northstar 800x600 G128", 800, 600, 128, d128, "viridis", "northstar", False)
    frame_case("northstar depth 800x600 G128", 800, 600, 128, d128, "magma", "northstar", True)
    frame_case("G=2 big tris 256x256", 256, 256, 2, None, "terrain", "shipped", False)
    frame_case("inside-terrain camera (clip path)", 400, 300, 64, d128, "viridis", "northstar", True,
               cam=((0.2, 0.3, 0.4), (1.0, 0.0, 1.0), (0, 1, 0), 70.0, 0.1, 100.0))
    frame_case("odd size 333x77 G=100 dem 61x17", 333, 77, 100, synthetic_dem(3, 64)[:17, :61].copy(), "terrain", "northstar", True)
    frame_case("micro: 256x192 G=1024 dem1024", 256, 192, 1024, synthetic_dem(4, 1024), "terrain", "northstar", True)
    frame_case("micro shipped: 256x192 G=1024", 256, 192, 1024, synthetic_dem(4, 1024), "terrain", "shipped", False)

def t_c3():
    dem = synthetic_dem(0, 4096)
    for mode, depth in (("northstar", False), ("shipped", False)):
        sc = frame_case(f"c3 1920x1080 G4096 {mode}", 1920, 1080, 4096, dem, "terrain", mode, depth)
        cams = vf.orbit_cameras(16)
        sc.render_batch_device(cams[:2])
        ms = sc.render_batch_device(cams)
        print(f"   c3 {mode}: 16 orbit frames device-only {ms:.2f} ms -> {16/ms*1e3:.1f} fps; last", sc.render_metrics(), flush=True)

if __name__ == "__main__":
    print(vf.enumerate_adapters())
    step("triangle", t_triangle); step("grid", t_grid); step("dem", t_dem)
    step("frames_small", t_frames_small)
    if "--c3" in sys.argv: step("c3", t_c3)
