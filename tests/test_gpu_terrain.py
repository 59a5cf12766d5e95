"""Terrain frame parity: CUDA kernels vs CPU oracle (see test_gpu_parity.py for the bars)."""
import os

import numpy as np
import pytest

from conftest import frame_diff

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIPPED_BAR = 0.999


def _render_both(ext, oracle, luts, W, H, G, dem, cmap, mode, depth, cam=None, auto=True, cls="Scene", sun=None):
    sc = getattr(ext, cls)(W, H, grid=G, colormap=cmap)
    if dem is not None:
        sc.set_height_from_r32f(dem)
    elif cls == "Scene":
        dem = np.array([[0.0, 0.25], [0.5, 0.75]], np.float32)         # src/scene/mod.rs:142-189
    else:
        dem = np.zeros((1, 1), np.float32)                             # src/terrain/mod.rs:342-378
    if cam is not None:
        sc.set_camera_look_at(*cam)
    hmin, hmax = -0.5, 0.5
    if mode == "northstar":
        sc.set_shading("northstar", depth=depth)
        if sun:
            sc.set_sun(*sun)
        sc.set_exposure(1.1)
        if auto:
            hmin, hmax = sc.set_auto_height_range()
            assert (np.float32(hmin), np.float32(hmax)) == oracle.h_range(dem)
    u = sc.debug_uniforms_f32()
    got = sc.render_rgba()
    gvis = sc.debug_visibility()
    ref, rvis = oracle.render_terrain(W, H, G, dem, u, luts[cmap], mode=oracle.MODE_NORTHSTAR if mode == "northstar" else oracle.MODE_SHIPPED,
                                      depth=depth, h_min=float(hmin), h_max=float(hmax), want_prim=True)
    return sc, got, ref, gvis, rvis


def _check(got, ref, gvis, rvis, mode):
    frac, mx = frame_diff(got, ref)
    cov_mismatch = int((gvis != rvis).sum())            # edge-coverage differences, reported
    if mode == "northstar":
        assert cov_mismatch == 0 and np.array_equal(got, ref), f"within1={frac} max={mx} coverage_mismatch={cov_mismatch}"
    else:
        assert frac >= SHIPPED_BAR, f"within1={frac} max={mx} coverage_mismatch={cov_mismatch}"
        assert cov_mismatch <= 1e-3 * gvis.size
    assert (rvis > 0).any() and (got[..., 3] == 255).all()


CASES = [
    # tag, W, H, G, dem(seed,n)|None, cmap, mode, depth
    ("spike default (c2 as shipped)", 800, 600, 128, None, "viridis", "shipped", False, "TerrainSpike"),
    ("scene dummy 2x2", 320, 240, 64, None, "viridis", "shipped", False, "Scene"),
    ("c2 scene dem128", 800, 600, 128, (1, 128), "viridis", "shipped", False, "Scene"),
    ("northstar", 800, 600, 128, (1, 128), "terrain", "northstar", False, "Scene"),
    ("northstar depth magma", 640, 360, 200, (2, 128), "magma", "northstar", True, "Scene"),
    ("G=2 huge triangles", 256, 256, 2, None, "terrain", "shipped", False, "Scene"),
    ("G=3 northstar", 123, 457, 3, (3, 16), "viridis", "northstar", True, "Scene"),
    ("width not multiple of 4", 333, 77, 100, (3, 64), "terrain", "northstar", True, "Scene"),
    ("micro triangles G=1024", 256, 192, 1024, (4, 1024), "terrain", "northstar", True, "Scene"),
    ("micro shipped G=1024", 256, 192, 1024, (4, 1024), "magma", "shipped", False, "Scene"),
    ("grid != dem (gather path)", 400, 300, 300, (5, 257), "viridis", "northstar", False, "Scene"),
    ("1x1 target", 1, 1, 16, (1, 16), "viridis", "shipped", False, "Scene"),
]


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_frame_parity(ext, oracle, luts, synthetic_dem, case):
    tag, W, H, G, d, cmap, mode, depth, cls = case
    dem = synthetic_dem(*d) if d else None
    if cls == "TerrainSpike":
        assert not hasattr(dem, "shape")
    _, got, ref, gvis, rvis = _render_both(ext, oracle, luts, W, H, G, dem, cmap, mode, depth, cls=cls, sun=(35.0, 120.0))
    _check(got, ref, gvis, rvis, mode)


def test_non_square_dem_and_ragged_rows(ext, oracle, luts, synthetic_dem):
    for shape in ((17, 61), (3, 255), (61, 17)):                       # tests/test_height_upload.py:114,145 sizes
        dem = np.ascontiguousarray(synthetic_dem(9, 256)[:shape[0], :shape[1]])
        _, got, ref, gvis, rvis = _render_both(ext, oracle, luts, 200, 150, 90, dem, "viridis", "northstar", True)
        _check(got, ref, gvis, rvis, "northstar")


def test_near_plane_clipping_path(ext, oracle, luts, synthetic_dem):
    dem = synthetic_dem(6, 64) * np.float32(0.2)
    cam = ((0.0, 0.15, 0.0), (1.0, 0.0, 1.0), (0.0, 1.0, 0.0), 90.0, 0.1, 100.0)   # eye just above the surface
    for mode, depth in (("northstar", True), ("shipped", False)):
        sc, got, ref, gvis, rvis = _render_both(ext, oracle, luts, 400, 300, 64, dem, "viridis", mode, depth, cam=cam)
        assert sc.render_metrics()["clipped_prims"] > 0, "camera did not exercise the Sutherland-Hodgman path"
        _check(got, ref, gvis, rvis, mode)


def test_speculative_binning_redo(ext, vf, oracle, luts, synthetic_dem):
    """k_binsort is launched only once a frame has queued primitives for it: a scene seen whole skips the
    stage, the first close-up is detected (k_raster reports the queue) and rendered again with the stage,
    also inside a batch whose first frames needed nothing.  Every frame equals the oracle's."""
    W, H, G = 320, 200, 64
    dem = synthetic_dem(6, 64) * np.float32(0.2)
    far = ((3.0, 2.0, 3.0), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), 45.0, 0.1, 100.0)
    near = ((0.0, 0.15, 0.0), (1.0, 0.0, 1.0), (0.0, 1.0, 0.0), 90.0, 0.1, 100.0)
    sc = ext.Scene(W, H, grid=G, colormap="viridis")
    sc.set_height_from_r32f(dem)
    sc.set_shading("northstar", depth=True)
    lo, hi = sc.set_auto_height_range()

    def ref(cam):
        sc.set_camera_look_at(*cam)
        return oracle.render_terrain(W, H, G, dem, sc.debug_uniforms_f32(), luts["viridis"], mode=oracle.MODE_NORTHSTAR,
                                     depth=True, h_min=float(lo), h_max=float(hi))
    r_far, r_near = ref(far), ref(near)
    for cam, want, bins in ((far, r_far, False), (near, r_near, True), (far, r_far, False), (near, r_near, True)):
        sc.set_camera_look_at(*cam)
        assert np.array_equal(sc.render_rgba(), want)
        assert (sc.render_metrics()["big_prims"] > 0) == bins
    fresh = ext.Scene(W, H, grid=G, colormap="viridis")                # batch: two quiet frames, then the close-up
    fresh.set_height_from_r32f(dem)
    fresh.set_shading("northstar", depth=True)
    fresh.set_height_range(lo, hi)
    batch = fresh.render_batch_rgba([far, far, near, far])
    for k, want in enumerate((r_far, r_far, r_near, r_far)):
        assert np.array_equal(batch[k], want), k


def _correct_projection_uniforms(oracle, W, H, eye, tgt, fovy, zn, zf):
    """The shipped gl_to_wgpu is transposed (src/camera.rs:14-21), which makes z <= w always true, so
    the far plane can only be reached with a correct [0,1]-depth projection supplied as raw uniforms."""
    D = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, .5, .5], [0, 0, 0, 1]], np.float32)
    return oracle.uniforms(oracle.look_at(eye, tgt, (0, 1, 0)), D @ oracle.perspective(fovy, W / H, zn, zf, "gl"))


@pytest.mark.parametrize("tag,G,cam", [
    ("far plane cuts the terrain", 48, ((1.4, 0.6, 1.4), (-1.5, 0.0, -1.5), 100.0, 0.05, 2.0)),
    ("guard band: huge triangles, 5 degree lens", 2, ((0.0, 2.0, 0.0), (0.3, 0.0, 0.3), 5.0, 0.1, 100.0)),
])
def test_far_plane_and_guard_band(ext, oracle, luts, synthetic_dem, tag, G, cam):
    dem = synthetic_dem(7, 64)
    W, H = 320, 200
    u = _correct_projection_uniforms(oracle, W, H, *cam)
    sc = ext.Scene(W, H, grid=G, colormap="magma")
    sc.set_height_from_r32f(dem)
    sc.set_shading("northstar", depth=True)
    sc.set_uniforms_raw(u)
    got, gvis = sc.render_rgba(), sc.debug_visibility()
    ref, rvis, cnt = oracle.render_terrain(W, H, G, dem, u, luts["magma"], mode=oracle.MODE_NORTHSTAR, depth=True,
                                           want_prim=True, want_counters=True)
    assert cnt["tris_clipped"] > 0 and sc.render_metrics()["clipped_prims"] == cnt["tris_clipped"], tag
    _check(got, ref, gvis, rvis, "northstar")


def test_draw_order_resolve_without_depth(ext, oracle, luts):
    """No depth buffer in the shipped pipeline (src/terrain/pipeline.rs:133): where front-facing
    triangles overlap, the one drawn LAST (highest primitive id) wins; with depth the nearest wins."""
    n = 64
    yy, xx = np.mgrid[0:n, 0:n]
    dem = (1.5 * np.exp(-((xx - 40) ** 2 + (yy - 24) ** 2) / 40.0)).astype(np.float32)   # a spike that self-occludes
    res = {}
    for depth in (False, True):
        _, got, ref, gvis, rvis = _render_both(ext, oracle, luts, 320, 240, n, dem, "viridis", "northstar", depth, auto=False,
                                               cam=((3.0, 0.2, -2.0), (0, 0.3, 0), (0, 1, 0), 40.0, 0.1, 100.0))
        _check(got, ref, gvis, rvis, "northstar")
        res[depth] = gvis
    differ = res[False] != res[True]
    assert differ.any(), "scene has no overlap: test is vacuous"
    assert (res[False][differ] > res[True][differ]).all()             # without depth the higher id wins


def test_golden_fixtures(ext, luts):
    g = np.load(os.path.join(ROOT, "tests", "golden", "frames.npz"))
    for name, cmap in (("shipped_viridis", "viridis"), ("northstar_terrain", "terrain"), ("northstar_depth_magma", "magma")):
        mode, depth, hmin, hmax = g[f"{name}_meta"]
        sc = ext.Scene(96, 64, grid=40, colormap=cmap)
        sc.set_height_from_r32f(g["dem48"])
        if int(mode) == 1:
            sc.set_shading("northstar", depth=bool(depth))
            sc.set_height_range(float(hmin), float(hmax))
        sc.set_uniforms_raw(g[f"{name}_u44"])
        got = sc.render_rgba()
        if int(mode) == 1:
            assert np.array_equal(got, g[f"{name}_rgba"])
        else:
            assert frame_diff(got, g[f"{name}_rgba"])[0] >= SHIPPED_BAR


def test_lut_unorm_path(ext, oracle, luts, synthetic_dem, monkeypatch):   # tests/test_colormap.py:140-159
    assert ext.Scene(64, 64).debug_lut_format() == "Rgba8UnormSrgb"
    monkeypatch.setenv("VF_FORCE_LUT_UNORM", "1")
    sc = ext.Scene(160, 120, grid=64, colormap="magma")
    assert sc.debug_lut_format() == "Rgba8Unorm"
    dem = synthetic_dem(1, 64)
    sc.set_height_from_r32f(dem)
    sc.set_shading("northstar")
    ref = oracle.render_terrain(160, 120, 64, dem, sc.debug_uniforms_f32(), luts["magma"], lut_unorm=True,
                                mode=oracle.MODE_NORTHSTAR, h_min=-0.5, h_max=0.5)
    assert np.array_equal(sc.render_rgba(), ref)


def test_batch_equals_single_frames(ext, vf, synthetic_dem):
    sc = ext.Scene(256, 144, grid=256, colormap="terrain")
    sc.set_height_from_r32f(synthetic_dem(0, 256))
    sc.set_shading("northstar", depth=True)
    cams = vf.orbit_cameras(5)
    batch = sc.render_batch_rgba(cams)
    assert batch.shape == (5, 144, 256, 4)
    for k, c in enumerate(cams):
        sc.set_camera_look_at(*c)
        assert np.array_equal(batch[k], sc.render_rgba()), k
    assert np.array_equal(sc.render_batch_rgba(sc.camera_blocks(cams)), batch)    # (N,44) uniform blocks accepted
    lo, frames = vf.render_orbit(sc, 5, rank=1, world=2)                # shard == slice of the full batch
    assert np.array_equal(frames, batch[lo:lo + frames.shape[0]]) and lo == 2


def test_pair_buffer_growth(ext, oracle, luts):
    """Tiny target + huge triangles: more (tile, primitive) pairs than the initial buffer once the
    grid is dense; the scene must grow the buffer and still match the oracle."""
    sc = ext.Scene(2048, 2048, grid=300, colormap="viridis")           # capacity = 4*W*H pairs
    sc.set_shading("northstar")
    sc.set_camera_look_at((0.05, 0.9, 0.05), (0, 0, 0), (0, 0, -1), 150.0, 0.1, 100.0)   # terrain fills the view
    got = sc.render_rgba()
    m = sc.render_metrics()
    assert m["bin_pairs"] <= m["bin_capacity"]
    dem = np.array([[0.0, 0.25], [0.5, 0.75]], np.float32)
    ref = oracle.render_terrain(2048, 2048, 300, dem, sc.debug_uniforms_f32(), luts["viridis"], mode=oracle.MODE_NORTHSTAR)
    assert np.array_equal(got, ref)


def test_queue_and_pair_buffers_grow_sync_and_async(ext, oracle, luts, synthetic_dem, monkeypatch):
    """VF_TEST_SMALL_BUFFERS starts a scene with 4 queue slots and 256 pairs: a close-up overflows
    both, the frame (or the whole pipelined batch) is rendered again after the buffers have grown,
    and the result is the oracle's.  Covers scene_check_overflow on the synchronous and the ticket path."""
    monkeypatch.setenv("VF_TEST_SMALL_BUFFERS", "1")
    W, H, G = 400, 320, 12                                              # 242 triangles of thousands of pixels each
    dem = synthetic_dem(9, G)
    cams = [((0.1, 0.35, 0.1), (0.8, 0.0, 0.6), (0, 1, 0), 70.0, 0.01, 50.0),
            ((3.0, 2.0, 3.0), (0, 0, 0), (0, 1, 0), 45.0, 0.1, 100.0),
            ((-0.2, 0.25, 0.3), (-1.0, 0.0, -0.4), (0, 1, 0), 90.0, 0.01, 50.0)]
    for depth in (False, True):
        sc = ext.Scene(W, H, grid=G, colormap="magma")
        sc.set_height_from_r32f(dem); sc.set_shading("northstar", depth=depth); sc.set_height_range(-0.3, 0.4)
        want = []
        for c in cams:
            sc.set_camera_look_at(*c)
            ref = oracle.render_terrain(W, H, G, dem, sc.debug_uniforms_f32(), luts["magma"], mode=oracle.MODE_NORTHSTAR,
                                        depth=depth, h_min=-0.3, h_max=0.4)
            want.append(ref)
        sc.set_camera_look_at(*cams[0])
        assert np.array_equal(sc.render_rgba(), want[0])                 # synchronous: grow + retry
        m = sc.render_metrics()
        assert m["bin_capacity"] > 256 and m["big_prims"] > 4 and m["bin_pairs"] <= m["bin_capacity"], m
        sc2 = ext.Scene(W, H, grid=G, colormap="magma")                 # fresh tiny buffers, pipelined path
        sc2.set_shading("northstar", depth=depth); sc2.set_height_range(-0.3, 0.4)
        sc2.set_height_from_r32f_async(dem)
        t1 = sc2.render_batch_rgba_async(cams)
        t2 = sc2.render_batch_rgba_async(cams[::-1])                    # submitted with the tiny buffers too
        a, b = sc2.collect(t1), sc2.collect(t2)                         # t1 grows them; t2 must still notice it overflowed
        for k in range(3):
            assert np.array_equal(a[k], want[k]), (depth, k)
            assert np.array_equal(b[k], want[2 - k]), (depth, k)
        # two height images in flight: the overflowed batch is redone from the image it was submitted with
        dem2 = np.ascontiguousarray(-dem)
        sc.set_height_from_r32f(dem2)
        sc.set_camera_look_at(*cams[0])
        want2 = sc.render_rgba().copy()
        assert not np.array_equal(want2, want[0])
        sc3 = ext.Scene(W, H, grid=G, colormap="magma")
        sc3.set_shading("northstar", depth=depth); sc3.set_height_range(-0.3, 0.4)
        sc3.set_height_from_r32f(dem)                                   # sizes the images; nothing rendered yet
        sc3.set_height_from_r32f_async(dem)
        t1 = sc3.render_batch_rgba_async(cams[:1])
        sc3.set_height_from_r32f_async(dem2)
        t2 = sc3.render_batch_rgba_async(cams[:1])
        sc3.set_shading("shipped", depth=not depth)                      # state changes after submission do not leak into a redo
        assert np.array_equal(sc3.collect(t1)[0], want[0]), depth
        assert np.array_equal(sc3.collect(t2)[0], want2), depth


# ---- reference API behaviour (tests/test_t31_integration.py, test_t41_scene.py, test_camera.py) ----
def test_uniform_lanes_and_camera_updates(ext, vf):
    spike = ext.TerrainSpike(256, 192, grid=64, colormap="viridis")
    u = spike.debug_uniforms_f32()
    assert isinstance(u, np.ndarray) and u.dtype == np.float32 and u.shape == (44,)
    assert tuple(u[36:40]) == (1.0, 1.0, 1.0, 0.0)
    spike.set_camera_look_at((1.0, 2.0, 3.0), (0, 0, 0), (0, 1, 0), 60.0, 0.1, 100.0)
    u2 = spike.debug_uniforms_f32()
    assert not np.allclose(u, u2)
    assert np.allclose(u2[:16].reshape(4, 4, order="F"), vf.camera_look_at((1.0, 2.0, 3.0), (0, 0, 0), (0, 1, 0)), rtol=1e-5, atol=1e-6)
    assert np.allclose(u2[16:32].reshape(4, 4, order="F"), vf.camera_perspective(60.0, 256 / 192, 0.1, 100.0, "wgpu"), rtol=1e-5, atol=1e-6)
    with pytest.raises(RuntimeError, match=r"fovy_deg must be finite and in \(0, 180\)"):
        spike.set_camera_look_at((0, 0, 3), (0, 0, 0), (0, 1, 0), 0.0, 0.1, 100.0)
    with pytest.raises(RuntimeError, match="up vector must not be colinear with view direction"):
        spike.set_camera_look_at((0, 0, 3), (0, 0, 0), (0, 0, -1), 45.0, 0.1, 100.0)
    t = ext.TerrainSpike(128, 96, grid=32)                             # tests/test_camera.py:298-328
    proj = t.debug_uniforms_f32()[16:32].reshape(4, 4, order="F")
    assert np.allclose(proj, vf.camera_perspective(45.0, 128 / 96, 0.1, 100.0, clip_space="wgpu"), atol=1e-6)


def test_png_outputs_and_height_upload_changes_image(ext, tmp_path):
    spike = ext.TerrainSpike(320, 240, grid=64, colormap="viridis")
    p = tmp_path / "spike.png"
    spike.render_png(str(p))
    assert p.exists() and p.stat().st_size > 4096                      # tests/test_t31_integration.py:30-36
    scn = ext.Scene(320, 240, grid=64, colormap="viridis")
    p1, p2 = tmp_path / "s1.png", tmp_path / "s2.png"
    scn.render_png(str(p1))
    h = (np.sin(np.linspace(0, 4 * np.pi, 128))[:, None] * np.cos(np.linspace(0, 4 * np.pi, 128))[None, :]).astype("float32") * 0.25
    scn.set_height_from_r32f(h)
    scn.render_png(str(p2))
    assert p1.stat().st_size > 4096 and p1.stat().st_size != p2.stat().st_size      # tests/test_t41_scene.py
    with pytest.raises(RuntimeError, match=r"height must be C-contiguous float32\[H,W\]"):
        scn.set_height_from_r32f(h.T)
    with pytest.raises(TypeError):
        scn.set_height_from_r32f(h.astype(np.float64))
    for name in ("viridis", "magma", "terrain"):                       # tests/test_colormap.py:67-92
        ext.Scene(64, 64, colormap=name)
    for bad in ("Viridis", "jet", ""):
        with pytest.raises(RuntimeError, match="colormap"):
            ext.Scene(64, 64, colormap=bad)
    assert ext.Scene(64, 64, grid=1).grid == 2                         # grid.max(2), src/scene/mod.rs:61


def test_renderer_draws_stored_terrain(vf, oracle, luts, synthetic_dem):
    """Additive Renderer.render_rgba() (named by tests/test_terrain_fs.py:17-23 of the reference)."""
    dem = synthetic_dem(8, 96)
    r = vf.Renderer(240, 160)
    r.add_terrain(dem, (1.0, 1.0), 2.0, "terrain")
    r.set_sun(35.0, 120.0)
    r.set_exposure(1.1)
    r.set_camera_look_at((3.0, 2.0, 3.0), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), 45.0, 0.1, 100.0)
    a = r.render_rgba()
    hmin, hmax = oracle.h_range(dem * np.float32(2.0))
    u = oracle.look_at_uniforms(240, 160, (3, 2, 3), sun=oracle.sun_dir(35.0, 120.0), exposure=1.1,
                                h_min=float(hmin), h_max=float(hmax), normalize_sun=False)
    ref = oracle.render_terrain(240, 160, 96, dem * np.float32(2.0), u, luts["terrain"], mode=oracle.MODE_NORTHSTAR, depth=True,
                                h_min=float(hmin), h_max=float(hmax))
    assert np.array_equal(a, ref)
    r.set_sun(35.0, 300.0)                                             # flipping the sun changes the shading
    assert not np.array_equal(a, r.render_rgba())


def test_renderer_framing_camera_spacing_and_device_resident_terrain(vf, oracle, luts):
    """Until set_camera_look_at is called the Renderer frames the terrain from the south (north = +X up, east = +Z to
    the right); spacing scales the plane (terrain.wgsl:58); float64 input, statistics and normalisation run on the
    device copy made by add_terrain and the frame equals the oracle's on that copy."""
    yy, xx = np.mgrid[0:96, 0:96]
    dem = (40.0 * np.exp(-(((xx - 47.5) ** 2 + (yy - 47.5) ** 2) / 96.0 ** 2) / (2 * 0.18 ** 2))).astype(np.float64)
    r = vf.Renderer(160, 120)
    r.add_terrain(dem, (30.0, 30.0), 1.0, "viridis")
    h32 = dem.astype(np.float32)
    lo, hi = r.height_range()
    olo, ohi = oracle.h_range(h32)
    assert (np.float32(lo), np.float32(hi)) == (olo, ohi)
    mn, mx, mean, std = r.terrain_stats()
    assert mn == h32.min() and mx == h32.max() and abs(mean - h32.mean()) < 1e-3 * abs(h32.mean()) and abs(std - h32.std()) < 1e-3 * h32.std()
    r.set_sun(45.0, 90.0)                                              # sun in the east
    east = r.render_rgba()
    R = max(45.0 * np.float32(1.41421356), 0.5 * (hi - lo))
    hmid = np.float32(0.5) * (np.float32(lo) + np.float32(hi))
    eye = (float(np.float32(-1.5) * np.float32(R)), float(hmid + np.float32(1.0) * np.float32(R)), 0.0)
    u = oracle.look_at_uniforms(160, 120, eye, (0.0, float(hmid), 0.0), (0, 1, 0), 45.0, float(np.float32(0.05) * np.float32(R)),
                                float(np.float32(12.0) * np.float32(R)), sun=oracle.sun_dir(45.0, 90.0), spacing=30.0,
                                h_min=float(lo), h_max=float(hi), normalize_sun=False)
    ref = oracle.render_terrain(160, 120, 96, h32, u, luts["viridis"], mode=oracle.MODE_NORTHSTAR, depth=True, h_min=float(lo), h_max=float(hi))
    assert np.array_equal(east, ref), f"{(east != ref).any(-1).sum()} pixels differ"
    r.set_sun(45.0, 270.0)
    west = r.render_rgba()
    lum = lambda im: im[..., :3].astype(np.float32) @ np.float32([0.2126, 0.7152, 0.0722])   # noqa: E731
    y, cx = 60, 80                 # east sun: the east flank (right of the summit) is brighter than under a west sun, and vice versa
    assert lum(east)[y, cx + 14] > lum(west)[y, cx + 14] and lum(west)[y, cx - 14] > lum(east)[y, cx - 14]
    r.normalize_terrain("minmax", (0.0, 1.0))                          # in place on the device copy; next frame sees it
    mn, mx, _, _ = r.terrain_stats()
    assert abs(mn) < 1e-6 and abs(mx - 1.0) < 1e-6
    assert not np.array_equal(west, r.render_rgba())
    m = r.render_metrics()
    assert m["d2h_bytes"] == 160 * 120 * 4 and m["vram_bytes"] > 0 and m["gpu_ms"] > 0


def test_render_metrics_and_report_environment(ext, vf, synthetic_dem):
    """render_metrics(): ROADMAP.md:1959-1967 (encode_ms, gpu_ms, readback_ms, total_ms, vram_bytes) plus byte counts;
    report_environment(): ROADMAP.md:1989-1999."""
    W, H, G = 320, 200, 256
    dem = synthetic_dem(3, G)
    sc = ext.Scene(W, H, grid=G, colormap="viridis")
    sc.set_height_from_r32f(dem)
    n = 5
    frames = sc.render_batch_rgba(vf.orbit_cameras(16)[:n])
    m = sc.render_metrics()
    for k in ("upload_ms", "gpu_ms", "readback_ms", "encode_ms", "total_ms", "h2d_bytes", "d2h_bytes", "vram_bytes", "frames"):
        assert k in m, k
    assert m["d2h_bytes"] == W * H * 4 * n == frames.nbytes and m["frames"] == n
    assert m["h2d_bytes"] == dem.nbytes and m["upload_ms"] > 0 and m["gpu_ms"] > 0 and m["readback_ms"] > 0
    assert m["total_ms"] >= m["gpu_ms"] * 0.5 and m["vram_bytes"] >= dem.nbytes + 2 * W * H * 4 + W * H * 8
    sc.render_batch_device(vf.orbit_cameras(16)[:2])
    m = sc.render_metrics()
    assert m["d2h_bytes"] == 0 and m["frames"] == 2 and m["readback_ms"] == 0
    tk = sc.render_batch_rgba_async(vf.orbit_cameras(16)[:3])
    sc.collect(tk)
    assert sc.render_metrics()["d2h_bytes"] == W * H * 4 * 3
    env = vf.report_environment()
    for k in ("os", "python", "package_version", "backend", "adapters", "cuda_driver_version", "cuda_runtime_version", "build"):
        assert k in env, k
    assert env["adapters"] and "B200" in env["adapters"][0]["name"] or env["adapters"][0]["backend"] == "CUDA"
    assert "sm_100a" in env["build"] and env["cuda_driver_version"]


def test_device_render_between_async_batch_and_collect(ext, vf, synthetic_dem):
    """A device-only render issued while an async batch still drains its frame slots to the host must not tear that
    batch (the slots wait for their copies whatever the later call does with its frames)."""
    W, H, G = 512, 384, 512
    sc = ext.Scene(W, H, grid=G, colormap="terrain")
    sc.set_height_from_r32f(synthetic_dem(6, G))
    sc.set_shading("northstar", depth=True)
    sc.set_auto_height_range()
    cams = vf.orbit_cameras(24)
    want = np.array(sc.render_batch_rgba(cams[:8]))
    for _ in range(3):
        tk = sc.render_batch_rgba_async(cams[:8])
        sc.render_batch_device(cams[8:16])
        sc.debug_visibility()
        assert np.array_equal(sc.collect(tk), want)


def test_tma_staging_matches_gather_path(ext, synthetic_dem, monkeypatch):
    """grid == DEM with identity texel mapping -> DEM tiles staged by TMA (cp.async.bulk.tensor);
    VF_DISABLE_TMA forces the __ldg gather path.  Both must give the same bytes."""
    dem = synthetic_dem(11, 512)
    frames = {}
    for disable in (False, True):
        if disable:
            monkeypatch.setenv("VF_DISABLE_TMA", "1")
        sc = ext.Scene(640, 360, grid=512, colormap="terrain")
        sc.set_height_from_r32f(dem)
        sc.set_shading("northstar", depth=True)
        frames[disable] = sc.render_rgba()
        assert sc.render_metrics()["tma"] == (0 if disable else 1)
    assert np.array_equal(frames[False], frames[True])
    sc = ext.Scene(64, 64, grid=100, colormap="terrain")        # grid != DEM: gather path
    sc.set_height_from_r32f(dem)
    sc.render_rgba()
    assert sc.render_metrics()["tma"] == 0


def test_pipelined_batches_equal_synchronous(ext, vf, synthetic_dem):
    """set_height_from_r32f_async / render_batch_rgba_async / collect: uploads alternate between two
    DEMs while earlier batches are still in flight; every batch must equal the synchronous result
    for the DEM that was current when it was submitted."""
    W, H, G = 320, 200, 256
    dems = [synthetic_dem(21, G), synthetic_dem(22, G)]
    cams = vf.orbit_cameras(12)[:5]
    sc = ext.Scene(W, H, grid=G, colormap="terrain")
    sc.set_shading("northstar", depth=True)
    sc.set_height_range(-0.3, 0.4)
    want = []
    for d in dems:
        sc.set_height_from_r32f(d)
        want.append(sc.render_batch_rgba(cams).copy())
    assert not np.array_equal(want[0], want[1])
    tickets = []
    for k in range(6):
        sc.set_height_from_r32f_async(dems[k & 1])
        tickets.append((k & 1, sc.render_batch_rgba_async(cams)))
        if len(tickets) == 3:                                   # three batches in flight at most here
            which, tk = tickets.pop(0)
            assert np.array_equal(sc.collect(tk), want[which]), k
    for which, tk in tickets:
        assert np.array_equal(sc.collect(tk), want[which])
    with pytest.raises(ValueError):
        sc.collect(0)
    # the synchronous path still works afterwards and sees the last uploaded image
    assert np.array_equal(sc.render_batch_rgba(cams), want[1])


# ---- PNG encoded on the device ----------------------------------------------------------------
def _png_chunks(data):
    import struct
    import zlib
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, out = 8, []
    while pos < len(data):
        n, typ = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        crc, = struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])
        assert crc == (zlib.crc32(typ + body) & 0xffffffff), f"bad CRC in {typ} chunk at {pos}"
        out.append((typ, body))
        pos += 12 + n
    assert pos == len(data) and out[0][0] == b"IHDR" and out[-1] == (b"IEND", b"")
    return out


def _png_decode(data, w, h):
    """Independent decoder: chunk CRCs, zlib (checks the Adler-32), Sub / None filters -> (h, w, 4) uint8."""
    import zlib
    chunks = _png_chunks(data)
    raw = np.frombuffer(zlib.decompress(b"".join(b for t, b in chunks if t == b"IDAT")), np.uint8).reshape(h, 4 * w + 1)
    assert set(np.unique(raw[:, 0])) <= {0, 1}
    px = raw[:, 1:].reshape(h, w, 4).astype(np.uint32)
    sub = raw[:, 0] == 1
    px[sub] = np.cumsum(px[sub], axis=1) & 255
    return px.astype(np.uint8)


@pytest.mark.parametrize("size", [(1, 1), (7, 3), (64, 64), (333, 77), (1920, 1080), (8192, 3)])
def test_device_png_encoder_arbitrary_images(ext, size):
    """Noise (every byte a literal: the 9-bits-per-byte worst case), flat colour (one long run), a gradient and
    period-4 patterns survive the device encoder byte for byte; PIL reads the files too."""
    from PIL import Image
    import io
    w, h = size
    rng = np.random.default_rng(w * 131 + h)
    yy, xx = np.mgrid[0:h, 0:w]
    images = {
        "noise": rng.integers(0, 256, (h, w, 4), dtype=np.uint8),
        "flat": np.full((h, w, 4), (39, 39, 48, 255), np.uint8),
        "gradient": np.stack([xx & 255, yy & 255, (xx + yy) & 255, np.full_like(xx, 255)], -1).astype(np.uint8),
        "mixed": np.where((xx // 37 + yy // 11)[..., None] % 2 == 0, rng.integers(0, 256, (h, w, 4)), 200).astype(np.uint8),
    }
    for name, img in images.items():
        data = ext.png_encode_device(img)
        assert np.array_equal(_png_decode(data, w, h), img), name
        assert np.array_equal(np.asarray(Image.open(io.BytesIO(data)).convert("RGBA")), img), name
    assert len(ext.png_encode_device(images["flat"])) < 0.02 * w * h * 4 + 600      # runs collapse to 13 bits per 258 bytes


def test_device_png_batch_equals_rgba_frames(ext, vf, synthetic_dem, tmp_path):
    """Scene.render_batch_png / render_batch_png_bytes / render_png: every file decodes to the bytes render_batch_rgba
    returns for the same camera; only the compressed bytes cross the bus (render_metrics)."""
    from PIL import Image
    W, H, G = 640, 360, 512
    sc = ext.Scene(W, H, grid=G, colormap="terrain")
    sc.set_height_from_r32f(synthetic_dem(7, G))
    sc.set_shading("northstar", depth=True)
    sc.set_auto_height_range()
    cams = vf.orbit_cameras(16)[:7]
    want = np.array(sc.render_batch_rgba(cams))
    blobs = sc.render_batch_png_bytes(cams)
    m = sc.render_metrics()
    assert m["d2h_bytes"] == sum(len(b) for b in blobs) < 0.5 * want.nbytes
    for k, b in enumerate(blobs):
        assert np.array_equal(_png_decode(b, W, H), want[k]), k
    paths = [str(tmp_path / f"f{k}.png") for k in range(len(cams))]
    sc.render_batch_png(paths, cams)
    for k, p in enumerate(paths):
        assert np.array_equal(np.asarray(Image.open(p).convert("RGBA")), want[k]), k
    sc.set_camera_look_at(*cams[3])
    sc.render_png(str(tmp_path / "one.png"))
    assert np.array_equal(np.asarray(Image.open(tmp_path / "one.png").convert("RGBA")), want[3])
    with pytest.raises(ValueError):
        sc.render_batch_png(paths[:2], cams)


def test_device_png_slots_grow_when_a_file_does_not_fit(ext, vf):
    """The host slots of a PNG batch start below the worst case; a batch of incompressible frames (a DEM of noise over a
    fine grid: every pixel a different colour) is repeated with larger slots behind the caller's back."""
    W, H, G = 512, 512, 1024
    rng = np.random.default_rng(3)
    sc = ext.Scene(W, H, grid=G, colormap="terrain")
    sc.set_height_from_r32f(rng.uniform(-0.05, 0.05, (G, G)).astype(np.float32))
    sc.set_shading("northstar", depth=True)
    sc.set_height_range(-0.05, 0.05)
    cams = [((0.0, 3.0, 0.01), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), 60.0, 0.1, 100.0)] * 3      # looking straight down: the frame is full
    want = np.array(sc.render_batch_rgba(cams))
    assert (want[0][..., :3] != want[0][0, 0, :3]).any(-1).mean() > 0.3
    blobs = sc.render_batch_png_bytes(cams)
    assert len(blobs[0]) > W * H * 4 // 4                   # larger than the first slot size
    for k, b in enumerate(blobs):
        assert np.array_equal(_png_decode(b, W, H), want[k])
    buf, sizes = sc.render_batch_png_raw(cams)
    assert np.array_equal(_png_decode(bytes(buf[1, :sizes[1]]), W, H), want[1])


def test_redo_of_a_batch_whose_image_is_no_longer_in_front(ext, oracle, luts, synthetic_dem):
    """Found by the API fuzzer in round 2: a pipelined batch that has to be rendered again (its frames queued primitives
    while the binning stage was left out) after a later upload has taken the front slot is redone from ITS image -- and
    with that image's height bounds (the per-task corner test of k_geom_march).  Two images with different height ranges,
    close-up cameras, several batches in flight, every batch compared with the oracle."""
    W, H, G = 395, 89, 753
    dems = [synthetic_dem(4, G) * np.float32(3.0), synthetic_dem(5, G)]
    cams = [((0.12, 0.2, -0.05), (0.6, -0.1, 0.4), (0.0, 1.0, 0.0), 90.0, 0.01, 100.0),
            ((4 * np.cos(1.0), 1.5, 4 * np.sin(1.0)), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), 45.0, 0.1, 100.0)]
    for depth in (False, True):
        sc = ext.Scene(W, H, grid=G, colormap="viridis")
        sc.set_shading("northstar", depth=depth)
        sc.set_height_range(-0.4, 0.5)
        blocks = sc.camera_blocks(cams)
        want = [np.stack([oracle.render_terrain(W, H, G, d, blocks[i], luts["viridis"], mode=oracle.MODE_NORTHSTAR, depth=depth,
                                                h_min=-0.4, h_max=0.5) for i in range(len(cams))]) for d in dems]
        pend = []
        for k in range(4):
            sc.set_height_from_r32f_async(dems[k & 1])
            pend.append((k & 1, sc.render_batch_rgba_async(cams)))
        assert sc.render_metrics()["big_prims"] >= 0
        for k, (which, tk) in enumerate(pend):
            got = sc.collect(tk)
            assert np.array_equal(got, want[which]), (depth, k, [int((got[i] != want[which][i]).any(-1).sum()) for i in range(len(cams))])


@pytest.mark.parametrize("size", [(640, 360), (322, 181)])
def test_list_shading_equals_tile_kernel(ext, vf, synthetic_dem, monkeypatch, size):
    """Frames that queue nothing for the binning stage are shaded by k_resolve_lean + k_shade_list (covered-pixel list,
    csrc/vf_raster_lean.cuh); VF_RASTER_FULL at scene creation keeps the tile kernel k_raster in every frame, and so does a
    width that is not a multiple of 4.  Same bytes either way, in both shading modes, with and without the depth test."""
    W, H = size
    dem = synthetic_dem(21, 384)
    cams = vf.orbit_cameras(5)
    frames = {}
    for full in (False, True):
        if full:
            monkeypatch.setenv("VF_RASTER_FULL", "1")
        out = []
        for mode, depth in (("northstar", False), ("northstar", True), ("shipped", False), ("shipped", True)):
            sc = ext.Scene(W, H, grid=384, colormap="viridis")
            sc.set_height_from_r32f(dem)
            sc.set_shading(mode, depth=depth); sc.set_sun(40.0, 200.0); sc.set_exposure(1.2)
            sc.set_auto_height_range()
            out.append(sc.render_batch_rgba(cams))
            vis = sc.debug_visibility()
            assert (vis > 0).any() and (vis == 0).any()
            out.append(vis)
        frames[full] = out
    for a, b in zip(frames[False], frames[True]):
        assert np.array_equal(a, b)


def test_png_batch_with_the_next_height_image_in_flight(ext, vf, synthetic_dem):
    """render_batch_png_raw(cameras, next_height=...) (vf_scene_render_batch_png_next): the batch is drawn from the current
    image while the NEXT call's image uploads; files equal those of the plain calls, image by image -- also when the first
    attempt's host slots are too small and the call repeats itself (the staged upload must stay pending, not become current)."""
    W, H, G = 320, 200, 256
    dems = [synthetic_dem(31 + k, G) for k in range(3)]
    cams = vf.orbit_cameras(4)
    ref = ext.Scene(W, H, grid=G, colormap="terrain")
    ref.set_shading("northstar", depth=True)
    want = []
    for d in dems:
        ref.set_height_from_r32f(d); ref.set_height_range(-0.3, 0.3)
        want.append(np.array(ref.render_batch_rgba(cams)))
    sc = ext.Scene(W, H, grid=G, colormap="terrain")
    sc.set_shading("northstar", depth=True); sc.set_height_range(-0.3, 0.3)
    sc.set_height_from_r32f(dems[0])
    for k in range(3):
        buf, sizes = sc.render_batch_png_raw(cams, next_height=dems[k + 1] if k < 2 else None)
        for f in range(len(cams)):
            assert np.array_equal(_png_decode(bytes(buf[f, :sizes[f]]), W, H), want[k][f]), (k, f)
    assert np.array_equal(sc.render_batch_rgba(cams), want[2])
    with pytest.raises(Exception):
        sc.render_batch_png_raw(cams, next_height=np.zeros((G, G + 4), np.float32))        # another size cannot be double-buffered
    # incompressible frames: the first attempt does not fit its slots and is repeated with the staged upload pending
    W, H, G = 512, 512, 1024
    rng = np.random.default_rng(5)
    noise = [rng.uniform(-0.05, 0.05, (G, G)).astype(np.float32) for _ in range(2)]
    down = [((0.0, 3.0, 0.01), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), 60.0, 0.1, 100.0)] * 2
    ref = ext.Scene(W, H, grid=G, colormap="terrain"); ref.set_shading("northstar", depth=False); ref.set_height_range(-0.05, 0.05)
    want = []
    for d in noise:
        ref.set_height_from_r32f(d); want.append(np.array(ref.render_batch_rgba(down)))
    sc = ext.Scene(W, H, grid=G, colormap="terrain"); sc.set_shading("northstar", depth=False); sc.set_height_range(-0.05, 0.05)
    sc.set_height_from_r32f(noise[0])
    buf, sizes = sc.render_batch_png_raw(down, next_height=noise[1])
    assert int(sizes[0]) > W * H * 4 // 4                                                 # larger than the first slot size: the call repeated itself
    assert np.array_equal(_png_decode(bytes(buf[1, :sizes[1]]), W, H), want[0][1])
    buf, sizes = sc.render_batch_png_raw(down)
    assert np.array_equal(_png_decode(bytes(buf[0, :sizes[0]]), W, H), want[1][0])


def test_geometry_pass_launched_behind_the_previous_frames_shading(ext, vf, synthetic_dem, monkeypatch):
    """In a run of list-shading frames k_geom_march(k+1) is a programmatic dependent of k_shade_list(k) and starts on the
    SMs that kernel's tail leaves idle (no memset node, no dependence between the two); VF_NO_PDL at scene creation keeps
    the plain stream order.  Same frames either way: long device-resident batches (where the overlap happens), batches with
    read-back, both modes, with and without depth, and a close-up in the middle that forces the binning stage back in."""
    W, H, G = 640, 360, 512
    dem = synthetic_dem(41, G)
    cams = vf.orbit_cameras(24)
    close = ((0.3, 0.4, 0.2), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), 70.0, 0.05, 50.0)
    mixed = cams[:7] + [close] + cams[7:14]
    out = {}
    for plain in (False, True):
        if plain:
            monkeypatch.setenv("VF_NO_PDL", "1")
        res = []
        for mode, depth in (("northstar", False), ("northstar", True), ("shipped", False)):
            sc = ext.Scene(W, H, grid=G, colormap="magma")
            sc.set_height_from_r32f(dem); sc.set_shading(mode, depth=depth); sc.set_auto_height_range()
            sc.render_batch_device(cams)                     # frames stay on the device: nothing but kernels in the stream
            sc.render_batch_device(cams)
            last2 = [sc.read_device_frame((len(cams) - 2) % 2), sc.read_device_frame((len(cams) - 1) % 2)]
            want2 = np.array(sc.render_batch_rgba(cams[-2:]))          # the same two cameras with read-back (events between frames)
            assert np.array_equal(last2[0], want2[0]) and np.array_equal(last2[1], want2[1])
            res += last2
            res.append(np.array(sc.render_batch_rgba(mixed)))
            res.append(sc.debug_visibility())
        out[plain] = res
    for a, b in zip(out[False], out[True]):
        assert np.array_equal(a, b)


def test_mode_and_depth_changes_between_memset_free_frames(ext, vf, synthetic_dem):
    """One scene, runs of list-shading frames (no memset node, geometry pass launched under the previous shading tail) with
    the depth test and the shading mode switched in between -- the visibility buffer is re-initialised, the key width
    changes -- and two scenes taking turns on the same stream: every batch equals the one a fresh scene renders."""
    W, H, G = 384, 216, 300
    dem = synthetic_dem(51, G)
    cams = vf.orbit_cameras(6)
    def fresh(mode, depth):
        sc = ext.Scene(W, H, grid=G, colormap="viridis")
        sc.set_height_from_r32f(dem); sc.set_shading(mode, depth=depth); sc.set_height_range(-0.3, 0.4)
        return np.array(sc.render_batch_rgba(cams))
    want = {(m, d): fresh(m, d) for m in ("northstar", "shipped") for d in (False, True)}
    a = ext.Scene(W, H, grid=G, colormap="viridis"); a.set_height_from_r32f(dem); a.set_height_range(-0.3, 0.4)
    b = ext.Scene(W, H, grid=G, colormap="viridis"); b.set_height_from_r32f(dem); b.set_height_range(-0.3, 0.4)
    seq = [("northstar", False), ("northstar", True), ("northstar", True), ("shipped", False), ("northstar", False), ("shipped", True)]
    for k, (m, d) in enumerate(seq):
        a.set_shading(m, depth=d)
        a.render_batch_device(cams)
        b.set_shading(*seq[-1 - k][:1], depth=seq[-1 - k][1])
        b.render_batch_device(cams[:3])
        got = [a.read_device_frame((len(cams) - 2) % 2), a.read_device_frame((len(cams) - 1) % 2)]
        assert np.array_equal(got[0], want[(m, d)][-2]) and np.array_equal(got[1], want[(m, d)][-1]), (k, m, d)
        assert np.array_equal(np.array(a.render_batch_rgba(cams)), want[(m, d)]), (k, m, d)
        assert np.array_equal(b.read_device_frame(0), want[seq[-1 - k]][2]), k
