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
    a = r.render_rgba()
    hmin, hmax = oracle.h_range(dem * np.float32(2.0))
    u = oracle.look_at_uniforms(240, 160, (3, 2, 3), sun=oracle.sun_dir(35.0, 120.0), exposure=1.1,
                                h_min=float(hmin), h_max=float(hmax), normalize_sun=False)
    ref = oracle.render_terrain(240, 160, 96, dem * np.float32(2.0), u, luts["terrain"], mode=oracle.MODE_NORTHSTAR, depth=True,
                                h_min=float(hmin), h_max=float(hmax))
    assert np.array_equal(a, ref)
    r.set_sun(35.0, 300.0)                                             # flipping the sun changes the shading
    assert not np.array_equal(a, r.render_rgba())


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
