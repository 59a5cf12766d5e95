"""Pins the CPU oracle against every golden vector / known answer the reference's own tests and
sources hold for the path.  No GPU needed.  Citations: reference file:line."""
import math

import numpy as np
import pytest


# ---- grid: tests/test_grid_generate.py ------------------------------------------------------
def test_grid_shapes_dtypes(oracle):                                   # :9-20
    xy, uv, idx = oracle.make_grid(4, 3, 2.0, 1.0)
    assert xy.shape == (12, 2) and uv.shape == (12, 2) and idx.shape == (36,)
    assert xy.dtype == np.float32 and uv.dtype == np.float32 and idx.dtype == np.uint32


def test_grid_uv_corners(oracle):                                      # :23-41
    _, uv, _ = oracle.make_grid(4, 3, 2.0, 1.0)
    for k, want in ((0, (0, 0)), (3, (1, 0)), (8, (0, 1)), (11, (1, 1))):
        assert tuple(uv[k]) == want


def test_grid_first_triangle_ccw(oracle):                              # :44-62
    xy, _, idx = oracle.make_grid(3, 3, 1.0, 1.0)
    p0, p1, p2 = xy[idx[0]], xy[idx[1]], xy[idx[2]]
    e1, e2 = p1 - p0, p2 - p0
    assert e1[0] * e2[1] - e1[1] * e2[0] > 0


def test_grid_256_index_count(oracle):                                 # :65-79
    xy, uv, idx = oracle.make_grid(256, 256, 1.0, 1.0)
    assert idx.shape == (255 * 255 * 6,) and idx.dtype == np.uint32 and xy.shape == (65536, 2)
    assert int(idx.max()) == 65535


def test_grid_centered_positions(oracle):                              # :107-121
    xy, _, _ = oracle.make_grid(3, 3, 2.0, 2.0)
    want = [[-2, -2], [0, -2], [2, -2], [-2, 0], [0, 0], [2, 0], [-2, 2], [0, 2], [2, 2]]
    assert np.array_equal(xy, np.array(want, np.float32))


def test_grid_index_pattern_rust_unit_tests(oracle):                   # src/terrain/mesh.rs:96-129
    _, _, idx = oracle.make_grid(3, 2, 1.0, 1.0)
    assert idx.tolist() == [0, 1, 3, 3, 1, 4, 1, 2, 4, 4, 2, 5]


def test_render_grid_matches_scene_inline_copy(oracle):               # src/scene/mod.rs:85-116
    verts, idx = oracle.render_grid_xyuv(4)
    assert np.allclose(verts[0], [-1.5, -1.5, 0, 0]) and np.allclose(verts[-1], [1.5, 1.5, 1, 1], atol=1e-6)
    assert idx[:6].tolist() == [0, 4, 1, 1, 4, 5]                       # [a,c,b,b,c,d]


# ---- camera: tests/test_camera.py -----------------------------------------------------------
def test_look_at_z_translation(oracle):                                # :56-68
    v = oracle.look_at((0, 0, 3), (0, 0, 0), (0, 1, 0))
    assert v.shape == (4, 4) and v.dtype == np.float32 and v.flags.c_contiguous
    assert abs(v[2, 3] + 3.0) < 1e-6


def test_gl_vs_wgpu_rows01_equal(oracle):                              # :113-126
    gl = oracle.perspective(45.0, 1.0, 0.1, 100.0, "gl")
    wg = oracle.perspective(45.0, 1.0, 0.1, 100.0, "wgpu")
    assert not np.allclose(gl, wg)
    assert np.allclose(gl[:2], wg[:2], rtol=1e-5, atol=1e-6)


def test_view_proj_composition(oracle):                                # :185-201
    args = ((0, 0, 3), (0, 0, 0), (0, 1, 0))
    vp = oracle.view_proj(*args, 45.0, 16 / 9, 0.1, 100.0)
    want = oracle.perspective(45.0, 16 / 9, 0.1, 100.0) @ oracle.look_at(*args)
    assert np.allclose(vp, want, rtol=1e-5, atol=1e-6)


def test_shipped_projection_known_answer(oracle):                      # SURVEY 8c, src/camera.rs:14-21
    p = oracle.perspective(45.0, 800 / 600, 0.1, 100.0)
    want = [[1.8106601, 0, 0, 0], [0, 2.4142137, 0, 0], [0, 0, -0.501001, -0.1001001], [0, 0, -1.501001, -0.1001001]]
    assert np.allclose(p, want, rtol=1e-6, atol=1e-6)
    v = oracle.look_at((3, 2, 3), (0, 0, 0), (0, 1, 0))
    assert abs(v[2, 3] + 4.6904154) < 1e-5


@pytest.mark.parametrize("args,msg", [
    (((float("inf"), 0, 3), (0, 0, 0), (0, 1, 0), 45.0, 0.1, 100.0), "eye/target/up components must be finite"),
    (((0, 0, 3), (0, 0, 0), (0, 0, -1), 45.0, 0.1, 100.0), "up vector must not be colinear with view direction"),
    (((0, 0, 3), (0, 0, 0), (0, 1, 0), 0.0, 0.1, 100.0), "fovy_deg must be finite and in (0, 180)"),
    (((0, 0, 3), (0, 0, 0), (0, 1, 0), 180.0, 0.1, 100.0), "fovy_deg must be finite and in (0, 180)"),
    (((0, 0, 3), (0, 0, 0), (0, 1, 0), 45.0, 0.0, 100.0), "znear must be finite and > 0"),
    (((0, 0, 3), (0, 0, 0), (0, 1, 0), 45.0, 0.1, 0.1), "zfar must be finite and > znear"),
])
def test_camera_validation_strings(oracle, args, msg):                 # src/camera.rs:24-30, tests :70-167
    assert oracle.validate_camera(*args) == msg
    assert oracle.validate_camera((0, 0, 3), (0, 0, 0), (0, 1, 0), 45.0, 0.1, 100.0) is None


# ---- uniforms: tests/test_t31_integration.py:13-28, src/terrain/mod.rs:698-707 --------------
def test_uniform_layout_defaults(oracle):
    for kind in ("spike", "scene"):
        u = oracle.default_uniforms(kind, 256, 192)
        assert u.shape == (44,) and u.dtype == np.float32 and u.nbytes == 176
        assert tuple(u[36:40]) == (1.0, 1.0, 1.0, 0.0) and not u[40:].any()
        proj = u[16:32].reshape(4, 4, order="F")                        # tests/test_camera.py:285-295
        assert np.allclose(proj, oracle.perspective(45.0, 256 / 192, 0.1, 100.0), atol=1e-6)
    s = oracle.default_uniforms("scene", 64, 64)[32:35]
    assert np.allclose(s, np.array([0.5, 0.8, 0.6]) / np.linalg.norm([0.5, 0.8, 0.6]), atol=1e-6)
    s = oracle.default_uniforms("spike", 64, 64)[32:35]
    assert np.allclose(s, np.array([0.5, 1.0, 0.3]) / np.linalg.norm([0.5, 1.0, 0.3]), atol=1e-6)


# ---- DEM statistics: tests/test_height_upload.py:20-49, test_dem_normalization.py -----------
def test_stats_ramp_times_two(oracle):
    h = (np.linspace(0, 1, 16, dtype=np.float32).reshape(4, 4) * np.float32(2.0))
    mn, mx, mean, std = oracle.dem_stats(h)
    assert mn == 0.0 and mx == 2.0 and abs(mean - 1.0) < 1e-6
    assert abs(std - np.linspace(0, 2, 16).std()) < 1e-3 * np.linspace(0, 2, 16).std()


def test_normalize_minmax_and_zscore(oracle):
    rng = np.random.default_rng(0)
    a = rng.normal(5, 3, (32, 32)).astype(np.float32)
    n = oracle.normalize(a, "minmax", rng=(10.0, 20.0))
    assert abs(n.min() - 10) < 1e-4 and abs(n.max() - 20) < 1e-4
    z = oracle.normalize(a, "zscore")
    assert abs(z.mean()) < 1e-4 and abs(z.std() - 1) < 1e-3


@pytest.mark.parametrize("n,i1,i99,step", [(16, 0, 15, 1), (1024, 10, 1013, 1), (128 * 128, 163, 16220, 1),
                                            (65536, 655, 64880, 1), (65537, 655, 64881, 1), (10 ** 6, 666, 66000, 15),
                                            (4096 * 4096, 655, 64880, 256)])
def test_percentile_index_arithmetic(oracle, n, i1, i99, step):       # src/terrain_stats.rs:23-33
    a = np.arange(n, dtype=np.float32)
    lo, hi = oracle.min_max(a, True)
    assert lo == np.float32(i1 * step) and hi == np.float32(i99 * step)


def test_percentile_is_order_statistic_of_strided_sample(oracle):
    rng = np.random.default_rng(7)
    a = rng.normal(size=300_000).astype(np.float32)
    step = a.size // 65536
    s = np.sort(a[::step], kind="stable")
    lo, hi = oracle.min_max(a, True)
    assert lo == s[int(np.float32(s.size) * np.float32(0.01))] and hi == s[int(np.float32(s.size) * np.float32(0.99))]
    assert oracle.min_max(a, False) == (a.min(), a.max())


def test_h_range_guard(oracle):                                        # src/renderer.rs:24-28
    lo, hi = oracle.h_range(np.full((8, 8), 3.0, np.float32))
    assert lo == np.float32(3.0) and hi == np.float32(3.0) + np.float32(1e-5)


def test_stable_sort_keeps_sign_of_zero(oracle):
    a = np.zeros(100, np.float32)
    a[::2] = -0.0
    lo, _ = oracle.min_max(a, True)              # index 1 of a stable sort of all-equal keys
    assert np.signbit(lo) == np.signbit(a[1])


# ---- colour known answers -------------------------------------------------------
def test_srgb_known_answers(oracle):
    assert [oracle.srgb_encode_u8(x) for x in (0.02, 0.02, 0.03)] == [39, 39, 48]   # terrain clear
    assert oracle.srgb_encode_u8(1.0) == 255 and oracle.srgb_encode_u8(0.2) == 124
    assert oracle.srgb_encode_u8(-1.0) == 0 and oracle.srgb_encode_u8(float("nan")) == 0
    for k in (0, 1, 17, 128, 254, 255):
        assert oracle.srgb_encode_u8(oracle.srgb8_to_linear(k)) == k


def test_lut_fixture_texels(luts):
    want = {"viridis": ((68, 1, 84), (32, 144, 140), (253, 231, 36)),
            "magma": ((0, 0, 3), (182, 54, 121), (251, 252, 191)),
            "terrain": ((0, 0, 127), (52, 177, 25), (255, 255, 255))}
    for name, (a, b, c) in want.items():
        t = luts[name].reshape(256, 4)
        assert tuple(t[0, :3]) == a and tuple(t[128, :3]) == b and tuple(t[255, :3]) == c
        assert (t[:, 3] == 255).all()


def test_lut_linearize_matches_formula(oracle, luts):                  # src/colormap/mod.rs:59-79
    out = oracle.lut_linearize_u8(luts["viridis"]).reshape(256, 4)
    s = luts["viridis"].reshape(256, 4)[:, :3].astype(np.float32) / np.float32(255)
    lin = np.where(s <= 0.04045, s / 12.92, ((s + 0.055) / 1.055) ** 2.4)
    assert np.abs(out[:, :3].astype(int) - np.floor(lin * 255 + 0.5).astype(int)).max() <= 1
    assert (out[:, 3] == 255).all()


# ---- rendering known answers ----------------------------------------------------------------
def test_triangle_known_pixels(oracle):                                # src/lib.rs:72-91, triangle.wgsl
    img = oracle.render_triangle(256, 256)
    assert img.shape == (256, 256, 4) and img.dtype == np.uint8
    assert tuple(img[0, 0]) == (255, 255, 255, 255) and tuple(img[250, 128]) == (255, 255, 255, 255)
    covered = (img[..., :3] != 255).any(-1).sum()
    assert abs(covered - 0.5 * (0.8 * 256) * (0.8 * 256)) < 400        # area of the triangle in pixels
    assert (img[..., 3] == 255).all()
    # bottom-left vertex is red (1,.2,.2) -> sRGB (255,124,124): the pixel just inside it is close
    r, g, b, _ = img[229, 27]
    assert r > 240 and abs(int(g) - 124) < 12 and abs(int(b) - 124) < 12


def test_terrain_clear_colour_and_origin_projection(oracle, luts):
    u = oracle.default_uniforms("scene", 800, 600)
    img, prim = oracle.render_terrain(800, 600, 128, np.zeros((1, 1), np.float32), u, luts["viridis"], want_prim=True)
    assert tuple(img[0, 0]) == (39, 39, 48, 255)                       # (0.02,0.02,0.03,1) -> sRGB8
    assert prim[300, 400] != 0                                         # world origin -> NDC (0,0): centre is covered
    assert (prim > 0).mean() > 0.05


def test_two_triangles_share_edge_without_gap_or_overlap(oracle, luts):
    # top-left rule: a full-screen flat quad (G=2) viewed head-on covers every pixel exactly once
    view = oracle.look_at((0, 5, 0), (0, 0, 0), (0, 0, -1))
    proj = np.array([[1 / 1.5, 0, 0, 0], [0, 1 / 1.5, 0, 0], [0, 0, -0.01, 0.5], [0, 0, 0, 1]], np.float32)
    u = oracle.uniforms(view, proj)
    img, prim, cnt = oracle.render_terrain(64, 64, 2, np.zeros((1, 1), np.float32), u, luts["viridis"],
                                           mode=oracle.MODE_NORTHSTAR, want_prim=True, want_counters=True)
    assert (prim > 0).all() and cnt["fragments"] == 64 * 64
