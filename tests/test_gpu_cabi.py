"""The drop-in boundary itself: libvfcuda.so driven through ctypes exactly as a foreign host
(Rust `extern "C"`, cgo, ...) would, with plain pointers and sizes."""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib(vf_build):
    L = C.CDLL(vf_build.lib_path())
    L.vf_last_error.restype = C.c_char_p
    assert L.vf_init(-1) == 0, L.vf_last_error()
    return L


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def test_cabi_triangle_grid_stats(lib, oracle):
    out = np.zeros((48, 64, 4), np.uint8)
    assert lib.vf_triangle_render(64, 48, _p(out, C.c_uint8)) == 0
    assert np.array_equal(out, oracle.render_triangle(64, 48))
    xy, uv, idx = np.zeros((20, 2), np.float32), np.zeros((20, 2), np.float32), np.zeros(72, np.uint32)
    assert lib.vf_grid_generate(5, 4, C.c_float(0.5), C.c_float(2.0), _p(xy, C.c_float), _p(uv, C.c_float), _p(idx, C.c_uint32)) == 0
    o = oracle.make_grid(5, 4, 0.5, 2.0)
    assert np.array_equal(xy, o[0]) and np.array_equal(uv, o[1]) and np.array_equal(idx, o[2])
    assert lib.vf_grid_generate(1, 4, C.c_float(1), C.c_float(1), _p(xy, C.c_float), _p(uv, C.c_float), _p(idx, C.c_uint32)) != 0
    assert b"nx and nz must be >= 2" in lib.vf_last_error()
    a = np.random.default_rng(0).normal(size=100_000).astype(np.float32)
    lo, hi = C.c_float(), C.c_float()
    assert lib.vf_dem_minmax(_p(a, C.c_float), C.c_size_t(a.size), 1, C.byref(lo), C.byref(hi)) == 0
    assert (np.float32(lo.value), np.float32(hi.value)) == oracle.min_max(a, True)
    st = np.zeros(4, np.float32)
    assert lib.vf_dem_stats(_p(a, C.c_float), C.c_size_t(a.size), _p(st, C.c_float)) == 0
    assert st[0] == a.min() and st[1] == a.max() and abs(st[2] - a.mean(dtype=np.float64)) < 1e-6 and abs(st[3] - a.std(dtype=np.float64)) < 1e-5


def test_cabi_scene_lifecycle(lib, oracle, luts, synthetic_dem):
    scene = C.c_void_p()
    assert lib.vf_scene_create(160, 120, 96, 1, 0, C.byref(scene)) == 0, lib.vf_last_error()
    dem = synthetic_dem(1, 96)
    assert lib.vf_scene_set_height_f32(scene, _p(dem, C.c_float), 96, 96) == 0
    u = oracle.look_at_uniforms(160, 120, (2.0, 2.5, -3.0))
    assert lib.vf_scene_set_uniforms(scene, _p(u, C.c_float)) == 0
    back = np.zeros(44, np.float32)
    assert lib.vf_scene_get_uniforms(scene, _p(back, C.c_float)) == 0 and np.array_equal(back, u)
    lo, hi = C.c_float(), C.c_float()
    assert lib.vf_scene_auto_height_range(scene, C.byref(lo), C.byref(hi)) == 0
    assert (np.float32(lo.value), np.float32(hi.value)) == oracle.h_range(dem)
    assert lib.vf_scene_set_mode(scene, 1, 1, lo, hi) == 0
    assert lib.vf_scene_set_mode(scene, 1, 1, C.c_float(1.0), C.c_float(1.0)) != 0 and b"min must be < max" in lib.vf_last_error()
    out = np.zeros((120, 160, 4), np.uint8)
    assert lib.vf_scene_render_rgba(scene, _p(out, C.c_uint8)) == 0, lib.vf_last_error()
    ref = oracle.render_terrain(160, 120, 96, dem, u, luts["magma"], mode=oracle.MODE_NORTHSTAR, depth=True,
                                h_min=lo.value, h_max=hi.value)
    assert np.array_equal(out, ref)
    us = np.stack([oracle.look_at_uniforms(160, 120, (3 * np.cos(t), 2.0, 3 * np.sin(t))) for t in (0.1, 1.3, 2.9)])
    batch = np.zeros((3, 120, 160, 4), np.uint8)
    assert lib.vf_scene_render_batch(scene, _p(us, C.c_float), 3, _p(batch, C.c_uint8)) == 0
    for k in range(3):
        r = oracle.render_terrain(160, 120, 96, dem, us[k], luts["magma"], mode=oracle.MODE_NORTHSTAR, depth=True,
                                  h_min=lo.value, h_max=hi.value)
        assert np.array_equal(batch[k], r)
    assert lib.vf_scene_render_device(scene, _p(us, C.c_float), 3) == 0
    vis = np.zeros((120, 160), np.uint32)
    assert lib.vf_scene_debug_visibility(scene, _p(vis, C.c_uint32)) == 0 and (vis > 0).any()
    lib.vf_scene_destroy(scene)
    bad = C.c_void_p()
    assert lib.vf_scene_create(0, 10, 8, 0, 0, C.byref(bad)) != 0 and lib.vf_scene_create(8, 8, 8, 9, 0, C.byref(bad)) != 0


def test_cabi_height_image_and_png(lib, tmp_path):
    h = np.random.default_rng(1).normal(size=(17, 61)).astype(np.float32)
    t = C.c_void_p()
    assert lib.vf_height_upload(_p(h, C.c_float), 61, 17, C.byref(t)) == 0
    patch = np.zeros((5, 7), np.float32)
    assert lib.vf_height_read_patch(t, 50, 10, 7, 5, _p(patch, C.c_float)) == 0 and np.array_equal(patch, h[10:15, 50:57])
    assert lib.vf_height_read_patch(t, 60, 0, 2, 1, _p(patch, C.c_float)) != 0
    lib.vf_height_destroy(t)
    img = np.random.default_rng(2).integers(0, 255, (40, 50, 4), dtype=np.uint8)
    p = str(tmp_path / "a.png").encode()
    assert lib.vf_png_write(p, _p(img, C.c_uint8), 50, 40) == 0
    from PIL import Image
    assert np.array_equal(np.asarray(Image.open(p.decode())), img)


def _scene_frame(lib, W, H, G, dem, u, mode, depth, lo, hi):
    scene = C.c_void_p()
    assert lib.vf_scene_create(W, H, G, 0, 0, C.byref(scene)) == 0, lib.vf_last_error()
    assert lib.vf_scene_set_height_f32(scene, _p(dem, C.c_float), dem.shape[1], dem.shape[0]) == 0
    assert lib.vf_scene_set_uniforms(scene, _p(u, C.c_float)) == 0
    assert lib.vf_scene_set_mode(scene, mode, depth, C.c_float(lo), C.c_float(hi)) == 0
    out = np.zeros((H, W, 4), np.uint8)
    assert lib.vf_scene_render_rgba(scene, _p(out, C.c_uint8)) == 0, lib.vf_last_error()
    lib.vf_scene_destroy(scene)
    return out


@pytest.mark.parametrize("kind", ["sheared", "ortho"])
def test_cabi_custom_matrices(lib, oracle, luts, synthetic_dem, kind):
    """Uniform blocks a foreign host may hand over that the camera API never produces: a sheared
    projection (general matrices: tiled k_geom) and a symmetric orthographic one (w == 1: the
    sparse form of k_geom_march).  Both bit-identical to the oracle in NORTHSTAR mode."""
    W, H, G = 200, 150, 128
    dem = synthetic_dem(5, G)
    u = oracle.look_at_uniforms(W, H, (2.5, 2.0, 2.0)).copy()
    if kind == "sheared":
        u[16 + 4] = np.float32(0.15)          # P[4]: x += 0.15 * y_view
        u[16 + 1] = np.float32(-0.05)         # P[1]
    else:
        P = np.zeros(16, np.float32)          # column-major symmetric orthographic, depth 0..1
        P[0], P[5], P[10], P[14], P[15] = 0.45, 0.6, -1.0 / 20.0, 0.0, 1.0
        u[16:32] = P
    for depth in (0, 1):
        got = _scene_frame(lib, W, H, G, dem, u, 1, depth, -0.3, 0.4)
        ref = oracle.render_terrain(W, H, G, dem, u, luts["viridis"], mode=oracle.MODE_NORTHSTAR, depth=bool(depth),
                                    h_min=-0.3, h_max=0.4)
        assert (ref[..., :3] != ref[0, 0, :3]).any(), "test camera sees nothing"
        assert np.array_equal(got, ref), f"{kind} depth={depth}: {(got != ref).any(-1).sum()} pixels differ"


def test_cabi_march_equals_tiled_geometry(lib, synthetic_dem, monkeypatch):
    """k_geom_march (persistent warp workers) and the tiled k_geom (VF_GEOM_TILED=1) write the same
    visibility: frames are byte-identical, fine grid and coarse grid, with and without depth."""
    for (W, H, G, seed) in [(320, 200, 512, 3), (160, 120, 9, 4), (257, 131, 100, 6)]:
        dem = synthetic_dem(seed, G)
        import vf_oracle as o
        u = o.look_at_uniforms(W, H, (3.0, 1.2, -2.0))
        for mode in (0, 1):
            for depth in (0, 1):
                monkeypatch.delenv("VF_GEOM_TILED", raising=False)
                a = _scene_frame(lib, W, H, G, dem, u, mode, depth, -0.3, 0.4)
                monkeypatch.setenv("VF_GEOM_TILED", "1")
                b = _scene_frame(lib, W, H, G, dem, u, mode, depth, -0.3, 0.4)
                monkeypatch.delenv("VF_GEOM_TILED", raising=False)
                assert np.array_equal(a, b), (W, H, G, mode, depth)


def test_cabi_division_helpers_equal_ieee_division(lib):
    """vf_selftest_div: the shading's division helpers (csrc/vf_device.cuh: the instruction sequence div.rn / rcp.rn
    take themselves for operands of ordinary magnitude, without the range check and with a shared reciprocal) against
    the plain correctly rounded division, on the device, bit for bit: 2^27 groups of random, special, window-edge and
    zero operands per seed (13 comparisons per group)."""
    lib.vf_selftest_div.argtypes = [C.c_uint64, C.c_uint32, C.POINTER(C.c_uint64)]
    for seed in (1, 20261105):
        bad = C.c_uint64(123)
        assert lib.vf_selftest_div(1 << 27, seed, C.byref(bad)) == 0, lib.vf_last_error()
        assert bad.value == 0
    assert lib.vf_selftest_div(0, 0, C.byref(bad)) == 0 and bad.value == 0
    assert lib.vf_selftest_div(16, 0, None) != 0


def test_cabi_srgb_encoder_equals_threshold_search_for_every_float(lib):
    """vf_selftest_srgb: the shading's sRGB8 encoder (approximate OETF + one correction step) against the binary search
    over the exact thresholds (whose codes tests/test_pixel_known_answers.py checks against the float64 OETF), for all 1 065 353 217
    floats of [0, 1] and 2^24 bit patterns outside (above 1, negative, infinities, NaNs)."""
    lib.vf_selftest_srgb.argtypes = [C.POINTER(C.c_uint64)]
    bad = C.c_uint64(7)
    assert lib.vf_selftest_srgb(C.byref(bad)) == 0, lib.vf_last_error()
    assert bad.value == 0
