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
