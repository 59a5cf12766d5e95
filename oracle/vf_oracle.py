"""ctypes loader for the CPU oracle (oracle/libvfo.so).

TEST INFRASTRUCTURE ONLY: import this from tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py, never from the product package.
Each wrapper cites the reference file:line its C implementation follows (see vfo.h).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libvfo.so")

MODE_SHIPPED = 0
MODE_NORTHSTAR = 1

CAMERA_ERRORS = {
    1: "eye/target/up components must be finite",
    2: "up vector must not be colinear with view direction",
    3: "fovy_deg must be finite and in (0, 180)",
    4: "znear must be finite and > 0",
    5: "zfar must be finite and > znear",
}


def build(force: bool = False) -> str:
    """Compile oracle/libvfo.so with the committed Makefile (gcc only)."""
    srcs = [os.path.join(_HERE, n) for n in ("vfo_host.c", "vfo_render.c", "vfo.h", "Makefile")]
    stale = (not os.path.exists(_LIB_PATH)) or any(
        os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-B", "libvfo.so"],
                              stdout=subprocess.DEVNULL)
    return _LIB_PATH


class _Frame(C.Structure):
    _fields_ = [
        ("width", C.c_uint32), ("height", C.c_uint32), ("grid", C.c_uint32),
        ("dem", C.POINTER(C.c_float)), ("dem_w", C.c_uint32), ("dem_h", C.c_uint32),
        ("uniforms", C.c_float * 44),
        ("lut_rgba", C.POINTER(C.c_uint8)), ("lut_unorm", C.c_int),
        ("mode", C.c_int), ("depth", C.c_int),
        ("h_min", C.c_float), ("h_max", C.c_float), ("threads", C.c_int),
    ]


class _Counters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in
                ("tris_total", "tris_front", "tris_sampled", "tris_clipped", "fragments")]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = C.CDLL(_LIB_PATH)
        fp, u32p, u8p = C.POINTER(C.c_float), C.POINTER(C.c_uint32), C.POINTER(C.c_uint8)
        L.vfo_make_grid.argtypes = [C.c_uint32, C.c_uint32, C.c_float, C.c_float, fp, fp, u32p]
        L.vfo_render_grid_xyuv.argtypes = [C.c_uint32, fp, u32p]
        L.vfo_min_max.argtypes = [fp, C.c_size_t, C.c_int, fp, fp]
        L.vfo_h_range.argtypes = [fp, C.c_size_t, fp, fp]
        L.vfo_dem_stats.argtypes = [fp, C.c_size_t, fp]
        L.vfo_normalize.argtypes = [fp, C.c_size_t, C.c_int, C.c_float, C.c_float, C.c_float, fp]
        L.vfo_look_at_rh.argtypes = [fp, fp, fp, fp]
        L.vfo_perspective_rh_gl.argtypes = [C.c_float] * 4 + [fp]
        L.vfo_perspective_wgpu.argtypes = [C.c_float] * 4 + [fp]
        L.vfo_mat4_mul.argtypes = [fp, fp, fp]
        L.vfo_validate_camera.argtypes = [fp, fp, fp, C.c_float, C.c_float, C.c_float]
        L.vfo_validate_camera.restype = C.c_int
        L.vfo_uniforms.argtypes = [fp, fp, fp] + [C.c_float] * 5 + [fp]
        L.vfo_sun_dir.argtypes = [C.c_float, C.c_float, fp]
        L.vfo_default_uniforms.argtypes = [C.c_int, C.c_uint32, C.c_uint32, fp]
        L.vfo_lut_linearize_u8.argtypes = [u8p, C.c_size_t, u8p]
        L.vfo_srgb_encode_u8.argtypes = [C.c_float]
        L.vfo_srgb_encode_u8.restype = C.c_uint8
        L.vfo_srgb8_to_linear.argtypes = [C.c_uint8]
        L.vfo_srgb8_to_linear.restype = C.c_float
        L.vfo_render_triangle.argtypes = [C.c_uint32, C.c_uint32, u8p]
        L.vfo_render_triangle.restype = C.c_int
        L.vfo_render_terrain.argtypes = [C.POINTER(_Frame), u8p, u32p, C.POINTER(_Counters)]
        L.vfo_render_terrain.restype = C.c_int
        L.vfo_max_threads.restype = C.c_int
        _lib = L
    return _lib


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _f3(v):
    return np.ascontiguousarray(np.asarray(v, dtype=np.float32).reshape(3))


# ---- grid ---------------------------------------------------------------------------
def make_grid(nx: int, nz: int, dx: float = 1.0, dy: float = 1.0):
    """src/terrain/mesh.rs:35-90,149-203 -> (XY (N,2) f32, UV (N,2) f32, idx (M,) u32)."""
    xy = np.empty((nx * nz, 2), np.float32)
    uv = np.empty((nx * nz, 2), np.float32)
    idx = np.empty(((nx - 1) * (nz - 1) * 6,), np.uint32)
    lib().vfo_make_grid(nx, nz, dx, dy, _fp(xy), _fp(uv),
                        idx.ctypes.data_as(C.POINTER(C.c_uint32)))
    return xy, uv, idx


def render_grid_xyuv(g: int):
    """src/terrain/mod.rs:553-598 -> (verts (G*G,4) f32 [x,z,u,v], idx u32)."""
    g = max(int(g), 2)
    verts = np.empty((g * g, 4), np.float32)
    idx = np.empty(((g - 1) * (g - 1) * 6,), np.uint32)
    lib().vfo_render_grid_xyuv(g, _fp(verts), idx.ctypes.data_as(C.POINTER(C.c_uint32)))
    return verts, idx


# ---- DEM statistics -----------------------------------------------------------------
def min_max(data, clamp: bool):
    """src/terrain_stats.rs:11-35."""
    a = np.ascontiguousarray(data, dtype=np.float32).ravel()
    lo, hi = C.c_float(), C.c_float()
    lib().vfo_min_max(_fp(a), a.size, int(bool(clamp)), C.byref(lo), C.byref(hi))
    return np.float32(lo.value), np.float32(hi.value)


def h_range(data):
    """src/renderer.rs:24-28."""
    a = np.ascontiguousarray(data, dtype=np.float32).ravel()
    lo, hi = C.c_float(), C.c_float()
    lib().vfo_h_range(_fp(a), a.size, C.byref(lo), C.byref(hi))
    return np.float32(lo.value), np.float32(hi.value)


def dem_stats(data):
    """src/lib.rs:905-932 -> (min, max, mean, std) as float32."""
    a = np.ascontiguousarray(data, dtype=np.float32).ravel()
    out = np.zeros(4, np.float32)
    lib().vfo_dem_stats(_fp(a), a.size, _fp(out))
    return tuple(np.float32(x) for x in out)


def normalize(data, mode: str, eps: float = 1e-8, rng=(0.0, 1.0)):
    """src/lib.rs:934-951 (returns a normalised copy)."""
    a = np.array(data, dtype=np.float32, order="C", copy=True)
    st = np.array(dem_stats(a), np.float32)
    lib().vfo_normalize(_fp(a), a.size, {"minmax": 0, "zscore": 1}[mode.lower()],
                        eps, rng[0], rng[1], _fp(st))
    return a


# ---- camera -------------------------------------------------------------------------
def _mat(colmajor):
    """column-major float[16] -> (4,4) array with M[row, col] (src/camera.rs:94-112)."""
    return np.ascontiguousarray(np.asarray(colmajor, np.float32).reshape(4, 4).T)


def look_at(eye, target, up):
    m = np.zeros(16, np.float32)
    lib().vfo_look_at_rh(_fp(_f3(eye)), _fp(_f3(target)), _fp(_f3(up)), _fp(m))
    return _mat(m)


def perspective(fovy_deg, aspect, znear, zfar, clip_space="wgpu"):
    m = np.zeros(16, np.float32)
    fovy = np.float32(fovy_deg) * np.float32(np.float32(np.pi) / np.float32(180.0))
    fn = lib().vfo_perspective_wgpu if clip_space == "wgpu" else lib().vfo_perspective_rh_gl
    fn(float(fovy), aspect, znear, zfar, _fp(m))
    return _mat(m)


def view_proj(eye, target, up, fovy_deg, aspect, znear, zfar, clip_space="wgpu"):
    v = np.ascontiguousarray(look_at(eye, target, up).T).ravel()
    p = np.ascontiguousarray(perspective(fovy_deg, aspect, znear, zfar, clip_space).T).ravel()
    o = np.zeros(16, np.float32)
    lib().vfo_mat4_mul(_fp(p), _fp(v), _fp(o))
    return _mat(o)


def validate_camera(eye, target, up, fovy_deg, znear, zfar):
    """Returns None or the exact error string of src/camera.rs:24-30."""
    code = lib().vfo_validate_camera(_fp(_f3(eye)), _fp(_f3(target)), _fp(_f3(up)),
                                     fovy_deg, znear, zfar)
    return CAMERA_ERRORS.get(code)


def sun_dir(elevation_deg, azimuth_deg):
    o = np.zeros(3, np.float32)
    lib().vfo_sun_dir(elevation_deg, azimuth_deg, _fp(o))
    return o


def default_uniforms(kind: str, width: int, height: int):
    """kind 'scene' | 'spike' -> 44 f32 (src/terrain/mod.rs:114-142)."""
    o = np.zeros(44, np.float32)
    lib().vfo_default_uniforms(0 if kind == "scene" else 1, width, height, _fp(o))
    return o


def uniforms(view_mat, proj_mat, sun=(0.5, 0.8, 0.6), exposure=1.0, spacing=1.0,
             h_min=-0.5, h_max=0.5, exaggeration=1.0, normalize_sun=True):
    """Pack 44 floats from (4,4) M[row,col] matrices."""
    v = np.ascontiguousarray(np.asarray(view_mat, np.float32).T).ravel()
    p = np.ascontiguousarray(np.asarray(proj_mat, np.float32).T).ravel()
    s = _f3(sun)
    if normalize_sun:
        s = (s / np.float32(np.sqrt(np.float32((s * s).sum())))).astype(np.float32)
    o = np.zeros(44, np.float32)
    lib().vfo_uniforms(_fp(v), _fp(p), _fp(s), exposure, spacing, h_min, h_max, exaggeration, _fp(o))
    return o


def look_at_uniforms(width, height, eye, target=(0, 0, 0), up=(0, 1, 0), fovy_deg=45.0,
                     znear=0.1, zfar=100.0, **kw):
    """What Scene.set_camera_look_at produces (src/scene/mod.rs:209-224)."""
    aspect = float(np.float32(width) / np.float32(height))
    return uniforms(look_at(eye, target, up), perspective(fovy_deg, aspect, znear, zfar), **kw)


# ---- colour -------------------------------------------------------------------------
def lut_linearize_u8(rgba):
    a = np.ascontiguousarray(rgba, dtype=np.uint8).ravel()
    o = np.empty_like(a)
    u8p = C.POINTER(C.c_uint8)
    lib().vfo_lut_linearize_u8(a.ctypes.data_as(u8p), a.size // 4, o.ctypes.data_as(u8p))
    return o


def srgb_encode_u8(x: float) -> int:
    return int(lib().vfo_srgb_encode_u8(float(np.float32(x))))


def srgb8_to_linear(v: int) -> np.float32:
    return np.float32(lib().vfo_srgb8_to_linear(int(v)))


# ---- rendering ----------------------------------------------------------------------
def render_triangle(w: int, h: int):
    """src/lib.rs:287-309 -> (H,W,4) uint8."""
    out = np.empty((h, w, 4), np.uint8)
    rc = lib().vfo_render_triangle(w, h, out.ctypes.data_as(C.POINTER(C.c_uint8)))
    assert rc == 0
    return out


def render_terrain(width, height, grid, dem, uniforms44, lut_rgba, *, lut_unorm=False,
                   mode=MODE_SHIPPED, depth=False, h_min=-0.5, h_max=0.5, threads=0,
                   want_prim=False, want_counters=False):
    """One terrain frame -> (H,W,4) uint8 [, prim (H,W) u32][, counters dict]."""
    dem = np.ascontiguousarray(dem, dtype=np.float32)
    lut = np.ascontiguousarray(lut_rgba, dtype=np.uint8).ravel()
    assert lut.size == 1024 and dem.ndim == 2
    f = _Frame()
    f.width, f.height, f.grid = width, height, max(int(grid), 2)
    f.dem = _fp(dem)
    f.dem_h, f.dem_w = dem.shape
    u = np.ascontiguousarray(uniforms44, dtype=np.float32).ravel()
    assert u.size == 44
    for i in range(44):
        f.uniforms[i] = float(u[i])
    f.lut_rgba = lut.ctypes.data_as(C.POINTER(C.c_uint8))
    f.lut_unorm, f.mode, f.depth = int(lut_unorm), int(mode), int(depth)
    f.h_min, f.h_max, f.threads = h_min, h_max, threads
    out = np.empty((height, width, 4), np.uint8)
    prim = np.empty((height, width), np.uint32) if want_prim else None
    cnt = _Counters()
    rc = lib().vfo_render_terrain(
        C.byref(f), out.ctypes.data_as(C.POINTER(C.c_uint8)),
        prim.ctypes.data_as(C.POINTER(C.c_uint32)) if want_prim else None, C.byref(cnt))
    if rc != 0:
        raise MemoryError("oracle allocation failed")
    res = [out]
    if want_prim:
        res.append(prim)
    if want_counters:
        res.append({n: int(getattr(cnt, n)) for n, _ in _Counters._fields_})
    return res[0] if len(res) == 1 else tuple(res)


def max_threads() -> int:
    return int(lib().vfo_max_threads())
