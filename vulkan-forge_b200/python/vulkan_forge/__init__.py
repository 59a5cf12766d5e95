"""vulkan_forge -- B200-native drop-in for the headless render path of milos-agathon/vulkan-forge.

Same public surface as the reference shim (python/vulkan_forge/__init__.py:46-178): Renderer,
TerrainSpike, render_triangle_rgba/png, make_terrain, colormap_supported, camera_look_at,
camera_perspective, camera_view_proj, dem_stats, dem_normalize, grid_generate, generate_grid,
__version__.  `Scene` is reached as vulkan_forge._vulkan_forge.Scene, as in the reference.

Underneath, `_vulkan_forge` is a C++ host module over libvfcuda.so: hand-written sm_100a CUDA
kernels behind the C ABI of include/vf_cuda.h.  There is no CPU fallback: importing works
anywhere, but constructing a renderer without a CUDA device raises RuntimeError.
"""
from __future__ import annotations

import numpy as _np

from . import _vulkan_forge as _ext          # fails loudly if the native module is not built
from ._validate import grid as _check_grid
from ._validate import png_path as _check_png
from ._validate import size_wh as _check_size

__version__ = "0.0.8"                        # reference Cargo.toml:3 / pyproject.toml:7

Renderer = _ext.Renderer
TerrainSpike = _ext.TerrainSpike             # the reference's wheels enable it (pyproject.toml:18)
Scene = _ext.Scene                           # additive re-export; the reference only has _vulkan_forge.Scene
colormap_supported = _ext.colormap_supported
camera_look_at = _ext.camera_look_at
camera_perspective = _ext.camera_perspective
camera_view_proj = _ext.camera_view_proj
enumerate_adapters = _ext.enumerate_adapters
device_probe = _ext.device_probe


def render_triangle_rgba(width: int, height: int):
    """(H, W, 4) uint8 of the deterministic triangle (reference __init__.py:54-58)."""
    w, h = _check_size(width, height)
    return Renderer(w, h).render_triangle_rgba()


def render_triangle_png(path, width: int, height: int) -> None:
    """Write the deterministic triangle to ``path`` (reference __init__.py:60-64)."""
    w, h = _check_size(width, height)
    target = _check_png(path)                # all argument errors surface before the device is touched
    Renderer(w, h).render_triangle_png(target)


def make_terrain(width: int, height: int, grid: int = 128):
    """TerrainSpike after Python-side validation (reference __init__.py:66-76)."""
    w, h = _check_size(width, height)
    return TerrainSpike(w, h, _check_grid(grid))


def dem_stats(heightmap):
    """(min, max, mean, std) of a 2-D float32/float64 C-contiguous array (reference :120-127)."""
    a = _np.asarray(heightmap)
    if a.ndim != 2 or a.dtype not in (_np.float32, _np.float64) or not a.flags["C_CONTIGUOUS"]:
        raise RuntimeError("heightmap must be 2-D float32/float64 and C-contiguous")
    a = a.astype(_np.float32, copy=False)
    return float(a.min()), float(a.max()), float(a.mean()), float(a.std(dtype=_np.float32))


def dem_normalize(heightmap, *, mode="minmax", out_range=(0.0, 1.0), eps=1e-8, return_stats=False):
    """numpy-side normalisation helper (reference :129-142)."""
    mn, mx, mean, std = dem_stats(heightmap)
    a = _np.asarray(heightmap).astype(_np.float32, copy=False)
    if mode == "minmax":
        lo, hi = (float(v) for v in out_range)
        scale = 0.0 if mx == mn else (hi - lo) / max(mx - mn, float(eps))
        out = (a - mn) * scale + lo
    elif mode == "zscore":
        out = (a - mean) / max(std, float(eps))
    else:
        raise ValueError("mode must be 'minmax' or 'zscore'")
    return (out, (mn, mx, mean, std)) if return_stats else out


def grid_generate(nx: int, nz: int, spacing=(1.0, 1.0), origin="center"):
    """Regular grid mesh: (XY (N,2) f32, UV (N,2) f32, indices (M,) u32) (reference :153-169)."""
    return _ext.grid_generate(int(nx), int(nz), tuple(float(s) for s in spacing), str(origin))


generate_grid = grid_generate                # legacy alias (reference :172)


def report_environment() -> dict:
    """Environment report for bug reports and bench logs (planned in the reference, ROADMAP.md:1989-1999: OS, Python,
    package version, adapter name / backend / limits / features, library versions).  The backend here is CUDA; the
    adapter dictionaries have the shape of enumerate_adapters() (reference src/lib.rs:744-845)."""
    import os
    import platform
    import sys
    drv, rt = _ext.cuda_versions()
    fmt = lambda v: f"{v // 1000}.{(v % 1000) // 10}" if v else None   # noqa: E731
    try:
        adapters = list(enumerate_adapters())
    except RuntimeError:
        adapters = []
    return {
        "os": {"system": platform.system(), "release": platform.release(), "machine": platform.machine()},
        "python": sys.version.split()[0], "executable": sys.executable, "numpy": _np.__version__,
        "package": "vulkan_forge", "package_version": __version__, "backend": "CUDA (sm_100a kernels, no graphics API)",
        "adapters": adapters, "cuda_driver_version": fmt(drv), "cuda_runtime_version": fmt(rt),
        "native_library": os.path.join(os.path.dirname(os.path.abspath(__file__)), "libvfcuda.so"),
        "build": _ext.build_info(),
        "env": {k: os.environ[k] for k in ("CUDA_VISIBLE_DEVICES", "LOCAL_RANK", "VF_FORCE_LUT_UNORM", "VF_DISABLE_TMA") if k in os.environ},
    }

from .orbit import bind_host_to_gpu, orbit_cameras, render_orbit, shard_range  # noqa: E402  (additive multi-GPU helper)

__all__ = [
    "Renderer", "TerrainSpike", "render_triangle_rgba", "render_triangle_png", "make_terrain",
    "colormap_supported", "camera_look_at", "camera_perspective", "camera_view_proj",
    "dem_stats", "dem_normalize", "grid_generate", "generate_grid", "__version__",
    "Scene", "enumerate_adapters", "device_probe", "report_environment", "orbit_cameras", "render_orbit", "shard_range", "bind_host_to_gpu",
]
