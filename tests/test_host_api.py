"""Host-side drop-in surface, runnable without a GPU: names, validation, exception types and
message strings, camera math against the oracle, the C-ABI export list, the PNG writer.
Re-expresses tests/test_api_validation.py, test_camera.py, test_colormap.py:1-60,
test_grid_generate.py:82-104, test_tonemap.py:26-31 of the reference."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_public_names(vf, ext):
    for n in ("Renderer", "TerrainSpike", "render_triangle_rgba", "render_triangle_png", "make_terrain",
              "colormap_supported", "camera_look_at", "camera_perspective", "camera_view_proj", "dem_stats",
              "dem_normalize", "grid_generate", "generate_grid", "__version__"):
        assert hasattr(vf, n), n
        assert n in vf.__all__
    for n in ("Renderer", "TerrainSpike", "Scene", "enumerate_adapters", "device_probe", "grid_generate",
              "colormap_supported", "camera_look_at", "camera_perspective", "camera_view_proj"):
        assert hasattr(ext, n), n                                       # src/lib.rs:961-976
    for cls in (ext.Scene, ext.TerrainSpike):
        for m in ("render_png", "set_camera_look_at", "debug_uniforms_f32", "debug_lut_format", "render_rgba",
                  "set_height_from_r32f", "set_height_range", "set_auto_height_range", "set_sun", "set_exposure",
                  "set_spacing_exaggeration", "set_shading", "render_batch_rgba"):
            assert hasattr(cls, m), m
    for m in ("info", "render_triangle_rgba", "render_triangle_png", "add_terrain", "terrain_stats", "set_height_range",
              "set_sun", "set_exposure", "normalize_terrain", "upload_height_r32f", "debug_read_height_patch",
              "read_full_height_texture", "render_rgba"):
        assert hasattr(ext.Renderer, m), m
    assert vf.generate_grid is vf.grid_generate
    import vshade
    assert vshade.Renderer is vf.Renderer


def test_python_validation_raises_value_error(vf, tmp_path):           # tests/test_api_validation.py:15-38
    with pytest.raises(ValueError):
        vf.render_triangle_rgba(0, 16)
    with pytest.raises(ValueError):
        vf.render_triangle_rgba(16, -2)
    with pytest.raises(ValueError):
        vf.render_triangle_rgba(8193, 16)
    with pytest.raises(ValueError):
        vf.render_triangle_png(str(tmp_path / "x.jpg"), 16, 16)
    with pytest.raises(ValueError):
        vf.render_triangle_png(str(tmp_path / "missing_dir" / "x.png"), 16, 16)
    with pytest.raises(ValueError):
        vf.make_terrain(64, 64, grid=1)
    with pytest.raises(ValueError):
        vf.make_terrain(64, 64, grid=4097)


def test_grid_generate_validation_strings(vf):                         # tests/test_grid_generate.py:82-104
    with pytest.raises(ValueError, match="nx and nz must be >= 2"):
        vf.grid_generate(1, 3)
    with pytest.raises(ValueError, match="nx and nz must be >= 2"):
        vf.grid_generate(3, 1)
    for sp in ((0.0, 1.0), (1.0, -1.0), (float("inf"), 1.0)):
        with pytest.raises(ValueError, match="spacing components must be finite and > 0"):
            vf.grid_generate(3, 3, spacing=sp)
    with pytest.raises(ValueError, match="origin must be 'center'"):
        vf.grid_generate(3, 3, origin="corner")


def test_colormap_registry(vf):                                        # tests/test_colormap.py
    assert vf.colormap_supported() == ["viridis", "magma", "terrain"]


def test_camera_functions_match_oracle(vf, oracle):
    rng = np.random.default_rng(3)
    for _ in range(50):
        eye = tuple(float(x) for x in rng.uniform(-5, 5, 3))
        tgt = tuple(float(x) for x in rng.uniform(-1, 1, 3))
        fovy, asp = float(rng.uniform(10, 120)), float(rng.uniform(0.3, 3))
        zn = float(rng.uniform(0.01, 1))
        zf = zn + float(rng.uniform(1, 500))
        v = vf.camera_look_at(eye, tgt, (0, 1, 0))
        assert v.shape == (4, 4) and v.dtype == np.float32 and v.flags.c_contiguous
        assert np.array_equal(v, oracle.look_at(eye, tgt, (0, 1, 0)))
        for clip in ("wgpu", "gl"):
            assert np.array_equal(vf.camera_perspective(fovy, asp, zn, zf, clip), oracle.perspective(fovy, asp, zn, zf, clip))
            assert np.array_equal(vf.camera_view_proj(eye, tgt, (0, 1, 0), fovy, asp, zn, zf, clip),
                                  oracle.view_proj(eye, tgt, (0, 1, 0), fovy, asp, zn, zf, clip))
    assert np.array_equal(vf.camera_perspective(45.0, 1.0, 0.1, 100.0), vf.camera_perspective(45.0, 1.0, 0.1, 100.0, "wgpu"))
    assert abs(vf.camera_look_at((0, 0, 3), (0, 0, 0), (0, 1, 0))[2, 3] + 3.0) < 1e-6   # tests/test_camera.py:56-68


def test_camera_error_strings_are_runtime_errors(vf):                  # tests/test_camera.py:26-34,70-167
    ok = ((0.0, 0.0, 3.0), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0))
    with pytest.raises(RuntimeError, match=re.escape("eye/target/up components must be finite")):
        vf.camera_look_at((float("nan"), 0, 3), ok[1], ok[2])
    with pytest.raises(RuntimeError, match=re.escape("up vector must not be colinear with view direction")):
        vf.camera_look_at(ok[0], ok[1], (0.0, 0.0, -1.0))
    for f in (0.0, 180.0, -5.0, float("inf")):
        with pytest.raises(RuntimeError, match=re.escape("fovy_deg must be finite and in (0, 180)")):
            vf.camera_perspective(f, 1.0, 0.1, 100.0)
    with pytest.raises(RuntimeError, match=re.escape("aspect must be finite and > 0")):
        vf.camera_perspective(45.0, 0.0, 0.1, 100.0)
    with pytest.raises(RuntimeError, match=re.escape("znear must be finite and > 0")):
        vf.camera_perspective(45.0, 1.0, 0.0, 100.0)
    with pytest.raises(RuntimeError, match=re.escape("zfar must be finite and > znear")):
        vf.camera_perspective(45.0, 1.0, 1.0, 1.0)
    with pytest.raises(RuntimeError, match=re.escape("clip_space must be 'wgpu' or 'gl'")):
        vf.camera_perspective(45.0, 1.0, 0.1, 100.0, "dx")
    with pytest.raises(RuntimeError, match=re.escape("fovy_deg must be finite and in (0, 180)")):
        vf.camera_view_proj(*ok, 0.0, 1.0, 0.1, 100.0)
    with pytest.raises(RuntimeError, match=re.escape("up vector must not be colinear with view direction")):
        vf.camera_view_proj(ok[0], ok[1], (0.0, 0.0, -1.0), 45.0, 1.0, 0.1, 100.0)


def test_dem_python_helpers(vf):                                       # tests/test_dem_normalization.py
    x = np.linspace(-1, 1, 48, dtype=np.float32)
    y = np.linspace(-1, 1, 32, dtype=np.float32)
    X, Y = np.meshgrid(x, y)
    Z = (0.25 * np.sin(1.3 * X) + 0.25 * np.cos(1.1 * Y)).astype(np.float32)
    mn, mx, mean, std = vf.dem_stats(Z)
    assert mx > mn and std > 0
    N = vf.dem_normalize(Z, mode="minmax", out_range=(0.0, 1.0))
    assert N.shape == Z.shape and N.dtype == np.float32 and N.min() >= -1e-5 and N.max() <= 1 + 1e-5
    with pytest.raises(ValueError):
        vf.dem_normalize(Z, mode="nope")
    with pytest.raises(RuntimeError):
        vf.dem_stats(Z.T)


def test_c_abi_exports_every_declared_symbol(vf_build):
    """libvfcuda.so loads without a GPU and exports exactly what include/vf_cuda.h declares."""
    hdr = open(os.path.join(ROOT, "include", "vf_cuda.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = sorted(set(re.findall(r"\b(vf_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 30
    lib = ctypes.CDLL(vf_build.lib_path())
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    lib.vf_last_error.restype = ctypes.c_char_p
    assert isinstance(lib.vf_last_error(), bytes)
    out = (ctypes.c_uint8 * 1024)()
    assert lib.vf_colormap_rgba8(0, out) == 0 and lib.vf_colormap_rgba8(7, out) != 0
    assert b"colormap" in lib.vf_last_error()


def test_embedded_palettes_equal_golden_fixtures(ext, luts):
    for n in ("viridis", "magma", "terrain"):
        assert np.array_equal(ext.colormap_rgba8(n).ravel(), luts[n])
    with pytest.raises(RuntimeError, match="Unknown colormap 'jet'. Supported: viridis, magma, terrain"):
        ext.colormap_rgba8("jet")


def test_png_writer_roundtrip(ext, tmp_path):
    from PIL import Image
    rng = np.random.default_rng(1)
    for shape in ((7, 5), (240, 320), (1080, 1920)):
        img = rng.integers(0, 256, shape + (4,), dtype=np.uint8)
        img[: shape[0] // 2] = (img[: shape[0] // 2] // 64) * 64        # some compressible rows
        p = tmp_path / f"t{shape[0]}.png"
        ext.png_write(str(p), img)
        back = np.asarray(Image.open(p))
        assert back.shape == img.shape and np.array_equal(back, img)


@pytest.mark.skipif(_has_gpu(), reason="only meaningful without a CUDA device")
def test_no_cpu_fallback_without_gpu(vf, ext):
    """The product path must fail loudly when no CUDA device exists (no oracle / CPU fallback)."""
    with pytest.raises(RuntimeError, match="No suitable GPU adapter"):
        vf.Renderer(16, 16)
    with pytest.raises(RuntimeError, match="No suitable GPU adapter"):
        ext.Scene(32, 32)
    with pytest.raises(RuntimeError, match="No suitable GPU adapter"):
        vf.grid_generate(4, 4)
    with pytest.raises(RuntimeError, match="No suitable GPU adapter"):
        ext.dem_minmax(np.zeros(8, np.float32), True)
    assert ext.enumerate_adapters() == [] and ext.device_probe()["status"] == "unsupported"
    with pytest.raises(RuntimeError, match="Unknown colormap"):        # validation precedes device use
        ext.Scene(32, 32, colormap="jet")


def test_product_package_never_imports_oracle():
    pkg = os.path.join(ROOT, "vulkan-forge_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp", ".inc")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "vf_oracle" not in text and "libvfo" not in text and "vfo_" not in text, f


def test_integration_doc_matches_header():
    """INTEGRATION.md's Rust block binds exactly the functions include/vf_cuda.h declares, and its #[repr(C)] structs
    have the header's fields in the header's order with matching types (a host built from the document must not read a
    timing shifted by four bytes)."""
    import re
    hdr = open(os.path.join(ROOT, "include", "vf_cuda.h")).read()
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    rust = doc[doc.index("```rust"):doc.index("```\n", doc.index("```rust") + 8)]
    declared = set(re.findall(r"^(?:int|void|size_t|const char \*)\s*\*?(vf_\w+)\(", hdr, re.M))
    bound = set(re.findall(r"pub fn (vf_\w+)\(", rust))
    assert declared == bound, (sorted(declared - bound), sorted(bound - declared))
    ctype = {"uint32_t": "u32", "uint64_t": "u64", "int": "i32", "float": "f32", "char": "c_char"}

    def c_fields(name):
        body = re.search(r"typedef struct \{([^{}]*)\}\s*" + name + ";", hdr).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        out = []
        for decl in body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            ty, rest = decl.split(None, 1)
            for item in rest.split(","):
                m = re.match(r"\s*(\w+)(?:\[(\d+)\])?\s*$", item)
                out.append((m.group(1), ctype[ty] if not m.group(2) else f"[{ctype[ty]}; {m.group(2)}]"))
        return out

    def rust_fields(name):
        body = re.search(r"pub struct " + name + r" \{(.*?)\}", rust, re.S).group(1)
        return [(m.group(1), m.group(2).strip()) for m in re.finditer(r"pub (\w+): ([^,}]+)", body)]

    for c_name, r_name in (("vf_device_info_t", "VfDeviceInfo"), ("vf_frame_stats_t", "VfFrameStats"), ("vf_render_metrics_t", "VfRenderMetrics")):
        assert c_fields(c_name) == rust_fields(r_name), (c_name, c_fields(c_name), rust_fields(r_name))
