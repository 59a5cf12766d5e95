"""GPU parity tests proper: the CUDA path (through the Python drop-in -> C ABI) against the CPU
oracle on identical seeded inputs, against the committed golden fixtures, and at BASELINE.json's
full sizes.  Bars: grid / DEM order statistics / NORTHSTAR frames bit-exact; SHIPPED frames (device
sinf/cosf vs libm) within 1 LSB per channel on >= 99.9 % of pixels, coverage differences counted."""
import hashlib
import os

import numpy as np
import pytest

from conftest import frame_diff

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIPPED_BAR = 0.999   # north_star: RGBA8 within 1 LSB per channel on at least 99.9 % of pixels


# ---- triangle (BASELINE config 1) -----------------------------------------------------------
@pytest.mark.parametrize("w,h", [(16, 16), (64, 64), (256, 256), (512, 512), (32, 24), (37, 19), (1, 1), (1920, 1080)])
def test_triangle_bit_exact(vf, oracle, w, h):
    got = vf.render_triangle_rgba(w, h)
    assert got.shape == (h, w, 4) and got.dtype == np.uint8
    assert np.array_equal(got, oracle.render_triangle(w, h))


def test_triangle_determinism_and_golden(vf):                          # tests/test_determinism.py:9-18
    golden = np.load(os.path.join(ROOT, "tests", "golden", "frames.npz"))["triangle_48x32"]
    digests = {hashlib.sha256(vf.Renderer(48, 32).render_triangle_rgba().tobytes()).hexdigest() for _ in range(3)}
    assert digests == {hashlib.sha256(golden.tobytes()).hexdigest()}
    r = vf.Renderer(64, 64)
    a, b = r.render_triangle_rgba(), r.render_triangle_rgba()          # tests/edge_consistency.py
    assert ((a[..., :3] != 255).any(-1)).sum() == ((b[..., :3] != 255).any(-1)).sum() > 0
    assert r.info() == "Renderer 64x64, format=Rgba8UnormSrgb"


def test_triangle_png(vf, tmp_path):                                   # tests/test_api.py, smoke_test.py
    from PIL import Image
    p = tmp_path / "tri.png"
    vf.render_triangle_png(str(p), 128, 96)
    assert p.exists() and p.stat().st_size > 0
    assert np.array_equal(np.asarray(Image.open(p)), vf.render_triangle_rgba(128, 96))


# ---- grid (bit-exact) -------------------------------------------------------------------------
@pytest.mark.parametrize("nx,nz,sp", [(4, 3, (2.0, 1.0)), (3, 3, (2.0, 2.0)), (2, 2, (1.0, 1.0)), (256, 256, (1.0, 1.0)),
                                      (1024, 1024, (1.0, 1.0)), (257, 3, (0.1, 7.5)), (3, 1000, (1e-3, 1e3))])
def test_grid_generate_bit_exact(vf, oracle, nx, nz, sp):
    xy, uv, idx = vf.grid_generate(nx, nz, spacing=sp)
    oxy, ouv, oidx = oracle.make_grid(nx, nz, *sp)
    assert xy.dtype == np.float32 and uv.dtype == np.float32 and idx.dtype == np.uint32
    assert xy.shape == (nx * nz, 2) and idx.shape == ((nx - 1) * (nz - 1) * 6,)
    assert xy.tobytes() == oxy.tobytes() and uv.tobytes() == ouv.tobytes() and np.array_equal(idx, oidx)


def test_grid_reference_values(vf):                                    # tests/test_grid_generate.py:23-41,107-121
    xy, uv, idx = vf.grid_generate(4, 3, spacing=(2.0, 1.0))
    assert [tuple(uv[k]) for k in (0, 3, 8, 11)] == [(0, 0), (1, 0), (0, 1), (1, 1)]
    xy, _, idx = vf.grid_generate(3, 3, spacing=(2.0, 2.0))
    assert xy.tolist() == [[-2, -2], [0, -2], [2, -2], [-2, 0], [0, 0], [2, 0], [-2, 2], [0, 2], [2, 2]]
    p0, p1, p2 = xy[idx[0]], xy[idx[1]], xy[idx[2]]
    assert (p1 - p0)[0] * (p2 - p0)[1] - (p1 - p0)[1] * (p2 - p0)[0] > 0


# ---- DEM statistics -----------------------------------------------------------------------------
@pytest.mark.parametrize("n", [1, 2, 16, 1000, 128 * 128, 65536, 65537, 131071, 10 ** 6, 4096 * 4096])
def test_dem_minmax_bit_exact(ext, oracle, n):
    a = np.random.default_rng(n).normal(size=n).astype(np.float32)
    for clamp in (False, True):
        lo, hi = ext.dem_minmax(a, clamp)
        olo, ohi = oracle.min_max(a, clamp)
        assert np.float32(lo).tobytes() == olo.tobytes() and np.float32(hi).tobytes() == ohi.tobytes()


def test_dem_host_arrays_stream_in_chunks(ext, vf, oracle):
    """Host arrays larger than one 64 MiB chunk: min / max / statistics accumulate across chunks, the sign of a zero
    extreme is that of the first zero of the whole array, and the percentile call moves 256 KB, not the array."""
    import time
    n = (16 << 20) * 2 + 12345
    rng = np.random.default_rng(5)
    a = np.abs(rng.normal(size=n)).astype(np.float32) + np.float32(0.5)
    a[(16 << 20) + 7] = -0.0                      # first zero sits in the second chunk: min must be -0.0
    a[(16 << 20) * 2 + 99] = 0.0
    lo, hi = ext.dem_minmax(a, False)
    olo, ohi = oracle.min_max(a, False)
    assert np.float32(lo).tobytes() == olo.tobytes() and np.float32(hi).tobytes() == ohi.tobytes() and np.signbit(np.float32(lo))
    assert ext.dem_minmax(a, True) == tuple(float(x) for x in oracle.min_max(a, True))
    ext.dem_minmax(a, True)
    t0 = time.perf_counter()
    for _ in range(5):
        ext.dem_minmax(a, True)
    per_call_ms = (time.perf_counter() - t0) / 5 * 1e3
    assert per_call_ms < 3.0, f"percentile of a host array took {per_call_ms:.2f} ms: the whole array is being uploaded again?"


def test_dem_percentile_ties_zeros_and_outliers(ext, oracle):
    rng = np.random.default_rng(0)
    a = np.round(rng.normal(size=200_000), 1).astype(np.float32)       # heavy ties
    a[rng.choice(a.size, 2000, replace=False)] *= 1e6
    assert ext.dem_minmax(a, True) == tuple(float(x) for x in oracle.min_max(a, True))
    z = np.zeros(5000, np.float32)
    z[::3] = -0.0                                                      # stable-sort sign of zero
    lo, hi = ext.dem_minmax(z, True)
    olo, ohi = oracle.min_max(z, True)
    assert np.signbit(np.float32(lo)) == np.signbit(olo) and np.signbit(np.float32(hi)) == np.signbit(ohi)
    c = np.full(70_000, -7.25, np.float32)
    assert ext.dem_minmax(c, True) == (-7.25, -7.25)


def test_dem_minmax_sign_of_zero_extremes(ext, oracle):
    """`if v < lo { lo = v }` keeps the first element equal to the extreme and -0 == +0, so a zero
    extreme carries the sign of the first zero of the slice (found by tools/fuzz_parity.py dem)."""
    for first, n in ((-0.0, 81), (0.0, 81), (-0.0, 1_538_639), (0.0, 3)):
        z = np.where(np.arange(n) % 2 == 0, np.float32(first), np.float32(-first)).astype(np.float32)
        lo, hi = ext.dem_minmax(z, False)
        olo, ohi = oracle.min_max(z, False)
        assert np.float32(lo).tobytes() == olo.tobytes() and np.float32(hi).tobytes() == ohi.tobytes(), (first, n)
    a = np.array([3.0, 0.0, -0.0, 5.0], np.float32)          # min is +0 (first zero), max is 5
    assert np.float32(ext.dem_minmax(a, False)[0]).tobytes() == np.float32(0.0).tobytes()
    b = np.array([-3.0, -0.0, 0.0, -5.0], np.float32)        # max is -0 (first zero)
    assert np.float32(ext.dem_minmax(b, False)[1]).tobytes() == np.float32(-0.0).tobytes()


def test_renderer_terrain_stats_and_range(vf, oracle):                 # tests/test_height_upload.py:20-49
    r = vf.Renderer(32, 32)
    with pytest.raises(RuntimeError, match=r"no terrain uploaded; call add_terrain\(\) first"):
        r.terrain_stats()
    h = np.linspace(0, 1, 16, dtype=np.float32).reshape(4, 4)
    r.add_terrain(h, (1.0, 1.0), 2.0, "viridis")
    mn, mx, mean, std = r.terrain_stats()
    assert mn == 0.0 and mx == 2.0 and mean == pytest.approx(1.0, abs=1e-6)
    assert std == pytest.approx(np.linspace(0, 2, 16).std(), rel=1e-3)
    assert r.height_range() == tuple(float(x) for x in oracle.h_range(h * np.float32(2.0)))
    r.set_height_range(-5.0, 40.0)
    for bad in ((1.0, 1.0), (2.0, -3.0)):
        with pytest.raises(ValueError, match="min must be < max"):
            r.set_height_range(*bad)
    with pytest.raises(ValueError, match="min/max must be finite floats"):
        r.set_height_range(float("nan"), 1.0)
    r.add_terrain(h.astype(np.float64), spacing=(1.0, 1.0), exaggeration=1.0, colormap="magma")   # f64 accepted
    assert r.terrain_stats()[1] == 1.0
    r.set_sun(45.0, 30.0)
    r.set_exposure(1.25)
    with pytest.raises(ValueError, match="exposure must be > 0"):
        r.set_exposure(0.0)                                            # tests/test_tonemap.py:26-31
    with pytest.raises(ValueError, match="angles must be finite"):
        r.set_sun(float("inf"), 0.0)


def test_add_terrain_error_strings(vf):                                # src/lib.rs:344-408
    r = vf.Renderer(16, 16)
    h = np.zeros((4, 4), np.float32)
    with pytest.raises(RuntimeError, match="spacing components must be > 0"):
        r.add_terrain(h, (0.0, 1.0), 1.0, "viridis")
    with pytest.raises(RuntimeError, match="exaggeration must be > 0"):
        r.add_terrain(h, (1.0, 1.0), 0.0, "viridis")
    with pytest.raises(RuntimeError, match="heightmap must be a 2-D NumPy array of dtype float32 or float64"):
        r.add_terrain(h.astype(np.int32), (1.0, 1.0), 1.0, "viridis")
    with pytest.raises(RuntimeError, match="heightmap must be a 2-D NumPy array of dtype float32 or float64"):
        r.add_terrain(np.zeros(4, np.float32), (1.0, 1.0), 1.0, "viridis")
    with pytest.raises(RuntimeError, match=r"heightmap must be C-contiguous \(row-major\)"):
        r.add_terrain(np.zeros((4, 6), np.float64).T, (1.0, 1.0), 1.0, "viridis")
    with pytest.raises(RuntimeError, match="heightmap cannot be empty"):
        r.add_terrain(np.zeros((0, 4), np.float32), (1.0, 1.0), 1.0, "viridis")
    with pytest.raises(RuntimeError, match="Unknown colormap 'jet'. Supported: viridis, magma, terrain"):
        r.add_terrain(h, (1.0, 1.0), 1.0, "jet")
    with pytest.raises(TypeError):
        r.add_terrain(h)                                               # all four arguments are required


def test_normalize_terrain(vf, oracle):                                # tests/test_dem_normalization.py:23-31
    rng = np.random.default_rng(2)
    z = rng.normal(3, 2, (64, 48)).astype(np.float32)
    r = vf.Renderer(16, 16)
    r.add_terrain(z, (1.0, 1.0), 1.0, "viridis")
    r.normalize_terrain("MinMax", range=(10.0, 20.0))
    mn, mx, _, _ = r.terrain_stats()
    assert mn == pytest.approx(10.0, abs=1e-4) and mx == pytest.approx(20.0, abs=1e-4)
    r.upload_height_r32f()
    # min/max are exact on the device, so minmax normalisation is bit-exact vs the oracle
    assert np.array_equal(r.read_full_height_texture(), oracle.normalize(z, "minmax", rng=(10.0, 20.0)))
    r.add_terrain(z, (1.0, 1.0), 1.0, "viridis")
    r.normalize_terrain("zscore")
    _, _, mean, std = r.terrain_stats()
    assert abs(mean) < 1e-4 and std == pytest.approx(1.0, abs=1e-3)
    with pytest.raises(RuntimeError, match="mode must be 'minmax' or 'zscore'"):
        r.normalize_terrain("robust")


@pytest.mark.parametrize("shape", [(7, 5), (64, 48), (255, 3), (33, 33), (61, 17), (1, 1)])
def test_height_upload_roundtrip(vf, shape):                           # tests/test_height_upload.py:52-160
    h = np.random.default_rng(shape[0]).normal(size=shape).astype(np.float32)
    r = vf.Renderer(8, 8)
    assert not r.debug_read_height_patch(0, 0, 3, 2).any()             # zeros before any upload
    r.add_terrain(h, (1.0, 1.0), 1.0, "viridis")
    with pytest.raises(RuntimeError, match=r"no height texture uploaded; call upload_height_r32f\(\) first"):
        r.read_full_height_texture()
    r.upload_height_r32f()
    full = r.read_full_height_texture()
    assert full.shape == shape and full.dtype == np.float32 and np.array_equal(full, h)
    ph, pw = max(1, shape[0] // 2), max(1, shape[1] // 2)
    y0, x0 = shape[0] - ph, shape[1] - pw
    assert np.array_equal(r.debug_read_height_patch(x0, y0, pw, ph), h[y0:, x0:])
    with pytest.raises(RuntimeError, match="requested patch exceeds texture bounds in x"):
        r.debug_read_height_patch(1, 0, shape[1], 1)
    with pytest.raises(RuntimeError, match="requested patch exceeds texture bounds in y"):
        r.debug_read_height_patch(0, 1, 1, shape[0])
    with pytest.raises(RuntimeError, match="patch dimensions must be > 0"):
        r.debug_read_height_patch(0, 0, 0, 1)


# ---- differential fuzzing (tools/fuzz_parity.py) ---------------------------------------------
def _fuzz():
    import importlib.util
    spec = importlib.util.spec_from_file_location("vf_fuzz_parity", os.path.join(ROOT, "tools", "fuzz_parity.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


# Found by the fuzzer: grids of 530+ vertices a side give k_geom_march tasks of several rows; with a
# camera close to (or inside) the terrain the first vertex row of a task can lie outside the guard
# band, and the two-rows-at-a-time fast path used to continue from it without the clip state.
_FUZZ_REGRESSIONS = [
    dict(W=1, H=17, G=836, dw=835, dh=836, seed=1022241656, eye=(-0.20915687370489333, 1.3276651954749488, -0.1938512144016744),
         tgt=(-0.20815687370489333, 0.0, -0.1938512144016744), fovy=90.0, znear=0.1, zfar=3.0, spacing=2.0, exag=2.0, mode=1, depth=0,
         cmap="magma", lo=-0.5888416895866394, hi=0.5424208173751831),
    dict(W=16, H=568, G=1063, dw=1062, dh=1063, seed=870961174, eye=(-0.2371148881262813, 0.7558845256859157, 0.14162109310250104),
         tgt=(-0.9095650163295042, 0.0, -0.14357839549558027), fovy=90.0, znear=0.01, zfar=10.0, spacing=1.0, exag=2.0, mode=1, depth=1,
         cmap="terrain", lo=-0.5365256199836731, hi=0.5347423086166382),
    dict(W=36, H=29, G=817, dw=817, dh=817, seed=518997034, eye=(-0.06159753803867862, -2.076090922529322, 0.2585386545596223),
         tgt=(-0.06059753803867862, 0.0, 0.2585386545596223), fovy=45.0, znear=0.5, zfar=3.0, spacing=1.0, exag=2.0, mode=0, depth=0,
         cmap="terrain", lo=-0.5542551407814026, hi=0.501332772731781),
    dict(W=355, H=33, G=670, dw=671, dh=670, seed=604203175, eye=(-0.20128530605246475, 1.628495349921721, -0.3871352855286505),
         tgt=(-0.20028530605246475, 0.0, -0.3871352855286505), fovy=20.0, znear=0.01, zfar=100.0, spacing=0.5, exag=1.0, mode=1, depth=1,
         cmap="viridis", lo=-0.5893525013923645, hi=0.5419024953842163),
    dict(W=272, H=193, G=1019, dw=1019, dh=1019, seed=757362591, eye=(-0.21935821393244004, 0.4826241812021064, -0.028272396321713744),
         tgt=(-0.08587378354637576, 0.0, 0.9006280450266455), fovy=90.0, znear=0.1, zfar=10.0, spacing=1.0, exag=2.0, mode=1, depth=0,
         cmap="viridis", lo=-0.536085141658783, hi=0.5352432856559753),
    dict(W=14, H=264, G=946, dw=88, dh=278, seed=931067945, eye=(0.265642581102838, 0.5883513195032288, 0.37803053320399993),
         tgt=(0.3961539000131138, -0.3968453150433958, -0.6292692226925838), fovy=120.0, znear=0.01, zfar=3.0, spacing=0.5, exag=1.0, mode=0,
         depth=0, cmap="viridis", lo=-0.5523202548027039, hi=0.5304822454452515),
]


@pytest.mark.gpu
@pytest.mark.parametrize("case", _FUZZ_REGRESSIONS, ids=lambda c: f"G{c['G']}_{c['W']}x{c['H']}_m{c['mode']}d{c['depth']}")
def test_fuzz_regressions(case):
    fz = _fuzz()
    ok, detail, nothing = fz.check_case(case, fz.dem_for(case))
    assert not nothing, "regression camera sees nothing"
    assert ok, detail


@pytest.mark.gpu
def test_fuzz_short_run():
    """20 s of seeded random frames (sizes, grids up to 2600, DEM shapes, cameras inside / grazing /
    looking away, camera / sheared / orthographic matrices, both modes, depth on and off)."""
    fz = _fuzz()
    rng = np.random.default_rng(20240607)
    import time
    t0, n, bad = time.time(), 0, []
    while time.time() - t0 < 20.0:
        c, dem = fz.random_case(rng)
        ok, detail, _ = fz.check_case(c, dem)
        n += 1
        if not ok:
            bad.append((detail, c))
    assert n >= 50, f"only {n} frames in 20 s"
    assert not bad, f"{len(bad)} of {n} frames differ; first: {bad[0]}"
