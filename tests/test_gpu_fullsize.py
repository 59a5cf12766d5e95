"""BASELINE.json's full-size configurations: CUDA vs oracle frame comparison where the oracle
finishes in seconds (c3), plus size-independent properties (determinism, shard equality,
clear-colour invariants, percentile of the full DEM)."""
import hashlib

import numpy as np
import pytest

from conftest import frame_diff

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dem4096(synthetic_dem):
    return synthetic_dem(0, 4096)


def test_c3_1080p_dem4096_both_modes(ext, oracle, luts, dem4096):
    """config 3: Scene 1920x1080, 4096^2 DEM, G=4096, terrain LUT; SHIPPED and NORTHSTAR."""
    sc = ext.Scene(1920, 1080, grid=4096, colormap="terrain")
    sc.set_height_from_r32f(dem4096)
    got = sc.render_rgba()
    ref = oracle.render_terrain(1920, 1080, 4096, dem4096, sc.debug_uniforms_f32(), luts["terrain"])
    frac, mx = frame_diff(got, ref)
    assert frac >= 0.999, (frac, mx)
    sc.set_shading("northstar", depth=False)
    sc.set_sun(35.0, 120.0)
    sc.set_exposure(1.1)
    lo, hi = sc.set_auto_height_range()
    assert (np.float32(lo), np.float32(hi)) == oracle.h_range(dem4096)
    got = sc.render_rgba()
    ref = oracle.render_terrain(1920, 1080, 4096, dem4096, sc.debug_uniforms_f32(), luts["terrain"],
                                mode=oracle.MODE_NORTHSTAR, h_min=float(lo), h_max=float(hi))
    assert np.array_equal(got, ref)
    assert hashlib.sha256(got.tobytes()).digest() == hashlib.sha256(sc.render_rgba().tobytes()).digest()
    assert tuple(got[0, 0]) == (39, 39, 48, 255)


def test_c4_orbit_shards_equal_whole(ext, vf, dem4096):
    """config 4 (subsampled to 8 frames): rendering the orbit in 1, 2 or 4 shards gives the same bytes."""
    sc = ext.Scene(1024, 1024, grid=4096, colormap="terrain")
    sc.set_height_from_r32f(dem4096)
    sc.set_shading("northstar", depth=True)
    sc.set_auto_height_range()
    _, whole = vf.render_orbit(sc, 8, rank=0, world=1)
    for world in (2, 4):
        parts = [vf.render_orbit(sc, 8, rank=r, world=world) for r in range(world)]
        assert np.array_equal(np.concatenate([p[1] for p in parts]), whole)
        assert [p[0] for p in parts] == [vf.shard_range(8, r, world)[0] for r in range(world)]
    assert len({hashlib.sha256(f.tobytes()).hexdigest() for f in whole}) == 8    # every camera differs


@pytest.mark.parametrize("depth", [False, True], ids=["no depth", "depth"])
def test_c4_orbit_vs_oracle(ext, vf, oracle, luts, dem4096, depth):
    """config 4: frames k = 0, 341, 683 of the 1024-frame orbit at 1024 x 1024 over the 4096^2 DEM (G = 4096),
    NORTHSTAR, byte-compared with the oracle (the config the multi-GPU split is defined on)."""
    sc = ext.Scene(1024, 1024, grid=4096, colormap="terrain")
    sc.set_height_from_r32f(dem4096)
    sc.set_shading("northstar", depth=depth)
    sc.set_sun(35.0, 120.0)
    sc.set_exposure(1.1)
    lo, hi = sc.set_auto_height_range()
    cams = vf.orbit_cameras(1024)
    ks = (0, 341, 683)
    blocks = sc.camera_blocks([cams[k] for k in ks])
    got = sc.render_batch_rgba(blocks)
    for n, k in enumerate(ks):
        ref = oracle.render_terrain(1024, 1024, 4096, dem4096, blocks[n], luts["terrain"], mode=oracle.MODE_NORTHSTAR, depth=depth,
                                    h_min=float(lo), h_max=float(hi))
        assert (ref[..., :3] != ref[0, 0, :3]).any(-1).mean() > 0.05, "camera sees nothing"
        assert np.array_equal(got[n], ref), f"frame {k}: {(got[n] != ref).any(-1).sum()} pixels differ"


def test_c5_4k_dem16384_depth(ext, oracle, luts, synthetic_dem):
    """config 5: 3840x2160, 16384^2 DEM, depth, magma, exaggeration 2, auto 1-99 %.  The oracle needs a
    few seconds per frame on many cores; compared in full."""
    dem = synthetic_dem(2, 16384)
    sc = ext.Scene(3840, 2160, grid=16384, colormap="magma")
    sc.set_height_from_r32f(dem)
    sc.set_shading("northstar", depth=True)
    sc.set_spacing_exaggeration(1.0, 2.0)
    lo, hi = sc.set_auto_height_range()
    assert (np.float32(lo), np.float32(hi)) == oracle.h_range(dem)
    got = sc.render_rgba()
    ref = oracle.render_terrain(3840, 2160, 16384, dem, sc.debug_uniforms_f32(), luts["magma"], mode=oracle.MODE_NORTHSTAR,
                                depth=True, h_min=float(lo), h_max=float(hi))
    assert np.array_equal(got, ref)
