"""Differential fuzzing of the terrain frame: CUDA path (through the C ABI) against the CPU oracle on
random frame sizes, grids, DEM shapes, cameras (far, near, inside the terrain, grazing), modes and
depth settings.  NORTHSTAR frames must be byte-identical; SHIPPED frames (sin/cos in the vertex and
fragment stage) within 1 LSB on >= 99.9 % of the pixels.  Test infrastructure: run on the GPU box,
  python tools/fuzz_parity.py [seconds] [seed]
prints one line per mismatch (with everything needed to replay it) and a summary; exit code 1 on
any mismatch."""
import ctypes as C
import json
import math
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vulkan-forge_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import vf_oracle as o
import vulkan_forge as vf
from vulkan_forge._synthetic import synthetic_dem

lib = C.CDLL(os.path.join(ROOT, "vulkan-forge_b200", "python", "vulkan_forge", "libvfcuda.so"))
lib.vf_last_error.restype = C.c_char_p
LUTS = {n: np.fromfile(os.path.join(ROOT, "tests", "golden", f"lut_{n}.rgba"), np.uint8) for n in ("viridis", "magma", "terrain")}
CMAP_ID = {"viridis": 0, "magma": 1, "terrain": 2}


def ptr(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def gpu_frame(W, H, G, cmap, dem, u, mode, depth, lo, hi):
    scene = C.c_void_p()
    assert lib.vf_scene_create(W, H, G, CMAP_ID[cmap], 0, C.byref(scene)) == 0, lib.vf_last_error()
    try:
        if dem is not None:
            assert lib.vf_scene_set_height_f32(scene, ptr(dem, C.c_float), dem.shape[1], dem.shape[0]) == 0, lib.vf_last_error()
        assert lib.vf_scene_set_uniforms(scene, ptr(u, C.c_float)) == 0
        assert lib.vf_scene_set_mode(scene, mode, depth, C.c_float(lo), C.c_float(hi)) == 0
        out = np.zeros((H, W, 4), np.uint8)
        assert lib.vf_scene_render_rgba(scene, ptr(out, C.c_uint8)) == 0, lib.vf_last_error()
        out2 = np.zeros((H, W, 4), np.uint8)       # second frame: the self-cleaning visibility buffer must be clean
        assert lib.vf_scene_render_rgba(scene, ptr(out2, C.c_uint8)) == 0, lib.vf_last_error()
        assert np.array_equal(out, out2), "second frame of the same scene differs from the first"
        return out
    finally:
        lib.vf_scene_destroy(scene)


def dem_for(c, rng=None):
    """The DEM of a recorded case: everything but the optional spikes / flattening is a function of the case."""
    base = synthetic_dem(c["seed"] % 7, max(c["dw"], c["dh"]))[:c["dh"], :c["dw"]]
    return np.ascontiguousarray(base * np.float32(c.get("scale", 1.0)))


def check_case(c, dem, threads=0):
    u = np.ascontiguousarray(uniforms_for(c), np.float32)
    if c.get("matrix") == "sheared":
        u[16 + 4] = np.float32(0.15); u[16 + 1] = np.float32(-0.05)
    elif c.get("matrix") == "ortho":
        P = np.zeros(16, np.float32)
        P[0], P[5], P[10], P[14], P[15] = 0.45, 0.6, -1.0 / 20.0, 0.0, 1.0
        u[16:32] = P
    ref = o.render_terrain(c["W"], c["H"], c["G"], dem, u, LUTS[c["cmap"]], mode=o.MODE_NORTHSTAR if c["mode"] else o.MODE_SHIPPED,
                           depth=bool(c["depth"]), h_min=c["lo"], h_max=c["hi"], threads=threads)
    got = gpu_frame(c["W"], c["H"], c["G"], c["cmap"], dem, u, c["mode"], c["depth"], c["lo"], c["hi"])
    empty = not (ref[..., :3] != ref[0, 0, :3]).any()
    if c["mode"] == 1:
        return np.array_equal(got, ref), f"{int((got != ref).any(-1).sum())} pixels differ", empty
    d = np.abs(got.astype(np.int16) - ref.astype(np.int16)).max(-1)
    ok = (d <= 1).mean() >= 0.999 if d.size >= 1000 else (d <= 1).sum() >= d.size - 1
    return bool(ok), f"within 1 LSB {float((d <= 1).mean()):.5f}, max {int(d.max())}", empty


def random_case(rng):
    W = int(rng.choice([rng.integers(1, 40), rng.integers(40, 400), rng.integers(400, 900)]))
    H = int(rng.choice([rng.integers(1, 40), rng.integers(40, 300), rng.integers(300, 700)]))
    G = int(rng.choice([rng.integers(2, 8), rng.integers(8, 70), rng.integers(70, 400), rng.integers(400, 1100), rng.integers(1100, 2600)]))
    kind = rng.integers(0, 4)
    if kind == 0:
        dw, dh = G, G                                   # identity mapping: TMA path when the pitch allows
    elif kind == 1:
        dw, dh = int(rng.integers(1, 300)), int(rng.integers(1, 300))
    elif kind == 2:
        dw, dh = G + int(rng.integers(-1, 2)), G
        dw = max(dw, 1)
    else:
        dw, dh = int(rng.integers(2, 1200)), int(rng.integers(2, 1200))
    seed = int(rng.integers(0, 1 << 30))
    scale = float(rng.choice([0.2, 1.0, 3.0]))
    dem = dem_for(dict(seed=seed, dw=dw, dh=dh, scale=scale))
    if rng.random() < 0.3:                               # spikes: needles, huge triangles
        k = int(rng.integers(1, 12))
        dem[rng.integers(0, dh, k), rng.integers(0, dw, k)] += rng.normal(0, 3.0, k).astype(np.float32)
    if rng.random() < 0.1:
        dem[:] = np.float32(rng.normal())                # flat
    style = rng.integers(0, 6)
    if style == 0:      # orbit-like
        a = rng.uniform(0, 2 * math.pi); r = rng.uniform(1.5, 8.0)
        eye = (r * math.cos(a), rng.uniform(0.3, 4.0), r * math.sin(a)); tgt = (0.0, 0.0, 0.0)
    elif style == 1:    # inside / just above the terrain: near-plane clipping
        eye = (rng.uniform(-1, 1), rng.uniform(-0.2, 0.6), rng.uniform(-1, 1)); tgt = (rng.uniform(-2, 2), rng.uniform(-0.5, 0.5), rng.uniform(-2, 2))
    elif style == 2:    # straight down / up
        eye = (rng.uniform(-0.5, 0.5), rng.uniform(1.0, 6.0) * rng.choice([-1, 1]), rng.uniform(-0.5, 0.5)); tgt = (eye[0] + 1e-3, 0.0, eye[2])
    elif style == 3:    # grazing
        eye = (rng.uniform(2, 6), rng.uniform(0.0, 0.15), rng.uniform(-1, 1)); tgt = (0.0, 0.05, 0.0)
    elif style == 4:    # looking away: (almost) nothing visible
        eye = (3.0, 2.0, 3.0); tgt = (6.0, 3.0, 6.0 + rng.uniform(-1, 1))
    else:               # close-up: big triangles (binned path)
        eye = (rng.uniform(-0.3, 0.3), rng.uniform(0.2, 0.8), rng.uniform(-0.3, 0.3)); tgt = (eye[0] + rng.uniform(-1, 1), 0.0, eye[2] + rng.uniform(-1, 1))
    fovy = float(rng.choice([20.0, 45.0, 90.0, 120.0]))
    znear = float(rng.choice([0.01, 0.1, 0.5])); zfar = float(rng.choice([10.0, 100.0, 3.0]))
    spacing = float(rng.choice([1.0, 1.0, 0.5, 2.0])); exag = float(rng.choice([1.0, 1.0, 2.0, 0.25]))
    mode = int(rng.integers(0, 2)); depth = int(rng.integers(0, 2))
    cmap = str(rng.choice(["viridis", "magma", "terrain"]))
    matrix = str(rng.choice(["camera"] * 6 + ["sheared", "ortho"]))
    lo = float(np.percentile(dem, 1)) - 1e-3; hi = float(np.percentile(dem, 99)) + 1e-3
    return dict(W=W, H=H, G=G, dw=dw, dh=dh, seed=seed, eye=tuple(map(float, eye)), tgt=tuple(map(float, tgt)), fovy=fovy,
                znear=znear, zfar=zfar, spacing=spacing, exag=exag, mode=mode, depth=depth, cmap=cmap, lo=lo, hi=hi,
                style=int(style), scale=scale, matrix=matrix), dem


def uniforms_for(c):
    kw = {}
    try:
        return o.look_at_uniforms(c["W"], c["H"], c["eye"], c["tgt"], (0, 1, 0), c["fovy"], c["znear"], c["zfar"],
                                  spacing=c["spacing"], exaggeration=c["exag"], **kw)
    except TypeError:
        return o.look_at_uniforms(c["W"], c["H"], c["eye"], c["tgt"], (0, 1, 0), c["fovy"], c["znear"], c["zfar"])


def random_dem_values(rng):
    """Arrays that stress the exact percentile select: ties, signed zeros, denormals, infinities,
    huge dynamic range, constant data, sizes around the radix-select pass boundaries."""
    n = int(rng.choice([rng.integers(1, 5), rng.integers(5, 300), rng.integers(300, 70000), rng.integers(70000, 3000000)]))
    kind = int(rng.integers(0, 8))
    if kind == 0:
        a = rng.normal(0, 1, n)
    elif kind == 1:
        a = rng.integers(-3, 4, n).astype(np.float64)                    # heavy ties
    elif kind == 2:
        a = np.where(rng.random(n) < 0.5, 0.0, -0.0)                     # signed zeros only
    elif kind == 3:
        a = rng.normal(0, 1, n) * 10.0 ** rng.integers(-40, 38, n)       # denormals .. near overflow
    elif kind == 4:
        a = np.full(n, rng.normal())
    elif kind == 5:
        a = rng.normal(0, 1, n); a[rng.integers(0, n, max(1, n // 50))] = rng.choice([np.inf, -np.inf, 1e30, -1e30])
    elif kind == 6:
        a = np.sort(rng.normal(0, 1, n))[:: int(rng.choice([1, -1]))]
    else:
        a = rng.integers(0, 2, n) * 1e-45 + rng.integers(0, 2, n) * -1e-45   # +-smallest denormal and zeros
    with np.errstate(over="ignore"):
        return np.ascontiguousarray(a, np.float32), kind


def png_decode(data, w, h):
    """chunk walk + zlib (checks the Adler-32) + Sub / None filters -> (h, w, 4) uint8"""
    import struct, zlib
    pos, idat = 8, []
    while pos < len(data):
        ln, typ = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + ln]
        assert zlib.crc32(typ + body) == struct.unpack(">I", data[pos + 8 + ln:pos + 12 + ln])[0]
        if typ == b"IDAT": idat.append(body)
        pos += 12 + ln
    raw = np.frombuffer(zlib.decompress(b"".join(idat)), np.uint8).reshape(h, 4 * w + 1)
    px = raw[:, 1:].reshape(h, w, 4).astype(np.uint32)
    sub = raw[:, 0] == 1
    px[sub] = np.cumsum(px[sub], axis=1) & 255
    return px.astype(np.uint8)


def fuzz_dem(budget, seed):
    """dem_minmax (min/max and the 1-99 % percentile pair of src/terrain_stats.rs:11-35) bit-for-bit."""
    ext = vf._vulkan_forge
    rng = np.random.default_rng(seed)
    t0 = time.time(); n = 0; bad = 0
    while time.time() - t0 < budget:
        a, kind = random_dem_values(rng)
        shape = (1, a.size) if rng.random() < 0.5 or a.size < 4 else (2, a.size // 2)
        a = a[:shape[0] * shape[1]].reshape(shape)
        for clamp in (False, True):
            got = ext.dem_minmax(a, clamp)
            ref = o.min_max(a, clamp)
            same = all(np.float32(g).tobytes() == np.float32(r).tobytes() for g, r in zip(got, ref))
            n += 1
            if not same:
                bad += 1
                print("DEM MISMATCH", dict(kind=kind, n=int(a.size), clamp=clamp, got=[float(x) for x in got], ref=[float(x) for x in ref], seed=seed), flush=True)
    print(f"fuzz dem: {n} calls in {time.time() - t0:.0f} s, {bad} mismatches, seed {seed}")
    return bad


def fuzz_api(budget, seed):
    """Python API: grid_generate and the triangle against the oracle; pipelined batches (async upload +
    async batch + collect, close-up cameras that overflow the primitive / pair buffers and make them
    grow) against the synchronous single-frame path of a second Scene."""
    ext = vf._vulkan_forge
    rng = np.random.default_rng(seed)
    t0 = time.time(); n = 0; bad = 0
    while time.time() - t0 < budget:
        what = int(rng.integers(0, 4))
        if what == 0:
            nx, nz = int(rng.integers(2, 400)), int(rng.integers(2, 400))
            sp = (float(rng.choice([1.0, 0.5, 2.0, 0.1, 3.7])), float(rng.choice([1.0, 0.25, 2.0, 1.3])))
            got = vf.grid_generate(nx, nz, spacing=sp)
            ref = o.make_grid(nx, nz, sp[0], sp[1])
            ok = all(np.array_equal(a, b) for a, b in zip(got, ref))
            desc = dict(what="grid", nx=nx, nz=nz, sp=sp)
        elif what == 1:
            w, h = int(rng.integers(1, 900)), int(rng.integers(1, 700))
            ok = np.array_equal(vf.render_triangle_rgba(w, h), o.render_triangle(w, h))
            desc = dict(what="triangle", w=w, h=h)
        else:
            W, H = int(rng.integers(8, 500)), int(rng.integers(8, 400))
            G = int(rng.choice([rng.integers(16, 300), rng.integers(300, 1600)]))
            depth = bool(rng.integers(0, 2)); mode = str(rng.choice(["northstar", "shipped"]))
            dem_par = [(int(rng.integers(0, 7)), float(rng.choice([1.0, 3.0]))) for _ in range(2)]
            Gd = G if rng.random() < 0.7 else int(rng.integers(8, 500))     # DEM size != grid: the vertex-height image is resampled per upload
            dems = [synthetic_dem(sd, Gd) * np.float32(sc_) for sd, sc_ in dem_par]
            cams = []
            for _ in range(int(rng.integers(1, 5))):
                if rng.random() < 0.6:      # close-up: most triangles are big -> queues and pair buffers grow
                    e = (float(rng.uniform(-0.3, 0.3)), float(rng.uniform(0.05, 0.6)), float(rng.uniform(-0.3, 0.3)))
                    t = (e[0] + float(rng.uniform(-1, 1)), float(rng.uniform(-0.3, 0.1)), e[2] + float(rng.uniform(-1, 1)))
                else:
                    a = float(rng.uniform(0, 6.28)); e = (4 * math.cos(a), float(rng.uniform(0.5, 3)), 4 * math.sin(a)); t = (0.0, 0.0, 0.0)
                cams.append((e, t, (0.0, 1.0, 0.0), float(rng.choice([45.0, 90.0])), float(rng.choice([0.01, 0.1])), 100.0))
            desc = dict(what="pipeline", W=W, H=H, G=G, Gd=Gd, depth=depth, mode=mode, dems=dem_par, cams=cams, failed=[])
            def scene():
                sc = ext.Scene(W, H, grid=G, colormap="viridis")
                sc.set_shading(mode, depth=depth); sc.set_height_range(-0.4, 0.5)
                return sc
            ref_sc, sc = scene(), scene()
            want = []
            for d in dems:
                ref_sc.set_height_from_r32f(d)
                fr = []
                for c in cams:
                    ref_sc.set_camera_look_at(*c)
                    fr.append(ref_sc.render_rgba().copy())
                want.append(np.stack(fr))
            ok = True
            pending = []
            for k in range(4):
                sc.set_height_from_r32f_async(dems[k & 1])
                pending.append((k & 1, sc.render_batch_rgba_async(cams)))
                if len(pending) == 3:
                    which, tk = pending.pop(0)
                    got = sc.collect(tk)
                    if not np.array_equal(got, want[which]):
                        ok = False; desc["failed"].append(("collect", k - 2, [int((got[i] != want[which][i]).any(-1).sum()) for i in range(len(cams))], dict(sc.render_metrics())["big_prims"]))
            for j, (which, tk) in enumerate(pending):
                got = sc.collect(tk)
                if not np.array_equal(got, want[which]):
                    ok = False; desc["failed"].append(("collect-tail", j, [int((got[i] != want[which][i]).any(-1).sum()) for i in range(len(cams))]))
            got = sc.render_batch_rgba(cams)
            if not np.array_equal(got, want[1]):
                ok = False; desc["failed"].append(("sync-batch", [int((got[i] != want[1][i]).any(-1).sum()) for i in range(len(cams))]))
            # PNG batches with the next image in flight (same-size images only), then frames left on the device
            # (runs of frames without a memset node, the geometry pass launched under the previous shading tail)
            if W <= 8192:
                for k in range(3):
                    buf, sizes = sc.render_batch_png_raw(cams, next_height=dems[k & 1] if k < 2 else None)
                    which = (k + 1) & 1                     # image drawn by call k: dems[1], dems[0], dems[1]
                    for i in range(len(cams)):
                        if not np.array_equal(png_decode(bytes(buf[i, :sizes[i]]), W, H), want[which][i]):
                            ok = False; desc["failed"].append(("png-next", k, i)); break
            sc.render_batch_device(cams * 3)
            last = sc.read_device_frame((3 * len(cams) - 1) % 2)
            if not np.array_equal(last, want[1][-1]):
                ok = False; desc["failed"].append(("device-batch", int((last != want[1][-1]).any(-1).sum())))
            if not ok:                                      # is the reference scene itself stable?
                ref_sc.set_height_from_r32f(dems[1])
                again = []
                for c in cams:
                    ref_sc.set_camera_look_at(*c)
                    again.append(ref_sc.render_rgba().copy())
                desc["ref_unstable"] = [int((again[i] != want[1][i]).any(-1).sum()) for i in range(len(cams))]
            del sc, ref_sc
        n += 1
        if not ok:
            bad += 1
            print("API MISMATCH", json.dumps(desc), flush=True)
    print(f"fuzz api: {n} cases in {time.time() - t0:.0f} s, {bad} mismatches, seed {seed}")
    return bad


def fuzz_renderer(budget, seed):
    """Renderer height path: add_terrain (f32 / f64, exaggeration) -> terrain_stats / height_range ->
    normalize_terrain(minmax) -> upload_height_r32f -> read back, against the oracle (min, max, range and
    the min-max normalised image bit-for-bit; mean / std to the reference tests' 1e-3)."""
    rng = np.random.default_rng(seed)
    t0 = time.time(); n = 0; bad = 0
    r = vf.Renderer(8, 8)
    while time.time() - t0 < budget:
        hh, ww = int(rng.choice([rng.integers(1, 6), rng.integers(6, 400)])), int(rng.choice([rng.integers(1, 6), rng.integers(6, 400)]))
        dt = np.float64 if rng.random() < 0.4 else np.float32
        kind = int(rng.integers(0, 4))
        if kind == 0:
            z = rng.normal(rng.normal(0, 5), 10.0 ** rng.uniform(-3, 3), (hh, ww))
        elif kind == 1:
            z = np.round(rng.normal(0, 2, (hh, ww)))
        elif kind == 2:
            z = np.full((hh, ww), rng.normal())
        else:
            z = rng.normal(0, 1, (hh, ww)); z.flat[rng.integers(0, z.size)] = 1e4
        z = np.ascontiguousarray(z, dt)
        ex = float(rng.choice([1.0, 2.0, 0.5, 3.3]))
        r.add_terrain(z, (1.0, 1.0), ex, "viridis")
        h32 = (z.astype(np.float32) * np.float32(ex)).astype(np.float32)      # src/lib.rs:362,384
        got = r.terrain_stats(); ref = o.dem_stats(h32)
        scale = max(1e-30, float(np.abs(h32).max()))
        ok = np.float32(got[0]).tobytes() == ref[0].tobytes() and np.float32(got[1]).tobytes() == ref[1].tobytes()
        # the reference (and the oracle) accumulate sequentially in f32: their own error grows with n
        tol = (1e-3 + 6e-8 * h32.size) * scale
        ok = ok and abs(got[2] - float(ref[2])) <= tol and abs(got[3] - float(ref[3])) <= tol + 1e-3 * float(ref[3])
        ok = ok and r.height_range() == tuple(float(x) for x in o.h_range(h32))
        desc = dict(what="stats", shape=(hh, ww), dtype=str(np.dtype(dt)), kind=kind, ex=ex, got=[float(x) for x in got], ref=[float(x) for x in ref])
        if ok:
            lo = float(rng.choice([0.0, -1.0, 10.0])); hi = lo + float(rng.choice([1.0, 10.0, 0.001]))
            r.normalize_terrain("minmax", range=(lo, hi))
            r.upload_height_r32f()
            full = r.read_full_height_texture()
            want = o.normalize(h32, "minmax", rng=(lo, hi))
            ok = np.array_equal(full, want)
            desc = dict(what="normalize", shape=(hh, ww), dtype=str(np.dtype(dt)), kind=kind, ex=ex, range=(lo, hi), diff=int((full != want).sum()))
            if ok and hh > 1 and ww > 1:
                x0, y0 = int(rng.integers(0, ww - 1)), int(rng.integers(0, hh - 1))
                pw, ph = int(rng.integers(1, ww - x0 + 1)), int(rng.integers(1, hh - y0 + 1))
                ok = np.array_equal(r.debug_read_height_patch(x0, y0, pw, ph), want[y0:y0 + ph, x0:x0 + pw])
                desc["what"] = "patch"
        n += 1
        if not ok:
            bad += 1
            print("RENDERER MISMATCH", json.dumps(desc), flush=True)
    print(f"fuzz renderer: {n} cases in {time.time() - t0:.0f} s, {bad} mismatches, seed {seed}")
    return bad


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "renderer":
        sys.exit(1 if fuzz_renderer(float(sys.argv[2]) if len(sys.argv) > 2 else 30.0, int(sys.argv[3]) if len(sys.argv) > 3 else 1) else 0)
    if len(sys.argv) > 1 and sys.argv[1] == "api":
        sys.exit(1 if fuzz_api(float(sys.argv[2]) if len(sys.argv) > 2 else 30.0, int(sys.argv[3]) if len(sys.argv) > 3 else 1) else 0)
    if len(sys.argv) > 1 and sys.argv[1] == "dem":
        sys.exit(1 if fuzz_dem(float(sys.argv[2]) if len(sys.argv) > 2 else 30.0, int(sys.argv[3]) if len(sys.argv) > 3 else 1) else 0)
    budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    rng = np.random.default_rng(seed)
    threads = len(os.sched_getaffinity(0))
    t0 = time.time(); n = 0; bad = 0; empty = 0; stats = {}
    while time.time() - t0 < budget:
        c, dem = random_case(rng)
        try:
            ok, detail, nothing = check_case(c, dem, threads)
        except AssertionError as e:
            bad += 1
            print("ERROR", e, json.dumps(c), flush=True)
            continue
        n += 1
        empty += int(nothing)
        key = (c["style"], c["mode"], c["depth"])
        stats[key] = stats.get(key, 0) + 1
        if not ok:
            bad += 1
            print("MISMATCH", detail, json.dumps(c), flush=True)
    print(f"fuzz: {n} frames in {time.time() - t0:.0f} s, {bad} mismatches, {empty} frames saw nothing, seed {seed}")
    print("cases per (camera style, mode, depth):", sorted(stats.items()))
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
