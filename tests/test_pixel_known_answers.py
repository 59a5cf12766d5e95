"""Pixel known answers written out in this file, in numpy float64 / Python integers, WITHOUT calling any code
of oracle/ or of the product: scenes simple enough that coverage and colour follow by hand from the
reference's shader text (src/shaders/terrain.wgsl:44-91, triangle.wgsl:1-23), its pipeline state
(src/terrain/pipeline.rs:107-140: CCW front, cull back, no blend) and the D3D11 / Vulkan rasterisation rules
(pixel centres at +0.5, top-left fill rule, 8 sub-pixel bits).

Each answer is checked twice: against the C oracle (CPU suite) and against the CUDA path (GPU suite), so the
oracle and the kernels are pinned by the same independent derivation rather than by each other.  The
reference itself holds no pixel vector ("pixel parity unpinned"), and its wgpu renderer cannot run on
these machines; this file is the strongest pixel pin available without a graphics stack.
"""
import math
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ---- the two implementations under test, behind one call ---------------------------------------------------
def _render_oracle(oracle, luts, W, H, G, dem, u, cmap, depth, hmin, hmax):
    img, prim = oracle.render_terrain(W, H, G, dem, u, luts[cmap], mode=oracle.MODE_NORTHSTAR, depth=depth,
                                      h_min=hmin, h_max=hmax, want_prim=True)
    return img, prim


def _render_cuda(ext, W, H, G, dem, u, cmap, depth, hmin, hmax):
    sc = ext.Scene(W, H, grid=G, colormap=cmap)
    sc.set_height_from_r32f(np.ascontiguousarray(dem, np.float32))
    sc.set_shading("northstar", depth=depth)
    sc.set_height_range(hmin, hmax)
    sc.set_uniforms_raw(u)                      # last: the setters above rebuild the block from the camera
    return sc.render_rgba(), sc.debug_visibility()


IMPLS = [pytest.param("oracle", id="oracle"), pytest.param("cuda", marks=pytest.mark.gpu, id="cuda")]


@pytest.fixture
def render(request, luts):
    def go(impl, *a):
        if impl == "oracle":
            return _render_oracle(request.getfixturevalue("oracle"), luts, *a)
        return _render_cuda(request.getfixturevalue("ext"), *a)
    return go


# ---- hand-built uniform block: top-down orthographic camera ------------------------------------------------
def topdown_ortho_uniforms(sx, sy, sun=(0.3, 0.8, 0.5), exposure=1.0, spacing=1.0, hmin=-0.5, hmax=0.5, exag=1.0):
    """44 floats (src/terrain/mod.rs:114-142).  Eye (0,5,0) looking down -Y with up = -Z, written out by hand:
    s = f x up = (1,0,0), u = s x f = (0,0,-1), so x_view = x, y_view = -z, z_view = y - 5; orthographic
    projection x_ndc = sx x_view, y_ndc = sy y_view, z_ndc = 0.5 - 0.01 z_view, w = 1."""
    V = np.array([[1, 0, 0, 0], [0, 0, -1, 0], [0, 1, 0, -5], [0, 0, 0, 1]], np.float64)        # M[row, col]
    P = np.array([[sx, 0, 0, 0], [0, sy, 0, 0], [0, 0, -0.01, 0.5], [0, 0, 0, 1]], np.float64)
    u = np.zeros(44, np.float32)
    u[0:16] = V.T.ravel()                       # column-major
    u[16:32] = P.T.ravel()
    u[32:35] = sun
    u[35], u[36], u[37], u[38] = exposure, spacing, hmax - hmin, exag
    return u


def screen_rect(sx, sy, W, H):
    """The plane [-1.5, 1.5]^2 under the camera above: world x -> screen X, world z -> screen Y (y_view = -z and
    the framebuffer's y axis points down, so the two flips cancel)."""
    X = [(sx * x * 0.5 + 0.5) * W for x in (-1.5, 1.5)]
    Y = [(sy * z * 0.5 + 0.5) * H for z in (-1.5, 1.5)]
    return X[0], X[1], Y[0], Y[1]


# ---- KAT 1: one cell, two triangles, coverage by the top-left rule ------------------------------------------
@pytest.mark.parametrize("impl", IMPLS)
@pytest.mark.parametrize("edge", [8.0, 8.5], ids=["edges on pixel boundaries", "edges through pixel centres"])
def test_one_cell_two_triangles_top_left_rule(render, impl, edge):
    """G = 2: vertices a (top-left), b (top-right), c (bottom-left), d (bottom-right) of a screen-aligned square;
    T0 = (a, c, b) is primitive 0, T1 = (b, c, d) primitive 1 (src/terrain/mod.rs:575-583).  The shared diagonal
    c-b runs through the centres of the pixels with px + py = 63.  For T1 it is the LEFT edge (the triangle lies to
    its right), so T1 owns those centres; T0 must not draw them.  With the square's outline on pixel centres
    (edge = 8.5 .. 55.5) the top and left outline pixels are inside, the bottom and right ones are not."""
    W = H = 64
    half = (32.0 - edge) / 32.0                 # |x_ndc| of the outline
    sx = half / 1.5
    x0, x1, y0, y1 = screen_rect(sx, sx, W, H)
    assert abs(x0 - edge) < 1e-9 and abs(x1 - (64 - edge)) < 1e-9 and abs(y0 - edge) < 1e-9
    _, prim = render(impl, W, H, 2, np.zeros((1, 1), np.float32), topdown_ortho_uniforms(sx, sx), "viridis", False, -0.5, 0.5)
    want = np.zeros((H, W), np.uint32)
    for py in range(H):
        for px in range(W):
            cx, cy = px + 0.5, py + 0.5
            inside = (x0 <= cx < x1) and (y0 <= cy < y1)          # left / top inclusive, right / bottom exclusive
            if inside:
                want[py, px] = 1 if cx + cy < 64.0 else 2          # on the diagonal (== 64): T1 (prim 1, stored + 1)
    assert np.array_equal(prim, want), f"{int((prim != want).sum())} pixels differ"
    n = int(56 - math.ceil(edge - 0.5)) if edge == 8.0 else 47
    assert (want > 0).sum() == n * n


# ---- KAT 2: colour of a height ramp under the orthographic camera -------------------------------------------
def srgb8_to_linear(v):
    s = v / 255.0
    return np.where(s <= 0.04045, s / 12.92, ((s + 0.055) / 1.055) ** 2.4)


def linear_to_srgb8(c):
    c = np.clip(c, 0.0, 1.0)
    e = np.where(c <= 0.0031308, 12.92 * c, 1.055 * c ** (1 / 2.4) - 0.055)
    return np.floor(e * 255.0 + 0.5).astype(np.int64)


def lut_sample(lut_lin, t):
    """256x1 texture, linear filter, clamp to edge, texels decoded to linear BEFORE the lerp (Rgba8UnormSrgb)."""
    x = t * 256.0 - 0.5
    i0 = np.floor(x)
    f = x - i0
    a = np.clip(i0, 0, 255).astype(int)
    b = np.clip(i0 + 1, 0, 255).astype(int)
    return lut_lin[a] * (1 - f)[..., None] + lut_lin[b] * f[..., None]


@pytest.mark.parametrize("impl", IMPLS)
@pytest.mark.parametrize("cmap,sun,exposure,exag", [("viridis", (0.3, 0.8, 0.5), 1.0, 1.0), ("terrain", (-0.6, 0.5, 0.2), 1.7, 2.5)])
def test_height_ramp_colour(render, luts, impl, cmap, sun, exposure, exag):
    """A 17 x 17 DEM that rises linearly along x, grid 17, NORTHSTAR shading, seen by the orthographic
    top-down camera.  Interpolation of a linear function is exact, so for the pixel at column px
        u = (x_world + 1.5) / 3,  h = h0 + 16 u dh,  t = (h exag' - hmin) / (hmax - hmin)     [exag' : see below]
        normal = normalize(-s da, s^2, 0),  da = (h[tx+1] - h[tx]) exag,  tx = floor(17 u) clamped (0 in the last column)
        rgb = reinhard(lut(t) (0.15 + 0.85 max(n.L, 0)) exposure),  out = sRGB8(rgb), alpha 255
    and uncovered pixels hold the clear colour (39, 39, 48, 255).  Float32 rounding may move a value sitting on a
    rounding boundary: the bar is <= 1 LSB everywhere and identical on >= 98 % of the pixels."""
    W, H, G = 128, 64, 17
    h0, dh = -0.2, 0.03
    dem = (h0 + dh * np.arange(G, dtype=np.float64))[None, :].repeat(G, 0).astype(np.float32)
    hmin, hmax = -0.25, 0.35
    spacing = 1.0
    u44 = topdown_ortho_uniforms(0.5, 0.5, sun=sun, exposure=exposure, spacing=spacing, hmin=hmin, hmax=hmax, exag=exag)
    img, prim = render(impl, W, H, G, dem, u44, cmap, False, hmin, hmax)

    x0, x1, y0, y1 = screen_rect(0.5, 0.5, W, H)                  # 16 .. 112, 8 .. 56: on pixel boundaries
    px = np.arange(W) + 0.5
    xw = (px / W * 2.0 - 1.0) / 0.5
    u = (xw + 1.5) / 3.0
    demf = dem[0].astype(np.float64)
    h = demf[0] + u * (G - 1) * (demf[1] - demf[0])               # what interpolating the texel heights gives
    t = np.clip((h - np.float32(hmin)) / (np.float64(np.float32(hmax)) - np.float32(hmin)), 0, 1)   # the varying is the raw texel height
    lut_lin = srgb8_to_linear(luts[cmap].reshape(256, 4)[:, :3].astype(np.float64))
    base = lut_sample(lut_lin, t)
    tx = np.clip(np.floor(u * G), 0, G - 1).astype(int)
    tx1 = np.minimum(tx + 1, G - 1)
    da = (demf[tx1] - demf[tx]) * exag
    n = np.stack([-spacing * da, np.full_like(da, spacing * spacing), np.zeros_like(da)], -1)
    n /= np.linalg.norm(n, axis=-1, keepdims=True)
    L = np.asarray(sun, np.float32).astype(np.float64)
    L /= np.linalg.norm(L)
    lam = np.clip(n @ L, 0, 1)
    col = base * (0.15 + 0.85 * lam)[:, None] * exposure
    rgb8 = linear_to_srgb8(col / (1.0 + col))

    want = np.empty((H, W, 4), np.int64)
    want[...] = (39, 39, 48, 255)                                 # clear (0.02, 0.02, 0.03, 1), src/scene/mod.rs:286
    cols = (px >= x0) & (px < x1)
    rows = (np.arange(H) + 0.5 >= y0) & (np.arange(H) + 0.5 < y1)
    want[np.ix_(rows, cols)] = np.concatenate([rgb8[cols], np.full((cols.sum(), 1), 255)], -1)[None]
    assert np.array_equal(prim > 0, np.outer(rows, cols)), "covered pixel set differs from the hand-derived rectangle"
    d = np.abs(img.astype(np.int64) - want)
    assert d.max() <= 1, f"max channel difference {d.max()}"
    assert (d.max(-1) == 0).mean() >= 0.98, f"only {(d.max(-1) == 0).mean():.4f} of the pixels identical"
    assert len({tuple(p) for p in img[rows][:, cols].reshape(-1, 4)}) > 40       # the ramp really produces a gradient


# ---- KAT 3: the triangle of render_triangle_rgba, coverage in exact integers, colour in float64 --------------
def triangle_expected(W, H):
    """src/lib.rs:72-91 + triangle.wgsl: NDC (-0.8,-0.8) red(1,.2,.2), (0.8,-0.8) green(.2,1,.2), (0,0.8) blue(.2,.2,1),
    clear white.  Vertices snapped to 1/256 px; a sample is inside when all three edge functions are positive, or
    zero on a top or left edge (D3D11 3.4.2: top = horizontal edge above the triangle, left = edge on its left side)."""
    ndc = [(-0.8, -0.8), (0.8, -0.8), (0.0, 0.8)]
    col = np.array([(1.0, 0.2, 0.2), (0.2, 1.0, 0.2), (0.2, 0.2, 1.0)], np.float32).astype(np.float64)
    pts = []
    for x, y in ndc:
        x, y = float(np.float32(x)), float(np.float32(y))
        pts.append((round((x * 0.5 + 0.5) * W * 256), round((-y * 0.5 + 0.5) * H * 256)))    # no .5 ties at these sizes
    (ax, ay), (bx, by), (cx, cy) = pts
    area = (bx - ax) * (cy - ay) - (by - ay) * (cx - ax)
    assert area != 0
    want = np.full((H, W, 4), 255, np.int64)
    cov = np.zeros((H, W), bool)
    centroid = ((ax + bx + cx) / 3.0, (ay + by + cy) / 3.0)

    def edge_fn(p, q, sx, sy):                   # > 0 on the side of the triangle's interior (sign fixed by the centroid)
        v = (q[1] - p[1]) * (sx - p[0]) - (q[0] - p[0]) * (sy - p[1])
        c = (q[1] - p[1]) * (centroid[0] - p[0]) - (q[0] - p[0]) * (centroid[1] - p[1])
        return v if c > 0 else -v

    def top_or_left(p, q):
        if p[1] == q[1]:                          # horizontal: top edge iff the interior is below it (y grows downwards)
            return centroid[1] > p[1]
        xe = p[0] + (q[0] - p[0]) * (centroid[1] - p[1]) / (q[1] - p[1])     # edge's x at the centroid's height
        return xe < centroid[0]                   # left edge: the interior is to its right
    edges = [(pts[1], pts[2]), (pts[2], pts[0]), (pts[0], pts[1])]
    tl = [top_or_left(p, q) for p, q in edges]
    for py in range(H):
        for px in range(W):
            sx, sy = px * 256 + 128, py * 256 + 128
            e = [edge_fn(p, q, sx, sy) for p, q in edges]
            if all(v > 0 or (v == 0 and t) for v, t in zip(e, tl)):
                cov[py, px] = True
                b = np.array(e, np.float64) / sum(e)                   # barycentric weight of the opposite vertex
                want[py, px, :3] = linear_to_srgb8(b @ col)
    return want, cov


@pytest.mark.parametrize("impl", IMPLS)
@pytest.mark.parametrize("size", [(64, 64), (100, 60)])
def test_triangle_frame(request, impl, size):
    W, H = size
    if impl == "oracle":
        img = request.getfixturevalue("oracle").render_triangle(W, H)
    else:
        img = request.getfixturevalue("vf").render_triangle_rgba(W, H)
    want, cov = triangle_expected(W, H)
    got_cov = (img[..., :3] != 255).any(-1)
    assert np.array_equal(got_cov, cov), f"coverage differs on {int((got_cov != cov).sum())} pixels"
    d = np.abs(img.astype(np.int64) - want)
    assert d.max() <= 1 and (d.max(-1) == 0).mean() >= 0.98, (int(d.max()), float((d.max(-1) == 0).mean()))


# ---- KAT 4: percentile values against numpy's stable sort of the strided sample (src/terrain_stats.rs:20-34) --
def numpy_percentile_range(a):
    a = np.asarray(a, np.float32).ravel()
    step = a.size // 65536 if a.size > 65536 else 1
    s = np.sort(a[::step], kind="stable")
    m = np.float32(s.size)
    return s[int(m * np.float32(0.01))], s[int(m * np.float32(0.99))]


@pytest.mark.parametrize("impl", IMPLS)
@pytest.mark.parametrize("n", [1, 7, 4096, 65536, 65537, 300_000, 1 << 22])
def test_percentile_values_vs_numpy(request, impl, n):
    rng = np.random.default_rng(n)
    a = (rng.normal(size=n) * rng.choice([0.1, 1.0, 50.0], size=n)).astype(np.float32)
    lo, hi = numpy_percentile_range(a)
    if impl == "oracle":
        got = request.getfixturevalue("oracle").min_max(a, True)
    else:
        got = request.getfixturevalue("ext").dem_minmax(a, True)
    assert (np.float32(got[0]), np.float32(got[1])) == (lo, hi)
