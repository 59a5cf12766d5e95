/* vfo_render.c -- CPU oracle, rendering (triangle + terrain).  TEST INFRASTRUCTURE ONLY.
 *
 * Restates: src/shaders/terrain.wgsl:39-91, src/shaders/triangle.wgsl:1-23, the fixed-function
 * state of src/terrain/pipeline.rs:107-140 and src/lib.rs:105-141 (TriangleList, CCW front,
 * cull Back, no depth, no blend, MSAA 1), the pass setup of src/scene/mod.rs:279-296 (clear
 * (0.02,0.02,0.03,1), full viewport, one indexed draw) and the implicit grid of
 * src/terrain/mod.rs:553-598.  Raster rules (unspecified by the reference, they live in
 * wgpu/driver): pixel centres at +0.5, 8 sub-pixel bits, exact integer edge functions,
 * top-left rule, clip to 0<=z<=w plus an 8x guard band in xy, perspective-correct varyings,
 * last-drawn-wins (no depth) or Less (depth).  See DESIGN.md "Arithmetic contract": every
 * float operation below is written in the exact order (and with the explicit fmaf calls) that
 * the CUDA kernels use, so that NORTHSTAR frames are bit-identical and SHIPPED frames differ
 * only through sinf/cosf.  Compile with -ffp-contract=off.
 */
#include "vfo.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define GUARD_K 8.0f
#define SUBPIX 256
#define HALFPIX 128

typedef struct { float c[4]; float a[5]; } vtx_t;            /* clip xyzw ; varyings u,v,h,x,z */
typedef struct { int32_t X, Y; float invw, zn; float a[5]; } svtx_t;
typedef struct { svtx_t v[3]; } subtri_t;

typedef struct {
    const vfo_frame *f;
    uint32_t G, W, H, Wd, Hd;
    float gm1, step, spacing, exag, hw, hh;
    const float *V, *P;
    float lut[256][3];
    float sunL[3];
} ctx_t;

int vfo_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ---- vertex stage: terrain.wgsl:44-66 on the implicit grid of terrain/mod.rs:558-570 ---- */
static inline int texel_index(float coord, uint32_t size)
{
    /* nearest, clamp-to-edge: floor(u*size) clamped */
    const float f = floorf(coord * (float)size);
    int t = (int)f;
    if (t < 0) t = 0;
    if (t > (int)size - 1) t = (int)size - 1;
    return t;
}

static void vertex_stage(const ctx_t *c, uint32_t i, uint32_t j, vtx_t *o)
{
    const float x = -1.5f + (float)i * c->step;          /* host-side, no FMA (Rust) */
    const float z = -1.5f + (float)j * c->step;
    const float u = (float)i / c->gm1;
    const float v = (float)j / c->gm1;
    const int tx = texel_index(u, c->Wd), ty = texel_index(v, c->Hd);
    const float h_tex = c->f->dem[(size_t)ty * c->Wd + tx];
    float h;
    if (c->f->mode == VFO_MODE_SHIPPED) {
        const float s = sinf(x * 1.3f), k = cosf(z * 1.1f);
        h = h_tex + (s * 0.25f + k * 0.25f);              /* analytic_height, wgsl:39-41,55 */
    } else {
        h = h_tex;
    }
    const float wx = x * c->spacing, wy = h * c->exag, wz = z * c->spacing;
    float vv[4];
    for (int r = 0; r < 4; ++r)   /* view * vec4(world,1): ((c3 + c0*wx) + c2*wz) + c1*wy */
        vv[r] = fmaf(c->V[4 + r], wy, fmaf(c->V[8 + r], wz, fmaf(c->V[r], wx, c->V[12 + r])));
    for (int r = 0; r < 4; ++r)   /* proj * v */
        o->c[r] = fmaf(c->P[12 + r], vv[3],
                       fmaf(c->P[8 + r], vv[2], fmaf(c->P[4 + r], vv[1], c->P[r] * vv[0])));
    o->a[0] = u; o->a[1] = v; o->a[2] = h; o->a[3] = x; o->a[4] = z;
}

/* ---- clip + viewport ---- */
static inline void to_screen(const ctx_t *c, const vtx_t *v, svtx_t *s)
{
    const float invw = 1.0f / v->c[3];
    const float xn = v->c[0] * invw, yn = v->c[1] * invw;
    const float sx = fmaf(xn, c->hw, c->hw);
    const float sy = fmaf(-yn, c->hh, c->hh);          /* framebuffer y runs down */
    s->X = (int32_t)lrintf(sx * (float)SUBPIX);
    s->Y = (int32_t)lrintf(sy * (float)SUBPIX);
    s->invw = invw;
    s->zn = v->c[2] * invw;
    memcpy(s->a, v->a, sizeof s->a);
}

static inline int vtx_inside_all(const vtx_t *v)
{
    const float x = v->c[0], y = v->c[1], z = v->c[2], w = v->c[3];
    const float g = GUARD_K * w;
    return w > 0.0f && z >= 0.0f && z <= w && x >= -g && x <= g && y >= -g && y <= g;
}

static inline float plane_dist(const vtx_t *v, int p)
{
    const float x = v->c[0], y = v->c[1], z = v->c[2], w = v->c[3];
    switch (p) {
    case 0: return z;                      /* near   z >= 0 */
    case 1: return w - z;                  /* far    z <= w */
    case 2: return fmaf(GUARD_K, w, x);    /* x >= -K w */
    case 3: return fmaf(GUARD_K, w, -x);   /* x <=  K w */
    case 4: return fmaf(GUARD_K, w, y);
    default: return fmaf(GUARD_K, w, -y);
    }
}

static inline void vtx_lerp(const vtx_t *p, const vtx_t *q, float t, vtx_t *o)
{
    for (int k = 0; k < 4; ++k) o->c[k] = fmaf(t, q->c[k] - p->c[k], p->c[k]);
    for (int k = 0; k < 5; ++k) o->a[k] = fmaf(t, q->a[k] - p->a[k], p->a[k]);
}

/* Returns number of screen-space sub-triangles (0..7) for one primitive.  *clipped is set when
 * the Sutherland-Hodgman path ran. */
static int prim_setup(const ctx_t *c, const vtx_t tri[3], subtri_t out[7], int *clipped)
{
    *clipped = 0;
    /* trivial reject against the view frustum */
    int ox0 = 1, ox1 = 1, oy0 = 1, oy1 = 1, oz0 = 1, oz1 = 1;
    for (int k = 0; k < 3; ++k) {
        const float x = tri[k].c[0], y = tri[k].c[1], z = tri[k].c[2], w = tri[k].c[3];
        ox0 &= (x < -w); ox1 &= (x > w);
        oy0 &= (y < -w); oy1 &= (y > w);
        oz0 &= (z < 0.0f); oz1 &= (z > w);
    }
    if (ox0 | ox1 | oy0 | oy1 | oz0 | oz1) return 0;
    if (vtx_inside_all(&tri[0]) && vtx_inside_all(&tri[1]) && vtx_inside_all(&tri[2])) {
        for (int k = 0; k < 3; ++k) to_screen(c, &tri[k], &out[0].v[k]);
        return 1;
    }
    *clipped = 1;
    vtx_t bufa[12], bufb[12];
    vtx_t *in = bufa, *outp = bufb;
    int n = 3;
    memcpy(in, tri, 3 * sizeof(vtx_t));
    for (int p = 0; p < 6 && n > 0; ++p) {
        int m = 0;
        for (int k = 0; k < n; ++k) {
            const vtx_t *prev = &in[(k + n - 1) % n], *cur = &in[k];
            const float dp = plane_dist(prev, p), dc = plane_dist(cur, p);
            const int pin = dp >= 0.0f, cin = dc >= 0.0f;
            if (cin) {
                if (!pin) { vtx_lerp(prev, cur, dp / (dp - dc), &outp[m]); ++m; }
                outp[m++] = *cur;
            } else if (pin) {
                vtx_lerp(prev, cur, dp / (dp - dc), &outp[m]); ++m;
            }
        }
        vtx_t *t = in; in = outp; outp = t;
        n = m;
    }
    if (n < 3) return 0;
    for (int k = 0; k < n; ++k)
        if (!(in[k].c[3] > 0.0f) || !isfinite(in[k].c[3])) return 0;
    svtx_t s[12];
    for (int k = 0; k < n; ++k) to_screen(c, &in[k], &s[k]);
    int cnt = 0;
    for (int k = 1; k + 1 < n; ++k) {           /* fan */
        out[cnt].v[0] = s[0]; out[cnt].v[1] = s[k]; out[cnt].v[2] = s[k + 1];
        ++cnt;
    }
    return cnt;
}

/* ---- coverage ---- */
static inline int64_t orient2d(const svtx_t *a, const svtx_t *b, int64_t px, int64_t py)
{
    return (int64_t)(b->X - a->X) * (py - a->Y) - (int64_t)(b->Y - a->Y) * (px - a->X);
}
static inline int top_left(const svtx_t *a, const svtx_t *b)
{
    /* front faces are counter-clockwise as displayed (negative area in y-down coordinates):
     * a left edge runs downwards, a top edge is horizontal and runs to the left */
    return (b->Y > a->Y) || (b->Y == a->Y && b->X < a->X);
}
static inline int64_t tri_area2(const subtri_t *t)
{
    return (int64_t)(t->v[1].X - t->v[0].X) * (t->v[2].Y - t->v[0].Y)
         - (int64_t)(t->v[1].Y - t->v[0].Y) * (t->v[2].X - t->v[0].X);
}
/* e[k] >= 0 inside; returns coverage of the sample at pixel (px,py) */
static inline int covers(const subtri_t *t, int px, int py, int64_t e[3])
{
    const int64_t sx = (int64_t)px * SUBPIX + HALFPIX, sy = (int64_t)py * SUBPIX + HALFPIX;
    e[0] = -orient2d(&t->v[1], &t->v[2], sx, sy);
    e[1] = -orient2d(&t->v[2], &t->v[0], sx, sy);
    e[2] = -orient2d(&t->v[0], &t->v[1], sx, sy);
    if (e[0] < 0 || e[1] < 0 || e[2] < 0) return 0;
    if (e[0] == 0 && !top_left(&t->v[1], &t->v[2])) return 0;
    if (e[1] == 0 && !top_left(&t->v[2], &t->v[0])) return 0;
    if (e[2] == 0 && !top_left(&t->v[0], &t->v[1])) return 0;
    return 1;
}
static inline int bbox_pixels(const subtri_t *t, uint32_t W, uint32_t H,
                              int *x0, int *x1, int *y0, int *y1)
{
    int32_t mnx = t->v[0].X, mxx = mnx, mny = t->v[0].Y, mxy = mny;
    for (int k = 1; k < 3; ++k) {
        if (t->v[k].X < mnx) mnx = t->v[k].X;
        if (t->v[k].X > mxx) mxx = t->v[k].X;
        if (t->v[k].Y < mny) mny = t->v[k].Y;
        if (t->v[k].Y > mxy) mxy = t->v[k].Y;
    }
    *x0 = (mnx - HALFPIX + SUBPIX - 1) >> 8;      /* ceil((min-128)/256), arithmetic shift */
    *x1 = (mxx - HALFPIX) >> 8;                   /* floor((max-128)/256) */
    *y0 = (mny - HALFPIX + SUBPIX - 1) >> 8;
    *y1 = (mxy - HALFPIX) >> 8;
    if (*x0 < 0) *x0 = 0;
    if (*y0 < 0) *y0 = 0;
    if (*x1 > (int)W - 1) *x1 = (int)W - 1;
    if (*y1 > (int)H - 1) *y1 = (int)H - 1;
    return *x0 <= *x1 && *y0 <= *y1;
}
static inline void barycentric(const int64_t e[3], float b[3])
{
    const float r = 1.0f / (float)(e[0] + e[1] + e[2]);
    b[0] = (float)e[0] * r; b[1] = (float)e[1] * r; b[2] = (float)e[2] * r;
}
static inline float depth_at(const subtri_t *t, const float b[3])
{
    float z = fmaf(b[2], t->v[2].zn, fmaf(b[1], t->v[1].zn, b[0] * t->v[0].zn));
    z = fmaxf(z, 0.0f);
    return fminf(z, 1.0f);
}

static inline void atomic_max_u32(uint32_t *p, uint32_t v)
{
    uint32_t cur = __atomic_load_n(p, __ATOMIC_RELAXED);
    while (cur < v && !__atomic_compare_exchange_n(p, &cur, v, 1, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) {}
}
static inline void atomic_min_u64(uint64_t *p, uint64_t v)
{
    uint64_t cur = __atomic_load_n(p, __ATOMIC_RELAXED);
    while (cur > v && !__atomic_compare_exchange_n(p, &cur, v, 1, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) {}
}

static inline void prim_vertex_ids(uint32_t G, uint32_t prim, uint32_t ij[3][2])
{
    const uint32_t cell = prim >> 1, gm = G - 1;
    const uint32_t j = cell / gm, i = cell % gm;
    if ((prim & 1u) == 0) {          /* (a,c,b) : terrain/mod.rs:581 */
        ij[0][0] = i;     ij[0][1] = j;
        ij[1][0] = i;     ij[1][1] = j + 1;
        ij[2][0] = i + 1; ij[2][1] = j;
    } else {                         /* (b,c,d) */
        ij[0][0] = i + 1; ij[0][1] = j;
        ij[1][0] = i;     ij[1][1] = j + 1;
        ij[2][0] = i + 1; ij[2][1] = j + 1;
    }
}

/* ---- fragment stage ---- */
static inline void v3_norm(float v[3])
{
    const float d = fmaf(v[2], v[2], fmaf(v[0], v[0], v[1] * v[1]));
    const float r = 1.0f / sqrtf(d);
    v[0] *= r; v[1] *= r; v[2] *= r;
}
static inline float clamp01(float x) { return fminf(fmaxf(x, 0.0f), 1.0f); }

static inline void lut_sample(const ctx_t *c, float t, float rgb[3])
{
    /* linear filter, clamp-to-edge, 256x1: texel space x = t*256 - 0.5 */
    const float x = t * 256.0f - 0.5f;
    const float fl = floorf(x);
    const float fr = x - fl;
    int i0 = (int)fl, i1 = i0 + 1;
    if (i0 < 0) i0 = 0;
    if (i0 > 255) i0 = 255;
    if (i1 < 0) i1 = 0;
    if (i1 > 255) i1 = 255;
    for (int k = 0; k < 3; ++k)
        rgb[k] = fmaf(fr, c->lut[i1][k] - c->lut[i0][k], c->lut[i0][k]);
}

static void fragment_stage(const ctx_t *c, const float a[5], uint8_t px[4])
{
    const float u = a[0], v = a[1], h = a[2], x = a[3], z = a[4];
    float rgb[3], n[3], lam, lut[3];
    if (c->f->mode == VFO_MODE_SHIPPED) {                 /* terrain.wgsl:69-91 */
        const float hr = fmaxf(c->f->uniforms[37], 1e-8f);
        const float t = clamp01(0.5f + h / (2.0f * hr));
        lut_sample(c, t, lut);
        const float dhdx = 1.3f * cosf(x * 1.3f) * 0.25f;
        const float dhdz = -1.1f * sinf(z * 1.1f) * 0.25f;
        n[0] = -dhdx; n[1] = 1.0f; n[2] = -dhdz;
        v3_norm(n);
        lam = clamp01(fmaf(n[2], c->sunL[2], fmaf(n[1], c->sunL[1], n[0] * c->sunL[0])));
        const float shade = fmaf(0.85f, lam, 0.15f);      /* mix(0.15, 1.0, lambert) */
        const float ex = c->f->uniforms[35];
        for (int k = 0; k < 3; ++k) rgb[k] = (lut[k] * ex) * shade;
    } else {                                              /* SURVEY A3 / ROADMAP.md:426-436 */
        const float t = clamp01((h - c->f->h_min) / (c->f->h_max - c->f->h_min));
        lut_sample(c, t, lut);
        const int tx = texel_index(u, c->Wd), ty = texel_index(v, c->Hd);
        const int tx1 = tx + 1 > (int)c->Wd - 1 ? (int)c->Wd - 1 : tx + 1;
        const int ty1 = ty + 1 > (int)c->Hd - 1 ? (int)c->Hd - 1 : ty + 1;
        const float h0 = c->f->dem[(size_t)ty * c->Wd + tx];
        const float hx = c->f->dem[(size_t)ty * c->Wd + tx1];
        const float hz = c->f->dem[(size_t)ty1 * c->Wd + tx];
        const float s = c->spacing;
        const float da = (hx - h0) * c->exag, db = (hz - h0) * c->exag;
        /* cross(dpz, dpx), dpx=(s,da,0), dpz=(0,db,s) in the Y-up frame of the vertex stage */
        n[0] = -(s * da); n[1] = s * s; n[2] = -(db * s);
        v3_norm(n);
        lam = clamp01(fmaf(n[2], c->sunL[2], fmaf(n[1], c->sunL[1], n[0] * c->sunL[0])));
        const float shade = fmaf(0.85f, lam, 0.15f);
        const float ex = c->f->uniforms[35];
        for (int k = 0; k < 3; ++k) {
            const float col = (lut[k] * shade) * ex;
            rgb[k] = col / (1.0f + col);                  /* Reinhard, linear */
        }
    }
    for (int k = 0; k < 3; ++k) px[k] = vfo_srgb_encode_u8(rgb[k]);
    px[3] = 255;
}

/* Shade pixel (px,py) from primitive `prim` (visibility-buffer resolve). */
static int shade_pixel(const ctx_t *c, uint32_t prim, int px, int py, uint8_t out[4])
{
    uint32_t ij[3][2];
    vtx_t tri[3];
    subtri_t sub[7];
    int clipped;
    prim_vertex_ids(c->G, prim, ij);
    for (int k = 0; k < 3; ++k) vertex_stage(c, ij[k][0], ij[k][1], &tri[k]);
    const int n = prim_setup(c, tri, sub, &clipped);
    for (int s = 0; s < n; ++s) {
        if (tri_area2(&sub[s]) >= 0) continue;
        int64_t e[3];
        if (!covers(&sub[s], px, py, e)) continue;
        float b[3], a[5];
        barycentric(e, b);
        const svtx_t *v = sub[s].v;
        const float iw = fmaf(b[2], v[2].invw, fmaf(b[1], v[1].invw, b[0] * v[0].invw));
        for (int k = 0; k < 5; ++k) {
            const float num = fmaf(b[2], v[2].a[k] * v[2].invw,
                              fmaf(b[1], v[1].a[k] * v[1].invw, b[0] * (v[0].a[k] * v[0].invw)));
            a[k] = num / iw;
        }
        fragment_stage(c, a, out);
        return 1;
    }
    return 0; /* cannot happen: the visibility pass found coverage with identical arithmetic */
}

static void ctx_init(ctx_t *c, const vfo_frame *f)
{
    memset(c, 0, sizeof *c);
    c->f = f;
    c->G = f->grid < 2 ? 2 : f->grid;
    c->W = f->width; c->H = f->height; c->Wd = f->dem_w; c->Hd = f->dem_h;
    c->gm1 = (float)c->G - 1.0f;
    c->step = (2.0f * 1.5f) / c->gm1;
    c->spacing = fmaxf(f->uniforms[36], 1e-8f);          /* terrain.wgsl:46 */
    c->exag = f->uniforms[38];
    c->hw = 0.5f * (float)c->W; c->hh = 0.5f * (float)c->H;
    c->V = f->uniforms; c->P = f->uniforms + 16;
    uint8_t lin[1024];
    if (f->lut_unorm) vfo_lut_linearize_u8(f->lut_rgba, 256, lin);
    for (int i = 0; i < 256; ++i)
        for (int k = 0; k < 3; ++k)
            c->lut[i][k] = f->lut_unorm ? (float)lin[4 * i + k] / 255.0f
                                        : vfo_srgb8_to_linear(f->lut_rgba[4 * i + k]);
    c->sunL[0] = f->uniforms[32]; c->sunL[1] = f->uniforms[33]; c->sunL[2] = f->uniforms[34];
    v3_norm(c->sunL);                                    /* terrain.wgsl:83 */
}

int vfo_render_terrain(const vfo_frame *f, uint8_t *out_rgba, uint32_t *out_prim,
                       vfo_counters *counters)
{
    ctx_t cx;
    ctx_init(&cx, f);
    const ctx_t *c = &cx;
    const size_t npix = (size_t)c->W * c->H;
    const int depth = f->depth;
    uint32_t *vis = NULL;
    uint64_t *zkey = NULL;
    if (depth) {
        zkey = (uint64_t *)malloc(npix * sizeof(uint64_t));
        if (!zkey) return -1;
        for (size_t i = 0; i < npix; ++i) zkey[i] = ~(uint64_t)0;
    } else {
        vis = (uint32_t *)calloc(npix, sizeof(uint32_t));
        if (!vis) return -1;
    }
    int threads = f->threads > 0 ? f->threads : vfo_max_threads();
    const uint32_t G = c->G, R = 8;                       /* rows of cells per task */
    const uint32_t nblocks = (G - 1 + R - 1) / R;
    uint64_t n_total = 0, n_front = 0, n_sampled = 0, n_clipped = 0, n_frag = 0;

#pragma omp parallel num_threads(threads) reduction(+ : n_total, n_front, n_sampled, n_clipped, n_frag)
    {
        vtx_t *rows = (vtx_t *)malloc((size_t)(R + 1) * G * sizeof(vtx_t));
#pragma omp for schedule(dynamic, 1)
        for (uint32_t blk = 0; blk < nblocks; ++blk) {
            const uint32_t j0 = blk * R;
            const uint32_t j1 = j0 + R < G - 1 ? j0 + R : G - 1;
            for (uint32_t j = j0; j <= j1; ++j)
                for (uint32_t i = 0; i < G; ++i) vertex_stage(c, i, j, &rows[(size_t)(j - j0) * G + i]);
            for (uint32_t j = j0; j < j1; ++j)
                for (uint32_t i = 0; i + 1 < G; ++i) {
                    const vtx_t *ra = &rows[(size_t)(j - j0) * G + i];
                    const vtx_t *rc = ra + G;
                    for (int k = 0; k < 2; ++k) {
                        const uint32_t prim = 2u * (j * (G - 1) + i) + (uint32_t)k;
                        vtx_t tri[3];
                        if (k == 0) { tri[0] = ra[0]; tri[1] = rc[0]; tri[2] = ra[1]; }
                        else        { tri[0] = ra[1]; tri[1] = rc[0]; tri[2] = rc[1]; }
                        subtri_t sub[7];
                        int clipped;
                        ++n_total;
                        const int n = prim_setup(c, tri, sub, &clipped);
                        n_clipped += (uint64_t)clipped;
                        int any_front = 0, any_sample = 0;
                        for (int s = 0; s < n; ++s) {
                            int x0, x1, y0, y1;
                            if (tri_area2(&sub[s]) >= 0) continue;   /* cull Back / degenerate */
                            any_front = 1;
                            if (!bbox_pixels(&sub[s], c->W, c->H, &x0, &x1, &y0, &y1)) continue;
                            for (int py = y0; py <= y1; ++py)
                                for (int px = x0; px <= x1; ++px) {
                                    int64_t e[3];
                                    if (!covers(&sub[s], px, py, e)) continue;
                                    any_sample = 1;
                                    ++n_frag;
                                    const size_t p = (size_t)py * c->W + px;
                                    if (depth) {
                                        float b[3];
                                        barycentric(e, b);
                                        const float zf = depth_at(&sub[s], b);
                                        if (!(zf < 1.0f)) continue;          /* Less vs clear 1.0 */
                                        uint32_t zb;
                                        memcpy(&zb, &zf, 4);
                                        atomic_min_u64(&zkey[p], ((uint64_t)zb << 32) | prim);
                                    } else {
                                        atomic_max_u32(&vis[p], prim + 1u);  /* last drawn wins */
                                    }
                                }
                        }
                        n_front += (uint64_t)any_front;
                        n_sampled += (uint64_t)any_sample;
                    }
                }
        }
        free(rows);
    }

    uint8_t clear[4];                                      /* scene/mod.rs:286 */
    clear[0] = vfo_srgb_encode_u8(0.02f); clear[1] = vfo_srgb_encode_u8(0.02f);
    clear[2] = vfo_srgb_encode_u8(0.03f); clear[3] = 255;
#pragma omp parallel for num_threads(threads) schedule(dynamic, 4)
    for (uint32_t py = 0; py < c->H; ++py)
        for (uint32_t px = 0; px < c->W; ++px) {
            const size_t p = (size_t)py * c->W + px;
            uint32_t id1;
            if (depth) id1 = zkey[p] == ~(uint64_t)0 ? 0u : (uint32_t)(zkey[p] & 0xffffffffu) + 1u;
            else id1 = vis[p];
            if (out_prim) out_prim[p] = id1;
            uint8_t *o = out_rgba + 4 * p;
            if (id1 == 0 || !shade_pixel(c, id1 - 1u, (int)px, (int)py, o)) memcpy(o, clear, 4);
        }
    if (counters) {
        counters->tris_total = n_total; counters->tris_front = n_front;
        counters->tris_sampled = n_sampled; counters->tris_clipped = n_clipped;
        counters->fragments = n_frag;
    }
    free(vis);
    free(zkey);
    return 0;
}

/* ---- triangle path: src/lib.rs:72-91 (geometry), :685-721 (clear WHITE, one draw),
 * src/shaders/triangle.wgsl:1-23 (pass-through position, colour varying) ---- */
int vfo_render_triangle(uint32_t w, uint32_t h, uint8_t *out)
{
    static const float pos[3][2] = { { -0.8f, -0.8f }, { 0.8f, -0.8f }, { 0.0f, 0.8f } };
    static const float col[3][3] = { { 1.0f, 0.2f, 0.2f }, { 0.2f, 1.0f, 0.2f }, { 0.2f, 0.2f, 1.0f } };
    ctx_t c;
    memset(&c, 0, sizeof c);
    c.W = w; c.H = h; c.hw = 0.5f * (float)w; c.hh = 0.5f * (float)h;
    subtri_t t;
    for (int k = 0; k < 3; ++k) {
        vtx_t v;
        memset(&v, 0, sizeof v);
        v.c[0] = pos[k][0]; v.c[1] = pos[k][1]; v.c[2] = 0.0f; v.c[3] = 1.0f;
        to_screen(&c, &v, &t.v[k]);
    }
    const int front = tri_area2(&t) < 0;
    for (uint32_t py = 0; py < h; ++py)
        for (uint32_t px = 0; px < w; ++px) {
            uint8_t *o = out + 4 * ((size_t)py * w + px);
            int64_t e[3];
            if (front && covers(&t, (int)px, (int)py, e)) {
                float b[3];
                barycentric(e, b);
                for (int k = 0; k < 3; ++k)
                    o[k] = vfo_srgb_encode_u8(fmaf(b[2], col[2][k], fmaf(b[1], col[1][k], b[0] * col[0][k])));
                o[3] = 255;
            } else {
                o[0] = o[1] = o[2] = o[3] = 255;           /* CLEAR_COLOR white, lib.rs:19 */
            }
        }
    return 0;
}
