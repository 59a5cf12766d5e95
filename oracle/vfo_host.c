/* vfo_host.c -- CPU oracle, host-side pieces (grid, DEM statistics, camera, uniforms, LUT).
 * TEST INFRASTRUCTURE ONLY (see vfo.h).  Compile with -ffp-contract=off: the reference's host
 * code is Rust, which never contracts a*b+c into an FMA. */
#include "vfo.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------
 * Public grid: src/terrain/mesh.rs:35-90.  Vertex (x,y): position (x*dx-cx, y*dy-cy) with
 * cx=(w-1)*0.5*dx; uv (x/(w-1), y/(h-1)); indices per cell [i0,i1,i2,i2,i1,i3].
 * The u16/u32 split (mesh.rs:62-89) is invisible to Python (widened at :195-198). */
void vfo_make_grid(uint32_t w, uint32_t h, float dx, float dy,
                   float *xy, float *uv, uint32_t *idx)
{
    const float wm1 = (float)w - 1.0f, hm1 = (float)h - 1.0f;
    const float cx = wm1 * 0.5f * dx;
    const float cy = hm1 * 0.5f * dy;
    size_t k = 0;
    for (uint32_t y = 0; y < h; ++y) {
        const float wy = (float)y * dy - cy;
        const float v = (float)y / hm1;
        for (uint32_t x = 0; x < w; ++x, ++k) {
            xy[2 * k + 0] = (float)x * dx - cx;
            xy[2 * k + 1] = wy;
            uv[2 * k + 0] = (float)x / wm1;
            uv[2 * k + 1] = v;
        }
    }
    size_t m = 0;
    for (uint32_t y = 0; y + 1 < h; ++y) {
        const uint32_t row = y * w;
        for (uint32_t x = 0; x + 1 < w; ++x) {
            const uint32_t i0 = row + x, i1 = i0 + 1, i2 = i0 + w, i3 = i2 + 1;
            idx[m++] = i0; idx[m++] = i1; idx[m++] = i2;
            idx[m++] = i2; idx[m++] = i1; idx[m++] = i3;
        }
    }
}

/* Render-path grid: src/terrain/mod.rs:553-598 (identical inline copy src/scene/mod.rs:85-116).
 * x,z in [-1.5,1.5]; indices [a,c,b, b,c,d]. */
void vfo_render_grid_xyuv(uint32_t g, float *verts, uint32_t *idx)
{
    const uint32_t n = g < 2 ? 2 : g;
    const float scale = 1.5f;
    const float step = (2.0f * scale) / ((float)n - 1.0f);
    size_t k = 0;
    for (uint32_t j = 0; j < n; ++j)
        for (uint32_t i = 0; i < n; ++i, ++k) {
            verts[4 * k + 0] = -scale + (float)i * step;
            verts[4 * k + 1] = -scale + (float)j * step;
            verts[4 * k + 2] = (float)i / ((float)n - 1.0f);
            verts[4 * k + 3] = (float)j / ((float)n - 1.0f);
        }
    if (!idx) return;
    size_t m = 0;
    for (uint32_t j = 0; j + 1 < n; ++j)
        for (uint32_t i = 0; i + 1 < n; ++i) {
            const uint32_t a = j * n + i, b = a + 1, c = a + n, d = c + 1;
            idx[m++] = a; idx[m++] = c; idx[m++] = b;
            idx[m++] = b; idx[m++] = c; idx[m++] = d;
        }
}

/* ------------------------------------------------------------------------------------
 * Stable merge sort with the comparator of src/terrain_stats.rs:31
 * (partial_cmp, unordered -> Equal; slice::sort_by is stable). */
static int cmp_lt(float a, float b) { return a < b; } /* "Less" only when strictly ordered */

static void merge_sort(float *a, float *tmp, size_t n)
{
    if (n < 2) return;
    const size_t h = n / 2;
    merge_sort(a, tmp, h);
    merge_sort(a + h, tmp, n - h);
    size_t i = 0, j = h, k = 0;
    while (i < h && j < n) {
        /* stable: take from the right run only if right < left strictly */
        if (cmp_lt(a[j], a[i])) tmp[k++] = a[j++];
        else tmp[k++] = a[i++];
    }
    while (i < h) tmp[k++] = a[i++];
    while (j < n) tmp[k++] = a[j++];
    memcpy(a, tmp, n * sizeof(float));
}

/* src/terrain_stats.rs:11-35 */
void vfo_min_max(const float *data, size_t n, int clamp, float *lo_out, float *hi_out)
{
    if (!clamp) {
        float lo = INFINITY, hi = -INFINITY;
        for (size_t i = 0; i < n; ++i) {
            const float v = data[i];
            if (v < lo) lo = v;
            if (v > hi) hi = v;
        }
        *lo_out = lo; *hi_out = hi;
        return;
    }
    const size_t SAMPLE = 65536;
    size_t m;
    float *buf;
    if (n > SAMPLE) {
        const size_t step = n / SAMPLE;          /* integer division, :25 */
        m = (n + step - 1) / step;               /* len of iter().step_by(step) */
        buf = (float *)malloc(m * sizeof(float));
        for (size_t k = 0, i = 0; i < n; i += step, ++k) buf[k] = data[i];
    } else {
        m = n;
        buf = (float *)malloc(m * sizeof(float));
        memcpy(buf, data, m * sizeof(float));
    }
    float *tmp = (float *)malloc(m * sizeof(float));
    merge_sort(buf, tmp, m);
    const size_t i1 = (size_t)((float)m * 0.01f);   /* (len as f32 * 0.01) as usize, :32 */
    const size_t i99 = (size_t)((float)m * 0.99f);  /* :33 */
    *lo_out = buf[i1];
    *hi_out = buf[i99];
    free(tmp);
    free(buf);
}

/* src/renderer.rs:24-28 */
void vfo_h_range(const float *data, size_t n, float *h_min, float *h_max)
{
    float lo, hi;
    vfo_min_max(data, n, 1, &lo, &hi);
    const float guard = lo + 1e-5f;
    *h_min = lo;
    *h_max = hi > guard ? hi : guard;   /* f32::max */
}

/* src/lib.rs:905-932 : sequential f32 accumulation, two-pass population variance */
void vfo_dem_stats(const float *d, size_t n, float out4[4])
{
    if (n == 0) { out4[0] = out4[1] = out4[2] = out4[3] = 0.0f; return; }
    float mn = d[0], mx = d[0], sum = 0.0f;
    for (size_t i = 0; i < n; ++i) {
        const float h = d[i];
        if (h < mn) mn = h;
        if (h > mx) mx = h;
        sum += h;
    }
    const float mean = sum / (float)n;
    float vs = 0.0f;
    for (size_t i = 0; i < n; ++i) {
        const float diff = d[i] - mean;
        vs += diff * diff;
    }
    out4[0] = mn; out4[1] = mx; out4[2] = mean; out4[3] = sqrtf(vs / (float)n);
}

/* src/lib.rs:934-951 */
void vfo_normalize(float *d, size_t n, int mode, float eps, float lo, float hi,
                   const float s[4])
{
    if (mode == 0) {
        float denom = fabsf(s[1] - s[0]);
        if (!(denom > eps)) denom = eps;          /* f32::max(abs, eps) */
        const float scale = (hi - lo) / denom;
        for (size_t i = 0; i < n; ++i) d[i] = (d[i] - s[0]) * scale + lo;
    } else {
        float denom = s[3] > eps ? s[3] : eps;
        for (size_t i = 0; i < n; ++i) d[i] = (d[i] - s[2]) / denom;
    }
}

/* ------------------------------------------------------------------------------------
 * glam 0.24.2 (Cargo.lock:297-298), not vendored: restated from its published source.
 * Vec3::normalize = v * (1/length); length = sqrt(x*x+y*y+z*z). */
static void v3_normalize(const float v[3], float o[3])
{
    const float len = sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    const float r = 1.0f / len;
    o[0] = v[0] * r; o[1] = v[1] * r; o[2] = v[2] * r;
}
static void v3_normalize_or_zero(const float v[3], float o[3])
{
    const float len = sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    const float r = 1.0f / len;
    if (isfinite(r) && r > 0.0f) { o[0] = v[0] * r; o[1] = v[1] * r; o[2] = v[2] * r; }
    else { o[0] = o[1] = o[2] = 0.0f; }
}
static void v3_cross(const float a[3], const float b[3], float o[3])
{
    o[0] = a[1] * b[2] - b[1] * a[2];
    o[1] = a[2] * b[0] - b[2] * a[0];
    o[2] = a[0] * b[1] - b[0] * a[1];
}
static float v3_dot(const float a[3], const float b[3])
{
    return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}

/* Mat4::look_at_rh(eye, center, up) == look_to_rh(eye, center-eye, up) */
void vfo_look_at_rh(const float eye[3], const float target[3], const float up[3], float m[16])
{
    float dir[3] = { target[0] - eye[0], target[1] - eye[1], target[2] - eye[2] };
    float f[3], s[3], u[3], t[3];
    v3_normalize(dir, f);
    v3_cross(f, up, t);
    v3_normalize(t, s);
    v3_cross(s, f, u);
    m[0] = s[0]; m[1] = u[0]; m[2] = -f[0]; m[3] = 0.0f;
    m[4] = s[1]; m[5] = u[1]; m[6] = -f[1]; m[7] = 0.0f;
    m[8] = s[2]; m[9] = u[2]; m[10] = -f[2]; m[11] = 0.0f;
    m[12] = -v3_dot(eye, s); m[13] = -v3_dot(eye, u); m[14] = v3_dot(eye, f); m[15] = 1.0f;
}

/* Mat4::perspective_rh_gl */
void vfo_perspective_rh_gl(float fovy, float aspect, float zn, float zf, float m[16])
{
    const float inv_length = 1.0f / (zn - zf);
    const float f = 1.0f / tanf(0.5f * fovy);
    const float a = f / aspect;
    const float b = (zn + zf) * inv_length;
    const float c = (2.0f * zn * zf) * inv_length;
    memset(m, 0, 16 * sizeof(float));
    m[0] = a; m[5] = f; m[10] = b; m[11] = -1.0f; m[14] = c;
}

/* Mat4 * Mat4: column k = a.x_axis*b[k].x + a.y_axis*b[k].y + a.z_axis*b[k].z + a.w_axis*b[k].w */
void vfo_mat4_mul(const float a[16], const float b[16], float o[16])
{
    float t[16];
    for (int c = 0; c < 4; ++c)
        for (int r = 0; r < 4; ++r)
            t[4 * c + r] = ((a[0 + r] * b[4 * c + 0] + a[4 + r] * b[4 * c + 1])
                            + a[8 + r] * b[4 * c + 2]) + a[12 + r] * b[4 * c + 3];
    memcpy(o, t, sizeof t);
}

/* src/camera.rs:14-21: from_cols_array(&[1,0,0,0, 0,1,0,0, 0,0,0.5,0.5, 0,0,0,1]) -- the array
 * is consumed COLUMN by column, so column 2 is (0,0,0.5,0.5): z' = 0.5 z, w' = 0.5 z + w.
 * Reproduced as written. */
void vfo_perspective_wgpu(float fovy, float aspect, float zn, float zf, float m[16])
{
    static const float g[16] = { 1, 0, 0, 0,  0, 1, 0, 0,  0, 0, 0.5f, 0.5f,  0, 0, 0, 1 };
    float p[16];
    vfo_perspective_rh_gl(fovy, aspect, zn, zf, p);
    vfo_mat4_mul(g, p, m);
}

/* src/camera.rs:224-240 (order of checks matters: first failure wins).
 * 1 vec finite, 2 up colinear, 3 fovy, 4 near, 5 far */
int vfo_validate_camera(const float eye[3], const float target[3], const float up[3],
                        float fovy_deg, float znear, float zfar)
{
    for (int i = 0; i < 3; ++i)
        if (!isfinite(eye[i])) return 1;
    for (int i = 0; i < 3; ++i)
        if (!isfinite(target[i])) return 1;
    for (int i = 0; i < 3; ++i)
        if (!isfinite(up[i])) return 1;
    float d[3] = { target[0] - eye[0], target[1] - eye[1], target[2] - eye[2] };
    float vd[3], un[3], cr[3];
    v3_normalize_or_zero(d, vd);
    v3_normalize_or_zero(up, un);
    v3_cross(vd, un, cr);
    if (v3_dot(cr, cr) < 1e-6f) return 2;
    if (!isfinite(fovy_deg) || fovy_deg <= 0.0f || fovy_deg >= 180.0f) return 3;
    if (!isfinite(znear) || znear <= 0.0f) return 4;
    if (!isfinite(zfar) || zfar <= znear) return 5;
    return 0;
}

/* src/terrain/mod.rs:124-142 */
void vfo_uniforms(const float view[16], const float proj[16], const float sun[3],
                  float exposure, float spacing, float h_min, float h_max,
                  float exaggeration, float o[44])
{
    memcpy(o, view, 64);
    memcpy(o + 16, proj, 64);
    o[32] = sun[0]; o[33] = sun[1]; o[34] = sun[2]; o[35] = exposure;
    o[36] = spacing; o[37] = h_max - h_min; o[38] = exaggeration; o[39] = 0.0f;
    o[40] = o[41] = o[42] = o[43] = 0.0f;
}

/* src/lib.rs:444-453 */
void vfo_sun_dir(float el_deg, float az_deg, float out[3])
{
    const float pi = 3.14159265358979323846f;
    const float el = el_deg * pi / 180.0f;
    const float az = az_deg * pi / 180.0f;
    const float se = sinf(el), ce = cosf(el), sa = sinf(az), ca = cosf(az);
    const float d[3] = { ce * ca, se, ce * sa };
    v3_normalize_or_zero(d, out);
}

/* Scene: src/scene/mod.rs:16-21,119-122 with Globals::default (src/terrain/mod.rs:188-199).
 * TerrainSpike: src/terrain/mod.rs:322-329,681-691. */
void vfo_default_uniforms(int kind, uint32_t width, uint32_t height, float out[44])
{
    const float eye[3] = { 3.0f, 2.0f, 3.0f }, tgt[3] = { 0, 0, 0 }, up[3] = { 0, 1, 0 };
    float view[16], proj[16], sun[3];
    vfo_look_at_rh(eye, tgt, up, view);
    const float fovy = 45.0f * (3.14159265358979323846f / 180.0f); /* f32::to_radians */
    vfo_perspective_wgpu(fovy, (float)width / (float)height, 0.1f, 100.0f, proj);
    const float s0[3] = { 0.5f, 0.8f, 0.6f }, s1[3] = { 0.5f, 1.0f, 0.3f };
    v3_normalize(kind == 0 ? s0 : s1, sun);
    vfo_uniforms(view, proj, sun, 1.0f, 1.0f, -0.5f, 0.5f, 1.0f, out);
}

/* ------------------------------------------------------------------------------------
 * Colour.  src/colormap/mod.rs:59-79 (CPU linearisation for the Rgba8Unorm LUT fallback). */
void vfo_lut_linearize_u8(const uint8_t *src, size_t n, uint8_t *out)
{
    for (size_t i = 0; i < n; ++i) {
        for (int c = 0; c < 3; ++c) {
            const float s = (float)src[4 * i + c] / 255.0f;
            float lin = s <= 0.04045f ? s / 12.92f : powf((s + 0.055f) / 1.055f, 2.4f);
            if (lin < 0.0f) lin = 0.0f;
            if (lin > 1.0f) lin = 1.0f;
            out[4 * i + c] = (uint8_t)(lin * 255.0f + 0.5f);
        }
        out[4 * i + 3] = src[4 * i + 3];
    }
}

/* Store to Rgba8UnormSrgb (fixed function; src/terrain/mod.rs:219): clamp, sRGB OETF,
 * round-half-up to 8 bits.  Evaluated in double so that the result is the correctly rounded
 * one for every float input (the device path uses exact per-code float thresholds). */
uint8_t vfo_srgb_encode_u8(float linear)
{
    double c = (double)linear;
    if (!(c > 0.0)) c = 0.0;       /* also maps NaN to 0 */
    if (c > 1.0) c = 1.0;
    const double e = c <= 0.0031308 ? 12.92 * c : 1.055 * pow(c, 1.0 / 2.4) - 0.055;
    int q = (int)floor(e * 255.0 + 0.5);
    if (q < 0) q = 0;
    if (q > 255) q = 255;
    return (uint8_t)q;
}

/* Sample from Rgba8UnormSrgb: sRGB EOTF of k/255, double then rounded to float. */
float vfo_srgb8_to_linear(uint8_t v)
{
    const double s = (double)v / 255.0;
    const double l = s <= 0.04045 ? s / 12.92 : pow((s + 0.055) / 1.055, 2.4);
    return (float)l;
}
