/* vfo.h -- CPU ORACLE for the vulkan-forge headless render path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 *
 * The reference (Rust + wgpu/WGSL) cannot be compiled or run here (no Rust toolchain, no
 * graphics stack), so this is a plain-C restatement.  Every function cites the reference
 * file:line it follows (paths relative to the reference repo root).
 *
 * Pinning status:
 *   - grid, camera, uniform layout, DEM stats, R32F handling: pinned by the reference's own
 *     test vectors (tests/golden/ fixtures, generated from the reference tests' literals).
 *   - percentile (min_max clamp=true): parity unpinned by reference tests (source-only);
 *     index arithmetic pinned by known answers derived from src/terrain_stats.rs:23-33.
 *   - pixels (rasterisation, sampling, sRGB): PARITY UNPINNED -- the reference asserts no
 *     pixel value anywhere and its GPU path (wgpu) is unavailable on this box.  The raster
 *     rules restated here are the D3D11/Vulkan ones (see DESIGN.md "Arithmetic contract").
 */
#ifndef VFO_H
#define VFO_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- grid ------------------------------------------------------------------------- */
/* src/terrain/mesh.rs:35-90 (make_grid) + :149-203 (grid_generate, always u32 to Python) */
void vfo_make_grid(uint32_t w, uint32_t h, float dx, float dy,
                   float *xy, float *uv, uint32_t *idx);
/* src/terrain/mod.rs:553-598 == src/scene/mod.rs:85-116 (render-path grid [x,z,u,v]) */
void vfo_render_grid_xyuv(uint32_t g, float *verts_xzuv, uint32_t *idx);

/* ---- DEM statistics --------------------------------------------------------------- */
/* src/terrain_stats.rs:11-35 */
void vfo_min_max(const float *data, size_t n, int clamp, float *lo, float *hi);
/* src/renderer.rs:24-28 */
void vfo_h_range(const float *data, size_t n, float *h_min, float *h_max);
/* src/lib.rs:905-932 ; out4 = min,max,mean,std */
void vfo_dem_stats(const float *data, size_t n, float out4[4]);
/* src/lib.rs:934-951 ; mode 0 = minmax, 1 = zscore */
void vfo_normalize(float *data, size_t n, int mode, float eps, float lo, float hi,
                   const float stats4[4]);

/* ---- camera (glam 0.24.2 restated; call sites src/camera.rs:133,160,219) ------------- */
/* all matrices column-major float[16]: m[4*c+r] */
void vfo_look_at_rh(const float eye[3], const float target[3], const float up[3], float m[16]);
void vfo_perspective_rh_gl(float fovy_rad, float aspect, float znear, float zfar, float m[16]);
/* src/camera.rs:14-21,218-221: gl_to_wgpu() * P with gl_to_wgpu as WRITTEN (transposed) */
void vfo_perspective_wgpu(float fovy_rad, float aspect, float znear, float zfar, float m[16]);
void vfo_mat4_mul(const float a[16], const float b[16], float out[16]);
/* src/camera.rs:224-240 ; returns 0 ok, else 1..7 = error id (see vfo_host.c) */
int vfo_validate_camera(const float eye[3], const float target[3], const float up[3],
                        float fovy_deg, float znear, float zfar);
/* src/terrain/mod.rs:114-142,203-214 */
void vfo_uniforms(const float view[16], const float proj[16], const float sun[3],
                  float exposure, float spacing, float h_min, float h_max,
                  float exaggeration, float out44[44]);
/* src/lib.rs:444-453 */
void vfo_sun_dir(float elevation_deg, float azimuth_deg, float out3[3]);
/* defaults: kind 0 = Scene (src/terrain/mod.rs:188-199), 1 = TerrainSpike (:325-327,689) */
void vfo_default_uniforms(int kind, uint32_t width, uint32_t height, float out44[44]);

/* ---- colour ----------------------------------------------------------------------- */
/* src/colormap/mod.rs:59-79 */
void vfo_lut_linearize_u8(const uint8_t *srgb_rgba, size_t n_texels, uint8_t *out_rgba);
uint8_t vfo_srgb_encode_u8(float linear);   /* store to Rgba8UnormSrgb */
float vfo_srgb8_to_linear(uint8_t v);       /* sample from Rgba8UnormSrgb */

/* ---- rendering -------------------------------------------------------------------- */
#define VFO_MODE_SHIPPED 0
#define VFO_MODE_NORTHSTAR 1

typedef struct {
    uint32_t width, height;      /* framebuffer */
    uint32_t grid;               /* vertices per side (>= 2) */
    const float *dem;            /* row-major float32[dem_h][dem_w] */
    uint32_t dem_w, dem_h;
    float uniforms[44];          /* TerrainUniforms as 44 f32 */
    const uint8_t *lut_rgba;     /* 256 texels RGBA8, sRGB encoded (PNG bytes) */
    int lut_unorm;               /* 1 = Rgba8Unorm path (CPU-linearised u8 table) */
    int mode;                    /* VFO_MODE_* */
    int depth;                   /* 1 = depth test Less (NORTHSTAR only by convention) */
    float h_min, h_max;          /* NORTHSTAR LUT range */
    int threads;                 /* OpenMP threads; <=0 -> all */
} vfo_frame;

typedef struct {
    uint64_t tris_total, tris_front, tris_sampled, tris_clipped, fragments;
} vfo_counters;

/* src/lib.rs:72-91,685-721 + src/shaders/triangle.wgsl:1-23 */
int vfo_render_triangle(uint32_t w, uint32_t h, uint8_t *out_rgba);
/* src/shaders/terrain.wgsl:39-91 + src/terrain/pipeline.rs:107-140 + src/scene/mod.rs:279-296.
 * out_prim (optional, may be NULL): per pixel 0 = clear colour, else primitive id + 1. */
int vfo_render_terrain(const vfo_frame *f, uint8_t *out_rgba, uint32_t *out_prim,
                       vfo_counters *counters);
int vfo_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
