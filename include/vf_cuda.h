/* vf_cuda.h -- C ABI of libvfcuda.so: the B200 (sm_100a) replacement for the GPU half of
 * vulkan-forge's headless render path (what the WGSL shaders under src/shaders plus wgpu do in the reference).
 *
 * This is the drop-in boundary.  A Rust host would bind it with `extern "C"` exactly as listed
 * (see INTEGRATION.md for the Rust/ctypes stubs); in this repository the host above it is the
 * C++/pybind11 module `_vulkan_forge` because no Rust toolchain exists on the build machines.
 *
 * Conventions
 *   - every function returns 0 on success, a non-zero VF_E_* code otherwise; the message is
 *     available from vf_last_error() (thread-local, valid until the next call on that thread);
 *   - no exception or panic crosses the boundary;
 *   - the caller owns and sizes every host buffer; inputs are borrowed for the call only;
 *   - handles are not thread-safe; every function synchronises before returning unless its
 *     name ends in _async;
 *   - there is NO CPU fallback: without a usable CUDA device every compute entry point fails
 *     with VF_E_NO_DEVICE.
 * Paths in "replaces:" comments are relative to the reference repository root.
 */
#ifndef VF_CUDA_H
#define VF_CUDA_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VF_OK 0
#define VF_E_NO_DEVICE 1   /* no CUDA device / driver ("No suitable GPU adapter") */
#define VF_E_CUDA 2        /* a CUDA runtime call or kernel failed */
#define VF_E_INVALID 3     /* invalid argument */
#define VF_E_OOM 4         /* device or pinned-host allocation failed */
#define VF_E_IO 5          /* file output failed */

#define VF_MODE_SHIPPED 0   /* bit-faithful restatement of src/shaders/terrain.wgsl */
#define VF_MODE_NORTHSTAR 1 /* DEM forward-difference normals, [h_min,h_max] LUT, Reinhard */

#define VF_LUT_VIRIDIS 0    /* order of src/colormap/mod.rs:7 SUPPORTED */
#define VF_LUT_MAGMA 1
#define VF_LUT_TERRAIN 2

typedef struct vf_scene vf_scene;   /* one terrain renderer: target + DEM + LUT + uniforms */
typedef struct vf_height vf_height; /* a device-resident R32F height image */

typedef struct {
    char name[256];
    int cc_major, cc_minor, sm_count;
    uint64_t total_mem_bytes;
    int pci_device_id, pci_bus_id, l2_bytes, max_smem_per_block;
} vf_device_info_t;

/* per-frame statistics of the last vf_scene_render_* call (last frame of a batch) */
typedef struct {
    uint32_t bin_pairs;      /* (tile, primitive) pairs produced by the binning kernel */
    uint32_t bin_capacity;   /* capacity of the pair buffer (grown automatically) */
    uint32_t clipped_prims;  /* primitives that took the Sutherland-Hodgman path */
    uint32_t tiles;          /* screen tiles (16x16 px) */
    uint32_t big_prims;      /* primitives routed to the binned path (large or clipped) */
    uint32_t tma;            /* 1 if the DEM tiles were staged with TMA (grid == DEM, identity mapping) */
    float ms_geom, ms_bin, ms_raster, ms_total; /* CUDA-event times, only when timing enabled */
} vf_frame_stats_t;

/* timing and traffic of a scene (planned in the reference as render_metrics(): ROADMAP.md:1959-1967) */
typedef struct {
    float upload_ms;     /* last height upload: H2D copy (+ vertex-height resample and bounds), CUDA events on its stream */
    float gpu_ms;        /* kernels of the last render call, all its frames (CUDA events on the render stream) */
    float readback_ms;   /* device-to-host copies of the last render call (CUDA events on the copy stream) */
    float encode_ms;     /* PNG encoding of the last render_png / render_batch_png call (device deflate + host framing) */
    float total_ms;      /* host wall clock of the last render call, submit to completion */
    uint32_t frames;     /* frames of the last render call */
    uint32_t reserved;
    uint64_t last_h2d_bytes, last_d2h_bytes;   /* bytes of the last upload / the last render call */
    uint64_t h2d_bytes, d2h_bytes;             /* cumulative since the scene was created */
    uint64_t vram_bytes;                       /* device memory the scene currently holds */
} vf_render_metrics_t;

/* ---- context ----------------------------------------------------------------------
 * replaces: WgpuContext::get() src/lib.rs:31-60 and the per-object device creation in
 * src/terrain/mod.rs:277-294, src/scene/mod.rs:63-70 */
int vf_init(int device);                       /* device < 0 : keep current / LOCAL_RANK */
const char *vf_last_error(void);
int vf_device_count(int *count);
int vf_device_info(int device, vf_device_info_t *out);   /* src/lib.rs:744-845 diagnostics */
/* pinned host memory for zero-staging readback (numpy arrays are built on top of it) */
int vf_host_alloc(size_t bytes, void **out);
int vf_host_free(void *p);
/* driver / runtime versions (as cudaDriverGetVersion reports them) and the flags libvfcuda.so was compiled with
 * (vulkan_forge.report_environment(), planned in the reference: ROADMAP.md:1989-1999) */
int vf_versions(int *driver, int *runtime);
const char *vf_build_info(void);

/* ---- triangle path ------------------------------------------------------------------
 * replaces: render_into_offscreen + copy_texture_to_rgba_unpadded, src/lib.rs:162-216,685-721,
 * and src/shaders/triangle.wgsl.  out: tight uint8[h][w][4]. */
int vf_triangle_render(int w, int h, uint8_t *out_rgba_host);

/* ---- public grid generator ------------------------------------------------------------
 * replaces: make_grid + grid_generate, src/terrain/mesh.rs:35-90,149-203.
 * xy, uv: float[nx*nz*2]; idx: uint32[6*(nx-1)*(nz-1)] (host buffers). */
int vf_grid_generate(uint32_t nx, uint32_t nz, float dx, float dy,
                     float *xy, float *uv, uint32_t *idx);

/* ---- DEM statistics -------------------------------------------------------------------
 * replaces: terrain_stats::min_max src/terrain_stats.rs:11-35 (clamp=1: exact 1/99 percentile of
 * the strided sample via device radix select), dem_stats_from_slice / normalize_in_place
 * src/lib.rs:905-951.  `data` is a HOST pointer of n floats. */
int vf_dem_minmax(const float *data, size_t n, int clamp, float *lo, float *hi);
int vf_dem_stats(const float *data, size_t n, float out4[4]);        /* min,max,mean,std */
/* mode 0 = minmax to [lo,hi], 1 = zscore; in place on the host array */
int vf_dem_normalize(float *data, size_t n, int mode, float eps, float lo, float hi);
/* f64 -> f32 cast and scale by exaggeration on the device (add_terrain, src/lib.rs:351-388) */
int vf_dem_ingest_f64(const double *src, size_t n, float scale, float *dst);
int vf_dem_ingest_f32(const float *src, size_t n, float scale, float *dst);

/* Allocation-free variant of vf_dem_minmax.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  d_data: n floats in DEVICE memory of the current device; d_out2: two floats (lo, hi) in device memory;
 * the kernels are enqueued on cuda_stream (a cudaStream_t, NULL = the legacy stream) and the call returns without
 * waiting.  Nothing is allocated: the scratch (accumulators + the <= 131 071 strided samples of clamp = 1) is one
 * static device array of 512.25 KiB inside the library, held by every GPU that loads it.  n > 4 294 967 295 is rejected
 * with VF_E_INVALID.  All calls share that array: calls that overlap on the GPU race -- serialise them (one stream). */
int vf_dem_minmax_noalloc_async(const float *d_data, size_t n, int clamp, float *d_out2, void *cuda_stream);

/* Arithmetic self-test (no counterpart in the reference: a diagnostic of this library).  The kernels replace most
 * IEEE divisions of the shading by the shorter instruction sequence the division itself takes for operands of ordinary
 * magnitude (csrc/vf_device.cuh); this runs n_groups groups of random and special operands through those helpers and
 * the plain division on the device and counts the results that differ in any bit.  *mismatches must come back 0. */
int vf_selftest_div(uint64_t n_groups, uint32_t seed, uint64_t *mismatches);
/* The same for the sRGB8 store: the shading's encoder (approximate OETF, then one correction against the exact
 * thresholds) against the binary search over those thresholds, for EVERY float in [0, 1] and 2^24 values outside. */
int vf_selftest_srgb(uint64_t *mismatches);

/* ---- device-resident height images ----------------------------------------------------------
 * replaces: the Vec<f32> that Renderer.add_terrain keeps on the host (src/lib.rs:351-418) and everything computed
 * from it (terrain_stats :424-429, normalize_terrain :478-493, upload_height_r32f :496-571).  The array crosses
 * the bus once (cast / scaled on the device); statistics, normalisation and scenes then work on the device copy. */
int vf_height_ingest_f32(const float *host, uint32_t w, uint32_t h, float scale, vf_height **out);
int vf_height_ingest_f64(const double *host, uint32_t w, uint32_t h, float scale, vf_height **out);
int vf_height_clone(const vf_height *src, vf_height **out);
int vf_height_minmax(const vf_height *t, int clamp, float *lo, float *hi);
int vf_height_stats(const vf_height *t, float out4[4]);              /* min,max,mean,std */
int vf_height_normalize(vf_height *t, int mode, float eps, float lo, float hi);

/* ---- R32F height image round trip -----------------------------------------------------
 * replaces: Renderer.upload_height_r32f / debug_read_height_patch / read_full_height_texture,
 * src/lib.rs:496-681 (256-byte row padding is an artefact of wgpu and is gone). */
int vf_height_upload(const float *host, uint32_t w, uint32_t h, vf_height **out);
int vf_height_size(const vf_height *t, uint32_t *w, uint32_t *h);
int vf_height_read_patch(const vf_height *t, uint32_t x, uint32_t y, uint32_t w, uint32_t h,
                         float *out_host);
void vf_height_destroy(vf_height *t);

/* ---- colormap -------------------------------------------------------------------------
 * replaces: colormap::decode_png_rgba8, src/colormap/mod.rs:50-56 (palette bytes embedded) */
int vf_colormap_rgba8(int lut_id, uint8_t out1024[1024]);

/* ---- terrain scene --------------------------------------------------------------------
 * replaces: Scene / TerrainSpike GPU state and draw, src/scene/mod.rs:60-335,
 * src/terrain/mod.rs:258-493, src/terrain/pipeline.rs:19-143, src/shaders/terrain.wgsl.
 * The render-path grid (src/terrain/mod.rs:553-598) is implicit: no vertex/index buffers. */
int vf_scene_create(int w, int h, uint32_t grid, int lut_id, int lut_unorm, vf_scene **out);
void vf_scene_destroy(vf_scene *s);
/* src/scene/mod.rs:227-276: copies the host array into a device image (pitch multiple of 16 B) */
int vf_scene_set_height_f32(vf_scene *s, const float *host, uint32_t w, uint32_t h);
/* the same from a device-resident image: a device-to-device copy, no PCIe crossing */
int vf_scene_set_height_image(vf_scene *s, const vf_height *t);
/* TerrainUniforms as 44 f32, src/terrain/mod.rs:114-142 (view, proj column-major, ...) */
int vf_scene_set_uniforms(vf_scene *s, const float u44[44]);
int vf_scene_get_uniforms(const vf_scene *s, float u44[44]);
int vf_scene_set_mode(vf_scene *s, int mode, int depth, float h_min, float h_max);
/* 1-99 % range of the scene's DEM on the device (+1e-5 guard, src/renderer.rs:24-28) */
int vf_scene_auto_height_range(vf_scene *s, float *h_min, float *h_max);
/* render one frame with the stored uniforms; out: tight uint8[h][w][4], row 0 = top */
int vf_scene_render_rgba(vf_scene *s, uint8_t *out_rgba_host);
/* n frames, one uniform block each, frames gathered to host (pinned memory recommended) */
int vf_scene_render_batch(vf_scene *s, const float *u44xn, int n, uint8_t *out_rgba_host);
/* Pipelined variants (batches of independent frames): the upload goes to a second
 * device image on its own stream while frames of the current image render; the batch call returns
 * once everything is enqueued.  Host buffers must be pinned and stay untouched until vf_scene_wait
 * (output) / the next vf_scene_wait after the batch that uses them (height).  At most 4 batches in flight.
 * A batch renders with the mode, height range and height image current when it was submitted; if one
 * of its frames needed more queue / pair space than the scene had at that moment, vf_scene_wait grows
 * the buffers and renders the batch again (synchronously, from that same image) before it returns. */
int vf_scene_set_height_f32_async(vf_scene *s, const float *host_pinned, uint32_t w, uint32_t h);
int vf_scene_render_batch_async(vf_scene *s, const float *u44xn, int n, uint8_t *out_rgba_host_pinned, int *ticket);
int vf_scene_wait(vf_scene *s, int ticket);
/* n frames rendered AND PNG-encoded on the device (Sub filter + fixed-Huffman deflate with one-pixel matches, one IDAT
 * chunk per band of rows, CRC-32 and Adler-32 computed on the device): only the finished files cross the bus.  File k
 * occupies out_host[k * cap_per_file ...] for sizes[k] bytes; size the slots with vf_png_file_bound (worst case, about
 * 1.13 x the raw frame).  replaces: render + readback + image::save, src/scene/mod.rs:279-335, for batches. */
size_t vf_png_file_bound(int w, int h);
int vf_scene_render_batch_png(vf_scene *s, const float *u44xn, int n, uint8_t *out_host, size_t cap_per_file, uint32_t *sizes);
/* The same with the NEXT batch's height image travelling meanwhile (DEM-tile jobs: SURVEY 8f-1 + 8f-2 together).  The batch
 * is drawn from the current image (an upload pending at entry becomes current first, as in every render call); then, if
 * next_height_host is not NULL, the upload of that image -- same size as the current one, pinned or pageable host memory that
 * stays untouched until the next render call returns -- starts on the upload stream and overlaps this batch's rendering,
 * encoding and read-back; it becomes current with the next render call.  flags: VF_PNG_KEEP_IMAGE = draw the current
 * image again and leave a pending upload pending (for repeating a call that failed because a file did not fit its slot). */
#define VF_PNG_KEEP_IMAGE 1
int vf_scene_render_batch_png_next(vf_scene *scene, const float *u44xn, int n, uint8_t *out_host, size_t cap_per_file, uint32_t *sizes,
                                   const float *next_height_host, uint32_t next_w, uint32_t next_h, int flags);
/* n frames left in device memory (no D2H); for kernel-only timing */
int vf_scene_render_device(vf_scene *s, const float *u44xn, int n);
/* a frame left on the device by a render call: frame k of a batch is rendered into device slot k % 2, and stays there until
 * another frame takes the slot.  Waits for the scene's streams; out_host: W * H * 4 bytes. */
int vf_scene_read_frame(vf_scene *s, int slot, uint8_t *out_host);
/* visibility of the last frame: per pixel 0 = clear colour, else primitive id + 1 */
int vf_scene_debug_visibility(vf_scene *s, uint32_t *out_host);
int vf_scene_enable_timing(vf_scene *s, int enabled);
int vf_scene_last_stats(const vf_scene *s, vf_frame_stats_t *out);
int vf_scene_metrics(const vf_scene *s, vf_render_metrics_t *out);
/* run on a caller-owned cudaStream_t (e.g. torch's current stream); NULL = private stream */
int vf_scene_set_stream(vf_scene *s, void *cuda_stream);

/* ---- PNG ------------------------------------------------------------------------------
 * replaces: image::RgbaImage::save, src/scene/mod.rs:331-333 (RGBA8, zlib deflate) */
int vf_png_write(const char *path, const uint8_t *rgba, int w, int h);
/* any host RGBA8 image through the device encoder of vf_scene_render_batch_png: the finished file's bytes in out_host */
int vf_png_encode_rgba(const uint8_t *rgba_host, int w, int h, uint8_t *out_host, size_t cap, uint32_t *size);

#ifdef __cplusplus
}
#endif
#endif
