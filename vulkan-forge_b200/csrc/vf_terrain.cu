// vf_terrain.cu -- vf_scene: host side of the terrain frame (C ABI in include/vf_cuda.h).
// replaces: Scene / TerrainSpike GPU state + draw + readback, src/scene/mod.rs:60-335,
// src/terrain/mod.rs:258-493, src/terrain/pipeline.rs:19-143.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <vector>

#include "vf_internal.h"
#include "vf_geom_march.cuh"
#include "vf_raster_lean.cuh"
#include "vf_png_dev.h"

// cuTensorMapEncodeTiled through the runtime's driver entry point (libcuda is not linked)
typedef CUresult (*vf_tmap_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                      const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                      CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static vf_tmap_encode_fn tmap_encoder() {
    static vf_tmap_encode_fn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult st;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &st) != cudaSuccess ||
            st != cudaDriverEntryPointSuccess) { (void)cudaGetLastError(); p = nullptr; }
        return (vf_tmap_encode_fn)p;
    }();
    return fn;
}

// launch configuration that depends on the device (SM count, opt-in shared memory, cooperative grid size):
// cached per device ordinal, not per process -- vf_init(device) may switch devices within one process
struct VfDevCache { int sm_count = 0; bool gm_attr = false; int binsort_ctas = 0; };
static VfDevCache g_dev_cache[64];

struct vf_scene {
    int device = 0;                      // CUDA device the scene's buffers, streams and events live on
    int W = 0, H = 0;
    uint32_t G = 0;
    int lut_id = 0, lut_unorm = 0;
    float u44[44];
    int mode = VF_MODE_SHIPPED, depth = 0;
    float h_min = -0.5f, h_max = 0.5f;
    // device state
    float* d_dem = nullptr; uint32_t Wd = 0, Hd = 0, pitch = 0;   // pitch in floats (mult. of 4)
    VfAxis* d_axis = nullptr;
    float4* d_lut = nullptr;
    uint8_t* d_frame = nullptr;          // n_slots frames of W*H*4 bytes
    int n_slots = 0;
    uint32_t* d_vis = nullptr;           // debug visibility (allocated on demand)
    uint32_t tiles_x = 0, tiles_y = 0, n_tiles = 0;
    uint32_t* d_tile_count = nullptr;    // [n_tiles] counts | [VF_CNT_WORDS] counters (one memset)
    uint32_t* d_counters = nullptr;      // = d_tile_count + n_tiles
    uint32_t* d_tile_off = nullptr;      // [n_tiles + 1] offsets | [2] high-water marks (pairs, queued prims)
    uint32_t* d_maxpairs = nullptr;      // = d_tile_off + n_tiles + 1
    uint2* d_segtab = nullptr;           // k_geom_march row segments (first row, rows), rebuilt when G changes
    uint32_t segtab_cap = 0, segtab_n = 0, segtab_G = 0, segtab_workers = 0, segtab_rows = 0, segtab_nfull = 0;
    unsigned long long* d_visbuf = nullptr;  // W*H visibility keys (L2-resident, self-cleaning)
    uint2* d_shade_list = nullptr;           // W*H entries: covered pixels of the frame (vf_raster_lean.cuh); nullptr: k_raster always
    bool counters_clean = false;             // the last frame took the list-shading pass, whose kernels leave the counter block zeroed
    bool early_geom_ok = true;               // k_geom_march as a programmatic dependent of the previous frame's k_shade_list (VF_NO_PDL at creation: off)
    uint32_t shade_parity = 0;               // which of the two list-length counters the next list-shading frame uses
    int visbuf_depth = -1;               // which 'empty' value the buffer currently holds
    uint32_t* d_big = nullptr; uint32_t big_capacity = 0;   // queue of large / clipped primitives
    alignas(64) CUtensorMap dem_map;     // 2-D tiled map over the DEM (box 36x33 floats, k_geom)
    alignas(64) CUtensorMap dem_map_gm[2];  // same tensor, one box per warp task of k_geom_march: [0] 36x31 floats, [1] 36x27 (depth)
    // Height of every grid vertex as one G x G image, h_grid[j][i] = dem[ty[j]][tx[i]], built once per upload when the
    // grid -> texel mapping is not the identity (grid != DEM size; or f32 rounding of u * W near the upper end of large
    // grids: 7 indices at G = 16384).  The geometry kernels then stage their texels with TMA boxes whatever the sizes;
    // nullptr = identity, the DEM itself is that image.  The fragment stage always reads the DEM proper.
    float* d_hgrid = nullptr; float* d_hgrid_back = nullptr; uint32_t hpitch = 0;
    // (min, max) of every 32 x 32 block of vertex heights of that image (k_hminmax): k_geom_march bounds a task's
    // heights with it and skips the per-vertex clip tests where the task's box is provably inside the guard band
    float2* d_hmm = nullptr; float2* d_hmm_back = nullptr; uint32_t hmm_nb = 0;
    bool use_tma = false;                // texels of the geometry pass are staged with TMA
    bool dem_identity = false;           // grid == DEM and every grid index fetches the texel of the same index
    bool geom_tiled = false;             // VF_GEOM_TILED at creation: tiled k_geom instead of k_geom_march (A/B runs)
    // back DEM image for vf_scene_set_height_f32_async: uploaded on upload_stream while frames of the
    // front image render; swapped in by the next submit
    float* d_dem_back = nullptr;
    alignas(64) CUtensorMap dem_map_b, dem_map_gm_b[2];
    bool back_valid = false, back_pending = false, front_released_once = false;
    cudaStream_t upload_stream = nullptr;
    cudaEvent_t ev_uploaded = nullptr, ev_front_released = nullptr;
    // batches in flight (vf_scene_render_batch_async): completion event on the copy stream, the
    // high-water marks of the batch in a pinned slot, and what is needed to redo it after growing a buffer
    struct Ticket {
        cudaEvent_t done = nullptr; bool busy = false; std::vector<float> u; int n = 0; uint8_t* out = nullptr;
        // state of the scene when the batch was submitted: an overflow is judged against THESE capacities
        // (a later batch may have grown the buffers since) and the batch is redone from THIS height image
        uint32_t cap_pairs = 0, cap_big = 0; const float* dem = nullptr; bool skipped_bins = false;
        int mode = 0, depth = 0; float h_min = 0.0f, h_max = 0.0f;
        bool resolved = false; int rc = 0;
    };
    Ticket tickets[4];
    // Speculative binning: frames of a terrain seen whole queue nothing for the sort-middle stage, and the
    // launch of k_binsort alone costs 6 us (3 % of a 1080p frame), so it is left out until a frame does
    // queue something (k_raster reports it in high_water[2]); that frame / batch is rendered again with
    // the stage in place, which then stays for bin_hold frames (doubling each time it proves necessary).
    uint32_t bin_hold = 0, bin_hold_len = 32;
    bool submit_skipped_bins = false;
    bool slot_used[2] = {false, false};
    uint32_t* d_flag = nullptr;
    uint32_t* d_tile_cursor = nullptr;   // [n_tiles]
    uint32_t* d_pair_tile = nullptr; uint32_t* d_pair_prim = nullptr; uint32_t* d_sorted = nullptr;
    uint32_t capacity = 0;
    uint32_t* h_counters = nullptr;      // pinned: 2*VF_CNT_WORDS words per batch slot (counters, high-water marks)
    uint32_t h_counters_last[VF_CNT_WORDS] = {0};
    cudaStream_t stream = nullptr; bool own_stream = false;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev_done[2] = {nullptr, nullptr}, ev_copied[2] = {nullptr, nullptr};
    // PNG output encoded on the device (vf_png_dev.cu): created on first use; VF_PNG_SLOTS finished files can wait
    // in HBM for their (size-exact) copies while later frames render
    VfPngDev* png = nullptr;
    uint32_t* h_png_size = nullptr;                       // pinned, one per file slot
    cudaEvent_t ev_png_done[4] = {nullptr, nullptr, nullptr, nullptr}, ev_png_copied[4] = {nullptr, nullptr, nullptr, nullptr};
    bool png_slot_used[4] = {false, false, false, false};
    float encode_ms = 0.0f;
    bool timing = false;
    vf_frame_stats_t stats;
    // render_metrics(): one set of timing events per synchronous call / per ticket, byte counters, wall clock
    struct Marks { cudaEvent_t g0 = nullptr, g1 = nullptr, r0 = nullptr, r1 = nullptr; bool copies = false; int frames = 0; uint64_t d2h = 0;
                   std::chrono::steady_clock::time_point t0; };
    Marks marks, ticket_marks[4];
    cudaEvent_t ev_up0 = nullptr, ev_up1 = nullptr; cudaStream_t up_stream_of_events = nullptr; bool upload_timed = false;
    uint64_t h2d_bytes = 0, d2h_bytes = 0, last_h2d_bytes = 0;
    vf_render_metrics_t metrics;
};

static const int VF_MAX_SLOTS = 2;

static int marks_init(vf_scene::Marks& m) {
    if (m.g0) return VF_OK;
    VF_CUDA(cudaEventCreate(&m.g0)); VF_CUDA(cudaEventCreate(&m.g1));
    VF_CUDA(cudaEventCreate(&m.r0)); VF_CUDA(cudaEventCreate(&m.r1));
    return VF_OK;
}
static void marks_free(vf_scene::Marks& m) {
    for (cudaEvent_t e : {m.g0, m.g1, m.r0, m.r1}) if (e) cudaEventDestroy(e);
    m.g0 = m.g1 = m.r0 = m.r1 = nullptr;
}
// device memory the scene holds right now
static uint64_t scene_vram_bytes(const vf_scene* s) {
    const uint64_t img = (uint64_t)s->pitch * s->Hd * 4, hg = (uint64_t)s->hpitch * s->G * 4, hm = (uint64_t)s->hmm_nb * s->hmm_nb * 8;
    uint64_t b = 0;
    b += s->d_dem ? img : 0; b += s->d_dem_back ? img : 0;
    b += s->d_hgrid ? hg : 0; b += s->d_hgrid_back ? hg : 0;
    b += s->d_hmm ? hm : 0; b += s->d_hmm_back ? hm : 0;
    b += (uint64_t)s->G * sizeof(VfAxis) + 256 * sizeof(float4);
    b += (uint64_t)s->n_slots * s->W * s->H * 4;                       // frame slots
    b += (uint64_t)s->W * s->H * 8;                                    // visibility buffer
    b += s->d_shade_list ? (uint64_t)s->W * s->H * 8 : 0;              // covered-pixel list
    b += s->d_vis ? (uint64_t)s->W * s->H * 4 : 0;
    b += ((uint64_t)s->n_tiles + VF_CNT_WORDS) * 4 + ((uint64_t)s->n_tiles + 5) * 4 + (uint64_t)s->n_tiles * 4;
    b += 3ull * s->capacity * 4 + (uint64_t)s->big_capacity * 4 + (uint64_t)s->segtab_cap * 8;
    b += vf_png_dev_bytes(s->png);
    return b;
}

static void scene_free(vf_scene* s) {
    if (!s) return;
    cudaFree(s->d_dem); cudaFree(s->d_axis); cudaFree(s->d_lut); cudaFree(s->d_frame);
    cudaFree(s->d_vis); cudaFree(s->d_tile_count); cudaFree(s->d_tile_off);
    cudaFree(s->d_tile_cursor); cudaFree(s->d_pair_tile); cudaFree(s->d_pair_prim);
    cudaFree(s->d_sorted); cudaFree(s->d_visbuf); cudaFree(s->d_shade_list); cudaFree(s->d_big); cudaFree(s->d_flag); cudaFree(s->d_segtab);
    cudaFree(s->d_dem_back); cudaFree(s->d_hgrid); cudaFree(s->d_hgrid_back); cudaFree(s->d_hmm); cudaFree(s->d_hmm_back);
    if (s->upload_stream) cudaStreamDestroy(s->upload_stream);
    if (s->ev_uploaded) cudaEventDestroy(s->ev_uploaded);
    if (s->ev_front_released) cudaEventDestroy(s->ev_front_released);
    for (auto& t : s->tickets) if (t.done) cudaEventDestroy(t.done);
    marks_free(s->marks);
    for (auto& m : s->ticket_marks) marks_free(m);
    vf_png_dev_destroy(s->png);
    if (s->h_png_size) cudaFreeHost(s->h_png_size);
    for (auto& e : s->ev_png_done) if (e) cudaEventDestroy(e);
    for (auto& e : s->ev_png_copied) if (e) cudaEventDestroy(e);
    if (s->ev_up0) cudaEventDestroy(s->ev_up0);
    if (s->ev_up1) cudaEventDestroy(s->ev_up1);
    if (s->h_counters) cudaFreeHost(s->h_counters);
    for (auto& e : s->ev) if (e) cudaEventDestroy(e);
    for (auto& e : s->ev_done) if (e) cudaEventDestroy(e);
    for (auto& e : s->ev_copied) if (e) cudaEventDestroy(e);
    if (s->own_stream && s->stream) cudaStreamDestroy(s->stream);
    if (s->copy_stream) cudaStreamDestroy(s->copy_stream);
    (void)cudaGetLastError();
    delete s;
}

static int scene_alloc_pairs(vf_scene* s, uint32_t capacity) {
    cudaFree(s->d_pair_tile); cudaFree(s->d_pair_prim); cudaFree(s->d_sorted);
    s->d_pair_tile = s->d_pair_prim = s->d_sorted = nullptr;
    VF_CUDA(cudaMalloc(&s->d_pair_tile, (size_t)capacity * 4));
    VF_CUDA(cudaMalloc(&s->d_pair_prim, (size_t)capacity * 4));
    VF_CUDA(cudaMalloc(&s->d_sorted, (size_t)capacity * 4));
    s->capacity = capacity;
    return VF_OK;
}

static int scene_alloc_big(vf_scene* s, uint32_t capacity) {
    cudaFree(s->d_big);
    s->d_big = nullptr;
    VF_CUDA(cudaMalloc(&s->d_big, (size_t)capacity * 4));
    s->big_capacity = capacity;
    return VF_OK;
}

// LUT as the fragment stage sees it: linear float per texel.  Rgba8UnormSrgb: exact sRGB EOTF
// of k/255 (what the sampler does); Rgba8Unorm: the lossy CPU table of
// src/colormap/mod.rs:59-79 divided by 255.
static void build_lut(const uint8_t* pal, int unorm, float4* out) {
    for (int i = 0; i < 256; ++i) {
        float c[3];
        for (int k = 0; k < 3; ++k) {
            const uint8_t v = pal[4 * i + k];
            if (unorm) {
                const float sr = (float)v / 255.0f;
                float lin = sr <= 0.04045f ? sr / 12.92f : powf((sr + 0.055f) / 1.055f, 2.4f);
                lin = std::min(std::max(lin, 0.0f), 1.0f);
                const uint8_t q = (uint8_t)(lin * 255.0f + 0.5f);
                c[k] = (float)q / 255.0f;
            } else {
                const double sd = (double)v / 255.0;
                c[k] = (float)(sd <= 0.04045 ? sd / 12.92 : pow((sd + 0.055) / 1.055, 2.4));
            }
        }
        out[i] = make_float4(c[0], c[1], c[2], 1.0f);
    }
}

static int scene_rebuild_axis(vf_scene* s) {
    const float gm1 = (float)s->G - 1.0f;
    const float step = (2.0f * 1.5f) / gm1;
    k_axis<<<(s->G + 255) / 256, 256, 0, s->stream>>>(s->d_axis, s->G, step, gm1, s->Wd, s->Hd);
    VF_CUDA(cudaGetLastError());
    return VF_OK;
}

// the two 2-D tiled maps over a DEM image: boxes of k_geom (36x33) and of k_geom_march (36x25)
static bool scene_encode_maps(float* d_dem, uint32_t w, uint32_t h, uint32_t pitch, CUtensorMap* m_tiled, CUtensorMap* m_march) {
    if (!tmap_encoder()) return false;
    const cuuint64_t gdim[2] = {w, h};
    const cuuint64_t gstride[1] = {(cuuint64_t)pitch * 4};
    const cuuint32_t estr[2] = {1, 1};
    const cuuint32_t boxes[3][2] = {{VF_TMA_BOX_X, VF_TMA_BOX_Y}, {VF_GM_BOX_X, VfGmDim<false>::VROWS}, {VF_GM_BOX_X, VfGmDim<true>::VROWS}};
    CUtensorMap* maps[3] = {m_tiled, m_march, m_march + 1};
    bool ok = true;
    for (int k = 0; k < 3; ++k)
        ok = ok && tmap_encoder()(maps[k], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, d_dem, gdim, gstride, boxes[k], estr,
                                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
    return ok;
}

static void scene_resolve_readers(vf_scene* s, const float* image);

// h_grid[j][i] = dem[ty[j]][tx[i]] (nearest / clamp sampler of src/scene/mod.rs:263-268 applied once per upload)
__global__ void __launch_bounds__(256)
k_resample(const VfAxis* __restrict__ axis, const float* __restrict__ dem, uint32_t dem_pitch, uint32_t G, uint32_t hpitch,
           float* __restrict__ hgrid) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= G) return;
    const int tx = axis[i].tx;
    for (uint32_t j = blockIdx.y; j < G; j += gridDim.y)
        hgrid[(size_t)j * hpitch + i] = __ldg(dem + (size_t)axis[j].ty * dem_pitch + tx);
}
static void launch_resample(const vf_scene* s, const float* dem, float* hgrid, cudaStream_t st);

// (min, max) of one 32 x 32 block of the G x G vertex-height image per CTA; a block holding a NaN or an infinity
// gets (-inf, +inf), which no task passes
__global__ void __launch_bounds__(256)
k_hminmax(const float* __restrict__ img, uint32_t pitch, uint32_t G, uint32_t nb, float2* __restrict__ out) {
    __shared__ float s_lo[8], s_hi[8];
    const uint32_t bi = blockIdx.x, bj = blockIdx.y, tid = threadIdx.x;
    float lo = __int_as_float(0x7f800000), hi = __int_as_float(0xff800000);
    bool bad = false;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t i = bi * 32u + (tid & 31u), j = bj * 32u + (tid >> 5) + 8u * k;
        if (i < G && j < G) {
            const float v = __ldg(img + (size_t)j * pitch + i);
            bad = bad || !isfinite(v);
            lo = fminf(lo, v); hi = fmaxf(hi, v);
        }
    }
    if (bad) { lo = __int_as_float(0xff800000); hi = __int_as_float(0x7f800000); }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, d));
        hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, d));
    }
    if ((tid & 31u) == 0u) { s_lo[tid >> 5] = lo; s_hi[tid >> 5] = hi; }
    __syncthreads();
    if (tid == 0) {
        for (int k = 1; k < 8; ++k) { lo = fminf(lo, s_lo[k]); hi = fmaxf(hi, s_hi[k]); }
        out[(size_t)bj * nb + bi] = make_float2(lo, hi);
    }
}
// the image the geometry pass stages its texels from (vertex heights): the resampled one, else the DEM itself
static void launch_hminmax(const vf_scene* s, const float* dem, const float* hgrid, float2* hmm, cudaStream_t st) {
    k_hminmax<<<dim3(s->hmm_nb, s->hmm_nb), 256, 0, st>>>(hgrid ? hgrid : dem, hgrid ? s->hpitch : s->pitch, s->G, s->hmm_nb, hmm);
}

// every entry point works on the scene's own device, whatever vf_init selected since its creation
static int scene_bind(const vf_scene* s) {
    if (!s) return vf_set_error(VF_E_INVALID, "null scene");
    VF_CUDA(cudaSetDevice(s->device));
    return VF_OK;
}

// `src`: w x h floats, rows `src_pitch` bytes apart, in host memory (kind = cudaMemcpyHostToDevice) or on the scene's
// device (cudaMemcpyDeviceToDevice: a vf_height image, no PCIe crossing)
static int scene_upload_dem(vf_scene* s, const float* host, uint32_t w, uint32_t h, size_t src_pitch = 0,
                            cudaMemcpyKind kind = cudaMemcpyHostToDevice) {
    if (!src_pitch) src_pitch = (size_t)w * 4;
    if (w == 0 || h == 0) return vf_set_error(VF_E_INVALID, "height image cannot be empty");
    const uint32_t pitch = (w + 3u) & ~3u;            // rows start 16-byte aligned (TMA-ready)
    if (int rcb = scene_bind(s)) return rcb;
    scene_resolve_readers(s, nullptr);                // batches in flight read the images replaced below
    if (s->upload_stream) VF_CUDA(cudaStreamSynchronize(s->upload_stream));
    s->back_pending = false;
    if (pitch != s->pitch || h != s->Hd || !s->d_dem) {
        VF_CUDA(cudaStreamSynchronize(s->stream));
        cudaFree(s->d_dem); cudaFree(s->d_dem_back);
        s->d_dem = nullptr; s->d_dem_back = nullptr; s->back_valid = false;
        VF_CUDA(cudaMalloc(&s->d_dem, (size_t)pitch * h * 4));
    }
    s->Wd = w; s->Hd = h; s->pitch = pitch;
    if (!s->ev_up0) { VF_CUDA(cudaEventCreate(&s->ev_up0)); VF_CUDA(cudaEventCreate(&s->ev_up1)); }
    VF_CUDA(cudaEventRecord(s->ev_up0, s->stream));
    VF_CUDA(cudaMemcpy2DAsync(s->d_dem, (size_t)pitch * 4, host, src_pitch, (size_t)w * 4, h, kind, s->stream));
    s->last_h2d_bytes = kind == cudaMemcpyHostToDevice ? (uint64_t)w * h * 4 : 0;
    s->h2d_bytes += s->last_h2d_bytes;
    int rc = scene_rebuild_axis(s);
    if (rc) return rc;
    // TMA staging of the geometry pass: over the DEM itself when every grid index fetches the texel of the same
    // index (checked on the device: f32 rounding of u*W can break it near the upper end at large sizes, SURVEY
    // 7.3-5), otherwise over the G x G image of vertex heights resampled here once
    s->use_tma = false;
    s->dem_identity = false;
    memset(&s->dem_map, 0, sizeof s->dem_map);
    memset(s->dem_map_gm, 0, sizeof s->dem_map_gm);
    const uint64_t hbytes = (uint64_t)((s->G + 3u) & ~3u) * s->G * 4u;
    if (tmap_encoder() && !getenv("VF_DISABLE_TMA") && hbytes <= (8ull << 30)) {
        uint32_t flag = 0u;
        if (s->G == w && s->G == h) {
            if (!s->d_flag) VF_CUDA(cudaMalloc(&s->d_flag, 4));
            const uint32_t one = 1u;
            VF_CUDA(cudaMemcpyAsync(s->d_flag, &one, 4, cudaMemcpyHostToDevice, s->stream));
            k_axis_identity<<<(s->G + 255) / 256, 256, 0, s->stream>>>(s->d_axis, s->G, s->d_flag);
            VF_CUDA(cudaMemcpyAsync(&flag, s->d_flag, 4, cudaMemcpyDeviceToHost, s->stream));
            VF_CUDA(cudaStreamSynchronize(s->stream));
        }
        s->dem_identity = flag == 1u;
        if (flag == 1u) {
            cudaFree(s->d_hgrid); cudaFree(s->d_hgrid_back);
            s->d_hgrid = s->d_hgrid_back = nullptr; s->hpitch = 0;
            s->use_tma = scene_encode_maps(s->d_dem, w, h, pitch, &s->dem_map, s->dem_map_gm);
        } else {
            const uint32_t hp = (s->G + 3u) & ~3u;
            if (!s->d_hgrid || s->hpitch != hp) {
                cudaFree(s->d_hgrid); cudaFree(s->d_hgrid_back);
                s->d_hgrid = s->d_hgrid_back = nullptr; s->back_valid = false;
                if (cudaMalloc(&s->d_hgrid, hbytes) != cudaSuccess) { (void)cudaGetLastError(); s->d_hgrid = nullptr; }
                s->hpitch = hp;
            }
            if (s->d_hgrid) {
                launch_resample(s, s->d_dem, s->d_hgrid, s->stream);
                VF_CUDA(cudaGetLastError());
                s->use_tma = scene_encode_maps(s->d_hgrid, s->G, s->G, hp, &s->dem_map, s->dem_map_gm);
            }
        }
    } else {
        cudaFree(s->d_hgrid); cudaFree(s->d_hgrid_back);
        s->d_hgrid = s->d_hgrid_back = nullptr; s->hpitch = 0;
    }
    if (s->use_tma) {
        const uint32_t nb = (s->G + 31u) / 32u;
        if (!s->d_hmm || s->hmm_nb != nb) {
            cudaFree(s->d_hmm); cudaFree(s->d_hmm_back);
            s->d_hmm = s->d_hmm_back = nullptr; s->hmm_nb = nb;
            VF_CUDA(cudaMalloc(&s->d_hmm, (size_t)nb * nb * sizeof(float2)));
        }
        launch_hminmax(s, s->d_dem, s->d_hgrid, s->d_hmm, s->stream);
        VF_CUDA(cudaGetLastError());
    } else {
        cudaFree(s->d_hmm); cudaFree(s->d_hmm_back);
        s->d_hmm = s->d_hmm_back = nullptr; s->hmm_nb = 0;
    }
    VF_CUDA(cudaEventRecord(s->ev_up1, s->stream));
    VF_CUDA(cudaStreamSynchronize(s->stream));
    s->upload_timed = true;
    return VF_OK;
}

static void launch_resample(const vf_scene* s, const float* dem, float* hgrid, cudaStream_t st) {
    const uint32_t gy = std::min<uint32_t>(s->G, 4096u);
    k_resample<<<dim3((s->G + 255) / 256, gy), 256, 0, st>>>(s->d_axis, dem, s->pitch, s->G, s->hpitch, hgrid);
}

extern "C" int vf_scene_create(int w, int h, uint32_t grid, int lut_id, int lut_unorm, vf_scene** out) {
    if (!out) return vf_set_error(VF_E_INVALID, "out is null");
    *out = nullptr;
    if (w <= 0 || h <= 0 || w > 8192 || h > 8192)   // 8192: python/vulkan_forge/_validate.py:6; keeps |X| < 2^24
        return vf_set_error(VF_E_INVALID, "width/height must be in 1..8192");
    if (grid < 2) grid = 2;                                    // src/scene/mod.rs:61
    if (grid > 32768) return vf_set_error(VF_E_INVALID, "grid must be <= 32768");
    const uint8_t* pal = vf_palette_bytes(lut_id);
    if (!pal) return vf_set_error(VF_E_INVALID, "unknown colormap id %d", lut_id);
    int rc = vf_require_device();
    if (rc) return rc;
    vf_scene* s = new vf_scene();
    memset(&s->stats, 0, sizeof s->stats);
    memset(&s->metrics, 0, sizeof s->metrics);
    memset(s->u44, 0, sizeof s->u44);
    s->W = w; s->H = h; s->G = grid; s->lut_id = lut_id; s->lut_unorm = lut_unorm;
    s->tiles_x = (uint32_t)(w + VF_TILE - 1) / VF_TILE;
    s->tiles_y = (uint32_t)(h + VF_TILE - 1) / VF_TILE;
    s->n_tiles = s->tiles_x * s->tiles_y;
    s->n_slots = VF_MAX_SLOTS;
    s->geom_tiled = getenv("VF_GEOM_TILED") != nullptr;
    if (cudaGetDevice(&s->device) != cudaSuccess) s->device = 0;
#define SC(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { scene_free(s); \
        return vf_set_error(e_ == cudaErrorMemoryAllocation ? VF_E_OOM : VF_E_CUDA, \
        "%s failed: %s", #call, cudaGetErrorString(e_)); } } while (0)
    SC(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    s->own_stream = true;
    SC(cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking));
    for (auto& e : s->ev) SC(cudaEventCreate(&e));
    for (auto& e : s->ev_done) SC(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto& e : s->ev_copied) SC(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    SC(cudaMalloc(&s->d_axis, (size_t)grid * sizeof(VfAxis)));
    SC(cudaMalloc(&s->d_lut, 256 * sizeof(float4)));
    SC(cudaMalloc(&s->d_frame, (size_t)s->n_slots * w * h * 4));
    SC(cudaMalloc(&s->d_tile_count, ((size_t)s->n_tiles + VF_CNT_WORDS) * 4));
    s->d_counters = s->d_tile_count + s->n_tiles;
    SC(cudaMalloc(&s->d_tile_off, ((size_t)s->n_tiles + 5) * 4));   // + 4 high-water marks
    s->d_maxpairs = s->d_tile_off + s->n_tiles + 1;
    SC(cudaMalloc(&s->d_tile_cursor, (size_t)s->n_tiles * 4));
    SC(cudaHostAlloc(&s->h_counters, 64 * VF_CNT_WORDS * 4, cudaHostAllocDefault));
#undef SC
    float4 lut[256];
    build_lut(pal, lut_unorm, lut);
    cudaError_t e = cudaMemcpy(s->d_lut, lut, sizeof lut, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { scene_free(s); return vf_set_error(VF_E_CUDA, "LUT upload: %s", cudaGetErrorString(e)); }
    // VF_TEST_SMALL_BUFFERS: tests start with tiny queues so that every close-up takes the grow-and-retry path
    const bool tiny = getenv("VF_TEST_SMALL_BUFFERS") != nullptr;
    const uint64_t cap = tiny ? 256 : std::max<uint64_t>(1u << 20, 4ull * (uint64_t)w * (uint64_t)h);
    rc = scene_alloc_pairs(s, (uint32_t)std::min<uint64_t>(cap, 1u << 30));
    if (rc) { scene_free(s); return rc; }
    rc = scene_alloc_big(s, tiny ? 4u : 1u << 20);
    if (rc) { scene_free(s); return rc; }
    if (cudaMalloc(&s->d_visbuf, (size_t)w * h * 8) != cudaSuccess) {
        scene_free(s);
        (void)cudaGetLastError();
        return vf_set_error(VF_E_OOM, "visibility buffer allocation failed");
    }
    s->early_geom_ok = getenv("VF_NO_PDL") == nullptr;
    if ((w & 3) == 0 && !getenv("VF_RASTER_FULL") && cudaMalloc(&s->d_shade_list, (size_t)w * h * sizeof(uint2)) != cudaSuccess) {
        scene_free(s);
        (void)cudaGetLastError();
        return vf_set_error(VF_E_OOM, "covered-pixel list allocation failed");
    }
    const float zero = 0.0f;                                   // 1x1 zero image until set_height
    rc = scene_upload_dem(s, &zero, 1, 1);
    if (rc) { scene_free(s); return rc; }
    *out = s;
    return VF_OK;
}

extern "C" void vf_scene_destroy(vf_scene* s) {
    if (!s) return;
    cudaSetDevice(s->device);
    cudaStreamSynchronize(s->stream);
    if (s->copy_stream) cudaStreamSynchronize(s->copy_stream);       // batches / uploads still in flight
    if (s->upload_stream) cudaStreamSynchronize(s->upload_stream);
    scene_free(s);
}

extern "C" int vf_scene_set_height_f32(vf_scene* s, const float* host, uint32_t w, uint32_t h) {
    if (!s || !host) return vf_set_error(VF_E_INVALID, "null argument");
    return scene_upload_dem(s, host, w, h);
}

extern "C" int vf_scene_set_height_image(vf_scene* s, const vf_height* t) {
    if (!s || !t) return vf_set_error(VF_E_INVALID, "null argument");
    if (t->device != s->device) return vf_set_error(VF_E_INVALID, "height image lives on device %d, the scene on device %d", t->device, s->device);
    return scene_upload_dem(s, t->d, t->w, t->h, t->pitch, cudaMemcpyDeviceToDevice);
}

extern "C" int vf_scene_set_uniforms(vf_scene* s, const float u44[44]) {
    if (!s || !u44) return vf_set_error(VF_E_INVALID, "null argument");
    memcpy(s->u44, u44, sizeof s->u44);
    return VF_OK;
}

extern "C" int vf_scene_get_uniforms(const vf_scene* s, float u44[44]) {
    if (!s || !u44) return vf_set_error(VF_E_INVALID, "null argument");
    memcpy(u44, s->u44, sizeof s->u44);
    return VF_OK;
}

extern "C" int vf_scene_set_mode(vf_scene* s, int mode, int depth, float h_min, float h_max) {
    if (!s) return vf_set_error(VF_E_INVALID, "null scene");
    if (mode != VF_MODE_SHIPPED && mode != VF_MODE_NORTHSTAR)
        return vf_set_error(VF_E_INVALID, "mode must be VF_MODE_SHIPPED or VF_MODE_NORTHSTAR");
    if (!isfinite(h_min) || !isfinite(h_max) || !(h_min < h_max))
        return vf_set_error(VF_E_INVALID, "min must be < max");
    s->mode = mode; s->depth = depth ? 1 : 0; s->h_min = h_min; s->h_max = h_max;
    return VF_OK;
}

extern "C" int vf_scene_enable_timing(vf_scene* s, int enabled) {
    if (!s) return vf_set_error(VF_E_INVALID, "null scene");
    s->timing = enabled != 0;
    return VF_OK;
}

extern "C" int vf_scene_last_stats(const vf_scene* s, vf_frame_stats_t* out) {
    if (!s || !out) return vf_set_error(VF_E_INVALID, "null argument");
    *out = s->stats;
    return VF_OK;
}

extern "C" int vf_scene_set_stream(vf_scene* s, void* cuda_stream) {
    if (int rcb = scene_bind(s)) return rcb;
    VF_CUDA(cudaStreamSynchronize(s->stream));
    if (s->own_stream) { cudaStreamDestroy(s->stream); s->own_stream = false; }
    if (cuda_stream) s->stream = (cudaStream_t)cuda_stream;
    else { VF_CUDA(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking)); s->own_stream = true; }
    return VF_OK;
}

// ---------------------------------------------------------------------------------------
// frame constants from the 44-float uniform block (src/terrain/mod.rs:114-142)
static void build_frame(const vf_scene* s, const float* u, VfFrame* f) {
    memcpy(f->V, u, 64);
    memcpy(f->P, u + 16, 64);
    // normalize(sun) with the contract's operation order (terrain.wgsl:83)
    const float sx = u[32], sy = u[33], sz = u[34];
    const float d = fmaf(sz, sz, fmaf(sx, sx, sy * sy));
    const float r = 1.0f / sqrtf(d);
    f->sunL[0] = sx * r; f->sunL[1] = sy * r; f->sunL[2] = sz * r;
    f->exposure = u[35];
    f->spacing = fmaxf(u[36], 1e-8f);                    // terrain.wgsl:46
    f->h_range = fmaxf(u[37], 1e-8f);                    // terrain.wgsl:71
    f->exag = u[38];
    f->h_min = s->h_min; f->h_max = s->h_max;
    f->gm1 = (float)s->G - 1.0f;
    f->step = (2.0f * 1.5f) / f->gm1;
    f->hw = 0.5f * (float)s->W; f->hh = 0.5f * (float)s->H;
    f->G = s->G; f->W = (uint32_t)s->W; f->H = (uint32_t)s->H;
    f->Wd = s->Wd; f->Hd = s->Hd; f->dem_pitch = s->pitch;
    f->tiles_x = s->tiles_x; f->tiles_y = s->tiles_y;
    f->mode = s->mode; f->depth = s->depth;
    const float* P = f->P; const float* V = f->V;
    f->gm_inv = s->G > 2 ? (uint32_t)((1ull << 32) / (s->G - 1)) : 0xffffffffu;
    f->sparse = (P[1] == 0 && P[2] == 0 && P[3] == 0 && P[4] == 0 && P[6] == 0 && P[7] == 0 && P[8] == 0 &&
                 P[9] == 0 && P[12] == 0 && P[13] == 0 && V[3] == 0 && V[7] == 0 && V[11] == 0 && V[15] == 1.0f) ? 1 : 0;
}

static uint32_t clear_rgba8() {
    const float* thr = vf_srgb_thresholds_host();
    auto enc = [&](float x) { uint32_t c = 0; for (int k = 1; k < 256; ++k) if (thr[k] <= x) c = k; return c; };
    // clear colour (0.02, 0.02, 0.03, 1): src/scene/mod.rs:286
    return enc(0.02f) | (enc(0.02f) << 8) | (enc(0.03f) << 16) | 0xff000000u;
}

// k_geom_march: persistent grid of warp workers; rows per task shrink on coarse grids so that every
// warp of the machine gets a few tasks
static int launch_geom_march(vf_scene* s, const VfFrame& f, bool early) {
    VfDevCache& dc = g_dev_cache[s->device & 63];
    const size_t smem_d = VF_GM_WARPS * sizeof(VfGmWarpT<true>), smem_n = VF_GM_WARPS * sizeof(VfGmWarpT<false>);
    if (!dc.gm_attr) {                   // cudaFuncSetAttribute applies to the current device only
        VF_CUDA(cudaDeviceGetAttribute(&dc.sm_count, cudaDevAttrMultiProcessorCount, s->device));
#define VF_GM_ATTR(T, D, S) VF_CUDA(cudaFuncSetAttribute(k_geom_march<T, D, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)((D) ? smem_d : smem_n)))
        VF_GM_ATTR(true, true, true);   VF_GM_ATTR(true, true, false);   VF_GM_ATTR(true, false, true);   VF_GM_ATTR(true, false, false);
        VF_GM_ATTR(false, true, true);  VF_GM_ATTR(false, true, false);  VF_GM_ATTR(false, false, true);  VF_GM_ATTR(false, false, false);
#undef VF_GM_ATTR
        dc.gm_attr = true;
    }
    const int sm_count = dc.sm_count;
    const uint32_t gm = s->G - 1;
    const uint32_t nstrips = (gm + VF_GM_QCOLS - 1) / VF_GM_QCOLS;
    const uint32_t per_sm = VF_GM_CTAS;     // CTAs per SM (launch bounds of k_geom_march)
    const uint32_t workers = (uint32_t)sm_count * per_sm * VF_GM_WARPS;
    uint32_t rows = (uint32_t)(((uint64_t)gm * nstrips) / (2ull * workers));
    rows = std::min<uint32_t>(std::max<uint32_t>(rows, 1u), f.depth ? VfGmDim<true>::ROWS : VfGmDim<false>::ROWS);
    // Row segments: full height while plenty of work is left, then tapering (guided self-scheduling:
    // a task is 1/(taper x workers) of what remains) so that the last tasks handed out are short.
    static const int taper = getenv("VF_GM_TAPER") ? atoi(getenv("VF_GM_TAPER")) : 1;
    static const int min_rows = getenv("VF_GM_MIN_ROWS") ? atoi(getenv("VF_GM_MIN_ROWS")) : 4;
    if (s->segtab_G != s->G || s->segtab_workers != workers || s->segtab_rows != rows) {
        std::vector<uint2> tab;
        const uint32_t lo = std::min<uint32_t>(std::max(min_rows, 1), rows);
        for (uint32_t j = 0; j < gm;) {
            uint32_t h = rows;
            if (taper > 0) h = (uint32_t)(((uint64_t)(gm - j) * nstrips) / ((uint64_t)taper * workers));
            h = std::min(std::min(std::max(h, lo), rows), gm - j);
            tab.push_back(make_uint2(j, h));
            j += h;
        }
        if (tab.size() > s->segtab_cap) {
            if (s->d_segtab) cudaFree(s->d_segtab);
            s->d_segtab = nullptr; s->segtab_cap = 0;
            VF_CUDA(cudaMalloc(&s->d_segtab, tab.size() * sizeof(uint2)));
            s->segtab_cap = (uint32_t)tab.size();
        }
        VF_CUDA(cudaStreamSynchronize(s->stream));      // frames in flight may still read the old table
        VF_CUDA(cudaMemcpy(s->d_segtab, tab.data(), tab.size() * sizeof(uint2), cudaMemcpyHostToDevice));
        s->segtab_n = (uint32_t)tab.size(); s->segtab_G = s->G; s->segtab_workers = workers; s->segtab_rows = rows;
        s->segtab_nfull = 0;                             // leading segments (k * rows, rows): decoded without the table
        while (s->segtab_nfull < tab.size() && tab[s->segtab_nfull].x == s->segtab_nfull * rows && tab[s->segtab_nfull].y == rows) ++s->segtab_nfull;
    }
    const uint32_t nsegs = s->segtab_n, ntasks = nstrips * nsegs;
    // the task dispenser counts in float (vf_geom_march.cuh): exact while every request stays below 2^24
    // (grid <= 32768: at most 1058 strips x 8192 segments)
    if ((uint64_t)ntasks + 2ull * workers + 2ull >= (1ull << 24)) return vf_set_error(VF_E_INVALID, "grid too fine for the task dispenser");
    const uint32_t ctas = std::min<uint32_t>((uint32_t)sm_count * per_sm, (ntasks + VF_GM_WARPS - 1) / VF_GM_WARPS);
    // profiling aid (builds with VF_NVCC_EXTRA=-DVF_GM_PROFILE only): VF_GM_DBG=<file> dumps [ctas*8][start,end] ns + per-task ns of the last launch
#ifdef VF_GM_PROFILE
    static const char* dbg_path = getenv("VF_GM_DBG");
#else
    static const char* dbg_path = nullptr;
#endif
    static unsigned long long* d_dbg = nullptr;
    static size_t dbg_words = 0;
    if (dbg_path) {
        const size_t need = 6 * (size_t)ctas * VF_GM_WARPS + ntasks;
        if (need > dbg_words) { if (d_dbg) cudaFree(d_dbg); VF_CUDA(cudaMalloc(&d_dbg, need * 8)); dbg_words = need; }
        VF_CUDA(cudaMemsetAsync(d_dbg, 0, need * 8, s->stream));
    }
    static const bool no_safe = getenv("VF_GM_NO_SAFE") != nullptr;      // A/B: per-vertex clip tests in every task
    // early: programmatic dependent launch behind k_shade_list of the previous frame (see scene_enqueue)
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(ctas); cfg.blockDim = dim3(VF_GM_WARPS * 32); cfg.stream = s->stream;
    cudaLaunchAttribute pdl[1];
    pdl[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; pdl[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = pdl; cfg.numAttrs = early ? 1u : 0u;
    const float2* hmm_arg = (s->use_tma && !no_safe) ? s->d_hmm : nullptr;
    unsigned long long* dbg_arg = dbg_path ? d_dbg : nullptr;
#define VF_GM_LAUNCH(T, D, S) do { cfg.dynamicSmemBytes = (D) ? smem_d : smem_n; \
        VF_CUDA(cudaLaunchKernelEx(&cfg, k_geom_march<T, D, S>, f, s->dem_map_gm[(D) ? 1 : 0], (const VfAxis*)s->d_axis, (const float*)s->d_dem, s->d_visbuf, s->d_big, s->big_capacity, \
                                   s->d_counters, (const uint2*)s->d_segtab, nsegs, s->segtab_nfull, s->segtab_rows, hmm_arg, s->hmm_nb, dbg_arg)); } while (0)
#define VF_GM_LAUNCH2(T, D) do { if (f.mode == VF_MODE_SHIPPED) VF_GM_LAUNCH(T, D, true); else VF_GM_LAUNCH(T, D, false); } while (0)
    if (s->use_tma) { if (f.depth) VF_GM_LAUNCH2(true, true); else VF_GM_LAUNCH2(true, false); }
    else            { if (f.depth) VF_GM_LAUNCH2(false, true); else VF_GM_LAUNCH2(false, false); }
#undef VF_GM_LAUNCH2
#undef VF_GM_LAUNCH
    if (dbg_path) {
        const size_t need = 6 * (size_t)ctas * VF_GM_WARPS + ntasks;
        std::vector<unsigned long long> h(need + 4);
        VF_CUDA(cudaStreamSynchronize(s->stream));
        VF_CUDA(cudaMemcpy(h.data() + 4, d_dbg, need * 8, cudaMemcpyDeviceToHost));
        h[0] = ctas * VF_GM_WARPS; h[1] = ntasks; h[2] = nstrips; h[3] = ((unsigned long long)nsegs << 32) | (rows << 8);
        static int dump_no = 0;
        char path[512];
        snprintf(path, sizeof path, "%s.%d", dbg_path, dump_no++);
        if (FILE* fp = fopen(path, "wb")) { fwrite(h.data(), 8, h.size(), fp); fclose(fp); }
    }
    return VF_OK;
}

// enqueue one frame into slot `slot` on s->stream
static int scene_enqueue(vf_scene* s, const float* u44, int slot, uint32_t* d_vis) {
    VfFrame f;
    build_frame(s, u44, &f);
    static const uint32_t clear = clear_rgba8();
    const float* thr = vf_srgb_thresholds_dev();
    if (!thr) return vf_set_error(VF_E_CUDA, "no sRGB table on device %d (vf_init has not run for it)", s->device);
    uint8_t* frame = s->d_frame + (size_t)slot * s->W * s->H * 4;
    // Counter block + tile counts start every frame at zero.  A frame that takes the list-shading pass leaves them that way
    // itself (tile counts untouched, k_shade_list resets the task / queue counters and the OTHER frame parity's list
    // length), so a run of such frames needs no memset node between k_shade_list(k) and k_geom_march(k+1).
    static const bool bin_always = getenv("VF_BIN_ALWAYS") != nullptr;
    const bool skip_bins = s->bin_hold == 0 && !bin_always;
    const bool lean = skip_bins && s->d_shade_list != nullptr;
    if (!(lean && s->counters_clean)) {
        VF_CUDA(cudaMemsetAsync(s->d_tile_count, 0, ((size_t)s->n_tiles + VF_CNT_WORDS) * 4, s->stream));
        s->shade_parity = 0;
    }
    // The kernel before this frame's k_geom_march in the stream is then k_shade_list of the previous frame, which the
    // geometry pass does not depend on (it reads the list; the visibility buffer was reset by k_resolve_lean before it, the
    // counters by its first CTA before it releases its dependents): k_geom_march may start on the SMs k_shade_list's tail
    // leaves idle (programmatic dependent launch).  Not with per-kernel timing events in between.
    const bool early_geom = s->early_geom_ok && lean && s->counters_clean && !s->timing;
    s->counters_clean = false;                                 // (until this frame's kernels are all in the stream)
    if (s->timing) VF_CUDA(cudaEventRecord(s->ev[0], s->stream));
    if (s->visbuf_depth != f.depth) {     // (re)initialise the self-cleaning visibility buffer
        VF_CUDA(cudaMemsetAsync(s->d_visbuf, f.depth ? 0xff : 0x00, (size_t)s->W * s->H * 8, s->stream));
        s->visbuf_depth = f.depth;
    }
    if (!f.sparse || s->geom_tiled) {   // general matrices; also kept for A/B runs
        const uint32_t gb = (s->G - 1 + VF_GEOM_TX - 1) / VF_GEOM_TX;
        if (s->use_tma)
            k_geom<true><<<dim3(gb, gb), VF_GEOM_THREADS, 0, s->stream>>>(
                f, s->dem_map, s->d_axis, s->d_dem, s->d_visbuf, s->d_big, s->big_capacity, s->d_counters);
        else
            k_geom<false><<<dim3(gb, gb), VF_GEOM_THREADS, 0, s->stream>>>(
                f, s->dem_map, s->d_axis, s->d_dem, s->d_visbuf, s->d_big, s->big_capacity, s->d_counters);
    } else {
        int rc = launch_geom_march(s, f, early_geom);
        if (rc) return rc;
    }
    if (s->timing) VF_CUDA(cudaEventRecord(s->ev[1], s->stream));
    if (skip_bins) {
        s->submit_skipped_bins = true;
    } else {   // one cooperative launch: bin -> scan -> scatter with grid barriers in between
        if (s->bin_hold) --s->bin_hold;
        int& binsort_ctas = g_dev_cache[s->device & 63].binsort_ctas;
        if (!binsort_ctas) {
            int sms = 0, per_sm = 0;
            VF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->device));
            VF_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_binsort, 256, 0));
            binsort_ctas = sms * std::min(std::max(per_sm, 1), 2);
        }
        uint32_t big_capacity = s->big_capacity, capacity = s->capacity;
        void* args[] = {&f, &s->d_axis, &s->d_dem, &s->d_big, &big_capacity, &s->d_tile_count, &s->d_tile_off,
                        &s->d_tile_cursor, &s->d_pair_tile, &s->d_pair_prim, &s->d_sorted, &s->d_counters, &capacity,
                        &s->d_maxpairs};
        VF_CUDA(cudaLaunchCooperativeKernel((const void*)k_binsort, dim3(binsort_ctas), dim3(256), args, 0, s->stream));
    }
    if (s->timing) VF_CUDA(cudaEventRecord(s->ev[2], s->stream));
    // vertex heights for the shading pass: the resampled image, the DEM itself when the grid maps onto it one to one,
    // else (TMA staging off) a gather through the axis table inside the kernel
    const float* hg = s->d_hgrid ? s->d_hgrid : (s->dem_identity ? s->d_dem : nullptr);
    const uint32_t hgp = s->d_hgrid ? s->hpitch : s->pitch;
#define VF_RASTER_LAUNCH(M, D) k_raster<M, D><<<dim3(s->tiles_x, s->tiles_y), 256, 0, s->stream>>>(f, s->d_axis, s->d_dem, hg, hgp, s->d_lut, thr, \
        s->d_visbuf, s->d_tile_off, s->d_sorted, s->d_counters, clear, frame, d_vis, s->d_maxpairs)
    // Frames that skip the binning stage take the two kernels without the merge / clipper paths (vf_raster_lean.cuh): a
    // streaming resolve that lists the covered pixels, then a dense shading pass over the list.  If the frame did queue
    // something, k_resolve_lean reports it like k_raster and the frame is rendered again.
    if (lean) {                                                // (VF_RASTER_FULL at scene creation: no list, k_raster in every frame)
        int& sms = g_dev_cache[s->device & 63].sm_count;
        if (!sms) VF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->device));
        const uint32_t w4 = s->W >> 2, quads = w4 * s->H, half = ((quads + 1u) / 2u + 255u) & ~255u;
        const uint32_t w4_inv = (uint32_t)(0x100000000ull / w4 > 0xffffffffull ? 0xffffffffull : 0x100000000ull / w4);
        if (f.depth) k_resolve_lean<true><<<half / 256u, 256, 0, s->stream>>>(s->W, s->H, w4_inv, half, s->d_visbuf, s->d_counters, s->shade_parity, clear, frame, d_vis, s->d_shade_list, s->d_maxpairs);
        else         k_resolve_lean<false><<<half / 256u, 256, 0, s->stream>>>(s->W, s->H, w4_inv, half, s->d_visbuf, s->d_counters, s->shade_parity, clear, frame, d_vis, s->d_shade_list, s->d_maxpairs);
        const int ctas = sms * VF_SHADE_CTAS;
        if (f.mode == VF_MODE_SHIPPED) k_shade_list<0><<<ctas, 256, 0, s->stream>>>(f, s->d_axis, s->d_dem, hg, hgp, s->d_lut, thr, s->d_counters, s->shade_parity, s->d_shade_list, clear, frame);
        else                           k_shade_list<1><<<ctas, 256, 0, s->stream>>>(f, s->d_axis, s->d_dem, hg, hgp, s->d_lut, thr, s->d_counters, s->shade_parity, s->d_shade_list, clear, frame);
        s->shade_parity ^= 1u;
    } else {
        if (f.mode == VF_MODE_SHIPPED) { if (f.depth) VF_RASTER_LAUNCH(0, true); else VF_RASTER_LAUNCH(0, false); }
        else                           { if (f.depth) VF_RASTER_LAUNCH(1, true); else VF_RASTER_LAUNCH(1, false); }
    }
#undef VF_RASTER_LAUNCH
    if (s->timing) VF_CUDA(cudaEventRecord(s->ev[3], s->stream));
    VF_CUDA(cudaGetLastError());
    s->counters_clean = lean;
    return VF_OK;
}

// after the call's work has completed: fold its marks into the metrics block
static void scene_finish_metrics(vf_scene* s, vf_scene::Marks& mk) {
    vf_render_metrics_t& m = s->metrics;
    m.gpu_ms = m.readback_ms = 0.0f;
    if (mk.g0 && cudaEventElapsedTime(&m.gpu_ms, mk.g0, mk.g1) != cudaSuccess) { (void)cudaGetLastError(); m.gpu_ms = 0.0f; }
    if (mk.copies && cudaEventElapsedTime(&m.readback_ms, mk.r0, mk.r1) != cudaSuccess) { (void)cudaGetLastError(); m.readback_ms = 0.0f; }
    m.total_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - mk.t0).count();
    m.frames = (uint32_t)mk.frames;
    m.last_d2h_bytes = mk.d2h;
    s->d2h_bytes += mk.d2h;
}

extern "C" int vf_scene_metrics(const vf_scene* sc, vf_render_metrics_t* out) {
    if (!sc || !out) return vf_set_error(VF_E_INVALID, "null argument");
    vf_scene* s = const_cast<vf_scene*>(sc);
    if (s->upload_timed && s->ev_up0 && cudaEventQuery(s->ev_up1) == cudaSuccess) {
        if (cudaEventElapsedTime(&s->metrics.upload_ms, s->ev_up0, s->ev_up1) != cudaSuccess) (void)cudaGetLastError();
    } else {
        (void)cudaGetLastError();
    }
    s->metrics.h2d_bytes = s->h2d_bytes; s->metrics.d2h_bytes = s->d2h_bytes; s->metrics.last_h2d_bytes = s->last_h2d_bytes;
    s->metrics.vram_bytes = scene_vram_bytes(s);
    *out = s->metrics;
    return VF_OK;
}

static int scene_finish_stats(vf_scene* s) {
    s->stats.bin_pairs = s->h_counters_last[VF_CNT_PAIRS];
    s->stats.clipped_prims = s->h_counters_last[VF_CNT_CLIPPED];
    s->stats.bin_capacity = s->capacity;
    s->stats.tma = s->use_tma ? 1u : 0u;
    s->stats.big_prims = s->h_counters_last[VF_CNT_BIG];
    s->stats.tiles = s->n_tiles;
    if (s->timing) {
        VF_CUDA(cudaEventElapsedTime(&s->stats.ms_geom, s->ev[0], s->ev[1]));
        VF_CUDA(cudaEventElapsedTime(&s->stats.ms_bin, s->ev[1], s->ev[2]));
        VF_CUDA(cudaEventElapsedTime(&s->stats.ms_raster, s->ev[2], s->ev[3]));
        VF_CUDA(cudaEventElapsedTime(&s->stats.ms_total, s->ev[0], s->ev[3]));
    }
    return VF_OK;
}

// frames enqueued from here on read the image uploaded by vf_scene_set_height_f32_async
static int scene_apply_pending_dem(vf_scene* s) {
    if (!s->back_pending) return VF_OK;
    VF_CUDA(cudaStreamWaitEvent(s->stream, s->ev_uploaded, 0));
    VF_CUDA(cudaEventRecord(s->ev_front_released, s->stream));   // every frame of the old image precedes this
    s->front_released_once = true;
    std::swap(s->d_dem, s->d_dem_back); std::swap(s->d_hgrid, s->d_hgrid_back); std::swap(s->d_hmm, s->d_hmm_back);
    std::swap(s->dem_map, s->dem_map_b); std::swap(s->dem_map_gm[0], s->dem_map_gm_b[0]); std::swap(s->dem_map_gm[1], s->dem_map_gm_b[1]);
    s->back_pending = false;
    return VF_OK;
}

// Enqueues n frames (+ their D2H copies when out_host is given) and the read-back of the batch's
// counters into pinned slot `hslot`.  Frame k is rendered into slot k%2 while the copy stream
// drains the other slot to the host.  Does not wait.
static const int VF_PNG_SLOTS = 3;
struct VfPngOut { uint8_t* out; size_t cap; uint32_t* sizes; };          // n files of at most `cap` bytes each, sizes[k] bytes used
// the height image the call AFTER this one renders from: its upload starts as soon as this call's own image is current
struct VfNextImage { const float* host; uint32_t w, h; };
extern "C" int vf_scene_set_height_f32_async(vf_scene* s, const float* host, uint32_t w, uint32_t h);

// file `j` of a PNG batch has been encoded into its slot: read its size, start the size-exact copy to the host
static int png_collect(vf_scene* s, const VfPngOut& po, int j, vf_scene::Marks& mk) {
    const int ps = j % VF_PNG_SLOTS;
    VF_CUDA(cudaEventSynchronize(s->ev_png_done[ps]));
    const uint32_t size = s->png->h_file_size[ps];            // stored by k_png_pack straight into mapped host memory
    if (size == 0xffffffffu || size > po.cap)
        return vf_set_error(VF_E_INVALID, "PNG file %d needs %u bytes, the output slot holds %zu (size it with vf_png_file_bound)", j, size, po.cap);
    po.sizes[j] = size;
    if (!mk.copies) { VF_CUDA(cudaEventRecord(mk.r0, s->copy_stream)); mk.copies = true; }
    VF_CUDA(cudaMemcpyAsync(po.out + (size_t)j * po.cap, s->png->d_files + (size_t)ps * s->png->file_cap, size, cudaMemcpyDeviceToHost, s->copy_stream));
    VF_CUDA(cudaEventRecord(s->ev_png_copied[ps], s->copy_stream));
    s->png_slot_used[ps] = true;
    mk.d2h += size;
    return VF_OK;
}

static int scene_submit(vf_scene* s, const float* u44xn, int n, uint8_t* out_host, uint32_t* d_vis, int hslot,
                        bool apply_pending, vf_scene::Marks& mk, const VfPngOut* png_out = nullptr, const VfNextImage* next = nullptr) {
    int rc = apply_pending ? scene_apply_pending_dem(s) : VF_OK;
    if (rc) return rc;
    if (next) {
        // Only an image of the current size takes the double-buffered path; another size would replace the image
        // this very batch is about to draw.
        if (s->back_pending || next->w != s->Wd || next->h != s->Hd || !s->d_dem)
            return vf_set_error(VF_E_INVALID, "the next height image must have the size of the current one (%u x %u), with no other upload pending", s->Wd, s->Hd);
        rc = vf_scene_set_height_f32_async(s, next->host, next->w, next->h);
        if (rc) return rc;
    }
    rc = marks_init(mk);
    if (rc) return rc;
    if (png_out && !s->png) {
        rc = vf_png_dev_create(s->W, s->H, VF_PNG_SLOTS, &s->png);
        if (rc) return rc;
        VF_CUDA(cudaHostAlloc(&s->h_png_size, 4 * sizeof(uint32_t), cudaHostAllocDefault));
        for (int k = 0; k < VF_PNG_SLOTS; ++k) {
            VF_CUDA(cudaEventCreateWithFlags(&s->ev_png_done[k], cudaEventDisableTiming));
            VF_CUDA(cudaEventCreateWithFlags(&s->ev_png_copied[k], cudaEventDisableTiming));
        }
    }
    const size_t fbytes = (size_t)s->W * s->H * 4;
    mk.t0 = std::chrono::steady_clock::now(); mk.copies = false; mk.frames = n; mk.d2h = out_host ? (uint64_t)n * fbytes : 0;
    VF_CUDA(cudaEventRecord(mk.g0, s->stream));
    VF_CUDA(cudaMemsetAsync(s->d_maxpairs, 0, 16, s->stream));
    s->submit_skipped_bins = false;
    for (int k = 0; k < n; ++k) {
        const int slot = k % VF_MAX_SLOTS;
        // a frame slot may still be draining to the host on copy_stream (an earlier batch, possibly one
        // submitted with vf_scene_render_batch_async): wait for that copy whatever this call does with its frames
        if (s->slot_used[slot]) VF_CUDA(cudaStreamWaitEvent(s->stream, s->ev_copied[slot], 0));
        rc = scene_enqueue(s, u44xn + (size_t)k * 44, slot, d_vis);
        if (rc) return rc;
        // (the previous frame's file is reported done only here, behind this frame's render kernels: nothing then stands
        //  between k_png_pack(k-1) and k_geom_march(k) in the stream, and the geometry pass -- a programmatic dependent that
        //  needs nothing from the encoder -- starts under it; the host collects files two frames late anyway)
        if (png_out && k > 0) VF_CUDA(cudaEventRecord(s->ev_png_done[(k - 1) % VF_PNG_SLOTS], s->stream));
        if (out_host) {
            VF_CUDA(cudaEventRecord(s->ev_done[slot], s->stream));
            VF_CUDA(cudaStreamWaitEvent(s->copy_stream, s->ev_done[slot], 0));
            if (!mk.copies) { VF_CUDA(cudaEventRecord(mk.r0, s->copy_stream)); mk.copies = true; }
            VF_CUDA(cudaMemcpyAsync(out_host + (size_t)k * fbytes, s->d_frame + (size_t)slot * fbytes,
                                    fbytes, cudaMemcpyDeviceToHost, s->copy_stream));
            VF_CUDA(cudaEventRecord(s->ev_copied[slot], s->copy_stream));
            s->slot_used[slot] = true;
        }
        if (png_out) {
            // frame k -> .png bytes in file slot k % VF_PNG_SLOTS (two kernels behind k_raster on the render stream), its
            // size into pinned memory; the file of frame k - (VF_PNG_SLOTS - 1) is collected meanwhile, so the host's
            // wait for a size never stalls the render stream
            const int ps = k % VF_PNG_SLOTS;
            if (s->png_slot_used[ps]) VF_CUDA(cudaStreamWaitEvent(s->stream, s->ev_png_copied[ps], 0));
            rc = vf_png_dev_encode(s->png, s->d_frame + (size_t)slot * fbytes, ps, s->stream);
            if (rc) return rc;
            if (k == n - 1) VF_CUDA(cudaEventRecord(s->ev_png_done[ps], s->stream));
            if (k >= VF_PNG_SLOTS - 1) { rc = png_collect(s, *png_out, k - (VF_PNG_SLOTS - 1), mk); if (rc) return rc; }
        }
    }
    if (png_out)
        for (int j = std::max(0, n - (VF_PNG_SLOTS - 1)); j < n; ++j) { rc = png_collect(s, *png_out, j, mk); if (rc) return rc; }
    VF_CUDA(cudaEventRecord(mk.g1, s->stream));
    if (mk.copies) VF_CUDA(cudaEventRecord(mk.r1, s->copy_stream));
    uint32_t* h = s->h_counters + (size_t)hslot * 2 * VF_CNT_WORDS;
    VF_CUDA(cudaMemcpyAsync(h, s->d_counters, VF_CNT_WORDS * 4, cudaMemcpyDeviceToHost, s->stream));
    VF_CUDA(cudaMemcpyAsync(h + VF_CNT_WORDS, s->d_maxpairs, 16, cudaMemcpyDeviceToHost, s->stream));
    return VF_OK;
}

// after the batch of pinned slot `hslot` has completed: 0 = fine, 1 = a buffer overflowed and was
// grown (the batch must be rendered again), < 0 = error code
static int scene_check_overflow(vf_scene* s, int hslot, bool skipped_bins) {
    const uint32_t* h = s->h_counters + (size_t)hslot * 2 * VF_CNT_WORDS + VF_CNT_WORDS;
    const uint32_t need = h[0], need_big = std::max(h[1], h[2]);
    if (h[2] > 0) {                                      // a frame queued primitives: keep the binning stage for a while
        if (skipped_bins) s->bin_hold_len = std::min<uint32_t>(s->bin_hold_len * 2u, 1u << 20);
        s->bin_hold = s->bin_hold_len;
    }
    const bool redo_bins = skipped_bins && h[2] > 0;
    if (need <= s->capacity && need_big <= s->big_capacity) return redo_bins ? 1 : 0;
    VF_CUDA(cudaStreamSynchronize(s->stream));           // nothing in flight may still use the buffers
    if (need > (1u << 30)) return -vf_set_error(VF_E_OOM, "binning needs %u pairs (limit 2^30)", need);
    int rc;
    if (need > s->capacity) {
        rc = scene_alloc_pairs(s, (uint32_t)std::min<uint64_t>((uint64_t)need + need / 4 + 1024, 1u << 30));
        if (rc) return -rc;
    }
    if (need_big > s->big_capacity) {
        rc = scene_alloc_big(s, need_big + need_big / 4 + 1024);
        if (rc) return -rc;
    }
    return 1;
}

// Renders n frames and waits; out_host may be null (frames stay on the device).  If a kernel asked
// for more queue / pair space than its buffer holds, the buffer is grown and the call is repeated
// (an overflowed frame leaves a clean visibility buffer: k_raster always runs to completion).
static int scene_render(vf_scene* s, const float* u44xn, int n, uint8_t* out_host, uint32_t* d_vis, bool apply_pending = true,
                        const VfPngOut* png_out = nullptr, const VfNextImage* next = nullptr) {
    if (!s || !u44xn || n <= 0) return vf_set_error(VF_E_INVALID, "invalid scene / frame count");
    int rc = scene_bind(s);
    if (rc) return rc;
    for (int attempt = 0; attempt < 8; ++attempt) {
        // (a repeated attempt draws the same image again: an upload staged by the first attempt stays pending)
        rc = scene_submit(s, u44xn, n, out_host, d_vis, 0, apply_pending && attempt == 0, s->marks, png_out, attempt == 0 ? next : nullptr);
        if (rc) {                                            // nothing of the failed call may still be in flight when the caller reacts
            cudaStreamSynchronize(s->stream); cudaStreamSynchronize(s->copy_stream);
            return rc;
        }
        VF_CUDA(cudaStreamSynchronize(s->stream));
        if (out_host || png_out) VF_CUDA(cudaStreamSynchronize(s->copy_stream));
        scene_finish_metrics(s, s->marks);
        memcpy(s->h_counters_last, s->h_counters, VF_CNT_WORDS * 4);
        const int ov = scene_check_overflow(s, 0, s->submit_skipped_bins);
        if (ov < 0) return -ov;
        if (ov == 0) return scene_finish_stats(s);
    }
    return vf_set_error(VF_E_CUDA, "pair buffer kept overflowing");
}

// ---- pipelined interface: uploads and batches overlap across calls -------------------------------
static void scene_swap_images(vf_scene* s) {
    std::swap(s->d_dem, s->d_dem_back); std::swap(s->d_hgrid, s->d_hgrid_back); std::swap(s->d_hmm, s->d_hmm_back);
    std::swap(s->dem_map, s->dem_map_b); std::swap(s->dem_map_gm[0], s->dem_map_gm_b[0]); std::swap(s->dem_map_gm[1], s->dem_map_gm_b[1]);
}

// Waits for batch `id`.  If one of its kernels ran out of queue / pair space (judged against the
// buffers the batch was submitted with), the buffers are grown and the batch is rendered again,
// synchronously, with the mode and from the height image it was submitted with.  That image is
// still intact: whatever would overwrite it resolves the batches that read it first.
static int ticket_resolve(vf_scene* s, int id) {
    vf_scene::Ticket& t = s->tickets[id];
    if (t.resolved) return t.rc;
    t.resolved = true;
    t.rc = VF_OK;
    if (cudaEventSynchronize(t.done) != cudaSuccess) return t.rc = vf_set_error(VF_E_CUDA, "batch %d: %s", id, cudaGetErrorString(cudaGetLastError()));
    scene_finish_metrics(s, s->ticket_marks[id]);
    const uint32_t* h = s->h_counters + (size_t)(1 + id) * 2 * VF_CNT_WORDS + VF_CNT_WORDS;
    if (h[0] <= t.cap_pairs && std::max(h[1], h[2]) <= t.cap_big && !(t.skipped_bins && h[2] > 0)) {
        if (h[2] > 0) s->bin_hold = s->bin_hold_len;
        return VF_OK;
    }
    const int ov = scene_check_overflow(s, 1 + id, t.skipped_bins);   // grows what is still too small (a later batch may have done it already)
    if (ov < 0) return t.rc = -ov;
    const bool other = t.dem != s->d_dem;
    if (other && (t.dem != s->d_dem_back || s->back_pending))
        return t.rc = vf_set_error(VF_E_INVALID, "batch %d overflowed and its height image is gone", id);
    const int mode = s->mode, depth = s->depth;
    const float h_min = s->h_min, h_max = s->h_max;
    s->mode = t.mode; s->depth = t.depth; s->h_min = t.h_min; s->h_max = t.h_max;
    if (other) scene_swap_images(s);
    t.rc = scene_render(s, t.u.data(), t.n, t.out, nullptr, false);
    if (other) scene_swap_images(s);
    s->mode = mode; s->depth = depth; s->h_min = h_min; s->h_max = h_max;
    return t.rc;
}

// before `image` is overwritten: finish (and if need be redo) the batches that read it
static void scene_resolve_readers(vf_scene* s, const float* image) {
    for (int k = 0; k < 4; ++k)
        if (s->tickets[k].busy && (!image || s->tickets[k].dem == image)) ticket_resolve(s, k);
}

extern "C" int vf_scene_set_height_f32_async(vf_scene* s, const float* host, uint32_t w, uint32_t h) {
    if (!s || !host) return vf_set_error(VF_E_INVALID, "null argument");
    if (int rcb = scene_bind(s)) return rcb;
    const uint32_t pitch = (w + 3u) & ~3u;
    if (w != s->Wd || h != s->Hd || pitch != s->pitch || !s->d_dem) return scene_upload_dem(s, host, w, h);   // new size: synchronous path
    if (!s->upload_stream) {
        VF_CUDA(cudaStreamCreateWithFlags(&s->upload_stream, cudaStreamNonBlocking));
        VF_CUDA(cudaEventCreateWithFlags(&s->ev_uploaded, cudaEventDisableTiming));
        VF_CUDA(cudaEventCreateWithFlags(&s->ev_front_released, cudaEventDisableTiming));
    }
    if (!s->d_dem_back) {
        VF_CUDA(cudaMalloc(&s->d_dem_back, (size_t)pitch * h * 4));
        s->back_valid = false;
    }
    if (s->d_hgrid && !s->d_hgrid_back) {
        VF_CUDA(cudaMalloc(&s->d_hgrid_back, (size_t)s->hpitch * s->G * 4));
        s->back_valid = false;
    }
    if (s->d_hmm && !s->d_hmm_back) VF_CUDA(cudaMalloc(&s->d_hmm_back, (size_t)s->hmm_nb * s->hmm_nb * sizeof(float2)));
    if (!s->back_valid) {
        const bool ok = !s->use_tma || (s->d_hgrid ? scene_encode_maps(s->d_hgrid_back, s->G, s->G, s->hpitch, &s->dem_map_b, s->dem_map_gm_b)
                                                   : scene_encode_maps(s->d_dem_back, w, h, pitch, &s->dem_map_b, s->dem_map_gm_b));
        if (!ok) return vf_set_error(VF_E_CUDA, "tensor map encoding failed for the back height image");
        s->back_valid = true;
    }
    scene_resolve_readers(s, s->d_dem_back);      // normally none: those batches were collected long ago
    // the back image was the front image of frames submitted before the last swap
    if (s->front_released_once) VF_CUDA(cudaStreamWaitEvent(s->upload_stream, s->ev_front_released, 0));
    if (!s->ev_up0) { VF_CUDA(cudaEventCreate(&s->ev_up0)); VF_CUDA(cudaEventCreate(&s->ev_up1)); }
    VF_CUDA(cudaEventRecord(s->ev_up0, s->upload_stream));
    s->last_h2d_bytes = (uint64_t)w * h * 4; s->h2d_bytes += s->last_h2d_bytes;
    VF_CUDA(cudaMemcpy2DAsync(s->d_dem_back, (size_t)pitch * 4, host, (size_t)w * 4, (size_t)w * 4, h,
                              cudaMemcpyHostToDevice, s->upload_stream));
    if (s->d_hgrid_back) {                        // vertex heights of the new image, behind its upload
        launch_resample(s, s->d_dem_back, s->d_hgrid_back, s->upload_stream);
        VF_CUDA(cudaGetLastError());
    }
    if (s->d_hmm_back) {
        launch_hminmax(s, s->d_dem_back, s->d_hgrid_back, s->d_hmm_back, s->upload_stream);
        VF_CUDA(cudaGetLastError());
    }
    VF_CUDA(cudaEventRecord(s->ev_up1, s->upload_stream));
    s->upload_timed = true;
    VF_CUDA(cudaEventRecord(s->ev_uploaded, s->upload_stream));
    static const bool dbg_sync = getenv("VF_DEBUG_SYNC_UPLOAD") != nullptr;      // debugging aid: no overlap of upload and frames
    if (dbg_sync) VF_CUDA(cudaStreamSynchronize(s->upload_stream));
    s->back_pending = true;
    return VF_OK;
}

extern "C" int vf_scene_render_batch_async(vf_scene* s, const float* u44xn, int n, uint8_t* out_rgba_host, int* ticket) {
    if (!s || !u44xn || n <= 0 || !out_rgba_host || !ticket) return vf_set_error(VF_E_INVALID, "invalid argument");
    int rc = scene_bind(s);
    if (rc) return rc;
    int id = -1;
    for (int k = 0; k < 4; ++k) if (!s->tickets[k].busy) { id = k; break; }
    if (id < 0) return vf_set_error(VF_E_INVALID, "4 batches already in flight: call vf_scene_wait first");
    vf_scene::Ticket& t = s->tickets[id];
    if (!t.done) VF_CUDA(cudaEventCreateWithFlags(&t.done, cudaEventDisableTiming));
    t.u.assign(u44xn, u44xn + (size_t)n * 44); t.n = n; t.out = out_rgba_host;
    rc = scene_submit(s, u44xn, n, out_rgba_host, nullptr, 1 + id, true, s->ticket_marks[id]);
    if (rc) return rc;
    t.cap_pairs = s->capacity; t.cap_big = s->big_capacity; t.dem = s->d_dem;       // after the swap inside scene_submit
    t.skipped_bins = s->submit_skipped_bins;
    t.mode = s->mode; t.depth = s->depth; t.h_min = s->h_min; t.h_max = s->h_max;
    t.resolved = false; t.rc = VF_OK;
    // done = counters read back on the render stream and every frame copied on the copy stream
    VF_CUDA(cudaEventRecord(s->ev_done[0], s->stream));           // reuse as a plain marker: ordered after the read-back
    VF_CUDA(cudaStreamWaitEvent(s->copy_stream, s->ev_done[0], 0));
    VF_CUDA(cudaEventRecord(t.done, s->copy_stream));
    t.busy = true;
    *ticket = id;
    return VF_OK;
}

extern "C" int vf_scene_wait(vf_scene* s, int ticket) {
    if (!s || ticket < 0 || ticket >= 4 || !s->tickets[ticket].busy) return vf_set_error(VF_E_INVALID, "invalid ticket");
    if (int rcb = scene_bind(s)) return rcb;
    const int rc = ticket_resolve(s, ticket);
    s->tickets[ticket].busy = false;
    return rc;
}

extern "C" int vf_scene_render_rgba(vf_scene* s, uint8_t* out_rgba_host) {
    if (!s || !out_rgba_host) return vf_set_error(VF_E_INVALID, "null argument");
    return scene_render(s, s->u44, 1, out_rgba_host, nullptr);
}

extern "C" int vf_scene_render_batch(vf_scene* s, const float* u44xn, int n, uint8_t* out_rgba_host) {
    if (!s || !out_rgba_host) return vf_set_error(VF_E_INVALID, "null argument");
    return scene_render(s, u44xn, n, out_rgba_host, nullptr);
}

extern "C" size_t vf_png_file_bound(int w, int h) { return (w > 0 && h > 0) ? vf_png_dev_bound(w, h) : 0; }

extern "C" int vf_scene_render_batch_png(vf_scene* s, const float* u44xn, int n, uint8_t* out_host, size_t cap_per_file, uint32_t* sizes) {
    if (!s || !out_host || !sizes) return vf_set_error(VF_E_INVALID, "null argument");
    const VfPngOut po = {out_host, cap_per_file, sizes};
    return scene_render(s, u44xn ? u44xn : s->u44, n, nullptr, nullptr, true, &po);
}

extern "C" int vf_scene_render_batch_png_next(vf_scene* s, const float* u44xn, int n, uint8_t* out_host, size_t cap_per_file, uint32_t* sizes,
                                             const float* next_height_host, uint32_t next_w, uint32_t next_h, int flags) {
    if (!s || !out_host || !sizes) return vf_set_error(VF_E_INVALID, "null argument");
    if (flags & ~VF_PNG_KEEP_IMAGE) return vf_set_error(VF_E_INVALID, "unknown flag");
    const VfPngOut po = {out_host, cap_per_file, sizes};
    const VfNextImage nx = {next_height_host, next_w, next_h};
    return scene_render(s, u44xn ? u44xn : s->u44, n, nullptr, nullptr, !(flags & VF_PNG_KEEP_IMAGE), &po, next_height_host ? &nx : nullptr);
}

extern "C" int vf_scene_render_device(vf_scene* s, const float* u44xn, int n) {
    return scene_render(s, u44xn ? u44xn : (s ? s->u44 : nullptr), n, nullptr, nullptr);
}

extern "C" int vf_scene_read_frame(vf_scene* s, int slot, uint8_t* out_host) {
    if (!s || !out_host) return vf_set_error(VF_E_INVALID, "null argument");
    if (slot < 0 || slot >= VF_MAX_SLOTS) return vf_set_error(VF_E_INVALID, "frame slot must be 0 or 1");
    if (int rc = scene_bind(s)) return rc;
    const size_t fbytes = (size_t)s->W * s->H * 4;
    VF_CUDA(cudaStreamSynchronize(s->stream));
    VF_CUDA(cudaStreamSynchronize(s->copy_stream));
    VF_CUDA(cudaMemcpy(out_host, s->d_frame + (size_t)slot * fbytes, fbytes, cudaMemcpyDeviceToHost));
    return VF_OK;
}

extern "C" int vf_scene_debug_visibility(vf_scene* s, uint32_t* out_host) {
    if (!s || !out_host) return vf_set_error(VF_E_INVALID, "null argument");
    const size_t bytes = (size_t)s->W * s->H * 4;
    if (!s->d_vis) VF_CUDA(cudaMalloc(&s->d_vis, bytes));
    int rc = scene_render(s, s->u44, 1, nullptr, s->d_vis);
    if (rc) return rc;
    VF_CUDA(cudaMemcpy(out_host, s->d_vis, bytes, cudaMemcpyDeviceToHost));
    return VF_OK;
}

// 1-99 % range of the scene's DEM (src/renderer.rs:24-28); implemented in vf_dem.cu
int vf_dem_minmax_device(const float* d_data, uint32_t w, uint32_t h, uint32_t pitch, int clamp,
                         float* lo, float* hi, cudaStream_t stream);

extern "C" int vf_scene_auto_height_range(vf_scene* s, float* h_min, float* h_max) {
    float lo, hi;
    int rc = scene_bind(s);
    if (rc) return rc;
    rc = scene_apply_pending_dem(s);
    if (rc) return rc;
    rc = vf_dem_minmax_device(s->d_dem, s->Wd, s->Hd, s->pitch, 1, &lo, &hi, s->stream);
    if (rc) return rc;
    const float guard = lo + 1e-5f;
    hi = hi > guard ? hi : guard;
    s->h_min = lo; s->h_max = hi;
    if (h_min) *h_min = lo;
    if (h_max) *h_max = hi;
    return VF_OK;
}
