// vf_terrain_kernels.cuh -- the terrain frame on sm_100a.
//
//   k_axis    : per-index tables of the implicit grid (x|z, u|v, texel column/row)
//   k_geom    : vertex stage once per vertex (smem), hierarchical cull (quad bbox -> triangle),
//               MICRO triangles (bbox <= VF_MICRO_MAX pixel centres, no clipping) are rasterised
//               on the spot into the L2-resident visibility buffer with one atomic per covered
//               sample; everything else (large or clipped primitives) goes to a queue      [dominant]
//   k_bin     : queue -> (tile, primitive) pairs, warp-scan compaction
//   k_scan / k_scatter : per-tile prefix sum and scatter
//   k_raster  : per 16x16 tile: merges the visibility buffer with the tile's binned primitives in
//               shared memory (integer edge functions, top-left rule, last-drawn-wins or depth
//               Less), then the fused fragment stage, exact sRGB8 and 16-byte RGBA8 stores
// replaces: vs_main/fs_main of src/shaders/terrain.wgsl and the fixed-function state of
// src/terrain/pipeline.rs:107-140.  Results are independent of scheduling: visibility is a max
// over primitive ids (no depth) or a min over (depth, id) keys (depth), both order-free.
#pragma once
#include <cuda.h>          // CUtensorMap (type only; the encoder is fetched through cudaGetDriverEntryPoint)
#include "vf_device.cuh"

// ---- TMA / mbarrier primitives (sm_90+; SASS: UTMALDG / SYNCS) ----------------------------
__device__ __forceinline__ uint32_t vf_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void vf_mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(vf_smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void vf_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(vf_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void vf_tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(vf_smem_u32(dst)), "l"(map), "r"(c0), "r"(c1), "r"(vf_smem_u32(bar)) : "memory");
}
// bounded wait: a mis-encoded tensor map must fault the launch, not hang the GPU
__device__ __forceinline__ void vf_mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done = 0;
    for (uint32_t spin = 0; !done; ++spin) {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done) : "r"(vf_smem_u32(bar)), "r"(parity) : "memory");
        if (spin > (1u << 24)) __trap();
    }
}
#define VF_TMA_BOX_X 36      // 33 texels + padding to a 16-byte multiple (144 B rows)
#define VF_TMA_BOX_Y 33

struct VfAxis { float coord; float uv; int tx; int ty; };   // 16 B per grid index

#define VF_GEOM_TX 32          // cells per CTA in x
#define VF_GEOM_TY 32          // cells per CTA in y
#define VF_GEOM_THREADS 256
#define VF_GEOM_VX (VF_GEOM_TX + 1)
#define VF_GEOM_VY (VF_GEOM_TY + 1)
#ifndef VF_MICRO_MAX
#define VF_MICRO_MAX 4         // pixel-centre count up to which one thread rasterises a triangle
#endif
#define VF_SERIAL_LINES 256    // one thread walks the runs of a triangle with at most this many lines of its
#define VF_SERIAL_AREA 4096    // shorter bbox axis and at most this many pixels of area; larger -> binned

// counter block (cleared per frame together with the tile counts)
enum { VF_CNT_PAIRS = 0, VF_CNT_CLIPPED = 1, VF_CNT_BIG = 2, VF_CNT_TASK = 3, VF_CNT_GRIDBAR = 4, VF_CNT_SHADE = 5 /* and 6: one list length per frame parity */, VF_CNT_WORDS = 8 };

// ---------------------------------------------------------------------------------------
__global__ void k_axis(VfAxis* __restrict__ axis, uint32_t G, float step, float gm1,
                       uint32_t Wd, uint32_t Hd) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= G) return;
    VfAxis a;
    a.coord = vf_plane_coord(i, step);
    a.uv = vf_uv_coord(i, gm1);
    a.tx = vf_texel_index(a.uv, Wd);
    a.ty = vf_texel_index(a.uv, Hd);
    axis[i] = a;
}

// flag[0] stays 1 only if every grid index maps to the texel of the same index (tx[i]==i, ty[i]==i)
__global__ void k_axis_identity(const VfAxis* __restrict__ axis, uint32_t G, uint32_t* __restrict__ flag) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < G && (axis[i].tx != (int)i || axis[i].ty != (int)i)) atomicAnd(flag, 0u);
}

__device__ __forceinline__ void vf_grid_vertex(const VfFrame& f, const VfAxis* __restrict__ axis,
                                               const float* __restrict__ dem, uint32_t i,
                                               uint32_t j, VfVtx& o) {
    const VfAxis ax = axis[i], az = axis[j];
    const float h_tex = __ldg(dem + (size_t)az.ty * f.dem_pitch + ax.tx);
    vf_vertex_stage(f, ax.coord, az.coord, ax.uv, az.uv, h_tex, o);
}

__device__ __forceinline__ void vf_prim_vertices(const VfFrame& f, const VfAxis* __restrict__ axis,
                                                 const float* __restrict__ dem, uint32_t prim,
                                                 VfVtx tri[3]) {
    uint32_t ij[3][2];
    vf_prim_vertex_ids(f.G, prim, ij);
#pragma unroll
    for (int k = 0; k < 3; ++k) vf_grid_vertex(f, axis, dem, ij[k][0], ij[k][1], tri[k]);
}

// no-depth resolve of a sample as one predicated reduction (no branch around it)
// (visibility buffer layout: without depth one 32-bit key per pixel, prim + 1, 0 = empty; with depth one 64-bit key,
// depth bits << 32 | prim, all ones = empty)
__device__ __forceinline__ void vf_vis_max_if(bool covered, uint32_t* key32, uint32_t prim) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %0, 0;\n\t@p red.global.max.u32 [%1], %2;\n\t}"
                 :: "r"((uint32_t)covered), "l"(key32), "r"(prim + 1u) : "memory");
}
// key of pixel p in the global visibility buffer, as the pointer vf_vis_update takes
__device__ __forceinline__ unsigned long long* vf_vis_key(unsigned long long* vis, bool depth, size_t p) {
    return depth ? vis + p : reinterpret_cast<unsigned long long*>(reinterpret_cast<uint32_t*>(vis) + p);
}
// visibility resolve on one sample (global or shared memory)
__device__ __forceinline__ void vf_vis_update(unsigned long long* key64, bool depth, uint32_t prim, float zf) {
    if (depth) atomicMin(key64, ((unsigned long long)__float_as_uint(zf) << 32) | prim);
    else atomicMax(reinterpret_cast<uint32_t*>(key64), prim + 1u);   // low word (little endian)
}

// Work of the run rasteriser follows lines + covered samples, not the bbox: a needle 40 pixels across
// and 400 long with 20 pixels of area is cheap, a fat triangle of the same bbox is not.
__device__ __forceinline__ bool vf_serial_ok(const VfSubTri& t, int bw, int bh) {
    return min(bw, bh) <= VF_SERIAL_LINES && (-vf_area2(t) >> 17) <= (int64_t)VF_SERIAL_AREA;   // area2 is in (1/256 px)^2 * 2
}

// ---------------------------------------------------------------------------------------
// Thread-serial rasterisation of one triangle (edges set up at the bbox origin (x0, y0)).  The
// needles next to DEM outliers are a fraction of a pixel wide, up to hundreds of pixels long and
// tilted, so the work must follow the covered samples, not the bbox: the thread walks the lines of
// the SHORTER bbox axis; on each the covered run along the longer axis is bounded in float
// (crossing of each edge, margin 1/64 px: the float error is < 2^-8 px for crossings inside the
// bbox) and the exact test decides every sample of the run.
// emit(px, py, e) receives the BIASED edge values (vf_edges_unbias restores them).
template <typename Emit>
__device__ __forceinline__ void vf_thin_raster(const VfEdges& s, int x0, int x1, int y0, int y1, Emit&& emit) {
    const bool cols = (x1 - x0) <= (y1 - y0);        // minor axis = x: walk columns, runs along y
    const int nmin = cols ? x1 - x0 + 1 : y1 - y0 + 1;
    const int nmaj = cols ? y1 - y0 + 1 : x1 - x0 + 1;
    int32_t amaj[3];
    float rmaj[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        amaj[i] = cols ? s.B[i] : s.A[i];
        rmaj[i] = amaj[i] != 0 ? __fdividef(-1.0f, (float)amaj[i] * (float)VF_SUBPIX) : 0.0f;
    }
    for (int m = 0; m < nmin; ++m) {
        int64_t base[3];
        float lo = 0.0f, hi = (float)(nmaj - 1);
        bool dead = false;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            base[i] = s.E[i] + (int64_t)(cols ? s.A[i] : s.B[i]) * (int64_t)(m * VF_SUBPIX);
            const float tc = (float)base[i] * rmaj[i];           // base + amaj*256*t = 0
            if (amaj[i] > 0) lo = fmaxf(lo, tc);
            else if (amaj[i] < 0) hi = fminf(hi, tc);
            else dead = dead || base[i] < 0;
        }
        if (dead) continue;
        int c0 = (int)ceilf(lo - 0.015625f), c1 = (int)floorf(hi + 0.015625f);
        c0 = max(c0, 0); c1 = min(c1, nmaj - 1);
        for (int c = c0; c <= c1; ++c) {
            int64_t ed[3];
#pragma unroll
            for (int i = 0; i < 3; ++i) ed[i] = base[i] + (int64_t)amaj[i] * (int64_t)(c * VF_SUBPIX);
            if ((ed[0] | ed[1] | ed[2]) >= 0) emit(cols ? x0 + m : x0 + c, cols ? y0 + c : y0 + m, ed);
        }
    }
}

// ---------------------------------------------------------------------------------------
// The same walk with a 32-bit inner loop (k_geom_march).  Requires vf_serial_ok (|area2| < 2^29).
//   run bounds : crossing of edge i with line m in float, tc_i(m) = fma(m, slope_i, tc0_i); its error
//                stays below (|tc0_i| + nmin |slope_i|) 2^-20 px, which is added to the 1/64 px margin
//                of that edge, so the run [c0, c1] never loses a covered sample;
//   samples    : a sample of the run is within margin_i px (along the run) of satisfying every
//                non-flat edge, so E_i >= -margin_i 256 |amaj_i| > -2^27, and E_i = |area2| - E_j - E_k
//                < 2^31: the low 32 bits of the biased edge values carry their signs, and the test
//                runs on wrapped 32-bit sums (one IADD per edge and step);
//   flat edges : an edge parallel to the run (amaj == 0) is constant on a line; its exact 64-bit
//                sign is taken once per line and it is left out of the 32-bit test.
// emit(covered, idx, e0, e1, e2) is called for EVERY sample of a run (so that it can be a predicated store instead
// of a branch) with the sample's index in the W-wide frame and the low words of the three biased edge values: for a
// covered sample they ARE the values (0 <= E_i < 2^31), a flat edge included (constant along its line), which is what
// the depth interpolation needs.
template <typename Emit>
__device__ __forceinline__ void vf_thin_raster32(const VfEdges& s, int x0, int x1, int y0, int y1, uint32_t W, Emit&& emit) {
    const bool cols = (x1 - x0) <= (y1 - y0);        // minor axis = x: walk columns, runs along y
    const uint32_t istep = cols ? W : 1u;
    const int nmin = cols ? x1 - x0 + 1 : y1 - y0 + 1;
    const int nmaj = cols ? y1 - y0 + 1 : x1 - x0 + 1;
    uint32_t e32[3], smin[3], smaj[3];                // low words: value at (line 0, sample 0), step per line / per sample
    float tc0[3], slope[3];
    bool lower[3], upper[3], flat = false;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const int32_t amaj = cols ? s.B[i] : s.A[i], amin = cols ? s.A[i] : s.B[i];
        lower[i] = amaj > 0; upper[i] = amaj < 0;
        flat = flat || amaj == 0;
        const float r = amaj != 0 ? __fdividef(-1.0f, __fmul_rn((float)amaj, (float)VF_SUBPIX)) : 0.0f;
        const float t0 = __fmul_rn((float)s.E[i], r), sl = __fmul_rn(__fmul_rn((float)amin, (float)VF_SUBPIX), r);
        const float mg = fmaf(fmaf((float)nmin, fabsf(sl), fabsf(t0)), 0x1p-20f, 0.015625f);
        tc0[i] = lower[i] ? __fsub_rn(t0, mg) : __fadd_rn(t0, mg);
        slope[i] = sl;
        // a flat edge takes part too: on a line that is not dead it is >= 0 and, like the others, below 2^31 wherever
        // the run is not empty, so its low word is exact and cannot change the sign of the OR
        e32[i] = (uint32_t)s.E[i];
        smin[i] = (uint32_t)amin * (uint32_t)VF_SUBPIX;
        smaj[i] = (uint32_t)amaj * (uint32_t)VF_SUBPIX;
    }
    for (int m = 0; m < nmin; ++m) {
        const float fm = (float)m;
        float lo = 0.0f, hi = (float)(nmaj - 1);
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            const float tc = fmaf(fm, slope[i], tc0[i]);
            if (lower[i]) lo = fmaxf(lo, tc);
            if (upper[i]) hi = fminf(hi, tc);
        }
        if (flat) {
            bool dead = false;
#pragma unroll
            for (int i = 0; i < 3; ++i)
                if (!lower[i] && !upper[i])
                    dead = dead || s.E[i] + (int64_t)(cols ? s.A[i] : s.B[i]) * (int64_t)(m * VF_SUBPIX) < 0;
            if (dead) continue;
        }
        // (an empty run has c0 > c1; a sample tested in vain could not hurt: the test is exact)
        const int c0 = __float2int_ru(lo), c1 = __float2int_rd(hi);      // 0 <= lo, hi <= nmaj - 1
        uint32_t ed0 = e32[0] + smin[0] * (uint32_t)m + smaj[0] * (uint32_t)c0;
        uint32_t ed1 = e32[1] + smin[1] * (uint32_t)m + smaj[1] * (uint32_t)c0;
        uint32_t ed2 = e32[2] + smin[2] * (uint32_t)m + smaj[2] * (uint32_t)c0;
        // sample index in the frame, stepped along the run (W per sample down a column, 1 along a row)
        uint32_t idx = cols ? (uint32_t)(y0 + c0) * W + (uint32_t)(x0 + m) : (uint32_t)(y0 + m) * W + (uint32_t)(x0 + c0);
        int c = c0;
        emit(c <= c1 && (int32_t)(ed0 | ed1 | ed2) >= 0, idx, ed0, ed1, ed2);    // most runs hold one sample or none: no branch
        while (++c <= c1) {
            ed0 += smaj[0]; ed1 += smaj[1]; ed2 += smaj[2];
            idx += istep;
            emit((int32_t)(ed0 | ed1 | ed2) >= 0, idx, ed0, ed1, ed2);
        }
    }
}

// ---------------------------------------------------------------------------------------
struct VfGeomSmem {
    alignas(128) float dem[VF_TMA_BOX_Y][VF_TMA_BOX_X];   // TMA destination (DEM tile of this CTA)
    VfAxis ax[VF_GEOM_VX], az[VF_GEOM_VY];    // axis entries of the tile's columns / rows
    alignas(8) uint64_t mbar;
    int2 xy[VF_GEOM_VY][VF_GEOM_VX];          // snapped screen position (valid when fl & 64)
    float zn[VF_GEOM_VY][VF_GEOM_VX];         // z/w
    uint8_t fl[VF_GEOM_VY][VF_GEOM_VX + 3];   // clip flags
    uint16_t queue[VF_GEOM_TX * VF_GEOM_TY];  // cells that survived the quad-level cull
    uint16_t nqueue[2 * VF_GEOM_TX * VF_GEOM_TY];  // non-micro triangles (needles) for the run rasteriser
    uint32_t qcount, ncount;
};

// TMA = true: grid == DEM and the nearest-texel mapping is the identity (checked on the host side
// by k_axis_identity), so the CTA's 33x33 texels are one 2-D box: a single cp.async.bulk.tensor
// (UTMALDG) stages them in shared memory while the threads stage the axis entries.  Otherwise the
// texels are gathered through the axis tables with coalesced __ldg row loads.
template <bool TMA>
__global__ void __launch_bounds__(VF_GEOM_THREADS, 4)
k_geom(const __grid_constant__ VfFrame f, const __grid_constant__ CUtensorMap dem_map,
       const VfAxis* __restrict__ axis, const float* __restrict__ dem,
       unsigned long long* __restrict__ vis, uint32_t* __restrict__ big_prims, uint32_t big_capacity,
       uint32_t* __restrict__ counters) {
    __shared__ VfGeomSmem sm;
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    const uint32_t i0 = blockIdx.x * VF_GEOM_TX, j0 = blockIdx.y * VF_GEOM_TY;
    if (tid == 0) {
        sm.qcount = 0u; sm.ncount = 0u;
        if (TMA) {
            vf_mbar_init(&sm.mbar, 1u);
            vf_mbar_expect_tx(&sm.mbar, VF_TMA_BOX_X * VF_TMA_BOX_Y * 4u);
            vf_tma_load_2d(&sm.dem[0][0], &dem_map, (int)i0, (int)j0, &sm.mbar);   // OOB texels read 0
        }
    }
    if (tid < VF_GEOM_VX) sm.ax[tid] = axis[min(i0 + tid, f.G - 1u)];
    else if (tid >= 64u && tid < 64u + VF_GEOM_VY) sm.az[tid - 64u] = axis[min(j0 + tid - 64u, f.G - 1u)];
    __syncthreads();
    if (TMA) vf_mbar_wait(&sm.mbar, 0u);

    // ---- phase 1: every vertex of the 33x33 block once ----
    for (uint32_t v = tid; v < VF_GEOM_VX * VF_GEOM_VY; v += VF_GEOM_THREADS) {
        const uint32_t lj = v / VF_GEOM_VX, li = v - lj * VF_GEOM_VX;
        const uint32_t i = i0 + li, j = j0 + lj;
        int2 xy = make_int2(0, 0);
        float zn = 0.0f;
        uint32_t fl = 63u;
        if (i < f.G && j < f.G) {
            VfVtx o;
            const VfAxis ax = sm.ax[li], az = sm.az[lj];
            const float h_tex = TMA ? sm.dem[lj][li] : __ldg(dem + (size_t)az.ty * f.dem_pitch + ax.tx);
            vf_vertex_stage(f, ax.coord, az.coord, ax.uv, az.uv, h_tex, o);
            fl = vf_clip_flags(o.c);
            if (fl & 64u) {
                VfSVtx s;
                vf_to_screen(f, o, s);
                xy = make_int2(s.X, s.Y);
                zn = s.zn;
            }
        }
        sm.xy[lj][li] = xy;
        sm.zn[lj][li] = zn;
        sm.fl[lj][li] = (uint8_t)fl;
    }
    __syncthreads();

    // ---- phase 2: quad-level cull (one test for two triangles) ----
    const uint32_t li = tid & 31u;
#pragma unroll 1
    for (uint32_t r = 0; r < VF_GEOM_TY / (VF_GEOM_THREADS / 32); ++r) {
        const uint32_t lj = (tid >> 5) + r * (VF_GEOM_THREADS / 32);
        bool keep = (i0 + li + 1 < f.G) && (j0 + lj + 1 < f.G);
        if (keep) {
            const uint32_t fa = sm.fl[lj][li], fb = sm.fl[lj][li + 1];
            const uint32_t fc = sm.fl[lj + 1][li], fd = sm.fl[lj + 1][li + 1];
            const uint32_t fand = fa & fb & fc & fd;
            if (fand & 63u) {
                keep = false;                               // all four outside one frustum plane
            } else if (fand & 64u) {
                const int2 pa = sm.xy[lj][li], pb = sm.xy[lj][li + 1];
                const int2 pc = sm.xy[lj + 1][li], pd = sm.xy[lj + 1][li + 1];
                int x0, x1, y0, y1;                         // no pixel centre in the quad's bbox ->
                keep = vf_bbox_pixels(min(min(pa.x, pb.x), min(pc.x, pd.x)), max(max(pa.x, pb.x), max(pc.x, pd.x)),
                                      min(min(pa.y, pb.y), min(pc.y, pd.y)), max(max(pa.y, pb.y), max(pc.y, pd.y)),
                                      f.W, f.H, x0, x1, y0, y1);   // neither triangle can cover a sample
            }
        }
        const uint32_t m = __ballot_sync(0xffffffffu, keep);
        if (m) {
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(&sm.qcount, (uint32_t)__popc(m));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (keep) sm.queue[base + __popc(m & ((1u << lane) - 1u))] = (uint16_t)(lj * VF_GEOM_TX + li);
        }
    }
    __syncthreads();

    // ---- phase 3: surviving triangles, densely packed over the CTA ----
    const uint32_t n_entries = 2u * sm.qcount;
    const bool depth = f.depth != 0;
    // triangle k of the queued cell `slot`: T0 = (a, c, b), T1 = (b, c, d): src/terrain/mod.rs:581
    auto load_tri = [&](uint32_t slot, uint32_t k, VfSubTri& t) -> uint32_t {
        const uint32_t q = sm.queue[slot];
        const uint32_t qj = q / VF_GEOM_TX, qi = q - qj * VF_GEOM_TX;
        const uint32_t ai = k ? qi + 1 : qi, aj = qj, bi = qi, bj = qj + 1, ci = qi + 1, cj = k ? qj + 1 : qj;
        t.v[0].X = sm.xy[aj][ai].x; t.v[0].Y = sm.xy[aj][ai].y; t.v[0].zn = sm.zn[aj][ai];
        t.v[1].X = sm.xy[bj][bi].x; t.v[1].Y = sm.xy[bj][bi].y; t.v[1].zn = sm.zn[bj][bi];
        t.v[2].X = sm.xy[cj][ci].x; t.v[2].Y = sm.xy[cj][ci].y; t.v[2].zn = sm.zn[cj][ci];
        return 2u * ((j0 + qj) * (f.G - 1u) + (i0 + qi)) + k;
    };
    auto emit_sample = [&](const VfSubTri& t, const VfEdges& es, uint32_t prim, int px, int py, int64_t ed[3]) {
        float zf = 0.0f;
        if (depth) {
            float b[3];
            vf_edges_unbias(es, ed);
            vf_barycentric(ed, b);
            zf = vf_depth_at(t, b);
            if (!(zf < 1.0f)) return;
        }
        vf_vis_update(vf_vis_key(vis, depth, (size_t)py * f.W + px), depth, prim, zf);
    };
    for (uint32_t base = 0; base < n_entries; base += VF_GEOM_THREADS) {
        const uint32_t e = base + tid;
        bool to_big = false, to_thin = false;
        uint32_t prim = 0;
        if (e < n_entries) {
            const uint32_t q = sm.queue[e >> 1], k = e & 1u;
            const uint32_t qj = q / VF_GEOM_TX, qi = q - qj * VF_GEOM_TX;
            const uint32_t ai = k ? qi + 1 : qi, aj = qj, bi = qi, bj = qj + 1, ci = qi + 1, cj = k ? qj + 1 : qj;
            const uint32_t f0 = sm.fl[aj][ai], f1 = sm.fl[bj][bi], f2 = sm.fl[cj][ci];
            if (!(f0 & f1 & f2 & 63u)) {
                if (!(f0 & f1 & f2 & 64u)) {
                    prim = 2u * ((j0 + qj) * (f.G - 1u) + (i0 + qi)) + k;
                    to_big = true;                                  // needs clipping
                } else {
                    VfSubTri t;
                    prim = load_tri(e >> 1, k, t);
                    int x0, x1, y0, y1;
                    if (vf_subtri_bbox(t, f.W, f.H, x0, x1, y0, y1) && vf_area2(t) < 0) {
                        const int bw = x1 - x0 + 1, bh = y1 - y0 + 1, npx = bw * bh;
                        if (npx <= VF_MICRO_MAX) {
                            VfEdges es;
                            vf_edges_setup(t, x0, y0, es);
                            for (int dy = 0; dy < bh; ++dy)
                                for (int dx = 0; dx < bw; ++dx) {
                                    int64_t ed[3];
                                    if (vf_edges_at(es, dx, dy, ed)) emit_sample(t, es, prim, x0 + dx, y0 + dy, ed);
                                }
                        } else if (vf_serial_ok(t, bw, bh)) {
                            to_thin = true;
                        } else {
                            to_big = true;
                        }
                    }
                }
            }
        }
        const uint32_t below = (1u << lane) - 1u;
        const uint32_t nm = __ballot_sync(0xffffffffu, to_thin);
        if (nm) {
            uint32_t nb = 0;
            if (lane == 0) nb = atomicAdd(&sm.ncount, (uint32_t)__popc(nm));
            nb = __shfl_sync(0xffffffffu, nb, 0);
            if (to_thin) sm.nqueue[nb + __popc(nm & below)] = (uint16_t)e;
        }
        const uint32_t m = __ballot_sync(0xffffffffu, to_big);
        if (m) {
            uint32_t gb = 0;
            if (lane == 0) gb = atomicAdd(&counters[VF_CNT_BIG], (uint32_t)__popc(m));
            gb = __shfl_sync(0xffffffffu, gb, 0);
            const uint32_t pos = gb + __popc(m & below);
            if (to_big && pos < big_capacity) big_prims[pos] = prim;
        }
    }
    __syncthreads();

    // ---- phase 4: non-micro triangles (needles next to DEM outliers, coarse grids), one thread each, packed ----
    const uint32_t n_thin = sm.ncount;
    for (uint32_t w = tid; w < n_thin; w += VF_GEOM_THREADS) {
        const uint32_t e = sm.nqueue[w];                   // (queue slot << 1) | k
        VfSubTri t;
        const uint32_t prim = load_tri(e >> 1, e & 1u, t);
        int x0, x1, y0, y1;
        vf_subtri_bbox(t, f.W, f.H, x0, x1, y0, y1);
        VfEdges es;
        vf_edges_setup(t, x0, y0, es);
        vf_thin_raster(es, x0, x1, y0, y1, [&](int px, int py, int64_t ed[3]) { emit_sample(t, es, prim, px, py, ed); });
    }
}

// ---------------------------------------------------------------------------------------
// k_bin: one thread per queued primitive (large bbox or needs clipping).  Rebuilds the
// primitive, takes the union pixel bbox of its front-facing sub-triangles and appends one
// (tile, prim) pair per overlapped tile; positions come from a warp prefix sum + one atomic.
__device__ __forceinline__ void
vf_bin_phase(const VfFrame& f, const VfAxis* __restrict__ axis, const float* __restrict__ dem,
             const uint32_t* __restrict__ big_prims, uint32_t big_capacity,
             uint32_t* __restrict__ tile_count, uint32_t* __restrict__ pair_tile,
             uint32_t* __restrict__ pair_prim, uint32_t* __restrict__ counters, uint32_t capacity) {
    const uint32_t n = min(counters[VF_CNT_BIG], big_capacity);
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t n_round = (n + 31u) & ~31u;            // whole warps stay in the loop together
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n_round; p += stride) {
        int tx0 = 0, tx1 = -1, ty0 = 0, ty1 = -1;
        uint32_t prim = 0;
        bool live = false;
        if (p < n) {
            prim = big_prims[p];
            VfVtx tri[3];
            vf_prim_vertices(f, axis, dem, prim, tri);
            int bx0 = 1 << 30, bx1 = -1, by0 = 1 << 30, by1 = -1;
            const bool clipped = vf_for_each_subtri(f, tri, [&](const VfSubTri& t) {
                int x0, x1, y0, y1;
                if (vf_area2(t) >= 0) return;
                if (!vf_subtri_bbox(t, f.W, f.H, x0, x1, y0, y1)) return;
                bx0 = min(bx0, x0); bx1 = max(bx1, x1);
                by0 = min(by0, y0); by1 = max(by1, y1);
            });
            if (clipped) atomicAdd(&counters[VF_CNT_CLIPPED], 1u);
            if (bx1 >= bx0 && by1 >= by0) {
                live = true;
                tx0 = bx0 / VF_TILE; tx1 = bx1 / VF_TILE;
                ty0 = by0 / VF_TILE; ty1 = by1 / VF_TILE;
            }
        }
        const uint32_t nt = live ? (uint32_t)((tx1 - tx0 + 1) * (ty1 - ty0 + 1)) : 0u;
        if (__ballot_sync(0xffffffffu, nt != 0u) == 0u) continue;
        uint32_t incl = nt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= (uint32_t)d) incl += t;
        }
        const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(&counters[VF_CNT_PAIRS], total);
        base = __shfl_sync(0xffffffffu, base, 0);
        uint32_t pos = base + incl - nt;
        for (int ty = ty0; ty <= ty1; ++ty)
            for (int tx = tx0; tx <= tx1; ++tx, ++pos) {
                if (pos < capacity) {
                    const uint32_t tile = (uint32_t)ty * f.tiles_x + (uint32_t)tx;
                    pair_tile[pos] = tile;
                    pair_prim[pos] = prim;
                    atomicAdd(&tile_count[tile], 1u);
                }
            }
    }
}

// ---------------------------------------------------------------------------------------
// exclusive prefix sum of tile_count[0..n) -> tile_off[0..n]; zeroes tile_cursor.  One CTA of
// NT threads.  Also records the high-water marks (pairs, queued primitives) used for overflow
// detection.
template <uint32_t NT>
__device__ __forceinline__ void
vf_scan_phase(const uint32_t* __restrict__ tile_count, uint32_t* __restrict__ tile_off,
              uint32_t* __restrict__ tile_cursor, uint32_t n, const uint32_t* __restrict__ counters,
              uint32_t* __restrict__ high_water) {
    __shared__ uint32_t warp_sum[32];
    __shared__ uint32_t carry;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, wid = tid >> 5;
    if (tid == 0) {
        carry = 0;
        atomicMax(&high_water[0], __ldcg(&counters[VF_CNT_PAIRS]));
        atomicMax(&high_water[1], __ldcg(&counters[VF_CNT_BIG]));
    }
    __syncthreads();
    if (__ldcg(&counters[VF_CNT_PAIRS]) == 0u) {                    // nothing binned: all offsets are zero
        for (uint32_t idx = tid; idx <= n; idx += NT) tile_off[idx] = 0u;
        return;
    }
    for (uint32_t base = 0; base < n; base += NT) {
        const uint32_t idx = base + tid;
        const uint32_t v = idx < n ? tile_count[idx] : 0u;
        uint32_t incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= (uint32_t)d) incl += t;
        }
        if (lane == 31) warp_sum[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            uint32_t w = lane < NT / 32u ? warp_sum[lane] : 0u;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, w, d);
                if (lane >= (uint32_t)d) w += t;
            }
            warp_sum[lane] = w;
        }
        __syncthreads();
        const uint32_t prev_warps = wid ? warp_sum[wid - 1] : 0u;
        const uint32_t c = carry;
        if (idx < n) {
            tile_off[idx] = c + prev_warps + incl - v;
            tile_cursor[idx] = 0u;
        }
        __syncthreads();
        if (tid == NT - 1u) carry = c + prev_warps + incl;
        __syncthreads();
    }
    if (tid == 0) tile_off[n] = carry;
}

__device__ __forceinline__ void
vf_scatter_phase(const uint32_t* __restrict__ pair_tile, const uint32_t* __restrict__ pair_prim,
                 const uint32_t* __restrict__ counters, uint32_t capacity,
                 const uint32_t* __restrict__ tile_off, uint32_t* __restrict__ tile_cursor,
                 uint32_t* __restrict__ sorted_prim) {
    const uint32_t n = min(__ldcg(&counters[VF_CNT_PAIRS]), capacity);
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
        const uint32_t tile = pair_tile[p];
        const uint32_t slot = atomicAdd(&tile_cursor[tile], 1u);
        sorted_prim[tile_off[tile] + slot] = pair_prim[p];
    }
}

// ---------------------------------------------------------------------------------------
// k_binsort: the sort-middle stage for the queued (large / clipped) primitives as ONE launch:
// bin -> grid barrier -> per-tile prefix sum (CTA 0) -> grid barrier -> scatter.  The grid is
// launched cooperatively (every CTA resident), the barrier is a monotonic counter in counters[].
// With an empty queue (fine grids seen whole) the three phases are a few loads each.
__device__ __forceinline__ void vf_grid_barrier(uint32_t* bar, uint32_t target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(bar, 1u);
        while (*reinterpret_cast<volatile uint32_t*>(bar) < target) __nanosleep(32);
        __threadfence();
    }
    __syncthreads();
}

__global__ void __launch_bounds__(256)
k_binsort(const __grid_constant__ VfFrame f, const VfAxis* __restrict__ axis, const float* __restrict__ dem,
          const uint32_t* __restrict__ big_prims, uint32_t big_capacity,
          uint32_t* __restrict__ tile_count, uint32_t* __restrict__ tile_off, uint32_t* __restrict__ tile_cursor,
          uint32_t* __restrict__ pair_tile, uint32_t* __restrict__ pair_prim, uint32_t* __restrict__ sorted_prim,
          uint32_t* __restrict__ counters, uint32_t capacity, uint32_t* __restrict__ high_water) {
    if (__ldcg(&counters[VF_CNT_BIG]) == 0u) return;     // empty queue (same answer in every CTA): k_raster sees no pairs
    vf_bin_phase(f, axis, dem, big_prims, big_capacity, tile_count, pair_tile, pair_prim, counters, capacity);
    vf_grid_barrier(&counters[VF_CNT_GRIDBAR], gridDim.x);
    if (blockIdx.x == 0) vf_scan_phase<256>(tile_count, tile_off, tile_cursor, f.tiles_x * f.tiles_y, counters, high_water);
    vf_grid_barrier(&counters[VF_CNT_GRIDBAR], 2u * gridDim.x);
    vf_scatter_phase(pair_tile, pair_prim, counters, capacity, tile_off, tile_cursor, sorted_prim);
}

// ---------------------------------------------------------------------------------------
// k_raster: one CTA per 16x16 screen tile, 256 threads.
//   merge   : the tile's slice of the visibility buffer (written by the geometry pass) is loaded and reset for the
//             next frame; if primitives were binned for this tile (large / clipped ones: rare), they are rebuilt and
//             resolved in shared memory first (vf_raster_merge, out of line) -- atomicMax(prim+1) (no depth: last
//             drawn wins, src/terrain/pipeline.rs:133) or atomicMin((depth<<32)|prim) (depth Less);
//   shading : thread = pixel.  The winning primitive's three vertices come straight from the vertex-height image and
//             the axis table (no clip machinery, no stack: a winner that does need clipping goes to vf_shade_clipped,
//             out of line), then perspective-correct varyings -- only the three the shading mode reads --, the fragment
//             stage, exact sRGB8 from an approximate guess fixed by the threshold table; 4 pixels per 16-byte store.
#ifndef VF_RASTER_CTAS
#define VF_RASTER_CTAS 6      // resident CTAs per SM the register budget is set for (4: 0.031 ms, 6: 0.025 ms, 8: 0.025 ms at c3)
#endif

static __device__ __noinline__ void
vf_raster_merge(const VfFrame& f, const VfAxis* __restrict__ axis, const float* __restrict__ dem,
                const uint32_t* __restrict__ sorted_prim, uint32_t beg, uint32_t end, int px0, int py0,
                unsigned long long* s_key) {
    const bool depth = f.depth != 0;
    for (uint32_t k = beg + threadIdx.x; k < end; k += 256) {
        const uint32_t prim = sorted_prim[k];
        VfVtx tri[3];
        vf_prim_vertices(f, axis, dem, prim, tri);
        vf_for_each_subtri(f, tri, [&](const VfSubTri& t) {
            if (vf_area2(t) >= 0) return;
            int x0, x1, y0, y1;
            if (!vf_subtri_bbox(t, f.W, f.H, x0, x1, y0, y1)) return;
            x0 = max(x0, px0); x1 = min(x1, px0 + VF_TILE - 1);
            y0 = max(y0, py0); y1 = min(y1, py0 + VF_TILE - 1);
            for (int qy = y0; qy <= y1; ++qy)
                for (int qx = x0; qx <= x1; ++qx) {
                    int64_t e[3];
                    if (!vf_covers(t, qx, qy, e)) continue;
                    float zf = 0.0f;
                    if (depth) {
                        float b[3];
                        vf_barycentric(e, b);
                        zf = vf_depth_at(t, b);
                        if (!(zf < 1.0f)) continue;
                    }
                    vf_vis_update(&s_key[(qy - py0) * VF_TILE + (qx - px0)], depth, prim, zf);
                }
        });
    }
}

// perspective-correct varyings of pixel (px, py) inside screen triangle t, then the fragment stage
template <int MODE>
__device__ __forceinline__ uint32_t vf_shade_at(const VfFrame& f, const VfSubTri& t, const int64_t e[3], const float* __restrict__ dem,
                                                const float4* __restrict__ lut, const float* __restrict__ thr) {
    float b[3], a[5];
    vf_barycentric(e, b);
    const float iw = fmaf(b[2], t.v[2].invw, fmaf(b[1], t.v[1].invw, __fmul_rn(b[0], t.v[0].invw)));
    // NORTHSTAR reads u, v, h; SHIPPED reads h, x, z (vf_fragment_stage): the others are never computed
    constexpr uint32_t used = MODE == 1 ? 0x07u : MODE == 0 ? 0x1cu : 0x1fu;
    float num[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        a[k] = 0.0f; num[k] = 0.0f;
        if (!((used >> k) & 1u)) continue;
        num[k] = fmaf(b[2], __fmul_rn(t.v[2].a[k], t.v[2].invw),
                 fmaf(b[1], __fmul_rn(t.v[1].a[k], t.v[1].invw),
                      __fmul_rn(b[0], __fmul_rn(t.v[0].a[k], t.v[0].invw))));
    }
    vf_div_by<5>(num, iw, a, used);                      // a[k] = num[k] / iw
    return vf_fragment_stage<MODE, true>(f, a, dem, lut, thr);
}

// the winner needs clipping (a vertex outside near / far / the guard band): the general path, kept out of line
template <int MODE>
static __device__ __noinline__ uint32_t
vf_shade_clipped(const VfFrame& f, const VfAxis* __restrict__ axis, const float* __restrict__ dem, const float4* __restrict__ lut,
                 const float* __restrict__ thr, uint32_t prim, int px, int py, uint32_t clear_rgba) {
    VfVtx tri[3];
    vf_prim_vertices(f, axis, dem, prim, tri);
    uint32_t rgba = clear_rgba;
    bool done = false;
    vf_for_each_subtri(f, tri, [&](const VfSubTri& t) {
        if (done || vf_area2(t) >= 0) return;
        int64_t e[3];
        if (!vf_covers(t, px, py, e)) return;
        rgba = vf_shade_at<MODE>(f, t, e, dem, lut, thr);
        done = true;
    });
    return rgba;
}

// Shading of a pixel from its winning primitive, in three steps so that k_raster_lean can fetch the heights ahead of time.
// cell of a primitive -> (column i, row j) without a division
__device__ __forceinline__ void vf_winner_cell(const VfFrame& f, uint32_t prim, uint32_t& i, uint32_t& j) {
    const uint32_t cell = prim >> 1, gm = f.G - 1u;
    j = __umulhi(cell, f.gm_inv); i = cell - j * gm;            // j <= cell / gm <= j + 2
    if (i >= gm) { i -= gm; ++j; }
    if (i >= gm) { i -= gm; ++j; }
}
// T0 = (a, c, b), T1 = (b, c, d) (src/terrain/mod.rs:575-583): vertex 0 = (i + k, j), 1 = (i, j + 1), 2 = (i + 1, j + k).
// Offsets of the three vertex heights in the G x G vertex-height image (pitch in floats).
__device__ __forceinline__ void vf_winner_height_offsets(uint32_t i, uint32_t j, uint32_t k, uint32_t hpitch, size_t off[3]) {
    const size_t o = (size_t)j * hpitch + i;
    off[0] = o + k; off[1] = o + hpitch; off[2] = o + (k ? hpitch + 1u : 1u);
}
// the same heights gathered from the DEM through the axis table (no vertex-height image)
__device__ __forceinline__ void vf_winner_heights_dem(const VfFrame& f, const VfAxis* __restrict__ axis, const float* __restrict__ dem,
                                                      uint32_t i, uint32_t j, uint32_t k, float h[3]) {
    const int tx0 = axis[i].tx, tx1 = axis[i + 1u].tx, ty0 = axis[j].ty, ty1 = axis[j + 1u].ty;
    h[0] = __ldg(dem + (size_t)ty0 * f.dem_pitch + (k ? tx1 : tx0));
    h[1] = __ldg(dem + (size_t)ty1 * f.dem_pitch + tx0);
    h[2] = __ldg(dem + (size_t)(k ? ty1 : ty0) * f.dem_pitch + tx1);
}
// The winner's three vertices from their heights and the axis table (no clip machinery, no stack), then the pixel's colour.
// false (rgba untouched): a vertex lies outside near / far / the guard band, the triangle needs the clipper.
template <int MODE>
__device__ __forceinline__ bool
vf_shade_winner_h(const VfFrame& f, const VfAxis* __restrict__ axis, const float* __restrict__ dem, const float4* __restrict__ lut,
                  const float* __restrict__ thr, uint32_t i, uint32_t j, uint32_t k, const float h[3], int px, int py, uint32_t& rgba) {
    const float2 ax0 = __ldg(reinterpret_cast<const float2*>(axis + i)), ax1 = __ldg(reinterpret_cast<const float2*>(axis + i + 1u));
    const float2 az0 = __ldg(reinterpret_cast<const float2*>(axis + j)), az1 = __ldg(reinterpret_cast<const float2*>(axis + j + 1u));
    VfVtx tri[3];
    const float2 axk = k ? ax1 : ax0, azk = k ? az1 : az0;
    vf_vertex_stage<MODE>(f, axk.x, az0.x, axk.y, az0.y, h[0], tri[0]);
    vf_vertex_stage<MODE>(f, ax0.x, az1.x, ax0.y, az1.y, h[1], tri[1]);
    vf_vertex_stage<MODE>(f, ax1.x, azk.x, ax1.y, azk.y, h[2], tri[2]);
    const uint32_t f0 = vf_clip_flags(tri[0].c), f1 = vf_clip_flags(tri[1].c), f2 = vf_clip_flags(tri[2].c);
    if (!(f0 & f1 & f2 & 64u)) return false;             // not every vertex inside near / far / the guard band
    VfSubTri t;
    const float w3[3] = {tri[0].c[3], tri[1].c[3], tri[2].c[3]};
    float iw3[3];
    vf_rcp3(w3, iw3);
#pragma unroll
    for (int v = 0; v < 3; ++v) vf_to_screen_w(f, tri[v], iw3[v], t.v[v]);
    const int32_t sx = px * VF_SUBPIX + VF_HALFPIX, sy = py * VF_SUBPIX + VF_HALFPIX;
    const int64_t e[3] = {vf_edge(t.v[1], t.v[2], sx, sy), vf_edge(t.v[2], t.v[0], sx, sy), vf_edge(t.v[0], t.v[1], sx, sy)};
    rgba = vf_shade_at<MODE>(f, t, e, dem, lut, thr);
    return true;
}
template <int MODE>
__device__ __forceinline__ bool
vf_shade_winner(const VfFrame& f, const VfAxis* __restrict__ axis, const float* __restrict__ dem, const float* __restrict__ hgrid,
                uint32_t hpitch, const float4* __restrict__ lut, const float* __restrict__ thr, uint32_t prim, int px, int py,
                uint32_t& rgba) {
    uint32_t i, j;
    const uint32_t k = prim & 1u;
    vf_winner_cell(f, prim, i, j);
    float h[3];
    if (hgrid != nullptr) {
        size_t off[3];
        vf_winner_height_offsets(i, j, k, hpitch, off);
        h[0] = __ldg(hgrid + off[0]); h[1] = __ldg(hgrid + off[1]); h[2] = __ldg(hgrid + off[2]);
    } else {
        vf_winner_heights_dem(f, axis, dem, i, j, k, h);
    }
    return vf_shade_winner_h<MODE>(f, axis, dem, lut, thr, i, j, k, h, px, py, rgba);
}

// hgrid: the G x G image of vertex heights (vf_terrain.cu: the resampled image, or the DEM itself when the grid maps
// to it one to one); nullptr: heights are gathered from the DEM through the axis table
template <int MODE, bool DEPTH>
__global__ void __launch_bounds__(256, VF_RASTER_CTAS)
k_raster(const __grid_constant__ VfFrame f, const VfAxis* __restrict__ axis,
         const float* __restrict__ dem, const float* __restrict__ hgrid, uint32_t hpitch, const float4* __restrict__ lut,
         const float* __restrict__ thr_g, unsigned long long* __restrict__ vis,
         const uint32_t* __restrict__ tile_off, const uint32_t* __restrict__ sorted_prim,
         const uint32_t* __restrict__ counters, uint32_t clear_rgba, uint8_t* __restrict__ out_rgba,
         uint32_t* __restrict__ out_vis, uint32_t* __restrict__ high_water) {
    __shared__ unsigned long long s_key[VF_TILE * VF_TILE];
    __shared__ float s_thr[257];
    const uint32_t tid = threadIdx.x;
    // high_water[2]: most primitives any frame of the batch queued for binning.  The host launches k_binsort
    // speculatively NOT at all while frames queue nothing (vf_terrain.cu); this is how it learns otherwise.
    if ((blockIdx.x | blockIdx.y | tid) == 0u) {
        const uint32_t nbig = counters[VF_CNT_BIG];
        if (nbig) atomicMax(&high_water[2], nbig);
    }
    unsigned long long key;
    {
        const uint32_t tyi = blockIdx.y, txi = blockIdx.x;  // grid = tiles_x x tiles_y
        const uint32_t tile = tyi * f.tiles_x + txi;
        const int px0 = (int)txi * VF_TILE, py0 = (int)tyi * VF_TILE;
        const bool binned = counters[VF_CNT_PAIRS] != 0u;  // no pairs: k_binsort left tile_off untouched
        const uint32_t beg = binned ? tile_off[tile] : 0u, end = binned ? tile_off[tile + 1] : 0u;
        const unsigned long long empty = DEPTH ? ~0ull : 0ull;
        const int px = px0 + (int)(tid & 15u), py = py0 + (int)(tid >> 4);
        s_thr[tid] = thr_g[tid];
        if (tid == 0u) s_thr[256] = __int_as_float(0x7f800000);
        key = empty;
        if (px < (int)f.W && py < (int)f.H) {               // self-cleaning: the key is reset for the next frame
            if (DEPTH) {
                unsigned long long* g = vis + (size_t)py * f.W + px;
                key = *g;
                if (key != empty) *g = empty;
            } else {
                uint32_t* g = reinterpret_cast<uint32_t*>(vis) + (size_t)py * f.W + px;
                key = *g;
                if (key != 0ull) *g = 0u;
            }
        }
        if (beg != end) {                                   // block-uniform, rare
            s_key[tid] = key;
            __syncthreads();
            vf_raster_merge(f, axis, dem, sorted_prim, beg, end, px0, py0, s_key);
            __syncthreads();
            uint32_t t2;                                    // (nothing but the key's address is needed after the call, and it is
            asm volatile("mov.u32 %0, %%tid.x;" : "=r"(t2));   // re-derived: no register has to survive the call)
            key = s_key[t2];
        } else {
            __syncthreads();                                // s_thr
        }
    }
    // Thread and tile indices are read again here (volatile: not merged with the reads above), so that the values
    // above die before the out-of-line call instead of being spilled around it on every path.
    uint32_t tid_b, ctax, ctay;
    asm volatile("mov.u32 %0, %%tid.x;" : "=r"(tid_b));
    asm volatile("mov.u32 %0, %%ctaid.x;" : "=r"(ctax));
    asm volatile("mov.u32 %0, %%ctaid.y;" : "=r"(ctay));
    const int px = (int)(ctax * VF_TILE + (tid_b & 15u)), py = (int)(ctay * VF_TILE + (tid_b >> 4));
    const bool inb = px < (int)f.W && py < (int)f.H;

    uint32_t id1;
    if (DEPTH) id1 = key == ~0ull ? 0u : (uint32_t)(key & 0xffffffffull) + 1u;
    else id1 = (uint32_t)key;
    uint32_t rgba = clear_rgba;
    if (id1 != 0u && inb) {
        const uint32_t prim = id1 - 1u;
        if (!vf_shade_winner<MODE>(f, axis, dem, hgrid, hpitch, lut, s_thr, prim, px, py, rgba))
            rgba = vf_shade_clipped<MODE>(f, axis, dem, lut, s_thr, prim, px, py, clear_rgba);
    }
    if (out_vis && inb) out_vis[(size_t)py * f.W + px] = id1;

    // 4 horizontally adjacent pixels -> one 16-byte store (a tile row is 64 contiguous bytes)
    const uint32_t r1 = __shfl_down_sync(0xffffffffu, rgba, 1);
    const uint32_t r2 = __shfl_down_sync(0xffffffffu, rgba, 2);
    const uint32_t r3 = __shfl_down_sync(0xffffffffu, rgba, 3);
    if ((f.W & 3u) == 0u) {
        if ((tid_b & 3u) == 0u && inb)
            *reinterpret_cast<uint4*>(out_rgba + ((size_t)py * f.W + px) * 4) = make_uint4(rgba, r1, r2, r3);
    } else if (inb) {
        *reinterpret_cast<uint32_t*>(out_rgba + ((size_t)py * f.W + px) * 4) = rgba;
    }
}
