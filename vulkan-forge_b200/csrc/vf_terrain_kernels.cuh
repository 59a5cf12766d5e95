// vf_terrain_kernels.cuh -- the terrain frame as four sm_100a kernels:
//   k_axis    : per-index tables of the implicit grid (x|z, u|v, texel column/row)
//   k_geom    : vertex stage + cull + tile binning (warp-scan compaction)      [dominant]
//   k_scan / k_scatter : per-tile prefix sum and scatter of (tile, primitive) pairs
//   k_raster  : per-tile coverage (integer edge functions, top-left), on-chip visibility /
//               depth resolve, fused fragment stage, sRGB encode, 16-byte RGBA8 stores
// replaces: vs_main/fs_main of src/shaders/terrain.wgsl and the fixed-function state of
// src/terrain/pipeline.rs:107-140 (see DESIGN.md for the data layout and rooflines).
#pragma once
#include "vf_device.cuh"

struct VfAxis { float coord; float uv; int tx; int ty; };   // 16 B per grid index

#define VF_GEOM_TX 32          // cells per CTA in x
#define VF_GEOM_TY 32          // cells per CTA in y
#define VF_GEOM_THREADS 256
#define VF_GEOM_VX (VF_GEOM_TX + 1)
#define VF_GEOM_VY (VF_GEOM_TY + 1)

// counters[0] = pairs requested, [1] = clipped primitives, [2] = pairs stored
enum { VF_CNT_PAIRS = 0, VF_CNT_CLIPPED = 1, VF_CNT_WORDS = 4 };

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

// ---------------------------------------------------------------------------------------
// k_geom: one CTA = 32x32 cells (33x33 vertices).  Phase 1 transforms the vertices once and
// keeps only what culling needs in shared memory (snapped XY + clip flags); phase 2 gives every
// thread 4 cells (8 triangles), rejects triangles that are outside, back-facing or contain no
// pixel centre, and appends (tile, primitive) pairs for the survivors with one atomic per warp.
struct VfGeomSmem {
    int2 xy[VF_GEOM_VY][VF_GEOM_VX];
    uint8_t fl[VF_GEOM_VY][VF_GEOM_VX + 3];
};

// Tile range of one triangle in the fast path.  Returns false if culled.
__device__ __forceinline__ bool vf_tri_tiles_fast(const VfFrame& f, int2 p0, int2 p1, int2 p2,
                                                  int& tx0, int& tx1, int& ty0, int& ty1) {
    int x0, x1, y0, y1;
    const int32_t mnx = min(p0.x, min(p1.x, p2.x)), mxx = max(p0.x, max(p1.x, p2.x));
    const int32_t mny = min(p0.y, min(p1.y, p2.y)), mxy = max(p0.y, max(p1.y, p2.y));
    if (!vf_bbox_pixels(mnx, mxx, mny, mxy, f.W, f.H, x0, x1, y0, y1)) return false;
    const int64_t area2 = (int64_t)(p1.x - p0.x) * (p2.y - p0.y) - (int64_t)(p1.y - p0.y) * (p2.x - p0.x);
    if (area2 >= 0) return false;                              // cull Back / degenerate
    tx0 = x0 / VF_TILE; tx1 = x1 / VF_TILE; ty0 = y0 / VF_TILE; ty1 = y1 / VF_TILE;
    return true;
}

__global__ void __launch_bounds__(VF_GEOM_THREADS)
k_geom(const __grid_constant__ VfFrame f, const VfAxis* __restrict__ axis,
       const float* __restrict__ dem, uint32_t* __restrict__ tile_count,
       uint32_t* __restrict__ pair_tile, uint32_t* __restrict__ pair_prim,
       uint32_t* __restrict__ counters, uint32_t capacity) {
    __shared__ VfGeomSmem sm;
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    const uint32_t i0 = blockIdx.x * VF_GEOM_TX, j0 = blockIdx.y * VF_GEOM_TY;

    // ---- phase 1: vertices ----
    for (uint32_t v = tid; v < VF_GEOM_VX * VF_GEOM_VY; v += VF_GEOM_THREADS) {
        const uint32_t lj = v / VF_GEOM_VX, li = v - lj * VF_GEOM_VX;
        const uint32_t i = i0 + li, j = j0 + lj;
        int2 xy = make_int2(0, 0);
        uint32_t fl = 63u;
        if (i < f.G && j < f.G) {
            VfVtx o;
            vf_grid_vertex(f, axis, dem, i, j, o);
            fl = vf_clip_flags(o.c);
            if (fl & 64u) {
                VfSVtx s;
                vf_to_screen(f, o, s);
                xy = make_int2(s.X, s.Y);
            }
        }
        sm.xy[lj][li] = xy;
        sm.fl[lj][li] = (uint8_t)fl;
    }
    __syncthreads();

    // ---- phase 2: cells ----
    const uint32_t li = tid & 31u;
#pragma unroll 1
    for (uint32_t r = 0; r < VF_GEOM_TY / (VF_GEOM_THREADS / 32); ++r) {
        const uint32_t lj = (tid >> 5) + r * (VF_GEOM_THREADS / 32);
        const uint32_t i = i0 + li, j = j0 + lj;
        const bool valid = (i + 1 < f.G) && (j + 1 < f.G);
        const int2 pa = sm.xy[lj][li], pb = sm.xy[lj][li + 1];
        const int2 pc = sm.xy[lj + 1][li], pd = sm.xy[lj + 1][li + 1];
        const uint32_t fa = sm.fl[lj][li], fb = sm.fl[lj][li + 1];
        const uint32_t fc = sm.fl[lj + 1][li], fd = sm.fl[lj + 1][li + 1];
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const uint32_t prim = 2u * (j * (f.G - 1u) + i) + (uint32_t)k;
            const uint32_t f0 = k == 0 ? fa : fb, f1 = fc, f2 = k == 0 ? fb : fd;
            const int2 p0 = k == 0 ? pa : pb, p1 = pc, p2 = k == 0 ? pb : pd;
            int tx0 = 0, tx1 = -1, ty0 = 0, ty1 = -1;
            bool live = false;
            if (valid && !(f0 & f1 & f2 & 63u)) {
                if (f0 & f1 & f2 & 64u) {
                    live = vf_tri_tiles_fast(f, p0, p1, p2, tx0, tx1, ty0, ty1);
                } else {
                    // clip path: redo the vertex stage (clip coordinates were not kept)
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
            }
            // ---- warp-aggregated append of (tile, prim) pairs ----
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
            if (nt) {
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
    }
}

// ---------------------------------------------------------------------------------------
// exclusive prefix sum of tile_count[0..n) -> tile_off[0..n]; zeroes tile_cursor. One CTA.
__global__ void __launch_bounds__(1024)
k_scan(const uint32_t* __restrict__ tile_count, uint32_t* __restrict__ tile_off,
       uint32_t* __restrict__ tile_cursor, uint32_t n, const uint32_t* __restrict__ counters,
       uint32_t* __restrict__ max_pairs) {
    __shared__ uint32_t warp_sum[32];
    __shared__ uint32_t carry;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, wid = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n; base += 1024) {
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
            uint32_t w = warp_sum[lane];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, w, d);
                if (lane >= (uint32_t)d) w += t;
            }
            warp_sum[lane] = w;          // inclusive over warps
        }
        __syncthreads();
        const uint32_t prev_warps = wid ? warp_sum[wid - 1] : 0u;
        const uint32_t c = carry;
        if (idx < n) {
            tile_off[idx] = c + prev_warps + incl - v;
            tile_cursor[idx] = 0u;
        }
        __syncthreads();
        if (tid == 1023) carry = c + prev_warps + incl;
        __syncthreads();
    }
    if (tid == 0) {
        tile_off[n] = carry;
        atomicMax(max_pairs, counters[VF_CNT_PAIRS]);   // high-water mark over a batch
    }
}

__global__ void k_scatter(const uint32_t* __restrict__ pair_tile, const uint32_t* __restrict__ pair_prim,
                          const uint32_t* __restrict__ counters, uint32_t capacity,
                          const uint32_t* __restrict__ tile_off, uint32_t* __restrict__ tile_cursor,
                          uint32_t* __restrict__ sorted_prim) {
    const uint32_t n = min(counters[VF_CNT_PAIRS], capacity);
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
        const uint32_t tile = pair_tile[p];
        const uint32_t slot = atomicAdd(&tile_cursor[tile], 1u);
        sorted_prim[tile_off[tile] + slot] = pair_prim[p];
    }
}

// ---------------------------------------------------------------------------------------
// k_raster: one CTA per 16x16 screen tile, 256 threads.
//   coverage phase: thread k takes the k-th primitive of the tile's list, rebuilds its
//   screen-space triangle(s) with the same arithmetic as k_geom and resolves visibility in
//   shared memory -- atomicMax(prim+1) (no depth: last drawn wins, src/terrain/pipeline.rs:133)
//   or atomicMin((depth<<32)|prim) (depth Less).
//   shading phase: thread = pixel; rebuilds the winning primitive, perspective-correct
//   varyings, fragment stage, exact sRGB8; 4 pixels are packed per 16-byte store.
__global__ void __launch_bounds__(256)
k_raster(const __grid_constant__ VfFrame f, const VfAxis* __restrict__ axis,
         const float* __restrict__ dem, const float4* __restrict__ lut,
         const float* __restrict__ thr_g, const uint32_t* __restrict__ tile_off,
         const uint32_t* __restrict__ sorted_prim, uint32_t clear_rgba,
         uint8_t* __restrict__ out_rgba, uint32_t* __restrict__ out_vis) {
    __shared__ unsigned long long s_key[VF_TILE * VF_TILE];
    __shared__ float s_thr[256];
    const uint32_t tid = threadIdx.x;
    const uint32_t tile = blockIdx.x;
    const uint32_t tyi = tile / f.tiles_x, txi = tile - tyi * f.tiles_x;
    const int px0 = (int)txi * VF_TILE, py0 = (int)tyi * VF_TILE;
    const uint32_t beg = tile_off[tile], end = tile_off[tile + 1];
    const bool depth = f.depth != 0;

    s_thr[tid] = thr_g[tid];
    s_key[tid] = depth ? ~0ull : 0ull;
    __syncthreads();

    for (uint32_t k = beg + tid; k < end; k += 256) {
        const uint32_t prim = sorted_prim[k];
        VfVtx tri[3];
        vf_prim_vertices(f, axis, dem, prim, tri);
        vf_for_each_subtri(f, tri, [&](const VfSubTri& t) {
            if (vf_area2(t) >= 0) return;
            int x0, x1, y0, y1;
            if (!vf_subtri_bbox(t, f.W, f.H, x0, x1, y0, y1)) return;
            x0 = max(x0, px0); x1 = min(x1, px0 + VF_TILE - 1);
            y0 = max(y0, py0); y1 = min(y1, py0 + VF_TILE - 1);
            for (int py = y0; py <= y1; ++py)
                for (int px = x0; px <= x1; ++px) {
                    int64_t e[3];
                    if (!vf_covers(t, px, py, e)) continue;
                    const int idx = (py - py0) * VF_TILE + (px - px0);
                    if (depth) {
                        float b[3];
                        vf_barycentric(e, b);
                        const float zf = vf_depth_at(t, b);
                        if (!(zf < 1.0f)) continue;
                        atomicMin(&s_key[idx], ((unsigned long long)__float_as_uint(zf) << 32) | prim);
                    } else {
                        atomicMax(&s_key[idx], (unsigned long long)(prim + 1u));
                    }
                }
        });
    }
    __syncthreads();

    const int lx = (int)(tid & 15u), ly = (int)(tid >> 4);
    const int px = px0 + lx, py = py0 + ly;
    const unsigned long long key = s_key[tid];
    uint32_t id1;
    if (depth) id1 = key == ~0ull ? 0u : (uint32_t)(key & 0xffffffffull) + 1u;
    else id1 = (uint32_t)key;
    uint32_t rgba = clear_rgba;
    const bool inb = px < (int)f.W && py < (int)f.H;
    if (id1 != 0u && inb) {
        VfVtx tri[3];
        vf_prim_vertices(f, axis, dem, id1 - 1u, tri);
        bool done = false;
        vf_for_each_subtri(f, tri, [&](const VfSubTri& t) {
            if (done || vf_area2(t) >= 0) return;
            int64_t e[3];
            if (!vf_covers(t, px, py, e)) return;
            float b[3], a[5];
            vf_barycentric(e, b);
            const float iw = fmaf(b[2], t.v[2].invw, fmaf(b[1], t.v[1].invw, __fmul_rn(b[0], t.v[0].invw)));
#pragma unroll
            for (int k = 0; k < 5; ++k) {
                const float num = fmaf(b[2], __fmul_rn(t.v[2].a[k], t.v[2].invw),
                                  fmaf(b[1], __fmul_rn(t.v[1].a[k], t.v[1].invw),
                                       __fmul_rn(b[0], __fmul_rn(t.v[0].a[k], t.v[0].invw))));
                a[k] = __fdiv_rn(num, iw);
            }
            rgba = vf_fragment_stage(f, a, dem, lut, s_thr);
            done = true;
        });
    }
    if (out_vis && inb) out_vis[(size_t)py * f.W + px] = id1;

    // 4 horizontally adjacent pixels -> one 16-byte store (rows are 16 px = 64 B per tile)
    const uint32_t r1 = __shfl_down_sync(0xffffffffu, rgba, 1);
    const uint32_t r2 = __shfl_down_sync(0xffffffffu, rgba, 2);
    const uint32_t r3 = __shfl_down_sync(0xffffffffu, rgba, 3);
    if ((f.W & 3u) == 0u) {
        if ((tid & 3u) == 0u && inb)
            *reinterpret_cast<uint4*>(out_rgba + ((size_t)py * f.W + px) * 4) = make_uint4(rgba, r1, r2, r3);
    } else if (inb) {
        *reinterpret_cast<uint32_t*>(out_rgba + ((size_t)py * f.W + px) * 4) = rgba;
    }
}
