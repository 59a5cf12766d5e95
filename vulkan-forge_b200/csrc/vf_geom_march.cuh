// vf_geom_march.cuh -- k_geom_march: the geometry pass of the terrain frame as a persistent,
// warp-autonomous streaming kernel (replaces vs_main of src/shaders/terrain.wgsl:44-66 plus the
// clip / cull / coverage state of src/terrain/pipeline.rs:107-140 for the implicit grid of
// src/terrain/mod.rs:553-598).
//
//   task     : 32 vertex columns x 29 vertex rows of the grid (31 x 28 cells); tasks are handed
//              out by one atomic counter, every warp is its own worker
//   staging  : the task's 29 x 32 DEM texels are one 2-D TMA box (cp.async.bulk.tensor, 36 texels
//              wide: a box must start on a 16-byte boundary in global memory, strips start every
//              31 columns) into the warp's private two-stage ring; the box of task n+1 is in
//              flight while task n runs
//   march    : lane = vertex column.  The warp walks down the rows; a lane keeps the previous row's
//              snapped vertex in registers and gets the right-hand neighbour by shuffle, so a cell
//              costs one vertex stage + two shuffles + one bbox test.  Cells whose bbox holds no
//              pixel centre (97 % at G = 4096, 1080p) die here.
//   queues   : surviving cells are compacted (ballot + popc) into the warp's shared-memory queue and
//              drained 32 triangles at a time: micro triangles are rasterised by their lane into the
//              L2-resident visibility buffer, longer ones are compacted again and drained 32 at a
//              time by the run rasteriser, large / clipped ones go to the global queue of k_bin.
// Results do not depend on scheduling: the visibility resolve is order-free (vf_vis_update).
#pragma once
#include "vf_terrain_kernels.cuh"

#define VF_GM_WARPS 8                        // warps per CTA (each independent)
#define VF_GM_QCOLS 31                       // cell columns per task (32 vertex columns)
#define VF_GM_ROWS 28                        // cell rows per task
#define VF_GM_BOX_X 36                       // TMA box width: 32 columns + up to 3 of alignment slack
#define VF_GM_STAGE_FLOATS 1056              // 29 x 36 floats rounded up to a multiple of 128 bytes
#define VF_GM_VROWS (VF_GM_ROWS + 1)
#define VF_GM_QCAP 48                        // cell queue: drained when >= 16 cells wait, a row adds <= 31
#define VF_GM_NCAP 64                        // long-triangle queue: drained in groups of 32
#define VF_CNT_TASK 3                        // counters[] slot of the task dispenser

struct VfGmWarp {
    alignas(128) float dem[2][VF_GM_STAGE_FLOATS];       // TMA ring: texels of the current / next task (rows of 36)
    alignas(8) uint64_t mbar[2];
    float wz[VF_GM_VROWS], zc[VF_GM_VROWS], kz[VF_GM_VROWS];   // per row: z*spacing, z, 0.25*cos(1.1 z)
    int ty[VF_GM_VROWS];                                 // per row: texel row
    int32_t qx[4][VF_GM_QCAP], qy[4][VF_GM_QCAP];        // queued cells: snapped a, b, c, d
    float qz[4][VF_GM_QCAP];
    uint32_t qcell[VF_GM_QCAP];                          // cell index | bit 31: rebuild from scratch (clip flags mixed)
    int32_t nx[3][VF_GM_NCAP], ny[3][VF_GM_NCAP];        // queued long triangles
    float nz[3][VF_GM_NCAP];
    uint32_t nprim[VF_GM_NCAP];
};

// Full per-triangle treatment of a cell whose vertices are not all inside the guard band: the
// vertex stage is redone from the grid (same code as k_bin / k_raster).  0: nothing to draw,
// 1: t holds the snapped triangle, 2: needs clipping (global queue).
static __device__ __noinline__ int vf_gm_rebuild(const VfFrame& f, const VfAxis* __restrict__ axis,
                                                 const float* __restrict__ dem, uint32_t prim, VfSubTri& t) {
    VfVtx tri[3];
    vf_prim_vertices(f, axis, dem, prim, tri);
    const uint32_t f0 = vf_clip_flags(tri[0].c), f1 = vf_clip_flags(tri[1].c), f2 = vf_clip_flags(tri[2].c);
    if (f0 & f1 & f2 & 63u) return 0;
    if (!(f0 & f1 & f2 & 64u)) return 2;
#pragma unroll
    for (int k = 0; k < 3; ++k) vf_to_screen(f, tri[k], t.v[k]);
    return 1;
}

// one covered sample -> visibility buffer
template <bool DEPTH>
__device__ __forceinline__ void vf_gm_emit(const VfFrame& f, unsigned long long* __restrict__ vis, const VfSubTri& t,
                                           const VfEdges& es, uint32_t prim, int px, int py, int64_t ed[3]) {
    float zf = 0.0f;
    if (DEPTH) {
        float b[3];
        vf_edges_unbias(es, ed);
        vf_barycentric(ed, b);
        zf = vf_depth_at(t, b);
        if (!(zf < 1.0f)) return;
    }
    vf_vis_update(vis + (size_t)py * f.W + px, DEPTH, prim, zf);
}

// drain groups of 32 long triangles (all of them when `all`)
template <bool DEPTH>
__device__ __forceinline__ void vf_gm_drain_long(const VfFrame& f, VfGmWarp& w, uint32_t& ncount, bool all, uint32_t lane,
                                                 unsigned long long* __restrict__ vis) {
    while (ncount >= 32u || (all && ncount > 0u)) {
        const uint32_t take = min(ncount, 32u), base = ncount - take;
        if (lane < take) {
            const uint32_t s = base + lane;
            VfSubTri t;
#pragma unroll
            for (int k = 0; k < 3; ++k) { t.v[k].X = w.nx[k][s]; t.v[k].Y = w.ny[k][s]; t.v[k].zn = w.nz[k][s]; }
            const uint32_t prim = w.nprim[s];
            int x0, x1, y0, y1;
            vf_subtri_bbox(t, f.W, f.H, x0, x1, y0, y1);
            VfEdges es;
            vf_edges_setup(t, x0, y0, es);
            vf_thin_raster(es, x0, x1, y0, y1, [&](int px, int py, int64_t ed[3]) {
                vf_gm_emit<DEPTH>(f, vis, t, es, prim, px, py, ed);
            });
        }
        ncount = base;
        __syncwarp();
    }
}

// drain the warp's cell queue: two triangles per cell, 32 triangles per pass
template <bool DEPTH>
__device__ __forceinline__ void vf_gm_drain_cells(const VfFrame& f, VfGmWarp& w, uint32_t& qcount, uint32_t& ncount,
                                                  uint32_t lane, const VfAxis* __restrict__ axis,
                                                  const float* __restrict__ dem, unsigned long long* __restrict__ vis,
                                                  uint32_t* __restrict__ big_prims, uint32_t big_capacity,
                                                  uint32_t* __restrict__ counters) {
    const uint32_t n_entries = 2u * qcount;
    const uint32_t below = (1u << lane) - 1u;
    for (uint32_t base = 0; base < n_entries; base += 32u) {
        const uint32_t e = base + lane;
        bool to_big = false, to_long = false;
        uint32_t prim = 0;
        VfSubTri t;
        if (e < n_entries) {
            const uint32_t slot = e >> 1, k = e & 1u;
            const uint32_t cell = w.qcell[slot];
            prim = 2u * (cell & 0x7fffffffu) + k;
            int st = 1;
            if (cell >> 31) {
                st = vf_gm_rebuild(f, axis, dem, prim, t);
            } else {                                     // T0 = (a, c, b), T1 = (b, c, d): src/terrain/mod.rs:581
                const uint32_t i0 = k, i2 = 1u + 2u * k;
                t.v[0].X = w.qx[i0][slot]; t.v[0].Y = w.qy[i0][slot]; t.v[0].zn = w.qz[i0][slot];
                t.v[1].X = w.qx[2][slot];  t.v[1].Y = w.qy[2][slot];  t.v[1].zn = w.qz[2][slot];
                t.v[2].X = w.qx[i2][slot]; t.v[2].Y = w.qy[i2][slot]; t.v[2].zn = w.qz[i2][slot];
            }
            if (st == 2) {
                to_big = true;                           // needs clipping
            } else if (st == 1) {
                int x0, x1, y0, y1;
                if (vf_subtri_bbox(t, f.W, f.H, x0, x1, y0, y1) && vf_area2(t) < 0) {
                    const int bw = x1 - x0 + 1, bh = y1 - y0 + 1, npx = bw * bh;
                    if (npx <= VF_MICRO_MAX) {
                        VfEdges es;
                        vf_edges_setup(t, x0, y0, es);
                        for (int dy = 0; dy < bh; ++dy)
                            for (int dx = 0; dx < bw; ++dx) {
                                int64_t ed[3];
                                if (vf_edges_at(es, dx, dy, ed)) vf_gm_emit<DEPTH>(f, vis, t, es, prim, x0 + dx, y0 + dy, ed);
                            }
                    } else if (npx <= VF_MEDIUM_MAX) {
                        to_long = true;
                    } else {
                        to_big = true;
                    }
                }
            }
        }
        const uint32_t nm = __ballot_sync(0xffffffffu, to_long);
        if (nm) {
            if (to_long) {
                const uint32_t s = ncount + __popc(nm & below);
#pragma unroll
                for (int k = 0; k < 3; ++k) { w.nx[k][s] = t.v[k].X; w.ny[k][s] = t.v[k].Y; w.nz[k][s] = t.v[k].zn; }
                w.nprim[s] = prim;
            }
            ncount += (uint32_t)__popc(nm);
        }
        const uint32_t m = __ballot_sync(0xffffffffu, to_big);
        if (m) {
            uint32_t gb = 0;
            if (lane == 0) gb = atomicAdd(&counters[VF_CNT_BIG], (uint32_t)__popc(m));
            gb = __shfl_sync(0xffffffffu, gb, 0);
            const uint32_t pos = gb + __popc(m & below);
            if (to_big && pos < big_capacity) big_prims[pos] = prim;
        }
        __syncwarp();
        vf_gm_drain_long<DEPTH>(f, w, ncount, false, lane, vis);
    }
    qcount = 0u;
    __syncwarp();
}

// ---------------------------------------------------------------------------------------
// TMA = true: grid == DEM with the identity texel mapping (k_axis_identity), the task's texels are
// one 29 x 36 box.  Otherwise texels are gathered through the axis tables (one coalesced __ldg
// per row, issued one row ahead).
template <bool TMA, bool DEPTH>
__global__ void __launch_bounds__(VF_GM_WARPS * 32, 2)
k_geom_march(const __grid_constant__ VfFrame f, const __grid_constant__ CUtensorMap dem_map,
             const VfAxis* __restrict__ axis, const float* __restrict__ dem,
             unsigned long long* __restrict__ vis, uint32_t* __restrict__ big_prims, uint32_t big_capacity,
             uint32_t* __restrict__ counters, uint32_t rows) {
    extern __shared__ __align__(128) unsigned char vf_gm_smem[];
    const uint32_t lane = threadIdx.x & 31u;
    VfGmWarp& w = reinterpret_cast<VfGmWarp*>(vf_gm_smem)[threadIdx.x >> 5];
    const uint32_t gm = f.G - 1u;
    const uint32_t nstrips = (gm + VF_GM_QCOLS - 1u) / VF_GM_QCOLS;
    const uint32_t nsegs = (gm + rows - 1u) / rows;          // rows (<= VF_GM_ROWS) cell rows per task
    const uint32_t ntasks = nstrips * nsegs;
    const uint32_t below = (1u << lane) - 1u;

    auto next_task = [&]() -> uint32_t {
        uint32_t t = 0;
        if (lane == 0) t = atomicAdd(&counters[VF_CNT_TASK], 1u);
        return __shfl_sync(0xffffffffu, t, 0);
    };
    auto issue = [&](uint32_t task, uint32_t stage) {       // lane 0: start the TMA of `task` into `stage`
        if (TMA && lane == 0 && task < ntasks) {
            const uint32_t strip = task % nstrips, seg = task / nstrips;
            vf_mbar_expect_tx(&w.mbar[stage], VF_GM_VROWS * VF_GM_BOX_X * 4u);
            vf_tma_load_2d(&w.dem[stage][0], &dem_map, (int)((strip * VF_GM_QCOLS) & ~3u), (int)(seg * rows), &w.mbar[stage]);
        }
    };
    if (TMA) {                                              // thread 0 initialises every warp's pair of barriers
        if (threadIdx.x == 0) {
            VfGmWarp* ws = reinterpret_cast<VfGmWarp*>(vf_gm_smem);
#pragma unroll
            for (int k = 0; k < VF_GM_WARPS; ++k) { vf_mbar_init(&ws[k].mbar[0], 1u); vf_mbar_init(&ws[k].mbar[1], 1u); }
        }
        __syncthreads();
    }
    uint32_t qcount = 0u, ncount = 0u, stage = 0u, parity = 0u;     // parity: bit s = phase of stage s
    uint32_t cur = next_task();
    issue(cur, 0u);

    while (cur < ntasks) {
        const uint32_t nxt = next_task();
        issue(nxt, stage ^ 1u);
        const uint32_t strip = cur % nstrips, seg = cur / nstrips;
        const uint32_t i = strip * VF_GM_QCOLS + lane, j0 = seg * rows;
        const uint32_t nrows = min(rows + 1u, f.G - j0);   // vertex rows of this task
        const bool colv = i < f.G;
        const bool cellv = lane < VF_GM_QCOLS && i + 1u < f.G;
        // ---- per-column and per-row constants ----
        const VfAxis ax = axis[min(i, gm)];
        const float wx = __fmul_rn(ax.coord, f.spacing);
        const float a0 = fmaf(f.V[0], wx, f.V[12]), a1 = fmaf(f.V[1], wx, f.V[13]), a2 = fmaf(f.V[2], wx, f.V[14]);
        const float sx25 = f.mode == 0 ? __fmul_rn(sinf(__fmul_rn(ax.coord, 1.3f)), 0.25f) : 0.0f;
        for (uint32_t r = lane; r < VF_GM_VROWS; r += 32u) {
            const VfAxis az = axis[min(j0 + r, gm)];
            w.wz[r] = __fmul_rn(az.coord, f.spacing);
            w.zc[r] = az.coord;
            w.ty[r] = az.ty;
            w.kz[r] = f.mode == 0 ? __fmul_rn(cosf(__fmul_rn(az.coord, 1.1f)), 0.25f) : 0.0f;
        }
        __syncwarp();
        if (TMA) { vf_mbar_wait(&w.mbar[stage], (parity >> stage) & 1u); parity ^= 1u << stage; }

        int32_t pX = 0, pY = 0, pXn = 0, pYn = 0;          // previous row: own vertex and right neighbour
        float pZ = 0.0f, pZn = 0.0f;
        uint32_t pst = 0u, pstn = 0u, pnin = 0u;           // clip state (0 = inside the guard band)
        float h_next = 0.0f;
        if (!TMA) h_next = __ldg(dem + (size_t)w.ty[0] * f.dem_pitch + ax.tx);
#pragma unroll 1
        for (uint32_t r = 0; r < nrows; ++r) {
            float h_tex;
            if (TMA) {
                h_tex = w.dem[stage][r * VF_GM_BOX_X + ((strip * VF_GM_QCOLS) & 3u) + lane];
            } else {
                h_tex = h_next;
                if (r + 1u < nrows) h_next = __ldg(dem + (size_t)w.ty[r + 1u] * f.dem_pitch + ax.tx);
            }
            // ---- vertex stage (terrain.wgsl:44-66), bit-identical to vf_vertex_stage ----
            float c[4];
            if (f.sparse) {
                const float h = f.mode == 0 ? __fadd_rn(h_tex, __fadd_rn(sx25, w.kz[r])) : h_tex;
                const float wy = __fmul_rn(h, f.exag), wz = w.wz[r];
                const float v0 = fmaf(f.V[4], wy, fmaf(f.V[8], wz, a0));
                const float v1 = fmaf(f.V[5], wy, fmaf(f.V[9], wz, a1));
                const float v2 = fmaf(f.V[6], wy, fmaf(f.V[10], wz, a2));
                c[0] = __fmul_rn(f.P[0], v0);
                c[1] = __fmul_rn(f.P[5], v1);
                c[2] = __fadd_rn(f.P[14], __fmul_rn(f.P[10], v2));
                c[3] = __fadd_rn(f.P[15], __fmul_rn(f.P[11], v2));
            } else {
                VfVtx o;
                vf_vertex_stage(f, ax.coord, w.zc[r], 0.0f, 0.0f, h_tex, o);
                c[0] = o.c[0]; c[1] = o.c[1]; c[2] = o.c[2]; c[3] = o.c[3];
            }
            const bool in = vf_clip_in(c) || !colv;
            // ---- viewport + snap (vf_to_screen) ----
            const float invw = __fdiv_rn(1.0f, c[3]);
            const float xn = __fmul_rn(c[0], invw), yn = __fmul_rn(c[1], invw);
            const int32_t X = __float2int_rn(__fmul_rn(fmaf(xn, f.hw, f.hw), (float)VF_SUBPIX));
            const int32_t Y = __float2int_rn(__fmul_rn(fmaf(-yn, f.hh, f.hh), (float)VF_SUBPIX));
            const float Z = DEPTH ? __fmul_rn(c[2], invw) : 0.0f;
            const int32_t Xn = __shfl_down_sync(0xffffffffu, X, 1), Yn = __shfl_down_sync(0xffffffffu, Y, 1);
            float Zn = 0.0f;
            if (DEPTH) Zn = __shfl_down_sync(0xffffffffu, Z, 1);
            const uint32_t nin = __ballot_sync(0xffffffffu, !in);
            uint32_t st = 0u, stn = 0u;
            if (nin) {                                      // rare: some vertex of this row is outside the guard band
                st = in ? 0u : (128u | (vf_clip_flags(c) & 63u));
                stn = __shfl_down_sync(0xffffffffu, st, 1);
            }
            if (r > 0u) {
                bool keep = cellv, rebuild = false;
                if ((nin | pnin) != 0u) {
                    const uint32_t any = st | stn | pst | pstn;
                    if (any) {
                        rebuild = true;
                        if (st & stn & pst & pstn & 63u) keep = false;      // all four outside one frustum plane
                    }
                }
                if (keep && !rebuild) {
                    int x0, x1, y0, y1;                     // no pixel centre inside the cell's bbox -> neither triangle draws
                    keep = vf_bbox_pixels(min(min(pX, pXn), min(X, Xn)), max(max(pX, pXn), max(X, Xn)),
                                          min(min(pY, pYn), min(Y, Yn)), max(max(pY, pYn), max(Y, Yn)),
                                          f.W, f.H, x0, x1, y0, y1);
                }
                const uint32_t m = __ballot_sync(0xffffffffu, keep);
                if (m) {
                    if (keep) {
                        const uint32_t s = qcount + __popc(m & below);
                        w.qx[0][s] = pX;  w.qy[0][s] = pY;
                        w.qx[1][s] = pXn; w.qy[1][s] = pYn;
                        w.qx[2][s] = X;   w.qy[2][s] = Y;
                        w.qx[3][s] = Xn;  w.qy[3][s] = Yn;
                        if (DEPTH) { w.qz[0][s] = pZ; w.qz[1][s] = pZn; w.qz[2][s] = Z; w.qz[3][s] = Zn; }
                        w.qcell[s] = ((j0 + r - 1u) * gm + i) | (rebuild ? 0x80000000u : 0u);
                    }
                    qcount += (uint32_t)__popc(m);
                    __syncwarp();
                    if (qcount >= 16u)
                        vf_gm_drain_cells<DEPTH>(f, w, qcount, ncount, lane, axis, dem, vis, big_prims, big_capacity, counters);
                }
            }
            pX = X; pY = Y; pXn = Xn; pYn = Yn; pZ = Z; pZn = Zn;
            pst = st; pstn = stn; pnin = nin;
        }
        __syncwarp();                                       // all reads of this stage and of the row tables are done
        cur = nxt;
        stage ^= 1u;
    }
    vf_gm_drain_cells<DEPTH>(f, w, qcount, ncount, lane, axis, dem, vis, big_prims, big_capacity, counters);
    vf_gm_drain_long<DEPTH>(f, w, ncount, true, lane, vis);
}
