// vf_geom_split.cuh -- the geometry pass of the terrain frame as three occupancy-matched kernels
// (replaces vs_main of src/shaders/terrain.wgsl:44-66 plus the clip / cull / coverage state of
// src/terrain/pipeline.rs:107-140 for the implicit grid of src/terrain/mod.rs:553-598):
//
//   k_geom_cull  : the streaming part of k_geom_march (vf_geom_march.cuh) on its own: persistent warp
//                  workers, TMA ring per warp, two vertex rows per iteration in registers, one
//                  "does the cell's bbox hold a pixel centre" test per cell.  Survivors (7.5 % of the
//                  cells at G = 4096, 1080p) are compacted in a shared-memory ring per warp (snapped
//                  a, b, c, d + cell id) and flushed 32 at a time to a global structure-of-arrays
//                  queue (one atomic + 9 coalesced 128-byte stores).  No rasterisation code in this
//                  kernel: ~56 registers, so 24-32 warps per SM hide its latencies.
//   k_cells      : one lane per queued cell, both triangles in straight-line code: bbox, back-face;
//                  micro triangles are rasterised into the L2-resident visibility buffer, longer
//                  ones go to the needle queue (snapped vertices + prim id), large / clipped ones to
//                  the queue of k_binsort.  Cells with a vertex outside the guard band carry a flag
//                  and are rebuilt from the grid (vf_grid_vertex, the code of k_binsort / k_raster).
//   k_needles    : one lane per queued long triangle: the run rasteriser (vf_thin_raster).
// The queues (36 / 28 bytes per entry, + z planes with depth) are written and read once and mostly
// stay in L2.  Results do not depend on scheduling (vf_vis_update is order-free).
#pragma once
#include "vf_geom_march.cuh"

#define VF_GC_WARPS 8                        // warps per CTA of k_geom_cull (each independent)
#define VF_GC_ROWS 16                        // cell rows per task
#define VF_GC_VROWS (VF_GC_ROWS + 1)
#define VF_GC_STAGE_FLOATS 640               // 17 x 36 floats rounded up to a multiple of 128 bytes
#define VF_GC_QCAP 64                        // cell ring: flushed in groups of 32, a row adds <= 31

// global queues, structure of arrays: plane k of entry e is at [k * cap + e]
struct VfCellQ   { uint32_t* id;   int32_t* xy; float* z; uint32_t cap; };   // xy: ax ay bx by cx cy dx dy; z: a b c d (depth)
struct VfNeedleQ { uint32_t* prim; int32_t* xy; float* z; uint32_t cap; };   // xy: x0 y0 x1 y1 x2 y2; z: 3 planes (depth)

template <bool DEPTH>
struct VfGcWarpT {
    alignas(128) float dem[2][VF_GC_STAGE_FLOATS];       // TMA ring: texels of the current / next task (rows of 36)
    alignas(8) uint64_t mbar[2];
    float wz[VF_GC_VROWS], kz[VF_GC_VROWS];              // per row: z*spacing, 0.25*cos(1.1 z)
    int ty[VF_GC_VROWS];                                 // per row: texel row
    int32_t qxy[8][VF_GC_QCAP];                          // surviving cells: snapped a, b, c, d
    float qz[DEPTH ? 4 : 1][DEPTH ? VF_GC_QCAP : 1];
    uint32_t ids[VF_GC_QCAP];                            // cell id | bit 31: rebuild from the grid (clip flags mixed)
};

// ---------------------------------------------------------------------------------------
// Requires f.sparse (perspective projection, affine view: what the camera API produces).  TMA = true:
// grid == DEM with the identity texel mapping, the task's texels are one 17 x 36 box; otherwise they
// are gathered through the axis tables.
template <bool TMA, bool DEPTH, bool SHIPPED>
__global__ void __launch_bounds__(VF_GC_WARPS * 32, 3)
k_geom_cull(const __grid_constant__ VfFrame f, const __grid_constant__ CUtensorMap dem_map,
            const VfAxis* __restrict__ axis, const float* __restrict__ dem,
            const VfCellQ cq, uint32_t* __restrict__ counters, uint32_t rows) {
    extern __shared__ __align__(128) unsigned char vf_gm_smem[];
    const uint32_t lane = threadIdx.x & 31u;
    VfGcWarpT<DEPTH>& w = reinterpret_cast<VfGcWarpT<DEPTH>*>(vf_gm_smem)[threadIdx.x >> 5];
    const uint32_t gm = f.G - 1u;
    const uint32_t nstrips = (gm + VF_GM_QCOLS - 1u) / VF_GM_QCOLS;
    const uint32_t nsegs = (gm + rows - 1u) / rows;          // rows (<= VF_GC_ROWS) cell rows per task
    const uint32_t ntasks = nstrips * nsegs;
    const uint32_t below = (1u << lane) - 1u;

    // the dispenser's answer takes an L2 round trip: it is asked for one task before it is needed
    auto ask_task = [&]() -> uint32_t { return lane == 0 ? atomicAdd(&counters[VF_CNT_TASK], 1u) : 0u; };
    auto got_task = [&](uint32_t t) -> uint32_t { return __shfl_sync(0xffffffffu, t, 0); };
    auto issue = [&](uint32_t task, uint32_t stage) {       // lane 0: start the TMA of `task` into `stage`
        if (TMA && lane == 0 && task < ntasks) {
            const uint32_t strip = task % nstrips, seg = task / nstrips;
            vf_mbar_expect_tx(&w.mbar[stage], VF_GC_VROWS * VF_GM_BOX_X * 4u);
            vf_tma_load_2d(&w.dem[stage][0], &dem_map, (int)((strip * VF_GM_QCOLS) & ~3u), (int)(seg * rows), &w.mbar[stage]);
        }
    };
    if (TMA) {                                              // thread 0 initialises every warp's pair of barriers
        if (threadIdx.x == 0) {
            VfGcWarpT<DEPTH>* ws = reinterpret_cast<VfGcWarpT<DEPTH>*>(vf_gm_smem);
#pragma unroll
            for (int k = 0; k < VF_GC_WARPS; ++k) { vf_mbar_init(&ws[k].mbar[0], 1u); vf_mbar_init(&ws[k].mbar[1], 1u); }
        }
        __syncthreads();
    }
    uint32_t qhead = 0u, qn = 0u;                           // id ring
    // 32 cells (all of them when `all`) -> global queue: one atomic, coalesced plane stores
    auto flush = [&](bool all) {
        while (qn >= 32u || (all && qn > 0u)) {
            const uint32_t take = min(qn, 32u);
            uint32_t base = 0u;
            if (lane == 0) base = atomicAdd(&counters[VF_CNT_CELLS], take);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (lane < take && base + lane < cq.cap) {
                const uint32_t sl = (qhead + lane) & (VF_GC_QCAP - 1u);
                const size_t e = base + lane;
                cq.id[e] = w.ids[sl];
#pragma unroll
                for (int k = 0; k < 8; ++k) cq.xy[(size_t)k * cq.cap + e] = w.qxy[k][sl];
                if (DEPTH) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) cq.z[(size_t)k * cq.cap + e] = w.qz[k][sl];
                }
            }
            qhead = (qhead + take) & (VF_GC_QCAP - 1u);
            qn -= take;
        }
    };
    uint32_t stage = 0u, parity = 0u;                       // parity: bit s = phase of stage s
    uint32_t cur = got_task(ask_task());
    issue(cur, 0u);
    uint32_t nxt = got_task(ask_task());
    // frame constants of the row loop, pinned in registers (otherwise every use re-reads the constant bank)
    float V4 = f.V[4], V5 = f.V[5], V6 = f.V[6], V8 = f.V[8], V9 = f.V[9], V10 = f.V[10];
    float P0 = f.P[0], P5 = f.P[5], P10 = f.P[10], P11 = f.P[11], P14 = f.P[14], P15 = f.P[15];
    float exag = f.exag, hw = f.hw, hh = f.hh;
#define VF_PIN(x) asm volatile("" : "+f"(x))
    VF_PIN(V4); VF_PIN(V5); VF_PIN(V6); VF_PIN(V8); VF_PIN(V9); VF_PIN(V10);
    VF_PIN(P0); VF_PIN(P5); VF_PIN(P10); VF_PIN(P11); VF_PIN(P14); VF_PIN(P15);
    VF_PIN(exag); VF_PIN(hw); VF_PIN(hh);
#undef VF_PIN

    while (cur < ntasks) {
        issue(nxt, stage ^ 1u);
        const uint32_t asked = ask_task();                  // becomes `nxt` at the end of this task
        const uint32_t strip = cur % nstrips, seg = cur / nstrips;
        const uint32_t i = strip * VF_GM_QCOLS + lane, j0 = seg * rows;
        const uint32_t nrows = min(rows + 1u, f.G - j0);    // vertex rows of this task
        const bool colv = i < f.G;
        const bool cellv = lane < VF_GM_QCOLS && i + 1u < f.G;
        // ---- per-column and per-row constants ----
        const VfAxis ax = axis[min(i, gm)];
        const float wx = __fmul_rn(ax.coord, f.spacing);
        const float a0 = fmaf(f.V[0], wx, f.V[12]), a1 = fmaf(f.V[1], wx, f.V[13]), a2 = fmaf(f.V[2], wx, f.V[14]);
        const float sx25 = SHIPPED ? __fmul_rn(sinf(__fmul_rn(ax.coord, 1.3f)), 0.25f) : 0.0f;
        if (lane < VF_GC_VROWS) {
            const VfAxis az = axis[min(j0 + lane, gm)];
            w.wz[lane] = __fmul_rn(az.coord, f.spacing);
            w.ty[lane] = az.ty;
            if (SHIPPED) w.kz[lane] = __fmul_rn(cosf(__fmul_rn(az.coord, 1.1f)), 0.25f);
        }
        __syncwarp();
        if (TMA) { vf_mbar_wait(&w.mbar[stage], (parity >> stage) & 1u); parity ^= 1u << stage; }

        int32_t pX = 0, pY = 0, pXn = 0, pYn = 0;          // previous row: own vertex and right neighbour
        float pZ = 0.0f, pZn = 0.0f;
        uint32_t pst = 0u, pstn = 0u, pnin = 0u;           // clip state of the previous row (0 = inside the guard band)
        const float* texel = &w.dem[stage][((strip * VF_GM_QCOLS) & 3u) + lane];
        const float* gtexel = dem + ax.tx;
        const uint32_t cell0 = (j0 - 1u) * gm + i;          // cell index of (row r-1, column i) is cell0 + r*gm
        uint32_t r = 0u;
        auto load_h = [&](uint32_t rr) -> float {
            return TMA ? texel[rr * VF_GM_BOX_X] : __ldg(gtexel + (size_t)w.ty[rr] * f.dem_pitch);
        };
        // vertex stage (terrain.wgsl:44-66), bit-identical to the sparse form of vf_vertex_stage
        auto clip_pos = [&](uint32_t rr, float h_tex, float c[4]) {
            const float h = SHIPPED ? __fadd_rn(h_tex, __fadd_rn(sx25, w.kz[rr])) : h_tex;
            const float wy = __fmul_rn(h, exag), wz = w.wz[rr];
            const float v0 = fmaf(V4, wy, fmaf(V8, wz, a0));
            const float v1 = fmaf(V5, wy, fmaf(V9, wz, a1));
            const float v2 = fmaf(V6, wy, fmaf(V10, wz, a2));
            c[0] = __fmul_rn(P0, v0); c[1] = __fmul_rn(P5, v1);
            c[2] = __fadd_rn(P14, __fmul_rn(P10, v2)); c[3] = __fadd_rn(P15, __fmul_rn(P11, v2));
        };
        // viewport + snap (vf_to_screen) given 1/w
        auto snap = [&](const float c[4], float invw, int32_t& X, int32_t& Y, float& Z) {
            const float xn = __fmul_rn(c[0], invw), yn = __fmul_rn(c[1], invw);
            X = __float2int_rn(__fmul_rn(fmaf(xn, hw, hw), (float)VF_SUBPIX));
            Y = __float2int_rn(__fmul_rn(fmaf(-yn, hh, hh), (float)VF_SUBPIX));
            Z = DEPTH ? __fmul_rn(c[2], invw) : 0.0f;
        };
        auto append = [&](uint32_t m, bool keep, uint32_t cell, int32_t aX, int32_t aY, float aZ, int32_t bX, int32_t bY, float bZ,
                          int32_t cX, int32_t cY, float cZ, int32_t dX, int32_t dY, float dZ) {
            if (keep) {
                const uint32_t sl = (qhead + qn + __popc(m & below)) & (VF_GC_QCAP - 1u);
                w.qxy[0][sl] = aX; w.qxy[1][sl] = aY; w.qxy[2][sl] = bX; w.qxy[3][sl] = bY;
                w.qxy[4][sl] = cX; w.qxy[5][sl] = cY; w.qxy[6][sl] = dX; w.qxy[7][sl] = dY;
                if (DEPTH) { w.qz[0][sl] = aZ; w.qz[1][sl] = bZ; w.qz[2][sl] = cZ; w.qz[3][sl] = dZ; }
                w.ids[sl] = cell;
            }
            qn += (uint32_t)__popc(m);
        };

        // One vertex row, general form: tracks the clip state of the vertices.  A cell with a vertex
        // outside the guard band survives unless all four are outside one frustum plane (k_cells
        // redoes the full per-triangle treatment).
        auto row = [&]() {
            float c[4];
            clip_pos(r, load_h(r), c);
            const bool in = vf_clip_in(c) || !colv;
            int32_t X, Y;
            float Z;
            snap(c, __frcp_rn(c[3]), X, Y, Z);
            const int32_t Xn = __shfl_down_sync(0xffffffffu, X, 1), Yn = __shfl_down_sync(0xffffffffu, Y, 1);
            float Zn = 0.0f;
            if (DEPTH) Zn = __shfl_down_sync(0xffffffffu, Z, 1);
            uint32_t rebuild = 0u;
            // no pixel centre inside the cell's bbox -> neither triangle draws (off-screen cells die in k_cells)
            bool keep = cellv && r > 0u &&
                        vf_span_has_centre(min(min(pX, pXn), min(X, Xn)), max(max(pX, pXn), max(X, Xn))) &&
                        vf_span_has_centre(min(min(pY, pYn), min(Y, Yn)), max(max(pY, pYn), max(Y, Yn)));
            const uint32_t nin = __ballot_sync(0xffffffffu, !in);
            uint32_t st = 0u, stn = 0u;
            if ((nin | pnin) != 0u) {                       // a vertex of this or the previous row is outside the guard band
                if (nin) {
                    st = in ? 0u : (128u | (vf_clip_flags(c) & 63u));
                    stn = __shfl_down_sync(0xffffffffu, st, 1);
                }
                if (st | stn | pst | pstn) {
                    rebuild = 0x80000000u;
                    keep = cellv && r > 0u && !(st & stn & pst & pstn & 63u);   // not all four outside one frustum plane
                }
            }
            pst = st; pstn = stn; pnin = nin;
            const uint32_t m = __ballot_sync(0xffffffffu, keep);
            if (m) append(m, keep, (cell0 + r * gm) | rebuild, pX, pY, pZ, pXn, pYn, pZn, X, Y, Z, Xn, Yn, Zn);
            pX = X; pY = Y; pXn = Xn; pYn = Yn; pZ = Z; pZn = Zn;
            ++r;
        };

        // Two vertex rows at once, straight-line (the two vertex stages interleave): taken while
        // every vertex is inside the guard band with w in [2^-100, 2^100], where rcp.rn is
        // MUFU.RCP + one Newton step.  Returns false (nothing consumed) when that does not hold.
        auto row_pair = [&]() -> bool {
            float ca[4], cb[4];
            clip_pos(r, load_h(r), ca);
            clip_pos(r + 1u, load_h(r + 1u), cb);
            const float ga = __fmul_rn(VF_GUARD_K, ca[3]), gb = __fmul_rn(VF_GUARD_K, cb[3]);
            const bool oka = (ca[3] >= 0x1p-100f) & (ca[3] <= 0x1p100f) & (ca[2] >= 0.0f) & (ca[2] <= ca[3]) &
                             (fabsf(ca[0]) <= ga) & (fabsf(ca[1]) <= ga);
            const bool okb = (cb[3] >= 0x1p-100f) & (cb[3] <= 0x1p100f) & (cb[2] >= 0.0f) & (cb[2] <= cb[3]) &
                             (fabsf(cb[0]) <= gb) & (fabsf(cb[1]) <= gb);
            if (__any_sync(0xffffffffu, colv & !(oka & okb))) return false;
            float ra, rb;
            asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(ra) : "f"(ca[3]));
            asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rb) : "f"(cb[3]));
            ra = fmaf(ra, -fmaf(ca[3], ra, -1.0f), ra);
            rb = fmaf(rb, -fmaf(cb[3], rb, -1.0f), rb);
            int32_t XA, YA, XB, YB;
            float ZA, ZB;
            snap(ca, ra, XA, YA, ZA);
            snap(cb, rb, XB, YB, ZB);
            const int32_t XAn = __shfl_down_sync(0xffffffffu, XA, 1), YAn = __shfl_down_sync(0xffffffffu, YA, 1);
            const int32_t XBn = __shfl_down_sync(0xffffffffu, XB, 1), YBn = __shfl_down_sync(0xffffffffu, YB, 1);
            float ZAn = 0.0f, ZBn = 0.0f;
            if (DEPTH) { ZAn = __shfl_down_sync(0xffffffffu, ZA, 1); ZBn = __shfl_down_sync(0xffffffffu, ZB, 1); }
            const int32_t eAx0 = min(XA, XAn), eAx1 = max(XA, XAn), eAy0 = min(YA, YAn), eAy1 = max(YA, YAn);   // edge of row A
            const bool keepA = cellv &
                vf_span_has_centre(min(min(pX, pXn), eAx0), max(max(pX, pXn), eAx1)) &
                vf_span_has_centre(min(min(pY, pYn), eAy0), max(max(pY, pYn), eAy1));
            const bool keepB = cellv &
                vf_span_has_centre(min(min(XB, XBn), eAx0), max(max(XB, XBn), eAx1)) &
                vf_span_has_centre(min(min(YB, YBn), eAy0), max(max(YB, YBn), eAy1));
            const uint32_t mA = __ballot_sync(0xffffffffu, keepA), mB = __ballot_sync(0xffffffffu, keepB);
            if (mA) append(mA, keepA, cell0 + r * gm, pX, pY, pZ, pXn, pYn, pZn, XA, YA, ZA, XAn, YAn, ZAn);
            if (qn + (uint32_t)__popc(mB) > VF_GC_QCAP) {       // ring full (coarse grids): row B is redone after the flush
                pX = XA; pY = YA; pXn = XAn; pYn = YAn; pZ = ZA; pZn = ZAn;
                r += 1u;
                return true;
            }
            if (mB) append(mB, keepB, cell0 + (r + 1u) * gm, XA, YA, ZA, XAn, YAn, ZAn, XB, YB, ZB, XBn, YBn, ZBn);
            pX = XB; pY = YB; pXn = XBn; pYn = YBn; pZ = ZB; pZn = ZBn;
            r += 2u;
            return true;
        };

        row();                                              // row 0 primes the previous-row registers
        bool fast = true;                                   // no vertex outside the guard band met in this task so far
        while (r < nrows) {
            if (fast) {
#pragma unroll 1
                while (r + 1u < nrows && qn <= VF_GC_QCAP - 32u)
                    if (!row_pair()) { fast = false; break; }
            }
            if (!fast || r + 1u == nrows) {
#pragma unroll 1
                while (r < nrows && qn <= VF_GC_QCAP - 32u) row();
            }
            __syncwarp();
            flush(false);
        }
        __syncwarp();                                       // all reads of this stage and of the row tables are done
        cur = nxt;
        nxt = got_task(asked);
        stage ^= 1u;
    }
    __syncwarp();
    flush(true);
}

// ---------------------------------------------------------------------------------------
// k_cells: lane = queued cell (whole warps stay in the loop together).  Flagged cells (a vertex
// outside the guard band) are rebuilt with vf_grid_vertex / vf_to_screen -- the code k_binsort and
// k_raster use.  T0 = (a, c, b), T1 = (b, c, d): src/terrain/mod.rs:581.
template <bool DEPTH>
__global__ void __launch_bounds__(256)
k_cells(const __grid_constant__ VfFrame f, const VfAxis* __restrict__ axis, const float* __restrict__ dem,
        const VfCellQ cq, unsigned long long* __restrict__ vis, const VfNeedleQ nq,
        uint32_t* __restrict__ big_prims, uint32_t big_capacity, uint32_t* __restrict__ counters,
        uint32_t* __restrict__ high_water) {
    const uint32_t lane = threadIdx.x & 31u, below = (1u << lane) - 1u;
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicMax(&high_water[2], __ldcg(&counters[VF_CNT_CELLS]));
    const uint32_t n = min(__ldcg(&counters[VF_CNT_CELLS]), cq.cap);
    const uint32_t n_round = (n + 31u) & ~31u;
    const uint32_t gm = f.G - 1u;
    for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < n_round; idx += gridDim.x * blockDim.x) {
        bool to_long[2] = {false, false}, to_big[2] = {false, false};
        uint32_t cell = 0u;
        int32_t X[4], Y[4];
        float Z[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        if (idx < n) {
            const uint32_t id = __ldcs(cq.id + idx);
            cell = id & 0x7fffffffu;
            uint32_t fl[4] = {64u, 64u, 64u, 64u};
            if (id >> 31) {                                  // rebuild from the grid with full clip flags
                const uint32_t j = cell / gm, i = cell - j * gm;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    VfVtx v;
                    VfSVtx sv;
                    vf_grid_vertex(f, axis, dem, i + (q & 1), j + (q >> 1), v);
                    fl[q] = vf_clip_flags(v.c);
                    vf_to_screen(f, v, sv);                  // meaningful only when fl & 64
                    X[q] = sv.X; Y[q] = sv.Y; Z[q] = sv.zn;
                }
            } else {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    X[q] = __ldcs(cq.xy + (size_t)(2 * q) * cq.cap + idx);
                    Y[q] = __ldcs(cq.xy + (size_t)(2 * q + 1) * cq.cap + idx);
                    if (DEPTH) Z[q] = __ldcs(cq.z + (size_t)q * cq.cap + idx);
                }
            }
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const int i0 = k, i1 = 2, i2 = 1 + 2 * k;
                const uint32_t f0 = fl[i0], f1 = fl[i1], f2 = fl[i2];
                if (f0 & f1 & f2 & 63u) continue;            // trivially outside one frustum plane
                if (!(f0 & f1 & f2 & 64u)) { to_big[k] = true; continue; }   // needs clipping
                VfSubTri t;
                t.v[0].X = X[i0]; t.v[0].Y = Y[i0]; t.v[0].zn = Z[i0];
                t.v[1].X = X[i1]; t.v[1].Y = Y[i1]; t.v[1].zn = Z[i1];
                t.v[2].X = X[i2]; t.v[2].Y = Y[i2]; t.v[2].zn = Z[i2];
                int x0, x1, y0, y1;
                if (!vf_subtri_bbox(t, f.W, f.H, x0, x1, y0, y1) || vf_area2(t) >= 0) continue;
                const int bw = x1 - x0 + 1, bh = y1 - y0 + 1, npx = bw * bh;
                if (npx <= VF_MICRO_MAX) {
                    VfEdges es;
                    vf_edges_setup(t, x0, y0, es);
                    for (int dy = 0; dy < bh; ++dy)
                        for (int dx = 0; dx < bw; ++dx) {
                            int64_t ed[3];
                            if (vf_edges_at(es, dx, dy, ed)) vf_gm_emit<DEPTH>(f, vis, t, es, 2u * cell + k, x0 + dx, y0 + dy, ed);
                        }
                } else if (vf_serial_ok(t, bw, bh)) {
                    to_long[k] = true;
                } else {
                    to_big[k] = true;
                }
            }
        }
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const uint32_t nm = __ballot_sync(0xffffffffu, to_long[k]);
            if (nm) {
                uint32_t nb = 0;
                if (lane == 0) nb = atomicAdd(&counters[VF_CNT_NEEDLES], (uint32_t)__popc(nm));
                nb = __shfl_sync(0xffffffffu, nb, 0);
                const size_t pos = nb + __popc(nm & below);
                if (to_long[k] && pos < nq.cap) {
                    const int vi[3] = {k, 2, 1 + 2 * k};
                    nq.prim[pos] = 2u * cell + k;
#pragma unroll
                    for (int v = 0; v < 3; ++v) {
                        nq.xy[(size_t)(2 * v) * nq.cap + pos] = X[vi[v]];
                        nq.xy[(size_t)(2 * v + 1) * nq.cap + pos] = Y[vi[v]];
                        if (DEPTH) nq.z[(size_t)v * nq.cap + pos] = Z[vi[v]];
                    }
                }
            }
            const uint32_t m = __ballot_sync(0xffffffffu, to_big[k]);
            if (m) {
                uint32_t gb = 0;
                if (lane == 0) gb = atomicAdd(&counters[VF_CNT_BIG], (uint32_t)__popc(m));
                gb = __shfl_sync(0xffffffffu, gb, 0);
                const uint32_t pos = gb + __popc(m & below);
                if (to_big[k] && pos < big_capacity) big_prims[pos] = 2u * cell + k;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------
// k_needles: lane = queued long triangle (every vertex inside the guard band, front-facing, bbox on
// screen: checked by k_cells); thread-serial run rasteriser along the shorter bbox axis.
template <bool DEPTH>
__global__ void __launch_bounds__(256)
k_needles(const __grid_constant__ VfFrame f, const VfNeedleQ nq, unsigned long long* __restrict__ vis,
          const uint32_t* __restrict__ counters, uint32_t* __restrict__ high_water) {
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicMax(&high_water[3], __ldcg(&counters[VF_CNT_NEEDLES]));
    const uint32_t n = min(__ldcg(&counters[VF_CNT_NEEDLES]), nq.cap);
    for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += gridDim.x * blockDim.x) {
        const uint32_t prim = __ldcs(nq.prim + idx);
        VfSubTri t;
#pragma unroll
        for (int v = 0; v < 3; ++v) {
            t.v[v].X = __ldcs(nq.xy + (size_t)(2 * v) * nq.cap + idx);
            t.v[v].Y = __ldcs(nq.xy + (size_t)(2 * v + 1) * nq.cap + idx);
            t.v[v].zn = DEPTH ? __ldcs(nq.z + (size_t)v * nq.cap + idx) : 0.0f;
        }
        int x0, x1, y0, y1;
        vf_subtri_bbox(t, f.W, f.H, x0, x1, y0, y1);
        VfEdges es;
        vf_edges_setup(t, x0, y0, es);
        vf_thin_raster(es, x0, x1, y0, y1, [&](int px, int py, int64_t ed[3]) {
            vf_gm_emit<DEPTH>(f, vis, t, es, prim, px, py, ed);
        });
    }
}
