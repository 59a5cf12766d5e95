// vf_geom_march.cuh -- k_geom_march: the geometry pass of the terrain frame as a persistent,
// warp-autonomous streaming kernel (replaces vs_main of src/shaders/terrain.wgsl:44-66 plus the
// clip / cull / coverage state of src/terrain/pipeline.rs:107-140 for the implicit grid of
// src/terrain/mod.rs:553-598).
//
//   task     : 32 vertex columns x up to 31 vertex rows of the grid (31 x 30 cells; 27 rows with depth).  Worker k
//              starts on tasks k and k + workers, the rest are handed out in order by one (float) atomic counter;
//              the row segments get shorter towards the end of the grid (host-built table), which keeps the
//              finish times of the 2368 workers close together (VF_GM_DBG + tools/gm_tasks.py show them)
//   ahead    : id, place in the grid, texels and axis entries of a task are requested during the task before it
//   staging  : the task's DEM texels are one 2-D TMA box (cp.async.bulk.tensor, 36 texels wide: a box must start
//              on a 16-byte boundary in global memory, strips start every 31 columns) into the warp's private
//              two-stage ring
//   march    : lane = vertex column.  The warp walks down the rows, two at a time in packed f32x2 arithmetic; a
//              lane keeps the previous row's snapped vertex in registers and gets the right-hand neighbour by
//              shuffle, so a cell costs one vertex stage + two shuffles + one bbox test.  Cells whose bbox holds
//              no pixel centre (92.5 % at G = 4096, 1080p) die here.
//   queues   : surviving cells are compacted (ballot + popc, predicated stores) into the warp's shared-memory
//              ring and drained 16 cells = 32 triangles at a time (always full warps, straight-line code):
//              micro triangles are covered in 32-bit arithmetic by their lane with predicated reductions into
//              the L2-resident visibility buffer, longer ones are compacted again into a second ring and
//              drained 32 at a time by the 32-bit run rasteriser, large / clipped ones go to the global queue
//              of k_binsort.
// Results do not depend on scheduling: the visibility resolve is order-free (vf_vis_update).
#pragma once
#include <type_traits>
#include "vf_terrain_kernels.cuh"

#ifndef VF_GM_WARPS
#define VF_GM_WARPS 8                        // warps per CTA (each independent)
#endif
#ifndef VF_GM_CTAS
#define VF_GM_CTAS 2                         // resident CTAs per SM the register budget is set for
#endif
#define VF_GM_QCOLS 31                       // cell columns per task (32 vertex columns)
#ifndef VF_GM_ROWS_N
#define VF_GM_ROWS_N 30                      // cell rows per task; with depth the rings carry z as well and the
#endif
#ifndef VF_GM_ROWS_D
#define VF_GM_ROWS_D 26                      // staging ring gets shorter (16 warps x ~14 KB fill the SM's shared memory)
#endif
#define VF_GM_BOX_X 36                       // TMA box width: 32 columns + up to 3 of alignment slack
template <bool DEPTH>
struct VfGmDim {
    static constexpr uint32_t ROWS = DEPTH ? VF_GM_ROWS_D : VF_GM_ROWS_N;     // <= 31: per-row tables are filled by lane
    static constexpr uint32_t VROWS = ROWS + 1u;
    static constexpr uint32_t STAGE_FLOATS = (VROWS * VF_GM_BOX_X + 31u) / 32u * 32u;   // a box, rounded up to 128 bytes
};
#define VF_GM_QCAP 64                        // cell ring: a row adds <= 31, rows run while <= 32 wait
#define VF_GM_NCAP 64                        // long-triangle ring: drained in groups of 32

template <bool DEPTH>
struct VfGmWarpT {
    alignas(128) float dem[2][VfGmDim<DEPTH>::STAGE_FLOATS];   // TMA ring: texels of the current / next task (rows of 36)
    alignas(8) uint64_t mbar[2];
    float wz[VfGmDim<DEPTH>::VROWS], kz[VfGmDim<DEPTH>::VROWS];   // per row: z*spacing, 0.25*cos(1.1 z)
    int ty[VfGmDim<DEPTH>::VROWS];                       // per row: texel row
    alignas(16) int4 qv[2][VF_GM_QCAP];                  // queued cells: raw snapped (x, y) of (a, b) and of (c, d), see vf_gm_raw
    float qz[DEPTH ? 4 : 1][DEPTH ? VF_GM_QCAP : 1];     // depth only
    uint32_t qcell[VF_GM_QCAP];                          // cell index | bit 31: rebuild from scratch (clip flags mixed)
    alignas(8) int2 nv[3][VF_GM_NCAP];                   // queued long triangles: snapped (X, Y) of the vertices
    float nz[DEPTH ? 3 : 1][DEPTH ? VF_GM_NCAP : 1];     // depth only
    uint32_t nprim[VF_GM_NCAP];
};

// ring state, warp-uniform registers; crosses the call to vf_gm_drain_cells packed in one word so
// that it never lives in local memory
struct VfGmQueues {
    uint32_t qhead, qn, nhead, nn;
    __device__ __forceinline__ uint32_t pack() const { return qhead | (qn << 8) | (nhead << 16) | (nn << 24); }
    __device__ __forceinline__ void unpack(uint32_t v) { qhead = v & 255u; qn = (v >> 8) & 255u; nhead = (v >> 16) & 255u; nn = v >> 24; }
};

// Snap without a float -> int conversion, and the cell cull without a bounding box.
//   raw     : for a vertex inside the fast guard band (|X|, |Y| <= 2^22 - 256 sub-pixels) the march keeps
//             raw = float_as_int(-xa + VF_GM_MAGIC), xa = fma(xn, 256 hw, 256 hw) the screen position in sub-pixels.
//             The sum lies in [2^23, 2^24) where floats are the integers, so the FADD rounds -xa to nearest-even
//             exactly as cvt.rni does (the constant is even), and raw = bits(MAGIC) + (-X - 128): X = VF_GM_RAWC - raw.
//   key     : bits 8..23 of raw are floor((-X - 128) / 256) + 2^14, the index of the last pixel centre
//             (sub-pixel 128 mod 256) at or below -X.  Two vertices with the same index have no centre c with
//             mn <= c < mx between them.  A centre ON the right / bottom extreme of a front-facing triangle is
//             never covered (its two edges there run up, resp. the bottom edge runs right: not top-left), so a
//             cell whose four vertices share the x index or share the y index draws nothing.  One PRMT packs the
//             two indices of a vertex; (Ka^Kb)|(Ka^Kc)|(Ka^Kd) has a zero half exactly when one axis is shared.
#define VF_GM_MAGIC 12582784.0f                       // 1.5 * 2^23 - 128
#define VF_GM_RAWC  0x4B3FFF80                        // bits(VF_GM_MAGIC) - 128 : X = VF_GM_RAWC - raw
#define VF_GM_FAST_PX 16383.0f                        // fast guard band: |X| <= 16383 px keeps the sum inside the binade
__device__ __forceinline__ uint32_t vf_gm_key(int32_t rx, int32_t ry) {
    uint32_t k;
    asm("prmt.b32 %0, %1, %2, 0x6521;" : "=r"(k) : "r"(rx), "r"(ry));     // bytes 1,2 of rx | bytes 1,2 of ry
    return k;
}
// (Ka^Kb)|(Ka^Kc)|(Ka^Kd)
__device__ __forceinline__ uint32_t vf_gm_keydiff(uint32_t ka, uint32_t kb, uint32_t kc, uint32_t kd) {
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0x7e;" : "=r"(d) : "r"(ka), "r"(kb), "r"(kc));          // (a^b)|(a^c) == not all three equal
    asm("lop3.b32 %0, %1, %2, %3, 0xf6;" : "=r"(d) : "r"(d), "r"(ka), "r"(kd));           // d | (a^d)
    return d;
}
// both halves of the key difference non-zero (and the lane owns a cell): LOP3 with a predicate result + one ISETP
__device__ __forceinline__ bool vf_gm_cell_live(uint32_t d, bool cellv) {
    uint32_t r;
    asm("{\n\t.reg .pred p, q;\n\t.reg .b32 t;\n\t"
        "setp.ne.u32 q, %2, 0;\n\t"
        "lop3.and.b32 t|p, %1, 0xffff, 0, 0xC0, q;\n\t"          // p = ((d & 0xffff) != 0) && cellv
        "setp.gt.u32.and p, %1, 0xffff, p;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(r) : "r"(d), "r"((uint32_t)cellv));
    return r != 0u;
}

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

// One covered sample of the depth variant -> visibility buffer.  e0..e2 are the BIASED edge values of the sample
// (they fit 32 bits: 0 <= E_i < 2^31 on a covered sample of a micro or serial triangle).  vf_barycentric divides
// by e0 + e1 + e2, which is twice the triangle's area at every sample: rinv = 1 / float(-area2) is taken once per
// triangle (vf_gm_rinv) and the weights are the same floats as vf_barycentric's.
__device__ __forceinline__ float vf_gm_rinv(int64_t area2) { return vf_rcp1((float)(-area2)); }
__device__ __forceinline__ void vf_gm_emit_depth(unsigned long long* __restrict__ key64, const VfSubTri& t, float rinv, uint32_t nbias,
                                                 uint32_t prim, uint32_t e0, uint32_t e1, uint32_t e2) {
    float b[3];
    b[0] = __fmul_rn((float)(int32_t)(e0 + (nbias & 1u)), rinv);                 // vf_edges_unbias + vf_barycentric
    b[1] = __fmul_rn((float)(int32_t)(e1 + ((nbias >> 1) & 1u)), rinv);
    b[2] = __fmul_rn((float)(int32_t)(e2 + ((nbias >> 2) & 1u)), rinv);
    const float zf = vf_depth_at(t, b);
    if (!(zf < 1.0f)) return;
    vf_vis_update(key64, true, prim, zf);
}

// drain groups of 32 long triangles (whatever is left when `all`)
template <bool DEPTH>
__device__ __forceinline__ void vf_gm_drain_long(const VfFrame& f, VfGmWarpT<DEPTH>& w, VfGmQueues& q, bool all, uint32_t lane,
                                                 unsigned long long* __restrict__ vis) {
    while (q.nn >= 32u || (all && q.nn > 0u)) {
        const uint32_t take = min(q.nn, 32u);
        if (lane < take) {
            const uint32_t s = (q.nhead + lane) & (VF_GM_NCAP - 1u);
            VfSubTri t;
#pragma unroll
            for (int k = 0; k < 3; ++k) { const int2 v = w.nv[k][s]; t.v[k].X = v.x; t.v[k].Y = v.y; t.v[k].zn = DEPTH ? w.nz[k][s] : 0.0f; }
            const uint32_t prim = w.nprim[s];
            int x0, x1, y0, y1;
            vf_subtri_bbox(t, f.W, f.H, x0, x1, y0, y1);
            VfEdges es;
            vf_edges_setup(t, x0, y0, es);
            const float rinv = DEPTH ? vf_gm_rinv(vf_area2(t)) : 0.0f;
            vf_thin_raster32(es, x0, x1, y0, y1, f.W, [&](bool covered, uint32_t idx, uint32_t e0, uint32_t e1, uint32_t e2) {
                if (!DEPTH) { vf_vis_max_if(covered, reinterpret_cast<uint32_t*>(vis) + idx, prim); return; }
                if (covered) vf_gm_emit_depth(vis + idx, t, rinv, es.nbias, prim, e0, e1, e2);
            });
        }
        q.nhead = (q.nhead + take) & (VF_GM_NCAP - 1u);
        q.nn -= take;
        __syncwarp();
    }
}

// drain the warp's cell ring 16 cells (two triangles each) at a time; the remainder too when `all`
template <bool DEPTH>
__device__ __noinline__ uint32_t vf_gm_drain_cells(const VfFrame& f, VfGmWarpT<DEPTH>& w, uint32_t qpacked, bool all,
                                               const VfAxis* __restrict__ axis, const float* __restrict__ dem,
                                               unsigned long long* __restrict__ vis, uint32_t* __restrict__ big_prims,
                                               uint32_t big_capacity, uint32_t* __restrict__ counters) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t below = (1u << lane) - 1u;
    VfGmQueues q;
    q.unpack(qpacked);
    while (q.qn >= 16u || (all && q.qn > 0u)) {
        const uint32_t take = min(q.qn, 16u);
        bool to_big = false, to_long = false;
        uint32_t prim = 0;
        VfSubTri t;
        if (lane < 2u * take) {
            const uint32_t slot = (q.qhead + (lane >> 1)) & (VF_GM_QCAP - 1u), k = lane & 1u;
            const uint32_t cell = w.qcell[slot];
            prim = 2u * (cell & 0x7fffffffu) + k;
            int st = 1;
            if (cell >> 31) {
                VfSubTri tr;                             // address escapes into the call: lives on the stack
                st = vf_gm_rebuild(f, axis, dem, prim, tr);
#pragma unroll
                for (int v = 0; v < 3; ++v) { t.v[v].X = tr.v[v].X; t.v[v].Y = tr.v[v].Y; t.v[v].zn = tr.v[v].zn; }
            } else {                                     // T0 = (a, c, b), T1 = (b, c, d): src/terrain/mod.rs:581
                // ring entry: qv[0] = raw (a.x, a.y, b.x, b.y), qv[1] = raw (c.x, c.y, d.x, d.y); X = VF_GM_RAWC - raw
                const int2* ab = reinterpret_cast<const int2*>(&w.qv[0][slot]);
                const int2* cd = reinterpret_cast<const int2*>(&w.qv[1][slot]);
                const int2 v0 = ab[k], v1 = cd[0], v2 = k ? cd[1] : ab[1];
                t.v[0].X = VF_GM_RAWC - v0.x; t.v[0].Y = VF_GM_RAWC - v0.y;
                t.v[1].X = VF_GM_RAWC - v1.x; t.v[1].Y = VF_GM_RAWC - v1.y;
                t.v[2].X = VF_GM_RAWC - v2.x; t.v[2].Y = VF_GM_RAWC - v2.y;
                const uint32_t i0 = k, i2 = 1u + 2u * k;
                if (DEPTH) { t.v[0].zn = w.qz[i0][slot]; t.v[1].zn = w.qz[2][slot]; t.v[2].zn = w.qz[i2][slot]; }
            }
            // Classification and micro coverage in straight-line code: a divergent region costs the warp its
            // instructions whatever the number of lanes inside, so the set-up runs for every lane and only
            // the visibility updates are predicated (branch resolution was a sixth of this kernel's stalls).
            int x0, x1, y0, y1;
            const int32_t mnx = min(t.v[0].X, min(t.v[1].X, t.v[2].X)), mxx = max(t.v[0].X, max(t.v[1].X, t.v[2].X));
            const int32_t mny = min(t.v[0].Y, min(t.v[1].Y, t.v[2].Y)), mxy = max(t.v[0].Y, max(t.v[1].Y, t.v[2].Y));
            const bool inside = vf_bbox_pixels(mnx, mxx, mny, mxy, f.W, f.H, x0, x1, y0, y1);
            const int64_t area2 = vf_area2(t);
            const bool ok = st == 1 && inside && area2 < 0;
            const int bw = x1 - x0 + 1, bh = y1 - y0 + 1, npx = bw * bh;
            // Micro triangle that is small itself (not a large one with a few on-screen samples): its extent
            // is below 2^12 sub-pixels, the reference centre lies inside its bbox, so every edge value stays
            // below 2^26 and the coverage test runs in 32 bits (unsigned products: other lanes may wrap).
            const bool micro = ok && npx <= VF_MICRO_MAX && (((uint32_t)mxx - (uint32_t)mnx) | ((uint32_t)mxy - (uint32_t)mny)) < 4096u;
            const bool serial = min(bw, bh) <= VF_SERIAL_LINES && (-area2 >> 17) <= (int64_t)VF_SERIAL_AREA;   // vf_serial_ok
            to_long = ok && !micro && serial;
            to_big = st == 2 || (ok && !micro && !serial);
            const int32_t sx = x0 * VF_SUBPIX + VF_HALFPIX, sy = y0 * VF_SUBPIX + VF_HALFPIX;
            int32_t A[3], B[3], E[3];
            uint32_t nbias = 0u;
#pragma unroll
            for (int i = 0; i < 3; ++i) {                   // vf_edges_setup in 32 bits
                const VfSVtx& a = t.v[(i + 1) % 3];
                const VfSVtx& b = t.v[(i + 2) % 3];
                A[i] = b.Y - a.Y; B[i] = a.X - b.X;
                // vf_top_left(a, b) on the differences (|X|, |Y| < 2^23: they cannot wrap): b.Y > a.Y, or level and b.X < a.X
                const bool tl = (A[i] > 0) | ((A[i] == 0) & (B[i] > 0));
                nbias |= tl ? 0u : (1u << i);
                E[i] = (int32_t)((uint32_t)A[i] * (uint32_t)(sx - a.X) + (uint32_t)B[i] * (uint32_t)(sy - a.Y)) - (tl ? 0 : 1);
            }
            // the (at most four) samples: the first one without a loop, the others along the row, then down
            int32_t e0 = E[0], e1 = E[1], e2 = E[2];
            uint32_t idx = (uint32_t)y0 * f.W + (uint32_t)x0;                   // sample index in the frame
            const float rinv = DEPTH ? vf_gm_rinv(area2) : 0.0f;
            if (!DEPTH) {
                vf_vis_max_if(micro && (e0 | e1 | e2) >= 0, reinterpret_cast<uint32_t*>(vis) + idx, prim);
            } else if (micro && (e0 | e1 | e2) >= 0) {
                vf_gm_emit_depth(vis + idx, t, rinv, nbias, prim, (uint32_t)e0, (uint32_t)e1, (uint32_t)e2);
            }
            if (micro && npx > 1) {
                const int32_t a0 = A[0] * VF_SUBPIX, a1 = A[1] * VF_SUBPIX, a2 = A[2] * VF_SUBPIX;
                const int32_t r0 = B[0] * VF_SUBPIX - (bw - 1) * a0, r1 = B[1] * VF_SUBPIX - (bw - 1) * a1,
                              r2 = B[2] * VF_SUBPIX - (bw - 1) * a2;
                const uint32_t idx_wrap = f.W - (uint32_t)(bw - 1);
                int dx = 0, dy = 0;
                for (int n = 1; n < npx; ++n) {
                    const bool wrap = dx + 1 == bw;
                    e0 += wrap ? r0 : a0; e1 += wrap ? r1 : a1; e2 += wrap ? r2 : a2;
                    idx += wrap ? idx_wrap : 1u;
                    dx = wrap ? 0 : dx + 1;
                    dy += wrap ? 1 : 0;
                    const bool covered = (e0 | e1 | e2) >= 0;
                    if (!DEPTH) {
                        vf_vis_max_if(covered, reinterpret_cast<uint32_t*>(vis) + idx, prim);
                    } else if (covered) {
                        vf_gm_emit_depth(vis + idx, t, rinv, nbias, prim, (uint32_t)e0, (uint32_t)e1, (uint32_t)e2);
                    }
                }
            }
        }
        q.qhead = (q.qhead + take) & (VF_GM_QCAP - 1u);
        q.qn -= take;
        const uint32_t nm = __ballot_sync(0xffffffffu, to_long);
        {   // branch-free append to the ring of long triangles (predicated stores, as in the march)
            static_assert(VF_GM_NCAP * sizeof(int2) == 512, "offsets of the predicated stores below");
            const uint32_t s = (q.nhead + q.nn + __popc(nm & below)) & (VF_GM_NCAP - 1u);
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %0, 0;\n\t"
                         "@p st.shared.v2.b32 [%1], {%3, %4};\n\t@p st.shared.v2.b32 [%1+512], {%5, %6};\n\t"
                         "@p st.shared.v2.b32 [%1+1024], {%7, %8};\n\t@p st.shared.b32 [%2], %9;\n\t}"
                         :: "r"((uint32_t)to_long), "r"(vf_smem_u32(&w.nv[0][0]) + s * 8u), "r"(vf_smem_u32(&w.nprim[0]) + s * 4u),
                            "r"(t.v[0].X), "r"(t.v[0].Y), "r"(t.v[1].X), "r"(t.v[1].Y), "r"(t.v[2].X), "r"(t.v[2].Y), "r"(prim) : "memory");
            if (DEPTH && to_long) {
#pragma unroll
                for (int k = 0; k < 3; ++k) w.nz[k][s] = t.v[k].zn;
            }
            q.nn += (uint32_t)__popc(nm);
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
        vf_gm_drain_long<DEPTH>(f, w, q, false, lane, vis);
    }
    return q.pack();
}

// true when some pixel centre (sub-pixel 128 mod 256) lies in [mn, mx]: floor((mx-128)/256) differs
// from floor((mn-129)/256), i.e. the two differ somewhere above bit 7
__device__ __forceinline__ bool vf_span_has_centre(int32_t mn, int32_t mx) {
    return (uint32_t)((mn - (VF_HALFPIX + 1)) ^ (mx - VF_HALFPIX)) > (uint32_t)(VF_SUBPIX - 1);
}

// ---------------------------------------------------------------------------------------
// Requires f.sparse (perspective projection, affine view: what the camera API produces); other
// matrices take the tiled k_geom.  TMA = true: grid == DEM with the identity texel mapping
// (k_axis_identity), the task's texels are one 25 x 36 box.  Otherwise texels are gathered through
// the axis tables (one coalesced __ldg per row, issued one row ahead).
template <bool TMA, bool DEPTH, bool SHIPPED>
__global__ void __launch_bounds__(VF_GM_WARPS * 32, VF_GM_CTAS)
k_geom_march(const __grid_constant__ VfFrame f, const __grid_constant__ CUtensorMap dem_map,
             const VfAxis* __restrict__ axis, const float* __restrict__ dem,
             unsigned long long* __restrict__ vis, uint32_t* __restrict__ big_prims, uint32_t big_capacity,
             uint32_t* __restrict__ counters, const uint2* __restrict__ segtab, uint32_t nsegs, uint32_t nfull, uint32_t rows_full,
             const float2* __restrict__ hmm, uint32_t hmm_nb, unsigned long long* __restrict__ dbg) {
    extern __shared__ __align__(128) unsigned char vf_gm_smem[];
#ifndef VF_GM_PROFILE                                       // the timing dump costs registers: only in -DVF_GM_PROFILE builds
    dbg = nullptr;
#endif
    constexpr uint32_t VROWS = VfGmDim<DEPTH>::VROWS;
    const uint32_t lane = threadIdx.x & 31u;
    VfGmWarpT<DEPTH>& w = reinterpret_cast<VfGmWarpT<DEPTH>*>(vf_gm_smem)[threadIdx.x >> 5];
    const uint32_t gm = f.G - 1u;
    const uint32_t nstrips = (gm + VF_GM_QCOLS - 1u) / VF_GM_QCOLS;
    // A task = one strip of one row segment.  segtab[seg] = (first cell row, cell rows <= VfGmDim::ROWS):
    // the host tapers the segment height towards the end of the grid (vf_terrain.cu), and the
    // dispenser hands tasks out in order, so the short ones come last and the workers finish together.
    const uint32_t ntasks = nstrips * nsegs;
    auto decode = [&](uint32_t task, uint32_t& strip, uint32_t& j0, uint32_t& nr) {
        const uint32_t seg = task / nstrips;
        strip = task - seg * nstrips;
        if (seg < nfull) {                                  // the leading segments all have the full height: no table
            j0 = seg * rows_full; nr = rows_full;           // look-up (an L2 round trip at the start of every task)
        } else {
            const uint2 e = __ldg(segtab + seg);
            j0 = e.x; nr = e.y;
        }
    };
    const uint32_t below = (1u << lane) - 1u;
    const uint32_t qv_base = vf_smem_u32(&w.qv[0][0]), qcell_base = vf_smem_u32(&w.qcell[0]);
    // fast guard band per axis: the 8x guard band of the clipper, narrowed where 9 x half the frame would exceed
    // VF_GM_FAST_PX pixels (frames wider than 3640 px) so that the raw snap stays inside its binade
    float gfx = fminf(VF_GUARD_K, __fmul_rn(__fsub_rn(__fdiv_rn(VF_GM_FAST_PX, f.hw), 1.0f), 0.999f));
    float gfy = fminf(VF_GUARD_K, __fmul_rn(__fsub_rn(__fdiv_rn(VF_GM_FAST_PX, f.hh), 1.0f), 0.999f));

    // the dispenser's answer takes an L2 round trip: it is asked for one task before it is needed
    // The counter is kept as a FLOAT (exact below 2^24 tasks; the host falls back to more rows per task
    // above that): ptxas turns every integer atomic add, from atomicAdd() or from asm, into a
    // warp-aggregated atomic whose closing shuffle waits for the answer at once, which would put
    // the round trip back on the critical path.
    auto ask_task = [&]() -> uint32_t {
        float t = 0.0f;
        if (lane == 0) asm volatile("atom.global.add.f32 %0, [%1], 0f3F800000;" : "=f"(t) : "l"(counters + VF_CNT_TASK));
        return __float_as_uint(t);
    };
    // Worker k starts with tasks k and k + workers (no round trip, no burst of requests at the start of the
    // launch); the counter hands out the rest.
    const uint32_t nworkers = gridDim.x * VF_GM_WARPS, wid = blockIdx.x * VF_GM_WARPS + (threadIdx.x >> 5);
    auto got_task = [&](uint32_t t) -> uint32_t { return 2u * nworkers + (uint32_t)__uint_as_float(__shfl_sync(0xffffffffu, t, 0)); };
    auto issue = [&](bool valid, uint32_t strip, uint32_t j0, uint32_t stage) {   // lane 0: start the TMA of a task into `stage`
        if (TMA && lane == 0 && valid) {
            // The stage was read (generic proxy) by the task before the current one; the __syncwarp() that closes
            // every task orders those reads before this point, the proxy fence orders them before the bulk write.
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            vf_mbar_expect_tx(&w.mbar[stage], VROWS * VF_GM_BOX_X * 4u);
            vf_tma_load_2d(&w.dem[stage][0], &dem_map, (int)((strip * VF_GM_QCOLS) & ~3u), (int)j0, &w.mbar[stage]);
        }
    };
    if (TMA) {                                              // thread 0 initialises every warp's pair of barriers
        if (threadIdx.x == 0) {
            VfGmWarpT<DEPTH>* ws = reinterpret_cast<VfGmWarpT<DEPTH>*>(vf_gm_smem);
#pragma unroll
            for (int k = 0; k < VF_GM_WARPS; ++k) { vf_mbar_init(&ws[k].mbar[0], 1u); vf_mbar_init(&ws[k].mbar[1], 1u); }
        }
        __syncthreads();
    }
    VfGmQueues q = {0u, 0u, 0u, 0u};
    uint32_t stage = 0u, parity = 0u;                       // parity: bit s = phase of stage s
    // Everything a task needs at its start is prepared one task ahead: its id (the dispenser's answer takes an
    // L2 round trip), its place in the grid (a division and the segment table), its texels (TMA) and the axis
    // entries of its columns / rows.  (asm volatile: a plain load is sunk by the compiler to its first use at
    // the start of the next task, where its latency shows.)
    uint32_t cur = wid, nxt = nworkers + wid;
    uint32_t strip = 0u, j0 = 0u, nr = 0u;                  // place of `cur`
    float ax_coord = 0.0f, az_coord = 0.0f;                 // axis entries of `cur`: column of this lane, row `lane`
    int ax_tx = 0, az_ty = 0;
    auto fetch_axes = [&](uint32_t s0, uint32_t r0) {
        const VfAxis* pa = axis + min(s0 * VF_GM_QCOLS + lane, gm);
        const VfAxis* pz = axis + min(r0 + min(lane, VROWS - 1u), gm);
        asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(ax_coord) : "l"(&pa->coord));
        asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(az_coord) : "l"(&pz->coord));
        if (!TMA) {
            asm volatile("ld.global.nc.s32 %0, [%1];" : "=r"(ax_tx) : "l"(&pa->tx));
            asm volatile("ld.global.nc.s32 %0, [%1];" : "=r"(az_ty) : "l"(&pz->ty));
        }
    };
    // Height bounds of a task (vf_terrain.cu: k_hminmax, (min, max) of every 32 x 32 block of vertex heights): lanes
    // 0..3 fetch the <= 2 x 2 blocks the task's 32 x 31 vertices touch, one task ahead like the axis entries.
    float mm_lo = __int_as_float(0x7f800000), mm_hi = __int_as_float(0xff800000);
    auto fetch_mm = [&](uint32_t s0, uint32_t r0, uint32_t rows) {
        if (hmm != nullptr && lane < 4u) {
            const uint32_t c0 = s0 * VF_GM_QCOLS;
            const uint32_t bi = min((c0 >> 5) + (lane & 1u), min(c0 + 31u, gm) >> 5);
            const uint32_t bj = min((r0 >> 5) + (lane >> 1), min(r0 + rows, gm) >> 5);
            asm volatile("ld.global.nc.v2.f32 {%0, %1}, [%2];" : "=f"(mm_lo), "=f"(mm_hi) : "l"(hmm + (size_t)bj * hmm_nb + bi));
        }
    };
    if (cur < ntasks) {
        decode(cur, strip, j0, nr);
        issue(true, strip, j0, 0u);
        fetch_axes(strip, j0);
        fetch_mm(strip, j0, nr);
    }
    // magnitude of the view / projection entries the vertex stage uses: scale of the safety margin below
    float vp_scale;
    {
        float vmax = 1.0f, pmax = 1.0f;
#pragma unroll
        for (int k = 0; k < 15; ++k) if ((k & 3) != 3) vmax = fmaxf(vmax, fabsf(f.V[k]));
        pmax = fmaxf(pmax, fmaxf(fmaxf(fabsf(f.P[0]), fabsf(f.P[5])), fmaxf(fmaxf(fabsf(f.P[10]), fabsf(f.P[11])), fmaxf(fabsf(f.P[14]), fabsf(f.P[15])))));
        vp_scale = __fmul_rn(__fmul_rn(vmax, pmax), 1e-4f);
    }
    // frame constants of the row loop, pinned in registers (otherwise every use re-reads the constant bank)
    float V4 = f.V[4], V5 = f.V[5], V6 = f.V[6], V8 = f.V[8], V9 = f.V[9], V10 = f.V[10];
    float P0 = f.P[0], P5 = f.P[5], P10 = f.P[10], P11 = f.P[11], P14 = f.P[14], P15 = f.P[15];
    float exag = f.exag, hw = __fmul_rn(f.hw, (float)VF_SUBPIX), hh = __fmul_rn(f.hh, (float)VF_SUBPIX);   // in sub-pixels
#define VF_PIN(x) asm volatile("" : "+f"(x))
    VF_PIN(V4); VF_PIN(V5); VF_PIN(V6); VF_PIN(V8); VF_PIN(V9); VF_PIN(V10);
    VF_PIN(P0); VF_PIN(P5); VF_PIN(P10); VF_PIN(P11); VF_PIN(P14); VF_PIN(P15);
    VF_PIN(exag); VF_PIN(hw); VF_PIN(hh);
    VF_PIN(gfx); VF_PIN(gfy);
#undef VF_PIN

    unsigned long long t_task = 0ull;
    if (dbg) { asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_task)); if (lane == 0) dbg[2u * (blockIdx.x * VF_GM_WARPS + (threadIdx.x >> 5))] = t_task; }
    while (cur < ntasks) {
        // The axis entries fetched during the previous task are consumed BEFORE this task's requests go out
        // (asm volatile keeps the order): issued after them, their first use shares a scoreboard with the new
        // long-latency operations and waits for those round trips (4 % of the kernel's stall samples).
        float wx, wz_lane;
        asm volatile("mul.rn.f32 %0, %1, %2;" : "=f"(wx) : "f"(ax_coord), "f"(f.spacing));
        asm volatile("mul.rn.f32 %0, %1, %2;" : "=f"(wz_lane) : "f"(az_coord), "f"(f.spacing));
        // Is every vertex of this task inside near / far / the fast guard band with w in rcp's fast range?  Clip
        // coordinates are affine in (x, height, z), the conditions are linear in clip coordinates, so they hold on the
        // task's box [x0, x1] x [hmin, hmax] x [z0, z1] if they hold at its 8 corners -- checked by lanes 0..7 with a
        // margin far above the rounding error of the float evaluation (1e-4 of the magnitude of the terms).  A safe
        // task runs the row pairs without the per-vertex tests.
        bool task_safe = false;
        if (hmm != nullptr) {
            float lo = mm_lo, hi = mm_hi;
            lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, 1)); hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, 1));
            lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, 2)); hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, 2));
            lo = __shfl_sync(0xffffffffu, lo, 0); hi = __shfl_sync(0xffffffffu, hi, 0);
            const float cxw = __shfl_sync(0xffffffffu, wx, (lane & 1u) ? 31u : 0u);
            const float czw = __shfl_sync(0xffffffffu, wz_lane, (lane & 2u) ? min(nr, VROWS - 1u) : 0u);
            float hc = (lane & 4u) ? hi : lo;
            if (SHIPPED) hc = __fadd_rn(hc, (lane & 4u) ? 0.5f : -0.5f);          // analytic term: |0.25 sin + 0.25 cos| <= 0.5
            const float cwy = __fmul_rn(hc, f.exag);
            const float v0 = fmaf(f.V[4], cwy, fmaf(f.V[8], czw, fmaf(f.V[0], cxw, f.V[12])));
            const float v1 = fmaf(f.V[5], cwy, fmaf(f.V[9], czw, fmaf(f.V[1], cxw, f.V[13])));
            const float v2 = fmaf(f.V[6], cwy, fmaf(f.V[10], czw, fmaf(f.V[2], cxw, f.V[14])));
            const float c0 = __fmul_rn(f.P[0], v0), c1 = __fmul_rn(f.P[5], v1);
            const float c2 = __fadd_rn(f.P[14], __fmul_rn(f.P[10], v2)), c3 = __fadd_rn(f.P[15], __fmul_rn(f.P[11], v2));
            const float mg = __fmul_rn(vp_scale, __fadd_rn(__fadd_rn(fabsf(cxw), fabsf(cwy)), __fadd_rn(fabsf(czw), 1.0f)));
            const bool okc = (c3 >= __fadd_rn(mg, 1e-30f)) & (c3 <= 1e30f) & (c2 >= mg) & (__fsub_rn(c3, c2) >= mg) &
                             (__fsub_rn(__fmul_rn(__fmul_rn(gfx, 0.999f), c3), fabsf(c0)) >= mg) &
                             (__fsub_rn(__fmul_rn(__fmul_rn(gfy, 0.999f), c3), fabsf(c1)) >= mg) & (hi < 1e30f) & (lo > -1e30f);
            task_safe = (__ballot_sync(0xffffffffu, okc) & 0xffu) == 0xffu;
        }
        uint32_t nstrip = 0u, nj0 = 0u, nnr = 0u;           // place of `nxt`
        if (nxt < ntasks) decode(nxt, nstrip, nj0, nnr);
        issue(nxt < ntasks, nstrip, nj0, stage ^ 1u);
        const uint32_t asked = ask_task();                  // becomes `nxt` at the end of this task
        const uint32_t i = strip * VF_GM_QCOLS + lane;
        const uint32_t nrows = min(nr + 1u, f.G - j0);      // vertex rows of this task
        const bool colv = i < f.G;
        const bool cellv = lane < VF_GM_QCOLS && i + 1u < f.G;
        // ---- per-column and per-row constants ----
        const float a0 = fmaf(f.V[0], wx, f.V[12]), a1 = fmaf(f.V[1], wx, f.V[13]), a2 = fmaf(f.V[2], wx, f.V[14]);
        const float sx25 = SHIPPED ? __fmul_rn(sinf(__fmul_rn(ax_coord, 1.3f)), 0.25f) : 0.0f;
        if (lane < VROWS) {
            w.wz[lane] = wz_lane;
            w.ty[lane] = az_ty;
            if (SHIPPED) w.kz[lane] = __fmul_rn(cosf(__fmul_rn(az_coord, 1.1f)), 0.25f);
        }
        __syncwarp();
        unsigned long long tp0 = 0ull, tp1 = 0ull, tp2 = 0ull;
        if (dbg) { asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tp0) :: "memory"); tp0 += (wx != 12345.0f ? 0ull : 1ull); }
        if (TMA) { vf_mbar_wait(&w.mbar[stage], (parity >> stage) & 1u); parity ^= 1u << stage; }
        if (dbg) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tp1) :: "memory");

        // previous row: raw snapped position (vf_gm_raw) of the lane's vertex and of its right neighbour, their keys
        int32_t pX = 0, pY = 0, pXn = 0, pYn = 0;
        uint32_t pK = 0u, pKn = 0u;
        float pZ = 0.0f, pZn = 0.0f;
        uint32_t pst = 0u, pstn = 0u, pnin = 0u;           // clip state of the previous row (0 = inside the fast guard band)
        const float* texel = &w.dem[stage][((strip * VF_GM_QCOLS) & 3u) + lane];
        const float* gtexel = dem + ax_tx;
        if (nxt < ntasks) { fetch_axes(nstrip, nj0); fetch_mm(nstrip, nj0, nnr); }   // used at the start of the next task
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
        // The same for two rows at once in packed f32x2 arithmetic (FFMA2 / FMUL2 / FADD2 of sm_100: one
        // instruction per pair, IEEE round-to-nearest in each half, so the results are those of clip_pos).
        auto pk = [](float lo, float hi) -> unsigned long long {
            unsigned long long r;
            asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
            return r;
        };
        auto fma2 = [](unsigned long long a, unsigned long long b, unsigned long long c) -> unsigned long long {
            unsigned long long r;
            asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
            return r;
        };
        auto mul2 = [](unsigned long long a, unsigned long long b) -> unsigned long long {
            unsigned long long r;
            asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
            return r;
        };
        auto add2 = [](unsigned long long a, unsigned long long b) -> unsigned long long {
            unsigned long long r;
            asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
            return r;
        };
        auto lo32 = [](unsigned long long v) -> float { float a, b; asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); return a; };
        auto hi32 = [](unsigned long long v) -> float { float a, b; asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); return b; };
        // Two rows at once.  x and y of ONE vertex travel as a pair (clip x|y, then the raw snapped X|Y: adjacent
        // registers, which is what the 16-byte ring stores and the key's PRMT want); z and w pair up ACROSS the two rows.
        // Every component sees the operations of clip_pos in the same order.
        auto clip_pos2 = [&](uint32_t rr, unsigned long long& cxyA, unsigned long long& cxyB, float& czA, float& cwA, float& czB, float& cwB) {
            if (SHIPPED) {
                float ca[4], cb[4];
                clip_pos(rr, load_h(rr), ca); clip_pos(rr + 1u, load_h(rr + 1u), cb);
                cxyA = pk(ca[0], ca[1]); cxyB = pk(cb[0], cb[1]); czA = ca[2]; cwA = ca[3]; czB = cb[2]; cwB = cb[3];
                return;
            }
            const unsigned long long wy = mul2(pk(load_h(rr), load_h(rr + 1u)), pk(exag, exag));
            const float wyA = lo32(wy), wyB = hi32(wy), wzA = w.wz[rr], wzB = w.wz[rr + 1u];
            const unsigned long long V45 = pk(V4, V5), V89 = pk(V8, V9), a01 = pk(a0, a1);
            const unsigned long long v01A = fma2(V45, pk(wyA, wyA), fma2(V89, pk(wzA, wzA), a01));
            const unsigned long long v01B = fma2(V45, pk(wyB, wyB), fma2(V89, pk(wzB, wzB), a01));
            const unsigned long long v2 = fma2(pk(V6, V6), wy, fma2(pk(V10, V10), pk(wzA, wzB), pk(a2, a2)));
            cxyA = mul2(pk(P0, P5), v01A); cxyB = mul2(pk(P0, P5), v01B);
            // (the additions stay scalar: ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into one FFMA2, which
            // would round P14 + P10 v2 once instead of twice)
            const unsigned long long m2 = mul2(pk(P10, P10), v2), m3 = mul2(pk(P11, P11), v2);
            czA = __fadd_rn(P14, lo32(m2)); czB = __fadd_rn(P14, hi32(m2));
            cwA = __fadd_rn(P15, lo32(m3)); cwB = __fadd_rn(P15, hi32(m3));
        };
        // inside near / far and the fast guard band: the vertex can be projected and snapped in raw form
        auto fast_in = [&](const float c[4]) -> bool {
            return (c[3] > 0.0f) & (c[2] >= 0.0f) & (c[2] <= c[3]) &
                   (fabsf(c[0]) <= __fmul_rn(gfx, c[3])) & (fabsf(c[1]) <= __fmul_rn(gfy, c[3]));
        };
        // viewport + raw snap (vf_to_screen, see vf_gm_raw above) given 1/w.  fma(xn, hw, hw) * 256 == fma(xn, 256 hw,
        // 256 hw) (scaling by a power of two commutes with the rounding; 256 hw is exact, the sum is never subnormal),
        // and negating both constants negates the result exactly.
        auto snap = [&](const float c[4], float invw, int32_t& X, int32_t& Y, float& Z) {
            const float xn = __fmul_rn(c[0], invw), yn = __fmul_rn(c[1], invw);
            X = __float_as_int(__fadd_rn(fmaf(xn, -hw, -hw), VF_GM_MAGIC));      // raw of -X
            Y = __float_as_int(__fadd_rn(fmaf(yn, hh, -hh), VF_GM_MAGIC));       // raw of -Y  (Y = rint(fma(-yn, hh, hh)))
            Z = DEPTH ? __fmul_rn(c[2], invw) : 0.0f;
        };
        // Branch-free: every lane computes a slot, the stores are predicated (asm: the compiler would put
        // the stores behind a divergent branch, and branch resolution is a sixth of this kernel's stalls).
        static_assert(VF_GM_QCAP * sizeof(int4) == 1024, "offsets of the predicated stores below");
        auto append = [&](uint32_t m, bool keep, uint32_t cell, int32_t aX, int32_t aY, float aZ, int32_t bX, int32_t bY, float bZ,
                          int32_t cX, int32_t cY, float cZ, int32_t dX, int32_t dY, float dZ) {
            const uint32_t s4 = ((q.qhead + q.qn + __popc(m & below)) << 2) & (4u * VF_GM_QCAP - 4u);   // slot * 4
            const uint32_t s = s4 >> 2;
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %0, 0;\n\t"
                         "@p st.shared.v4.b32 [%1], {%3, %4, %5, %6};\n\t@p st.shared.v4.b32 [%1+1024], {%7, %8, %9, %10};\n\t"
                         "@p st.shared.b32 [%2], %11;\n\t}"
                         :: "r"((uint32_t)keep), "r"(qv_base + s4 * 4u), "r"(qcell_base + s4), "r"(aX), "r"(aY), "r"(bX), "r"(bY),
                            "r"(cX), "r"(cY), "r"(dX), "r"(dY), "r"(cell) : "memory");
            if (DEPTH && keep) { w.qz[0][s] = aZ; w.qz[1][s] = bZ; w.qz[2][s] = cZ; w.qz[3][s] = dZ; }
            q.qn += (uint32_t)__popc(m);
        };

        // One vertex row, general form: tracks the clip state of the vertices; cells with a vertex
        // outside the (fast) guard band are queued for a rebuild from scratch.
        auto row = [&]() {
            float c[4];
            clip_pos(r, load_h(r), c);
            const bool in = fast_in(c) || !colv;
            int32_t X, Y;
            float Z;
            snap(c, __frcp_rn(c[3]), X, Y, Z);
            const uint32_t K = vf_gm_key(X, Y);
            const int32_t Xn = __shfl_down_sync(0xffffffffu, X, 1), Yn = __shfl_down_sync(0xffffffffu, Y, 1);
            const uint32_t Kn = vf_gm_key(Xn, Yn);
            float Zn = 0.0f;
            if (DEPTH) Zn = __shfl_down_sync(0xffffffffu, Z, 1);
            // the four vertices share a pixel-centre index on one axis -> neither triangle draws (off-screen cells die in the drain)
            bool keep = vf_gm_cell_live(vf_gm_keydiff(pK, pKn, K, Kn), cellv && r > 0u);
            uint32_t rebuild = 0u;
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
            append(m, keep, (cell0 + r * gm) | rebuild, pX, pY, pZ, pXn, pYn, pZn, X, Y, Z, Xn, Yn, Zn);
            pX = X; pY = Y; pXn = Xn; pYn = Yn; pK = K; pKn = Kn; pZ = Z; pZn = Zn;
            ++r;
        };

        // Two vertex rows at once, straight-line (the two vertex stages interleave): taken while
        // every vertex is inside the fast guard band with w in [2^-100, 2^100], where rcp.rn is
        // MUFU.RCP + one Newton step.  Returns false (nothing consumed) when that does not hold.
        auto row_pair = [&](auto checked_tag) -> bool {
            constexpr bool CHECKED = decltype(checked_tag)::value;       // false: the task passed the corner test above
            unsigned long long cxyA, cxyB;
            float czA, cwA, czB, cwB;
            clip_pos2(r, cxyA, cxyB, czA, cwA, czB, cwB);
            if (CHECKED) {
                const unsigned long long g2 = pk(gfx, gfy);
                const unsigned long long gA = mul2(g2, pk(cwA, cwA)), gB = mul2(g2, pk(cwB, cwB));
                const bool oka = (cwA >= 0x1p-100f) & (cwA <= 0x1p100f) & (czA >= 0.0f) & (czA <= cwA) &
                                 (fabsf(lo32(cxyA)) <= lo32(gA)) & (fabsf(hi32(cxyA)) <= hi32(gA));
                const bool okb = (cwB >= 0x1p-100f) & (cwB <= 0x1p100f) & (czB >= 0.0f) & (czB <= cwB) &
                                 (fabsf(lo32(cxyB)) <= lo32(gB)) & (fabsf(hi32(cxyB)) <= hi32(gB));
                if (__any_sync(0xffffffffu, colv & !(oka & okb))) return false;
            }
            float ra, rb;
            asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(ra) : "f"(cwA));
            asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rb) : "f"(cwB));
            ra = fmaf(ra, -fmaf(cwA, ra, -1.0f), ra);
            rb = fmaf(rb, -fmaf(cwB, rb, -1.0f), rb);
            int32_t XA, YA, XB, YB;                         // raw (vf_gm_raw)
            {   // raw snap: (xn, yn) = clip (x, y) / w; (-X, -Y) = fma((xn, yn), (-hw, hh), (-hw, -hh)); + magic
                const unsigned long long sc = pk(-hw, hh), of = pk(-hw, -hh), mg = pk(VF_GM_MAGIC, VF_GM_MAGIC);
                const unsigned long long xyA = add2(fma2(mul2(cxyA, pk(ra, ra)), sc, of), mg);
                const unsigned long long xyB = add2(fma2(mul2(cxyB, pk(rb, rb)), sc, of), mg);
                asm("mov.b64 {%0, %1}, %2;" : "=r"(XA), "=r"(YA) : "l"(xyA));
                asm("mov.b64 {%0, %1}, %2;" : "=r"(XB), "=r"(YB) : "l"(xyB));
            }
            const float ZA = DEPTH ? __fmul_rn(czA, ra) : 0.0f, ZB = DEPTH ? __fmul_rn(czB, rb) : 0.0f;
            const uint32_t KA = vf_gm_key(XA, YA), KB = vf_gm_key(XB, YB);
            const int32_t XAn = __shfl_down_sync(0xffffffffu, XA, 1), YAn = __shfl_down_sync(0xffffffffu, YA, 1);
            const int32_t XBn = __shfl_down_sync(0xffffffffu, XB, 1), YBn = __shfl_down_sync(0xffffffffu, YB, 1);
            const uint32_t KAn = vf_gm_key(XAn, YAn), KBn = vf_gm_key(XBn, YBn);
            float ZAn = 0.0f, ZBn = 0.0f;
            if (DEPTH) { ZAn = __shfl_down_sync(0xffffffffu, ZA, 1); ZBn = __shfl_down_sync(0xffffffffu, ZB, 1); }
            const bool keepA = vf_gm_cell_live(vf_gm_keydiff(KA, KAn, pK, pKn), cellv);
            const bool keepB = vf_gm_cell_live(vf_gm_keydiff(KA, KAn, KB, KBn), cellv);
            const uint32_t mA = __ballot_sync(0xffffffffu, keepA), mB = __ballot_sync(0xffffffffu, keepB);
            append(mA, keepA, cell0 + r * gm, pX, pY, pZ, pXn, pYn, pZn, XA, YA, ZA, XAn, YAn, ZAn);
            if (q.qn + (uint32_t)__popc(mB) > VF_GM_QCAP) {      // ring full (coarse grids): row B is redone after the drain
                pX = XA; pY = YA; pXn = XAn; pYn = YAn; pK = KA; pKn = KAn; pZ = ZA; pZn = ZAn;
                r += 1u;
                return true;
            }
            append(mB, keepB, cell0 + (r + 1u) * gm, XA, YA, ZA, XAn, YAn, ZAn, XB, YB, ZB, XBn, YBn, ZBn);
            pX = XB; pY = YB; pXn = XBn; pYn = YBn; pK = KB; pKn = KBn; pZ = ZB; pZn = ZBn;
            r += 2u;
            return true;
        };

        // Four vertex rows at once, for tasks that passed the corner test (no per-vertex tests, so nothing can send the
        // rows back): the two pairs are independent up to the appends, which gives the scheduler two dependence chains
        // to interleave (the kernel waits on fixed-latency dependences with its 4 warps per scheduler).  Returns false,
        // with nothing consumed, when the ring could not take the four rows' cells.
        auto row_quad = [&]() -> bool {
            int32_t X[4], Y[4], Xn[4], Yn[4];
            float Z[4], Zn[4];
            uint32_t K[4], Kn[4];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                unsigned long long cxyA, cxyB;
                float czA, cwA, czB, cwB, ra, rb;
                clip_pos2(r + 2u * h, cxyA, cxyB, czA, cwA, czB, cwB);
                asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(ra) : "f"(cwA));
                asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rb) : "f"(cwB));
                ra = fmaf(ra, -fmaf(cwA, ra, -1.0f), ra);
                rb = fmaf(rb, -fmaf(cwB, rb, -1.0f), rb);
                const unsigned long long sc = pk(-hw, hh), of = pk(-hw, -hh), mg = pk(VF_GM_MAGIC, VF_GM_MAGIC);
                const unsigned long long xyA = add2(fma2(mul2(cxyA, pk(ra, ra)), sc, of), mg);
                const unsigned long long xyB = add2(fma2(mul2(cxyB, pk(rb, rb)), sc, of), mg);
                asm("mov.b64 {%0, %1}, %2;" : "=r"(X[2 * h]), "=r"(Y[2 * h]) : "l"(xyA));
                asm("mov.b64 {%0, %1}, %2;" : "=r"(X[2 * h + 1]), "=r"(Y[2 * h + 1]) : "l"(xyB));
                Z[2 * h] = DEPTH ? __fmul_rn(czA, ra) : 0.0f; Z[2 * h + 1] = DEPTH ? __fmul_rn(czB, rb) : 0.0f;
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                K[k] = vf_gm_key(X[k], Y[k]);
                Xn[k] = __shfl_down_sync(0xffffffffu, X[k], 1); Yn[k] = __shfl_down_sync(0xffffffffu, Y[k], 1);
                Kn[k] = vf_gm_key(Xn[k], Yn[k]);
                Zn[k] = DEPTH ? __shfl_down_sync(0xffffffffu, Z[k], 1) : 0.0f;
            }
            bool keep[4];
            uint32_t m[4];
            keep[0] = vf_gm_cell_live(vf_gm_keydiff(K[0], Kn[0], pK, pKn), cellv);
#pragma unroll
            for (int k = 1; k < 4; ++k) keep[k] = vf_gm_cell_live(vf_gm_keydiff(K[k], Kn[k], K[k - 1], Kn[k - 1]), cellv);
#pragma unroll
            for (int k = 0; k < 4; ++k) m[k] = __ballot_sync(0xffffffffu, keep[k]);
            if (q.qn + (uint32_t)(__popc(m[0]) + __popc(m[1]) + __popc(m[2]) + __popc(m[3])) > VF_GM_QCAP) return false;
            append(m[0], keep[0], cell0 + r * gm, pX, pY, pZ, pXn, pYn, pZn, X[0], Y[0], Z[0], Xn[0], Yn[0], Zn[0]);
#pragma unroll
            for (int k = 1; k < 4; ++k)
                append(m[k], keep[k], cell0 + (r + k) * gm, X[k - 1], Y[k - 1], Z[k - 1], Xn[k - 1], Yn[k - 1], Zn[k - 1],
                       X[k], Y[k], Z[k], Xn[k], Yn[k], Zn[k]);
            pX = X[3]; pY = Y[3]; pXn = Xn[3]; pYn = Yn[3]; pK = K[3]; pKn = Kn[3]; pZ = Z[3]; pZn = Zn[3];
            r += 4u;
            return true;
        };

        row();                                              // row 0 primes the previous-row registers
        if (dbg) {
            tp2 = (unsigned long long)pX & 1ull;            // the stamp waits for the row's results
            unsigned long long tq;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tq) :: "memory");
            tp2 = tq + tp2 - ((unsigned long long)pX & 1ull);
            if (lane == 0) {
                unsigned long long* acc = dbg + 2u * gridDim.x * VF_GM_WARPS + ntasks + 4u * (blockIdx.x * VF_GM_WARPS + (threadIdx.x >> 5));
                acc[0] += tp0 - t_task; acc[1] += tp1 - tp0; acc[2] += tp2 - tp1; acc[3] += 1ull;
            }
        }
        bool fast = true;                                   // no vertex outside the guard band met in this task so far
        while (r < nrows) {
            // row_pair assumes a clean previous row (it neither reads nor writes the clip state):
            // after a row with a vertex outside the guard band the general routine runs until a
            // row is clean again (pnin == 0 implies pst == pstn == 0)
            if (fast && pnin == 0u && r + 1u < nrows) {
                if (task_safe) {
#ifndef VF_GM_NO_QUAD
#pragma unroll 1
                    while (r + 3u < nrows && q.qn <= VF_GM_QCAP - 32u)
                        if (!row_quad()) row_pair(std::false_type{});       // ring nearly full: advance by a pair, drain below
#endif
#pragma unroll 1
                    while (r + 1u < nrows && q.qn <= VF_GM_QCAP - 32u) row_pair(std::false_type{});
                } else {
#pragma unroll 1
                    while (r + 1u < nrows && q.qn <= VF_GM_QCAP - 32u)
                        if (!row_pair(std::true_type{})) { fast = false; break; }
                }
            } else {
#pragma unroll 1
                while (r < nrows && q.qn <= VF_GM_QCAP - 32u) {
                    row();
                    if (fast && pnin == 0u && r + 1u < nrows) break;
                }
            }
            __syncwarp();
            if (q.qn >= 16u)
                q.unpack(vf_gm_drain_cells<DEPTH>(f, w, q.pack(), false, axis, dem, vis, big_prims, big_capacity, counters));
        }
        __syncwarp();                                       // all reads of this stage and of the row tables are done
        if (dbg) {                                          // profiling aid (VF_GM_DBG): duration of every task
            unsigned long long t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (lane == 0) dbg[2u * gridDim.x * VF_GM_WARPS + cur] = t1 - t_task;
            t_task = t1;
        }
        cur = nxt;
        strip = nstrip; j0 = nj0; nr = nnr;
        nxt = got_task(asked);
        stage ^= 1u;
    }
    q.unpack(vf_gm_drain_cells<DEPTH>(f, w, q.pack(), true, axis, dem, vis, big_prims, big_capacity, counters));
    vf_gm_drain_long<DEPTH>(f, w, q, true, lane, vis);
    if (dbg) {
        unsigned long long t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        if (lane == 0) dbg[2u * (blockIdx.x * VF_GM_WARPS + (threadIdx.x >> 5)) + 1u] = t1;
    }
}
