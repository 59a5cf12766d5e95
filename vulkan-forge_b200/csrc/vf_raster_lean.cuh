// vf_raster_lean.cuh -- k_resolve_lean + k_shade_list: the shading pass of a frame that queued nothing for the
// sort-middle stage, as a streaming pass and a dense one.
//
// k_raster (vf_terrain_kernels.cuh) carries two rare paths out of line: the merge of binned primitives into a tile's keys
// and the clipper for a winner with a vertex outside the guard band.  Both exist only for primitives that went through
// the global queue of large / clipped triangles (counters[VF_CNT_BIG] > 0): the geometry pass writes a key to the
// visibility buffer only for a triangle whose three vertices passed the very test vf_shade_winner_h repeats.  The host
// leaves k_binsort out while frames queue nothing (vf_terrain.cu, "speculative binning") and renders a frame again with
// the full pipeline when it turns out that one did; these kernels are what it launches in the first case.
//
// Why two kernels.  Shading one pixel is ~600 dependent instructions behind a key -> vertex heights -> normal taps chain of
// loads, several microseconds for a warp; an empty pixel is one load and one store.  A tile kernel (k_raster: CTA = tile,
// thread = pixel) pays a CTA's launch and its load round trip for every tile of sky, keeps the empty lanes of a
// half-covered warp waiting for the covered ones, and its duration is that of the warp with the most covered tiles
// (ncu: at c3 only 14 % of the pixel warps shade anything, yet 1080p took 22 us however the loads were prefetched).
//   k_resolve_lean : thread = two quads of 4 adjacent pixels, 16-byte accesses: takes the keys (and resets them for the next
//                    frame), writes the clear colour, and appends every covered pixel as (x | y << 16, primitive) to a
//                    list -- warp scan, CTA scan, one atomic per CTA.  Pure streaming.
//   k_shade_list   : persistent, thread = list entry: every lane of every warp shades (no empty lanes, equal work), no
//                    calls, no stack, and the SM's warps all carry the long chain at once.  It is the frame's last
//                    kernel: it leaves the counter block zeroed for the next frame (the list length alternates between
//                    two counters, one per frame parity), so that no memset node separates it from the next frame's
//                    k_geom_march, which the host launches as its programmatic dependent: nothing passes from this
//                    kernel to that one, and it starts on the SMs this kernel's tail leaves idle.
// Results are those of k_raster, which the parity tests compare with the oracle through either path (VF_BIN_ALWAYS=1
// keeps the binning stage and k_raster in every frame, VF_RASTER_FULL=1 keeps k_raster alone).
#pragma once
#include "vf_terrain_kernels.cuh"

#ifndef VF_SHADE_CTAS
#define VF_SHADE_CTAS 6
#endif

// requires W % 4 == 0 (the host takes k_raster otherwise).  list: W * H entries.  Thread t handles the pixel quads t and
// t + half (two independent 16-byte loads in flight; 1080p is then a single wave of CTAs); w4_inv = floor(2^32 / (W / 4)).
template <bool DEPTH>
__global__ void __launch_bounds__(256)
k_resolve_lean(uint32_t W, uint32_t H, uint32_t w4_inv, uint32_t half, unsigned long long* __restrict__ vis, uint32_t* __restrict__ counters,
               uint32_t parity, uint32_t clear_rgba, uint8_t* __restrict__ out_rgba, uint32_t* __restrict__ out_vis, uint2* __restrict__ list,
               uint32_t* __restrict__ high_water) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t g0 = blockIdx.x * 256u + threadIdx.x, w4 = W >> 2, total = w4 * H;
    // high_water[2]: most primitives any frame of the batch queued for binning -- how the host learns that this frame
    // has to be rendered again with the binning stage (what these launches write for it is then discarded)
    if (g0 == 0u) {
        const uint32_t nbig = counters[VF_CNT_BIG];
        if (nbig) atomicMax(&high_water[2], nbig);
    }
    uint32_t id[2][4], xy[2];                                      // primitive id + 1 (0: nothing drawn here); x | y << 16 of the quad
    bool live[2];
    uint4 ka[2];
    ulonglong2 kd[2][2];
#pragma unroll
    for (int q = 0; q < 2; ++q) {                                  // both loads first
        const uint32_t g = g0 + (uint32_t)q * half;
        live[q] = g0 < half && g < total;
        uint32_t y = __umulhi(g, w4_inv);
        if (g - y * w4 >= w4) ++y;
        const uint32_t x = (g - y * w4) << 2;
        xy[q] = x | (y << 16);
        if (live[q]) {
            const size_t p = (size_t)y * W + x;
            if (DEPTH) { const ulonglong2* k = reinterpret_cast<const ulonglong2*>(vis + p); kd[q][0] = k[0]; kd[q][1] = k[1]; }
            else ka[q] = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint32_t*>(vis) + p);
        }
    }
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        id[q][0] = id[q][1] = id[q][2] = id[q][3] = 0u;
        if (live[q]) {
            const size_t p = (size_t)(xy[q] >> 16) * W + (xy[q] & 0xffffu);
            if (DEPTH) {
                ulonglong2* k = reinterpret_cast<ulonglong2*>(vis + p);
                const ulonglong2 a = kd[q][0], b = kd[q][1];
                id[q][0] = a.x == ~0ull ? 0u : (uint32_t)a.x + 1u; id[q][1] = a.y == ~0ull ? 0u : (uint32_t)a.y + 1u;
                id[q][2] = b.x == ~0ull ? 0u : (uint32_t)b.x + 1u; id[q][3] = b.y == ~0ull ? 0u : (uint32_t)b.y + 1u;
                if (id[q][0] | id[q][1]) k[0] = make_ulonglong2(~0ull, ~0ull);   // self-cleaning: reset for the next frame
                if (id[q][2] | id[q][3]) k[1] = make_ulonglong2(~0ull, ~0ull);
            } else {
                const uint4 a = ka[q];
                id[q][0] = a.x; id[q][1] = a.y; id[q][2] = a.z; id[q][3] = a.w;
                if (a.x | a.y | a.z | a.w) *reinterpret_cast<uint4*>(reinterpret_cast<uint32_t*>(vis) + p) = make_uint4(0u, 0u, 0u, 0u);
            }
            *reinterpret_cast<uint4*>(out_rgba + p * 4) = make_uint4(clear_rgba, clear_rgba, clear_rgba, clear_rgba);
            if (out_vis) *reinterpret_cast<uint4*>(out_vis + p) = make_uint4(id[q][0], id[q][1], id[q][2], id[q][3]);
        }
    }
    // compaction: exclusive scan of the threads' covered-pixel counts over the warp, then over the CTA's eight warps, and
    // ONE atomic per CTA (atomics on one address are served one after the other by its L2 slice: with one per warp the
    // few thousand of a 1080p frame took longer than the streaming itself)
    __shared__ uint32_t s_warp[8], s_base;
    uint32_t mine = 0u;
#pragma unroll
    for (int q = 0; q < 2; ++q)
#pragma unroll
        for (int b = 0; b < 4; ++b) mine += id[q][b] ? 1u : 0u;
    uint32_t incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= (uint32_t)d) incl += v;
    }
    const uint32_t wid = threadIdx.x >> 5;
    if (lane == 31u) s_warp[wid] = incl;
    __syncthreads();
    if (threadIdx.x == 0u) {
        uint32_t tot = 0u;
#pragma unroll
        for (int w = 0; w < 8; ++w) { const uint32_t c = s_warp[w]; s_warp[w] = tot; tot += c; }
        s_base = tot ? atomicAdd(&counters[VF_CNT_SHADE + parity], tot) : 0u;
    }
    __syncthreads();
    if (mine == 0u) return;
    const uint32_t base = s_base + s_warp[wid];
    uint32_t pos = base + incl - mine;
#pragma unroll
    for (int q = 0; q < 2; ++q)
#pragma unroll
        for (int b = 0; b < 4; ++b)
            if (id[q][b]) list[pos++] = make_uint2(xy[q] + (uint32_t)b, id[q][b] - 1u);
}

template <int MODE>
__global__ void __launch_bounds__(256, VF_SHADE_CTAS)
k_shade_list(const __grid_constant__ VfFrame f, const VfAxis* __restrict__ axis, const float* __restrict__ dem,
             const float* __restrict__ hgrid, uint32_t hpitch, const float4* __restrict__ lut, const float* __restrict__ thr_g,
             uint32_t* __restrict__ counters, uint32_t parity, const uint2* __restrict__ list, uint32_t clear_rgba, uint8_t* __restrict__ out_rgba) {
    __shared__ float s_thr[257];
    const uint32_t tid = threadIdx.x;
    s_thr[tid] = thr_g[tid];
    if (tid == 0u) s_thr[256] = __int_as_float(0x7f800000);
    // The first entry is requested together with the length of the list, not after it (one L2 round trip less before the
    // first shading instruction; an entry past the end is stale or unwritten memory of the list, read and dropped).
    uint32_t n = blockIdx.x * 256u + tid;
    uint2 e = __ldg(list + min(n, f.W * f.H - 1u));
    const uint32_t count = min(__ldcg(&counters[VF_CNT_SHADE + parity]), f.W * f.H);
    // This is the frame's last kernel: it leaves the counter block as the next frame's first kernel wants it (the host
    // then skips the memset between them).  The geometry pass's task and queue counters have no reader left; this
    // frame's list length is still being read by other CTAs, so it is the OTHER parity's -- last read a frame ago, next
    // written after this kernel -- that is reset here.
    if ((blockIdx.x | tid) == 0u) {
        counters[VF_CNT_TASK] = 0u; counters[VF_CNT_BIG] = 0u; counters[VF_CNT_SHADE + (parity ^ 1u)] = 0u;
        __threadfence();
    }
    __syncthreads();
    // the next frame's geometry pass, if it was launched as a programmatic dependent, may start wherever SMs fall idle
    // (every CTA is past the counter reset above; nothing else of this kernel is read by it)
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    for (; n < count; n += gridDim.x * 256u, e = n < count ? __ldg(list + n) : e) {
        const int px = (int)(e.x & 0xffffu), py = (int)(e.x >> 16);
        uint32_t rgba = clear_rgba;
        if (!vf_shade_winner<MODE>(f, axis, dem, hgrid, hpitch, lut, s_thr, e.y, px, py, rgba)) rgba = clear_rgba;   // (needs the clipper:
        *reinterpret_cast<uint32_t*>(out_rgba + ((size_t)py * f.W + px) * 4) = rgba;                               //  cannot win here, see above)
    }
}
