// vf_misc.cu -- triangle path, public grid generator, R32F height image round trip.
#include <string.h>

#include "vf_device.cuh"
#include "vf_internal.h"

// ---------------------------------------------------------------------------------------
// k_triangle: the fixed triangle of src/lib.rs:72-91 through the same snap / edge-function /
// top-left / sRGB code as the terrain path; src/shaders/triangle.wgsl:1-23 (pass-through
// position, colour varying), clear WHITE (src/lib.rs:19,703).  One thread = 4 pixels of a row,
// one 16-byte store.
__global__ void __launch_bounds__(256)
k_triangle(uint32_t W, uint32_t H, const float* __restrict__ thr_g, uint8_t* __restrict__ out) {
    __shared__ float s_thr[256];
    s_thr[threadIdx.x] = thr_g[threadIdx.x];
    __syncthreads();
    const float pos[3][2] = {{-0.8f, -0.8f}, {0.8f, -0.8f}, {0.0f, 0.8f}};
    const float col[3][3] = {{1.0f, 0.2f, 0.2f}, {0.2f, 1.0f, 0.2f}, {0.2f, 0.2f, 1.0f}};
    VfFrame f;
    f.sparse = 0;
    f.hw = __fmul_rn(0.5f, (float)W); f.hh = __fmul_rn(0.5f, (float)H);
    VfSubTri t;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        VfVtx v;
        v.c[0] = pos[k][0]; v.c[1] = pos[k][1]; v.c[2] = 0.0f; v.c[3] = 1.0f;
#pragma unroll
        for (int a = 0; a < 5; ++a) v.a[a] = 0.0f;
        vf_to_screen(f, v, t.v[k]);
    }
    const bool front = vf_area2(t) < 0;
    const uint32_t groups_per_row = (W + 3u) / 4u;
    const size_t total = (size_t)groups_per_row * H;
    for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < total;
         g += (size_t)gridDim.x * blockDim.x) {
        const uint32_t py = (uint32_t)(g / groups_per_row);
        const uint32_t px0 = (uint32_t)(g - (size_t)py * groups_per_row) * 4u;
        uint32_t px4[4];
#pragma unroll
        for (uint32_t q = 0; q < 4; ++q) {
            const uint32_t px = px0 + q;
            uint32_t rgba = 0xffffffffu;
            int64_t e[3];
            if (front && px < W && vf_covers(t, (int)px, (int)py, e)) {
                float b[3];
                vf_barycentric(e, b);
                rgba = 0xff000000u;
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const float c = fmaf(b[2], col[2][k], fmaf(b[1], col[1][k], __fmul_rn(b[0], col[0][k])));
                    rgba |= vf_srgb_encode(s_thr, c) << (8 * k);
                }
            }
            px4[q] = rgba;
        }
        uint8_t* dst = out + ((size_t)py * W + px0) * 4;
        if ((W & 3u) == 0u) {
            *reinterpret_cast<uint4*>(dst) = make_uint4(px4[0], px4[1], px4[2], px4[3]);
        } else {
            for (uint32_t q = 0; q < 4 && px0 + q < W; ++q) reinterpret_cast<uint32_t*>(dst)[q] = px4[q];
        }
    }
}

extern "C" int vf_triangle_render(int w, int h, uint8_t* out_rgba_host) {
    if (w <= 0 || h <= 0 || !out_rgba_host) return vf_set_error(VF_E_INVALID, "invalid size / null output");
    int rc = vf_require_device();
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(vf_util_mutex());
    cudaStream_t st = vf_util_stream();
    const size_t bytes = (size_t)w * h * 4;
    uint8_t* d = (uint8_t*)vf_scratch(bytes);            // cached per device: no cudaMalloc / cudaFree per call
    if (!st || !d) return VF_E_CUDA;
    const size_t groups = (size_t)((w + 3) / 4) * h;
    const int blocks = (int)((groups + 255) / 256 > 148 * 8 ? 148 * 8 : (groups + 255) / 256);
    k_triangle<<<blocks, 256, 0, st>>>((uint32_t)w, (uint32_t)h, vf_srgb_thresholds_dev(), d);
    VF_CUDA(cudaGetLastError());
    VF_CUDA(cudaMemcpyAsync(out_rgba_host, d, bytes, cudaMemcpyDeviceToHost, st));
    VF_CUDA(cudaStreamSynchronize(st));
    return VF_OK;
}

// ---------------------------------------------------------------------------------------
// k_grid: src/terrain/mesh.rs:35-90.  Thread = one vertex; the thread of cell (x,y) also writes
// that cell's 6 indices.  Rust never contracts, so every mul/add/div is a separate rounding.
__global__ void __launch_bounds__(256)
k_grid(uint32_t w, uint32_t h, float dx, float dy, float2* __restrict__ xy, float2* __restrict__ uv,
       uint32_t* __restrict__ idx) {
    const float wm1 = __fsub_rn((float)w, 1.0f), hm1 = __fsub_rn((float)h, 1.0f);
    const float cx = __fmul_rn(__fmul_rn(wm1, 0.5f), dx);
    const float cy = __fmul_rn(__fmul_rn(hm1, 0.5f), dy);
    const size_t n = (size_t)w * h;
    for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < n;
         k += (size_t)gridDim.x * blockDim.x) {
        const uint32_t y = (uint32_t)(k / w), x = (uint32_t)(k - (size_t)y * w);
        xy[k] = make_float2(__fsub_rn(__fmul_rn((float)x, dx), cx), __fsub_rn(__fmul_rn((float)y, dy), cy));
        uv[k] = make_float2(__fdiv_rn((float)x, wm1), __fdiv_rn((float)y, hm1));
        if (x + 1 < w && y + 1 < h) {
            const uint32_t i0 = y * w + x, i1 = i0 + 1, i2 = i0 + w, i3 = i2 + 1;
            uint32_t* o = idx + ((size_t)y * (w - 1) + x) * 6;
            // 24 bytes per cell, 8-byte aligned: three 8-byte stores
            reinterpret_cast<uint2*>(o)[0] = make_uint2(i0, i1);
            reinterpret_cast<uint2*>(o)[1] = make_uint2(i2, i2);
            reinterpret_cast<uint2*>(o)[2] = make_uint2(i1, i3);
        }
    }
}

extern "C" int vf_grid_generate(uint32_t nx, uint32_t nz, float dx, float dy, float* xy, float* uv, uint32_t* idx) {
    if (nx < 2 || nz < 2) return vf_set_error(VF_E_INVALID, "nx and nz must be >= 2");
    if (!isfinite(dx) || !isfinite(dy) || dx <= 0.0f || dy <= 0.0f)
        return vf_set_error(VF_E_INVALID, "spacing components must be finite and > 0");
    if ((uint64_t)nx * nz > 0xffffffffull) return vf_set_error(VF_E_INVALID, "nx*nz exceeds u32 indices");
    if (!xy || !uv || !idx) return vf_set_error(VF_E_INVALID, "null output");
    int rc = vf_require_device();
    if (rc) return rc;
    const size_t n = (size_t)nx * nz, m = (size_t)(nx - 1) * (nz - 1) * 6;
    std::lock_guard<std::mutex> lk(vf_util_mutex());
    cudaStream_t st = vf_util_stream();
    const size_t o_uv = (n * 8 + 255) & ~(size_t)255, o_idx = o_uv + ((n * 8 + 255) & ~(size_t)255);
    char* d = (char*)vf_scratch(o_idx + m * 4);          // one cached allocation: xy | uv | indices
    if (!st || !d) return VF_E_CUDA;
    const size_t b = (n + 255) / 256;
    k_grid<<<(int)(b > 148 * 16 ? 148 * 16 : b), 256, 0, st>>>(nx, nz, dx, dy, (float2*)d, (float2*)(d + o_uv), (uint32_t*)(d + o_idx));
    VF_CUDA(cudaGetLastError());
    VF_CUDA(cudaMemcpyAsync(xy, d, n * 8, cudaMemcpyDeviceToHost, st));
    VF_CUDA(cudaMemcpyAsync(uv, d + o_uv, n * 8, cudaMemcpyDeviceToHost, st));
    VF_CUDA(cudaMemcpyAsync(idx, d + o_idx, m * 4, cudaMemcpyDeviceToHost, st));
    VF_CUDA(cudaStreamSynchronize(st));
    return VF_OK;
}

// ---------------------------------------------------------------------------------------
// R32F height image: src/lib.rs:496-681.  wgpu's 256-byte row padding has no equivalent here:
// the image is a pitched device allocation and patches come back with cudaMemcpy2D.
extern "C" int vf_height_upload(const float* host, uint32_t w, uint32_t h, vf_height** out) {
    if (!host || !out) return vf_set_error(VF_E_INVALID, "null argument");
    if (w == 0 || h == 0) return vf_set_error(VF_E_INVALID, "terrain dimensions are zero");
    int rc = vf_require_device();
    if (rc) return rc;
    vf_height* t = new vf_height();
    t->w = w; t->h = h;
    t->pitch = (((size_t)w * 4) + 15) & ~(size_t)15;
    if (cudaGetDevice(&t->device) != cudaSuccess) t->device = 0;
    cudaError_t e = cudaMalloc(&t->d, t->pitch * h);
    if (e == cudaSuccess)
        e = cudaMemcpy2D(t->d, t->pitch, host, (size_t)w * 4, (size_t)w * 4, h, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        cudaFree(t->d);
        delete t;
        (void)cudaGetLastError();
        return vf_set_error(e == cudaErrorMemoryAllocation ? VF_E_OOM : VF_E_CUDA, "height upload: %s", cudaGetErrorString(e));
    }
    *out = t;
    return VF_OK;
}

extern "C" int vf_height_size(const vf_height* t, uint32_t* w, uint32_t* h) {
    if (!t) return vf_set_error(VF_E_INVALID, "null height image");
    if (w) *w = t->w;
    if (h) *h = t->h;
    return VF_OK;
}

extern "C" int vf_height_read_patch(const vf_height* t, uint32_t x, uint32_t y, uint32_t w, uint32_t h, float* out_host) {
    if (!t || !out_host) return vf_set_error(VF_E_INVALID, "null argument");
    if (w == 0 || h == 0) return vf_set_error(VF_E_INVALID, "patch dimensions must be > 0");
    if ((uint64_t)x + w > t->w)
        return vf_set_error(VF_E_INVALID, "requested patch exceeds texture bounds in x: x+w (%u) > width (%u)", x + w, t->w);
    if ((uint64_t)y + h > t->h)
        return vf_set_error(VF_E_INVALID, "requested patch exceeds texture bounds in y: y+h (%u) > height (%u)", y + h, t->h);
    const char* src = (const char*)t->d + (size_t)y * t->pitch + (size_t)x * 4;
    VF_CUDA(cudaMemcpy2D(out_host, (size_t)w * 4, src, t->pitch, (size_t)w * 4, h, cudaMemcpyDeviceToHost));
    return VF_OK;
}

extern "C" void vf_height_destroy(vf_height* t) {
    if (!t) return;
    cudaFree(t->d);
    delete t;
}

// ---------------------------------------------------------------------------------------
// vf_selftest_div: the division helpers of vf_device.cuh (vf_rcp1 / vf_rcp3 / vf_div1 / vf_div_by / vf_div3) against the
// correctly rounded intrinsic they stand for, bit for bit, on the device.  Operands: a counter-based generator mixes
// plain random bit patterns (every exponent, subnormals, infinities, NaNs), values of the magnitudes the shading sees,
// values around the helpers' window limits, and exact zeros.  No reference counterpart: a diagnostic of this library.
__device__ __forceinline__ uint32_t vf_st_hash(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}
__device__ __forceinline__ float vf_st_operand(uint32_t h) {
    const uint32_t kind = h & 7u, bits = vf_st_hash(h ^ 0x9e3779b9u);
    if (kind < 2u) return __uint_as_float(bits);                                             // anything at all
    if (kind == 2u) {                                                                         // around the window limits
        const uint32_t e = (bits & 1u) ? 127u - 31u : 127u + 31u;
        return __uint_as_float((bits & 0x80000000u) | ((e - 1u + ((bits >> 1) % 3u)) << 23) | ((bits >> 3) & ((bits & 4u) ? 0x7fffffu : 0x3u)));
    }
    if (kind == 3u && (bits & 0xf0u) == 0u) return __uint_as_float(bits & 0x80000000u);       // a zero now and then
    const uint32_t e = 127u - 20u + (bits >> 8) % 40u;                                        // 2^-20 .. 2^20: the shading's range
    return __uint_as_float((bits & 0x80000000u) | (e << 23) | (vf_st_hash(bits) & 0x7fffffu));
}
__device__ __forceinline__ bool vf_st_same(float x, float y) {
    return __float_as_uint(x) == __float_as_uint(y) || (x != x && y != y);
}
__global__ void __launch_bounds__(256)
k_selftest_div(unsigned long long n, uint32_t seed, unsigned long long* __restrict__ bad) {
    unsigned long long mine = 0ull;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (unsigned long long)gridDim.x * blockDim.x) {
        const uint32_t h = vf_st_hash((uint32_t)i * 0x9e3779b1u + (uint32_t)(i >> 32) + seed);
        float a[5], b[3], q[5], want;
#pragma unroll
        for (int k = 0; k < 5; ++k) a[k] = vf_st_operand(vf_st_hash(h + 0x1000193u * (uint32_t)(k + 1)));
#pragma unroll
        for (int k = 0; k < 3; ++k) b[k] = vf_st_operand(vf_st_hash(h ^ (0x85ebca6bu * (uint32_t)(k + 7))));
        if (h & 0x100u) {                                        // groups that stay inside the window more often
#pragma unroll
            for (int k = 0; k < 5; ++k) a[k] = vf_st_operand((vf_st_hash(h + 31u * (uint32_t)k) & ~7u) | 4u);
#pragma unroll
            for (int k = 0; k < 3; ++k) b[k] = vf_st_operand((vf_st_hash(h + 77u * (uint32_t)k) & ~7u) | 5u);
        }
        uint32_t miss = 0u;
        miss += vf_st_same(vf_rcp1(b[0]), __fdiv_rn(1.0f, b[0])) ? 0u : 1u;
        miss += vf_st_same(vf_rcp1(b[0]), __frcp_rn(b[0])) ? 0u : 1u;
        float r3[3];
        vf_rcp3(b, r3);
#pragma unroll
        for (int k = 0; k < 3; ++k) miss += vf_st_same(r3[k], __fdiv_rn(1.0f, b[k])) ? 0u : 1u;
        miss += vf_st_same(vf_div1(a[0], b[1]), __fdiv_rn(a[0], b[1])) ? 0u : 1u;
#pragma unroll
        for (int k = 0; k < 5; ++k) q[k] = -1.0f;
        vf_div_by<5>(a, b[2], q, 0x1du);                         // numerators 0, 2, 3, 4
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            want = ((0x1du >> k) & 1u) ? __fdiv_rn(a[k], b[2]) : -1.0f;
            miss += vf_st_same(q[k], want) ? 0u : 1u;
        }
        vf_div3(a, b, q);
#pragma unroll
        for (int k = 0; k < 3; ++k) miss += vf_st_same(q[k], __fdiv_rn(a[k], b[k])) ? 0u : 1u;
        mine += miss;
    }
    if (mine) atomicAdd(bad, mine);
}

extern "C" int vf_selftest_div(uint64_t n_groups, uint32_t seed, uint64_t* mismatches) {
    if (!mismatches) return vf_set_error(VF_E_INVALID, "null output");
    int rc = vf_require_device();
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(vf_util_mutex());
    cudaStream_t st = vf_util_stream();
    unsigned long long* d = (unsigned long long*)vf_scratch(sizeof(unsigned long long));
    if (!st || !d) return VF_E_CUDA;
    VF_CUDA(cudaMemsetAsync(d, 0, sizeof(unsigned long long), st));
    if (n_groups) {
        k_selftest_div<<<148 * 8, 256, 0, st>>>((unsigned long long)n_groups, seed, d);
        VF_CUDA(cudaGetLastError());
    }
    unsigned long long hres = 0ull;
    VF_CUDA(cudaMemcpyAsync(&hres, d, sizeof(hres), cudaMemcpyDeviceToHost, st));
    VF_CUDA(cudaStreamSynchronize(st));
    *mismatches = (uint64_t)hres;
    return VF_OK;
}

// ---------------------------------------------------------------------------------------
// vf_selftest_srgb: vf_srgb_encode_fast (guess + one correction) against vf_srgb_encode (binary search over the exact
// thresholds) for every float from +0 to 1.0 inclusive, plus negative values, values above 1, infinities and NaNs.
__global__ void __launch_bounds__(256)
k_selftest_srgb(const float* __restrict__ thr_g, unsigned long long* __restrict__ bad) {
    __shared__ float s_thr[257];
    s_thr[threadIdx.x] = thr_g[threadIdx.x];
    if (threadIdx.x == 0) s_thr[256] = __int_as_float(0x7f800000);
    __syncthreads();
    unsigned long long mine = 0ull;
    const unsigned long long n_unit = 0x3f800000ull + 1ull, n_extra = 1ull << 24;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_unit + n_extra;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        // past the unit interval: bit patterns spread over (1, +inf], NaNs and the negative half
        const uint32_t bits = i < n_unit ? (uint32_t)i : 0x3f800001u + (uint32_t)(i - n_unit) * 193u;
        const float x = __uint_as_float(bits);
        mine += vf_srgb_encode_fast(s_thr, x) != vf_srgb_encode(s_thr, x) ? 1ull : 0ull;
    }
    if (mine) atomicAdd(bad, mine);
}

extern "C" int vf_selftest_srgb(uint64_t* mismatches) {
    if (!mismatches) return vf_set_error(VF_E_INVALID, "null output");
    int rc = vf_require_device();
    if (rc) return rc;
    const float* thr = vf_srgb_thresholds_dev();
    if (!thr) return VF_E_CUDA;
    std::lock_guard<std::mutex> lk(vf_util_mutex());
    cudaStream_t st = vf_util_stream();
    unsigned long long* d = (unsigned long long*)vf_scratch(sizeof(unsigned long long));
    if (!st || !d) return VF_E_CUDA;
    VF_CUDA(cudaMemsetAsync(d, 0, sizeof(unsigned long long), st));
    k_selftest_srgb<<<148 * 8, 256, 0, st>>>(thr, d);
    VF_CUDA(cudaGetLastError());
    unsigned long long hres = 0ull;
    VF_CUDA(cudaMemcpyAsync(&hres, d, sizeof(hres), cudaMemcpyDeviceToHost, st));
    VF_CUDA(cudaStreamSynchronize(st));
    *mismatches = (uint64_t)hres;
    return VF_OK;
}
