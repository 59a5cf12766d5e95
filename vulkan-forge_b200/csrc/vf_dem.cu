// vf_dem.cu -- DEM statistics on the device.
// replaces: terrain_stats::min_max (src/terrain_stats.rs:11-35), dem_stats_from_slice and
// normalize_in_place (src/lib.rs:905-951), the f32/f64 ingest loops of add_terrain
// (src/lib.rs:351-388).
//   k_dem_gather : strided sample (or all values) -> compact buffer        (HBM: 32-B sectors)
//   k_dem_select : exact k-th order statistics by 4-pass 8-bit radix select (one CTA, L2/smem)
//   k_dem_reduce : min / max / sum, then sum of squared deviations          (HBM streaming)
//   k_dem_normalize, k_dem_ingest_* : elementwise                           (HBM streaming)
#include <math.h>
#include <string.h>

#include "vf_internal.h"

// order-preserving key; -0.0 and +0.0 share one key because partial_cmp calls them Equal
__device__ __forceinline__ uint32_t vf_float_key(float v) {
    uint32_t b = __float_as_uint(v);
    if (v == 0.0f) b = 0u;
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float vf_key_float(uint32_t k) {
    const uint32_t b = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
    return __uint_as_float(b);
}

__global__ void k_dem_gather(const float* __restrict__ data, uint32_t w, uint32_t pitch,
                             size_t step, uint32_t m, float* __restrict__ vals) {
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < m; k += gridDim.x * blockDim.x) {
        const size_t idx = (size_t)k * step;
        const size_t r = idx / w, c = idx - r * w;
        vals[k] = __ldg(data + r * pitch + c);
    }
}

// Exact order statistics of vals[0..m) for the two ranks rank_in[0..1].  Semantics of a STABLE
// sort under partial_cmp: among equal keys the input order is kept, which only matters for the
// sign of a zero result (resolved in the last step).  One CTA of 1024 threads.
__global__ void __launch_bounds__(1024)
k_dem_select(const float* __restrict__ vals, uint32_t m, uint32_t rank0, uint32_t rank1,
             float* __restrict__ out2) {
    __shared__ uint32_t hist[2][256];
    __shared__ uint32_t s_prefix[2], s_rank[2];
    __shared__ uint32_t s_cnt[1024];
    const uint32_t tid = threadIdx.x;
    if (tid == 0) { s_prefix[0] = s_prefix[1] = 0u; s_rank[0] = rank0; s_rank[1] = rank1; }
    __syncthreads();
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = 24 - 8 * pass;
        const uint32_t mask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
        if (tid < 512) hist[tid >> 8][tid & 255u] = 0u;
        __syncthreads();
        const uint32_t p0 = s_prefix[0], p1 = s_prefix[1];
        for (uint32_t k = tid; k < m; k += 1024) {
            const uint32_t key = vf_float_key(vals[k]);
            const uint32_t bin = (key >> shift) & 255u;
            if ((key & mask) == p0) atomicAdd(&hist[0][bin], 1u);
            if ((key & mask) == p1) atomicAdd(&hist[1][bin], 1u);
        }
        __syncthreads();
        if (tid < 2) {
            uint32_t r = s_rank[tid], cum = 0u, b = 0u;
            for (; b < 256u; ++b) {
                const uint32_t c = hist[tid][b];
                if (cum + c > r) break;
                cum += c;
            }
            s_rank[tid] = r - cum;
            s_prefix[tid] |= b << shift;
        }
        __syncthreads();
    }
    // s_prefix = key of the answer, s_rank = index among equal keys in input order
    for (int q = 0; q < 2; ++q) {
        const uint32_t key = s_prefix[q];
        float ans = vf_key_float(key);
        if (key == 0x80000000u) {
            // zero: the sign is the one of the s_rank[q]-th zero of the sample, in input order
            const uint32_t chunk = (m + 1023u) / 1024u;
            const uint32_t b = min(tid * chunk, m), e = min(b + chunk, m);
            uint32_t c = 0;
            for (uint32_t k = b; k < e; ++k) c += (vals[k] == 0.0f) ? 1u : 0u;
            s_cnt[tid] = c;
            __syncthreads();
            if (tid == 0) {
                uint32_t want = s_rank[q], t = 0;
                while (t < 1023u && want >= s_cnt[t]) { want -= s_cnt[t]; ++t; }
                const uint32_t bb = min(t * chunk, m), ee = min(bb + chunk, m);
                for (uint32_t k = bb; k < ee; ++k)
                    if (vals[k] == 0.0f) { if (want == 0u) { ans = vals[k]; break; } --want; }
                out2[q] = ans;
            }
            __syncthreads();
        } else if (tid == 0) {
            out2[q] = ans;
        }
    }
}

// acc[0] = min key, acc[1] = max key (uint32), accd[0] = sum, accd[1] = sum (v-mean)^2
__global__ void __launch_bounds__(256)
k_dem_reduce(const float* __restrict__ data, uint32_t w, uint32_t h, uint32_t pitch, int phase,
             double mean, uint32_t* __restrict__ acc, double* __restrict__ accd) {
    __shared__ uint32_t s_mn[8], s_mx[8];
    __shared__ double s_sum[8];
    const size_t n = (size_t)w * h;
    uint32_t mn = 0xffffffffu, mx = 0u;
    double sum = 0.0;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n;
         idx += (size_t)gridDim.x * blockDim.x) {
        const size_t r = idx / w, c = idx - r * w;
        const float v = __ldg(data + r * pitch + c);
        if (phase == 0) {
            if (v == v) { const uint32_t k = vf_float_key(v); mn = min(mn, k); mx = max(mx, k); }
            sum += (double)v;
        } else {
            const double d = (double)v - mean;
            sum += d * d;
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, d));
        mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, d));
        sum += __shfl_xor_sync(0xffffffffu, sum, d);
    }
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    if (lane == 0) { s_mn[wid] = mn; s_mx[wid] = mx; s_sum[wid] = sum; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < 8; ++k) { mn = min(mn, s_mn[k]); mx = max(mx, s_mx[k]); sum += s_sum[k]; }
        if (phase == 0) { atomicMin(&acc[0], mn); atomicMax(&acc[1], mx); atomicAdd(&accd[0], sum); }
        else atomicAdd(&accd[1], sum);
    }
}

// index of the first +-0 of the image.  `if v < lo { lo = v }` (terrain_stats.rs:17, lib.rs:914) keeps
// the FIRST element that equals the extreme value, and -0 == +0: when an extreme is zero its sign is
// the sign of the first zero in index order, which an order-free reduction cannot know.
__global__ void __launch_bounds__(256)
k_dem_first_zero(const float* __restrict__ data, uint32_t w, uint32_t h, uint32_t pitch, unsigned long long* __restrict__ first) {
    const size_t n = (size_t)w * h;
    unsigned long long best = ~0ull;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t)gridDim.x * blockDim.x) {
        const size_t r = idx / w, c = idx - r * w;
        if (__ldg(data + r * pitch + c) == 0.0f) { best = idx; break; }      // a thread's indices ascend
    }
    if (best != ~0ull) atomicMin(first, best);
}

// replaces a zero extreme by the first zero of the image (sign included); out2 = (lo, hi)
static int fix_zero_sign(const float* d_data, uint32_t w, uint32_t h, uint32_t pitch, float out2[2], cudaStream_t stream) {
    if (out2[0] != 0.0f && out2[1] != 0.0f) return VF_OK;
    unsigned long long* d_first = nullptr;
    VF_CUDA(cudaMalloc(&d_first, 8));
    unsigned long long first = ~0ull;
    cudaError_t e = cudaMemcpyAsync(d_first, &first, 8, cudaMemcpyHostToDevice, stream);
    if (e == cudaSuccess) {
        const size_t n = (size_t)w * h;
        const size_t blocks = (n + 255) / 256;
        k_dem_first_zero<<<(int)(blocks > 148 * 16 ? 148 * 16 : blocks), 256, 0, stream>>>(d_data, w, h, pitch, d_first);
        e = cudaMemcpyAsync(&first, d_first, 8, cudaMemcpyDeviceToHost, stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    float z = 0.0f;
    if (e == cudaSuccess && first != ~0ull)
        e = cudaMemcpy(&z, d_data + (size_t)(first / w) * pitch + (size_t)(first % w), 4, cudaMemcpyDeviceToHost);
    cudaFree(d_first);
    VF_CUDA(e);
    if (out2[0] == 0.0f) out2[0] = z;
    if (out2[1] == 0.0f) out2[1] = z;
    return VF_OK;
}

__global__ void k_dem_keys_to_float(const uint32_t* __restrict__ acc, float* __restrict__ out2) {
    out2[0] = vf_key_float(acc[0]);
    out2[1] = vf_key_float(acc[1]);
}

__global__ void k_dem_normalize(float* __restrict__ d, size_t n, int mode, float a, float scale, float lo) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float v = d[i];
        // src/lib.rs:941 `(*v - min) * scale + lo` and :947 `(*v - mean) / denom` (no FMA in Rust)
        d[i] = mode == 0 ? __fadd_rn(__fmul_rn(__fsub_rn(v, a), scale), lo) : __fdiv_rn(__fsub_rn(v, a), scale);
    }
}
__global__ void k_dem_ingest_f64(const double* __restrict__ s, size_t n, float scale, float* __restrict__ d) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        d[i] = __fmul_rn((float)s[i], scale);      // src/lib.rs:384
}
__global__ void k_dem_ingest_f32(const float* s, size_t n, float scale, float* d) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        d[i] = __fmul_rn(s[i], scale);             // src/lib.rs:362
}

static inline int grid_for(size_t n, int threads) {
    const size_t b = (n + threads - 1) / threads;
    return (int)(b < 1 ? 1 : (b > 148 * 16 ? 148 * 16 : b));
}

struct DevBuf {
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
};

// (lo, hi) of a device-resident image.  clamp=1: percentile semantics of terrain_stats.rs:20-34.
int vf_dem_minmax_device(const float* d_data, uint32_t w, uint32_t h, uint32_t pitch, int clamp,
                         float* lo, float* hi, cudaStream_t stream) {
    const size_t n = (size_t)w * h;
    if (n == 0) return vf_set_error(VF_E_INVALID, "heightmap slice empty");
    DevBuf scratch;
    float out[2];
    if (!clamp) {
        VF_CUDA(cudaMalloc(&scratch.p, 64));
        uint32_t* acc = (uint32_t*)scratch.p;
        double* accd = (double*)((char*)scratch.p + 16);
        float* o2 = (float*)((char*)scratch.p + 48);
        const uint32_t init[4] = {0xffffffffu, 0u, 0u, 0u};
        VF_CUDA(cudaMemcpyAsync(acc, init, 16, cudaMemcpyHostToDevice, stream));
        VF_CUDA(cudaMemsetAsync(accd, 0, 32, stream));
        k_dem_reduce<<<grid_for(n, 256), 256, 0, stream>>>(d_data, w, h, pitch, 0, 0.0, acc, accd);
        k_dem_keys_to_float<<<1, 1, 0, stream>>>(acc, o2);
        VF_CUDA(cudaMemcpyAsync(out, o2, 8, cudaMemcpyDeviceToHost, stream));
        VF_CUDA(cudaStreamSynchronize(stream));
        const int rc = fix_zero_sign(d_data, w, h, pitch, out, stream);
        if (rc) return rc;
    } else {
        const size_t SAMPLE = 65536;
        size_t step = 1, m = n;
        if (n > SAMPLE) { step = n / SAMPLE; m = (n + step - 1) / step; }   // terrain_stats.rs:23-27
        const uint32_t r1 = (uint32_t)((float)m * 0.01f), r99 = (uint32_t)((float)m * 0.99f);  // :32-33
        VF_CUDA(cudaMalloc(&scratch.p, m * 4 + 16));
        float* vals = (float*)scratch.p;
        float* o2 = vals + m;
        k_dem_gather<<<grid_for(m, 256), 256, 0, stream>>>(d_data, w, pitch, step, (uint32_t)m, vals);
        k_dem_select<<<1, 1024, 0, stream>>>(vals, (uint32_t)m, r1, r99, o2);
        VF_CUDA(cudaMemcpyAsync(out, o2, 8, cudaMemcpyDeviceToHost, stream));
        VF_CUDA(cudaStreamSynchronize(stream));
    }
    VF_CUDA(cudaGetLastError());
    *lo = out[0]; *hi = out[1];
    return VF_OK;
}

static int upload(const void* host, size_t bytes, DevBuf& b) {
    VF_CUDA(cudaMalloc(&b.p, bytes ? bytes : 4));
    VF_CUDA(cudaMemcpy(b.p, host, bytes, cudaMemcpyHostToDevice));
    return VF_OK;
}

extern "C" int vf_dem_minmax(const float* data, size_t n, int clamp, float* lo, float* hi) {
    if (!data || !lo || !hi) return vf_set_error(VF_E_INVALID, "null argument");
    if (n == 0) return vf_set_error(VF_E_INVALID, "heightmap slice empty");
    if (n > 0xffffffffull) return vf_set_error(VF_E_INVALID, "more than 2^32 samples");
    int rc = vf_require_device();
    if (rc) return rc;
    DevBuf d;
    rc = upload(data, n * 4, d);
    if (rc) return rc;
    // present the flat slice as rows of <= 65536 columns... a 1-row image keeps indices identical
    return vf_dem_minmax_device((const float*)d.p, (uint32_t)n, 1, (uint32_t)n, clamp, lo, hi, 0);
}

static int stats_device(const float* d_data, uint32_t w, uint32_t h, uint32_t pitch, float out4[4]) {
    const size_t n = (size_t)w * h;
    DevBuf scratch;
    VF_CUDA(cudaMalloc(&scratch.p, 64));
    uint32_t* acc = (uint32_t*)scratch.p;
    double* accd = (double*)((char*)scratch.p + 16);
    const uint32_t init[4] = {0xffffffffu, 0u, 0u, 0u};
    VF_CUDA(cudaMemcpy(acc, init, 16, cudaMemcpyHostToDevice));
    VF_CUDA(cudaMemset(accd, 0, 32));
    k_dem_reduce<<<grid_for(n, 256), 256>>>(d_data, w, h, pitch, 0, 0.0, acc, accd);
    double hd[2];
    uint32_t hk[2];
    VF_CUDA(cudaMemcpy(hd, accd, 16, cudaMemcpyDeviceToHost));
    VF_CUDA(cudaMemcpy(hk, acc, 8, cudaMemcpyDeviceToHost));
    const double mean = hd[0] / (double)n;
    k_dem_reduce<<<grid_for(n, 256), 256>>>(d_data, w, h, pitch, 1, mean, acc, accd);
    VF_CUDA(cudaMemcpy(hd, accd, 16, cudaMemcpyDeviceToHost));
    VF_CUDA(cudaGetLastError());
    auto kf = [](uint32_t k) { uint32_t b = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k; float f; memcpy(&f, &b, 4); return f; };
    out4[0] = kf(hk[0]); out4[1] = kf(hk[1]);
    const int rcz = fix_zero_sign(d_data, w, h, pitch, out4, 0);
    if (rcz) return rcz;
    out4[2] = (float)mean;
    out4[3] = (float)sqrt(hd[1] / (double)n);
    return VF_OK;
}

extern "C" int vf_dem_stats(const float* data, size_t n, float out4[4]) {
    if (!data || !out4) return vf_set_error(VF_E_INVALID, "null argument");
    if (n == 0) { out4[0] = out4[1] = out4[2] = out4[3] = 0.0f; return VF_OK; }   // src/lib.rs:906-908
    if (n > 0xffffffffull) return vf_set_error(VF_E_INVALID, "more than 2^32 samples");
    int rc = vf_require_device();
    if (rc) return rc;
    DevBuf d;
    rc = upload(data, n * 4, d);
    if (rc) return rc;
    return stats_device((const float*)d.p, (uint32_t)n, 1, (uint32_t)n, out4);
}

extern "C" int vf_dem_normalize(float* data, size_t n, int mode, float eps, float lo, float hi) {
    if (!data) return vf_set_error(VF_E_INVALID, "null argument");
    if (mode != 0 && mode != 1) return vf_set_error(VF_E_INVALID, "mode must be 'minmax' or 'zscore'");
    if (n == 0) return VF_OK;
    if (n > 0xffffffffull) return vf_set_error(VF_E_INVALID, "more than 2^32 samples");
    int rc = vf_require_device();
    if (rc) return rc;
    DevBuf d;
    rc = upload(data, n * 4, d);
    if (rc) return rc;
    float st[4];
    rc = stats_device((const float*)d.p, (uint32_t)n, 1, (uint32_t)n, st);
    if (rc) return rc;
    float a, scale;
    if (mode == 0) {                                   // src/lib.rs:936-943
        float denom = fabsf(st[1] - st[0]);
        if (!(denom > eps)) denom = eps;
        scale = (hi - lo) / denom;
        a = st[0];
    } else {                                           // :944-949
        scale = st[3] > eps ? st[3] : eps;
        a = st[2];
    }
    k_dem_normalize<<<grid_for(n, 256), 256>>>((float*)d.p, n, mode, a, scale, lo);
    VF_CUDA(cudaGetLastError());
    VF_CUDA(cudaMemcpy(data, d.p, n * 4, cudaMemcpyDeviceToHost));
    return VF_OK;
}

extern "C" int vf_dem_ingest_f64(const double* src, size_t n, float scale, float* dst) {
    if (!src || !dst) return vf_set_error(VF_E_INVALID, "null argument");
    if (n == 0) return VF_OK;
    int rc = vf_require_device();
    if (rc) return rc;
    DevBuf s, d;
    rc = upload(src, n * 8, s);
    if (rc) return rc;
    VF_CUDA(cudaMalloc(&d.p, n * 4));
    k_dem_ingest_f64<<<grid_for(n, 256), 256>>>((const double*)s.p, n, scale, (float*)d.p);
    VF_CUDA(cudaGetLastError());
    VF_CUDA(cudaMemcpy(dst, d.p, n * 4, cudaMemcpyDeviceToHost));
    return VF_OK;
}

extern "C" int vf_dem_ingest_f32(const float* src, size_t n, float scale, float* dst) {
    if (!src || !dst) return vf_set_error(VF_E_INVALID, "null argument");
    if (n == 0) return VF_OK;
    int rc = vf_require_device();
    if (rc) return rc;
    DevBuf s;
    rc = upload(src, n * 4, s);
    if (rc) return rc;
    k_dem_ingest_f32<<<grid_for(n, 256), 256>>>((const float*)s.p, n, scale, (float*)s.p);
    VF_CUDA(cudaGetLastError());
    VF_CUDA(cudaMemcpy(dst, s.p, n * 4, cudaMemcpyDeviceToHost));
    return VF_OK;
}
