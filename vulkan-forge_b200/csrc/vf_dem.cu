// vf_dem.cu -- DEM statistics on the device.
// replaces: terrain_stats::min_max (src/terrain_stats.rs:11-35), dem_stats_from_slice and
// normalize_in_place (src/lib.rs:905-951), the f32/f64 ingest loops of add_terrain
// (src/lib.rs:351-388).
//   k_dem_gather : strided sample (or all values) -> compact buffer        (HBM: 32-B sectors)
//   k_dem_select : exact k-th order statistics by 4-pass 8-bit radix select (one CTA, L2/smem)
//   k_dem_moments: min / max / shifted sums / first zero in ONE pass        (HBM streaming)
//   k_dem_normalize, k_dem_ingest_f64, k_dem_scale : elementwise            (HBM streaming)
// Host arrays stream through a cached per-device scratch allocation on a utility stream (no cudaMalloc per call);
// device-resident images (vf_height) are processed in place.
#include <math.h>
#include <string.h>

#include <condition_variable>
#include <functional>
#include <thread>
#include <vector>

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
        // Heights of one terrain share their sign and exponent bits: in the first pass nearly every key falls into a handful
        // of bins, and one shared-memory atomic per value would serialise on them.  The lanes of a warp that hit the same
        // bin are found with match.any and their leader adds the count; while the two ranks still share their prefix (always
        // in the first pass) one histogram serves both.
        const bool same = p0 == p1;
        const uint32_t lane = tid & 31u;
        for (uint32_t k0 = 0; k0 < m; k0 += 1024) {          // (warp-uniform trip count: match.any needs the whole warp)
            const uint32_t k = k0 + tid;
            const uint32_t key = k < m ? vf_float_key(vals[k]) : 0u;
            const uint32_t bin = (key >> shift) & 255u;
            const bool in0 = k < m && (key & mask) == p0, in1 = !same && k < m && (key & mask) == p1;
            const uint32_t g0 = __match_any_sync(0xffffffffu, in0 ? bin : 256u);
            if (in0 && lane == (uint32_t)__ffs((int)g0) - 1u) atomicAdd(&hist[0][bin], (uint32_t)__popc(g0));
            if (!same) {
                const uint32_t g1 = __match_any_sync(0xffffffffu, in1 ? bin : 256u);
                if (in1 && lane == (uint32_t)__ffs((int)g1) - 1u) atomicAdd(&hist[1][bin], (uint32_t)__popc(g1));
            }
        }
        __syncthreads();
        if (same && tid < 256) hist[1][tid] = hist[0][tid];
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

// One pass over an image (or a chunk of a flat host array uploaded piecewise; `base` = index of its first element):
//   acc[0] = min key, acc[1] = max key (order-preserving uint32 keys; NaN skipped as `if v < lo` skips it),
//   accd[0] = sum (v - pivot), accd[1] = sum (v - pivot)^2 in double (mean and population variance follow; the shift
//   by a pivot value keeps the single pass free of cancellation),
//   first = index of the first +-0 of the slice.  `if v < lo { lo = v }` (terrain_stats.rs:17, lib.rs:914) keeps the
//   FIRST element that equals the extreme value, and -0 == +0: when an extreme is zero its sign is that of the first
//   zero in index order, which an order-free reduction cannot know.
__global__ void __launch_bounds__(256)
k_dem_moments(const float* __restrict__ data, uint32_t w, uint32_t h, uint32_t pitch, unsigned long long base, float pivot,
              uint32_t* __restrict__ acc, double* __restrict__ accd, unsigned long long* __restrict__ first) {
    __shared__ uint32_t s_mn[8], s_mx[8];
    __shared__ double s_sum[8], s_sq[8];
    const size_t n = (size_t)w * h;
    uint32_t mn = 0xffffffffu, mx = 0u;
    double sum = 0.0, sq = 0.0;
    unsigned long long zero_at = ~0ull;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t)gridDim.x * blockDim.x) {
        const size_t r = idx / w, c = idx - r * w;
        const float v = __ldg(data + r * pitch + c);
        if (v == v) { const uint32_t k = vf_float_key(v); mn = min(mn, k); mx = max(mx, k); }
        if (v == 0.0f && zero_at == ~0ull) zero_at = base + idx;          // a thread's indices ascend
        const double d = (double)v - (double)pivot;
        sum += d; sq += d * d;
    }
    if (zero_at != ~0ull) atomicMin(first, zero_at);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, d));
        mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, d));
        sum += __shfl_xor_sync(0xffffffffu, sum, d);
        sq += __shfl_xor_sync(0xffffffffu, sq, d);
    }
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    if (lane == 0) { s_mn[wid] = mn; s_mx[wid] = mx; s_sum[wid] = sum; s_sq[wid] = sq; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < 8; ++k) { mn = min(mn, s_mn[k]); mx = max(mx, s_mx[k]); sum += s_sum[k]; sq += s_sq[k]; }
        atomicMin(&acc[0], mn); atomicMax(&acc[1], mx); atomicAdd(&accd[0], sum); atomicAdd(&accd[1], sq);
    }
}

__global__ void k_dem_normalize(float* __restrict__ d, uint32_t w, uint32_t h, uint32_t pitch, int mode, float a, float scale, float lo) {
    const size_t n = (size_t)w * h;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t)gridDim.x * blockDim.x) {
        const size_t r = idx / w, c = idx - r * w;
        float* p = d + r * pitch + c;
        const float v = *p;
        // src/lib.rs:941 `(*v - min) * scale + lo` and :947 `(*v - mean) / denom` (no FMA in Rust)
        *p = mode == 0 ? __fadd_rn(__fmul_rn(__fsub_rn(v, a), scale), lo) : __fdiv_rn(__fsub_rn(v, a), scale);
    }
}
// f64 -> f32 cast and scale (src/lib.rs:384), f32 scale in place (src/lib.rs:362): rows of a pitched image
__global__ void k_dem_ingest_f64(const double* __restrict__ s, size_t n, uint32_t w, uint32_t pitch, size_t first, float scale, float* __restrict__ d) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const size_t idx = first + i, r = idx / w, c = idx - r * w;
        d[r * pitch + c] = __fmul_rn((float)s[i], scale);
    }
}
__global__ void k_dem_scale(float* __restrict__ d, uint32_t w, uint32_t h, uint32_t pitch, float scale) {
    const size_t n = (size_t)w * h;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t)gridDim.x * blockDim.x) {
        const size_t r = idx / w, c = idx - r * w;
        d[r * pitch + c] = __fmul_rn(d[r * pitch + c], scale);
    }
}

static inline int grid_for(size_t n, int threads) {
    const size_t b = (n + threads - 1) / threads;
    return (int)(b < 1 ? 1 : (b > 148 * 16 ? 148 * 16 : b));
}
static inline float key_to_float(uint32_t k) {
    const uint32_t b = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
    float f;
    memcpy(&f, &b, 4);
    return f;
}

// accumulator block at the head of the scratch allocation (256 bytes): acc[4] | accd[2] | first | out2
struct Moments { uint32_t acc[4]; double accd[2]; unsigned long long first; float out2[2]; };

static int moments_begin(Moments* d_m, cudaStream_t st) {
    Moments init;
    memset(&init, 0, sizeof init);
    init.acc[0] = 0xffffffffu; init.first = ~0ull;
    VF_CUDA(cudaMemcpyAsync(d_m, &init, sizeof init, cudaMemcpyHostToDevice, st));   // (pageable source: staged before the call returns)
    return VF_OK;
}
// min, max, mean, std from the accumulators; zero_value = the first zero of the data (sign included)
static void moments_finish(const Moments& m, size_t n, float pivot, float zero_value, float out4[4]) {
    out4[0] = key_to_float(m.acc[0]); out4[1] = key_to_float(m.acc[1]);
    if (out4[0] == 0.0f) out4[0] = zero_value;
    if (out4[1] == 0.0f) out4[1] = zero_value;
    const double md = m.accd[0] / (double)n;
    double var = m.accd[1] / (double)n - md * md;
    if (var < 0.0) var = 0.0;
    out4[2] = (float)((double)pivot + md);
    out4[3] = (float)sqrt(var);
}

// statistics of a device-resident image -> (min, max, mean, std)
static int stats_device(const float* d_data, uint32_t w, uint32_t h, uint32_t pitch, float out4[4], cudaStream_t st) {
    const size_t n = (size_t)w * h;
    Moments* d_m = (Moments*)vf_scratch(256);
    if (!d_m) return VF_E_OOM;
    float pivot = 0.0f;
    VF_CUDA(cudaMemcpyAsync(&pivot, d_data, 4, cudaMemcpyDeviceToHost, st));
    int rc = moments_begin(d_m, st);
    if (rc) return rc;
    VF_CUDA(cudaStreamSynchronize(st));
    if (!isfinite(pivot)) pivot = 0.0f;
    k_dem_moments<<<grid_for(n, 256), 256, 0, st>>>(d_data, w, h, pitch, 0ull, pivot, d_m->acc, d_m->accd, &d_m->first);
    Moments m;
    VF_CUDA(cudaMemcpyAsync(&m, d_m, sizeof m, cudaMemcpyDeviceToHost, st));
    VF_CUDA(cudaStreamSynchronize(st));
    VF_CUDA(cudaGetLastError());
    float z = 0.0f;
    if (m.first != ~0ull)
        VF_CUDA(cudaMemcpy(&z, d_data + (size_t)(m.first / w) * pitch + (size_t)(m.first % w), 4, cudaMemcpyDeviceToHost));
    moments_finish(m, n, pivot, z, out4);
    return VF_OK;
}

// exact order statistics of m gathered values already on the device (vals) -> out2 (host)
static int select_device(const float* d_vals, uint32_t m, float* d_out2, float out[2], cudaStream_t st) {
    const uint32_t r1 = (uint32_t)((float)m * 0.01f), r99 = (uint32_t)((float)m * 0.99f);      // terrain_stats.rs:32-33
    k_dem_select<<<1, 1024, 0, st>>>(d_vals, m, r1, r99, d_out2);
    VF_CUDA(cudaMemcpyAsync(out, d_out2, 8, cudaMemcpyDeviceToHost, st));
    VF_CUDA(cudaStreamSynchronize(st));
    VF_CUDA(cudaGetLastError());
    return VF_OK;
}
static void sample_plan(size_t n, size_t& step, size_t& m) {
    const size_t SAMPLE = 65536;
    step = 1; m = n;
    if (n > SAMPLE) { step = n / SAMPLE; m = (n + step - 1) / step; }            // terrain_stats.rs:23-27
}

// (lo, hi) of a device-resident image.  clamp=1: percentile semantics of terrain_stats.rs:20-34.
int vf_dem_minmax_device(const float* d_data, uint32_t w, uint32_t h, uint32_t pitch, int clamp,
                         float* lo, float* hi, cudaStream_t stream) {
    const size_t n = (size_t)w * h;
    if (n == 0) return vf_set_error(VF_E_INVALID, "heightmap slice empty");
    std::lock_guard<std::mutex> lk(vf_util_mutex());
    if (!clamp) {
        float st4[4];
        const int rc = stats_device(d_data, w, h, pitch, st4, stream);
        if (rc) return rc;
        *lo = st4[0]; *hi = st4[1];
        return VF_OK;
    }
    size_t step, m;
    sample_plan(n, step, m);
    char* sc = (char*)vf_scratch(256 + m * 4);
    if (!sc) return VF_E_OOM;
    float* vals = (float*)(sc + 256);
    float out[2];
    k_dem_gather<<<grid_for(m, 256), 256, 0, stream>>>(d_data, w, pitch, step, (uint32_t)m, vals);
    const int rc = select_device(vals, (uint32_t)m, (float*)sc, out, stream);
    if (rc) return rc;
    *lo = out[0]; *hi = out[1];
    return VF_OK;
}

// ---- host-array entry points ---------------------------------------------------------------------------------
// A flat host array is streamed through the scratch allocation in chunks (64 MiB): nothing is allocated per call,
// and an array larger than the device's free memory still works.
static const size_t kChunkFloats = 16u << 20;
static int vf_dem_stats_locked(const float* data, size_t n, float out4[4], cudaStream_t st);

// Helper threads for the strided host gather of vf_dem_minmax(clamp = 1): parked on a condition variable between calls
// (creating threads per call cost more than the gather they shared), started on first use, joined at unload.
namespace {
class VfGatherPool {
public:
    explicit VfGatherPool(int helpers) {
        for (int t = 0; t < helpers; ++t) th_.emplace_back([this, t] { loop(t); });
    }
    ~VfGatherPool() {
        { std::lock_guard<std::mutex> lk(mu_); stop_ = true; ++gen_; }
        go_.notify_all();
        for (auto& t : th_) t.join();
    }
    int parts() const { return (int)th_.size() + 1; }
    // f(part) for part = 0 .. parts() - 1, the last one on the calling thread; returns when all are done
    void run(const std::function<void(int)>& f) {
        { std::lock_guard<std::mutex> lk(mu_); job_ = &f; pending_ = (int)th_.size(); ++gen_; }
        go_.notify_all();
        f((int)th_.size());
        std::unique_lock<std::mutex> lk(mu_);
        done_.wait(lk, [this] { return pending_ == 0; });
        job_ = nullptr;
    }
private:
    void loop(int t) {
        uint64_t seen = 0;
        for (;;) {
            const std::function<void(int)>* f;
            {
                std::unique_lock<std::mutex> lk(mu_);
                go_.wait(lk, [&] { return gen_ != seen; });
                seen = gen_;
                if (stop_) return;
                f = job_;
            }
            (*f)(t);
            { std::lock_guard<std::mutex> lk(mu_); if (--pending_ == 0) done_.notify_one(); }
        }
    }
    std::vector<std::thread> th_;
    std::mutex mu_;
    std::condition_variable go_, done_;
    const std::function<void(int)>* job_ = nullptr;
    uint64_t gen_ = 0;
    int pending_ = 0;
    bool stop_ = false;
};
VfGatherPool& vf_gather_pool() {
    static VfGatherPool pool([] {
        const char* e = getenv("VF_GATHER_THREADS");              // tuning aid: total threads of the gather, caller included
        const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
        const unsigned want = e ? (unsigned)std::max(1, atoi(e)) : 8u;
        return (int)(std::min(want, hw) - 1u);
    }());
    return pool;
}
}   // namespace

extern "C" int vf_dem_minmax(const float* data, size_t n, int clamp, float* lo, float* hi) {
    if (!data || !lo || !hi) return vf_set_error(VF_E_INVALID, "null argument");
    if (n == 0) return vf_set_error(VF_E_INVALID, "heightmap slice empty");
    int rc = vf_require_device();
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(vf_util_mutex());
    cudaStream_t st = vf_util_stream();
    if (!st) return VF_E_CUDA;
    if (clamp) {
        // terrain_stats.rs:23-27: the statistic is defined on the strided sample data[0], data[step], ...: those
        // <= 65 537 values are staged in pinned memory (a memory-bound gather, no arithmetic) and 256 KB cross the bus
        // instead of the whole array; the order statistics themselves are selected on the device.
        size_t step, m;
        sample_plan(n, step, m);
        float* stage = (float*)vf_pinned_scratch(m * 4);
        char* sc = (char*)vf_scratch(256 + m * 4);
        if (!stage || !sc) return VF_E_OOM;
        // (every sample is a cache and, at large strides, a TLB miss: software prefetch keeps ~32 of them in flight per
        // thread, up to eight parked threads share the work on large arrays)
        auto gather = [&](size_t k0, size_t k1) {
            for (size_t k = k0; k < k1; ++k) {
                if (k + 32 < k1) __builtin_prefetch(data + (k + 32) * step, 0, 0);
                stage[k] = data[k * step];
            }
        };
        if (step >= 64 && m >= 16384) {
            VfGatherPool& pool = vf_gather_pool();
            const size_t parts = (size_t)pool.parts();
            pool.run([&](int part) { gather(m * (size_t)part / parts, m * ((size_t)part + 1) / parts); });
        } else {
            gather(0, m);
        }
        float* vals = (float*)(sc + 256);
        VF_CUDA(cudaMemcpyAsync(vals, stage, m * 4, cudaMemcpyHostToDevice, st));
        float out[2];
        rc = select_device(vals, (uint32_t)m, (float*)sc, out, st);
        if (rc) return rc;
        *lo = out[0]; *hi = out[1];
        return VF_OK;
    }
    float st4[4];
    rc = vf_dem_stats_locked(data, n, st4, st);
    if (rc) return rc;
    *lo = st4[0]; *hi = st4[1];
    return VF_OK;
}

// min / max / mean / std of a flat host array, chunked (the util mutex is held by the caller)
static int vf_dem_stats_locked(const float* data, size_t n, float out4[4], cudaStream_t st) {
    const size_t chunk = n < kChunkFloats ? n : kChunkFloats;
    char* sc = (char*)vf_scratch(256 + chunk * 4);
    if (!sc) return VF_E_OOM;
    Moments* d_m = (Moments*)sc;
    float* buf = (float*)(sc + 256);
    int rc = moments_begin(d_m, st);
    if (rc) return rc;
    const float pivot = isfinite(data[0]) ? data[0] : 0.0f;
    for (size_t off = 0; off < n; off += chunk) {
        const size_t c = n - off < chunk ? n - off : chunk;
        VF_CUDA(cudaMemcpyAsync(buf, data + off, c * 4, cudaMemcpyHostToDevice, st));      // stream order protects `buf`
        k_dem_moments<<<grid_for(c, 256), 256, 0, st>>>(buf, (uint32_t)c, 1u, (uint32_t)c, (unsigned long long)off, pivot,
                                                        d_m->acc, d_m->accd, &d_m->first);
    }
    Moments m;
    VF_CUDA(cudaMemcpyAsync(&m, d_m, sizeof m, cudaMemcpyDeviceToHost, st));
    VF_CUDA(cudaStreamSynchronize(st));
    VF_CUDA(cudaGetLastError());
    moments_finish(m, n, pivot, m.first != ~0ull ? data[m.first] : 0.0f, out4);
    return VF_OK;
}

extern "C" int vf_dem_stats(const float* data, size_t n, float out4[4]) {
    if (!data || !out4) return vf_set_error(VF_E_INVALID, "null argument");
    if (n == 0) { out4[0] = out4[1] = out4[2] = out4[3] = 0.0f; return VF_OK; }   // src/lib.rs:906-908
    int rc = vf_require_device();
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(vf_util_mutex());
    cudaStream_t st = vf_util_stream();
    if (!st) return VF_E_CUDA;
    return vf_dem_stats_locked(data, n, out4, st);
}

// normalisation parameters of src/lib.rs:936-949 from (min, max, mean, std)
static void normalize_params(const float st[4], int mode, float eps, float lo, float hi, float& a, float& scale) {
    if (mode == 0) {
        float denom = fabsf(st[1] - st[0]);
        if (!(denom > eps)) denom = eps;
        scale = (hi - lo) / denom;
        a = st[0];
    } else {
        scale = st[3] > eps ? st[3] : eps;
        a = st[2];
    }
}

extern "C" int vf_dem_normalize(float* data, size_t n, int mode, float eps, float lo, float hi) {
    if (!data) return vf_set_error(VF_E_INVALID, "null argument");
    if (mode != 0 && mode != 1) return vf_set_error(VF_E_INVALID, "mode must be 'minmax' or 'zscore'");
    if (n == 0) return VF_OK;
    int rc = vf_require_device();
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(vf_util_mutex());
    cudaStream_t st = vf_util_stream();
    if (!st) return VF_E_CUDA;
    float s4[4];
    rc = vf_dem_stats_locked(data, n, s4, st);
    if (rc) return rc;
    float a, scale;
    normalize_params(s4, mode, eps, lo, hi, a, scale);
    const size_t chunk = n < kChunkFloats ? n : kChunkFloats;
    char* sc = (char*)vf_scratch(256 + chunk * 4);
    if (!sc) return VF_E_OOM;
    float* buf = (float*)(sc + 256);
    for (size_t off = 0; off < n; off += chunk) {
        const size_t c = n - off < chunk ? n - off : chunk;
        VF_CUDA(cudaMemcpyAsync(buf, data + off, c * 4, cudaMemcpyHostToDevice, st));
        k_dem_normalize<<<grid_for(c, 256), 256, 0, st>>>(buf, (uint32_t)c, 1u, (uint32_t)c, mode, a, scale, lo);
        VF_CUDA(cudaMemcpyAsync(data + off, buf, c * 4, cudaMemcpyDeviceToHost, st));
    }
    VF_CUDA(cudaStreamSynchronize(st));
    VF_CUDA(cudaGetLastError());
    return VF_OK;
}

// ---- device-resident height images (vf_height): add_terrain without bouncing through the host ---------------------
static int height_alloc(uint32_t w, uint32_t h, vf_height** out) {
    vf_height* t = new vf_height();
    t->w = w; t->h = h;
    t->pitch = (((size_t)w * 4) + 15) & ~(size_t)15;
    if (cudaGetDevice(&t->device) != cudaSuccess) t->device = 0;
    const cudaError_t e = cudaMalloc(&t->d, t->pitch * h);
    if (e != cudaSuccess) {
        delete t;
        (void)cudaGetLastError();
        return vf_set_error(e == cudaErrorMemoryAllocation ? VF_E_OOM : VF_E_CUDA, "height image allocation: %s", cudaGetErrorString(e));
    }
    *out = t;
    return VF_OK;
}

extern "C" int vf_height_ingest_f32(const float* host, uint32_t w, uint32_t h, float scale, vf_height** out) {
    if (!host || !out) return vf_set_error(VF_E_INVALID, "null argument");
    if (w == 0 || h == 0) return vf_set_error(VF_E_INVALID, "heightmap cannot be empty");
    int rc = vf_require_device();
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(vf_util_mutex());
    cudaStream_t st = vf_util_stream();
    if (!st) return VF_E_CUDA;
    vf_height* t = nullptr;
    rc = height_alloc(w, h, &t);
    if (rc) return rc;
    cudaError_t e = cudaMemcpy2DAsync(t->d, t->pitch, host, (size_t)w * 4, (size_t)w * 4, h, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess && scale != 1.0f) {
        k_dem_scale<<<grid_for((size_t)w * h, 256), 256, 0, st>>>(t->d, w, h, (uint32_t)(t->pitch / 4), scale);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { vf_height_destroy(t); return vf_set_error(VF_E_CUDA, "height ingest: %s", cudaGetErrorString(e)); }
    *out = t;
    return VF_OK;
}

extern "C" int vf_height_ingest_f64(const double* host, uint32_t w, uint32_t h, float scale, vf_height** out) {
    if (!host || !out) return vf_set_error(VF_E_INVALID, "null argument");
    if (w == 0 || h == 0) return vf_set_error(VF_E_INVALID, "heightmap cannot be empty");
    int rc = vf_require_device();
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(vf_util_mutex());
    cudaStream_t st = vf_util_stream();
    if (!st) return VF_E_CUDA;
    vf_height* t = nullptr;
    rc = height_alloc(w, h, &t);
    if (rc) return rc;
    const size_t n = (size_t)w * h, chunk = n < kChunkFloats / 2 ? n : kChunkFloats / 2;      // doubles: 64 MiB chunks
    double* buf = (double*)vf_scratch(chunk * 8);
    if (!buf) { vf_height_destroy(t); return VF_E_OOM; }
    cudaError_t e = cudaSuccess;
    for (size_t off = 0; off < n && e == cudaSuccess; off += chunk) {
        const size_t c = n - off < chunk ? n - off : chunk;
        e = cudaMemcpyAsync(buf, host + off, c * 8, cudaMemcpyHostToDevice, st);
        if (e == cudaSuccess) {
            k_dem_ingest_f64<<<grid_for(c, 256), 256, 0, st>>>(buf, c, w, (uint32_t)(t->pitch / 4), off, scale, t->d);
            e = cudaGetLastError();
        }
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { vf_height_destroy(t); return vf_set_error(VF_E_CUDA, "height ingest: %s", cudaGetErrorString(e)); }
    *out = t;
    return VF_OK;
}

extern "C" int vf_height_clone(const vf_height* src, vf_height** out) {
    if (!src || !out) return vf_set_error(VF_E_INVALID, "null argument");
    VF_CUDA(cudaSetDevice(src->device));
    vf_height* t = nullptr;
    int rc = height_alloc(src->w, src->h, &t);
    if (rc) return rc;
    const cudaError_t e = cudaMemcpy(t->d, src->d, src->pitch * src->h, cudaMemcpyDeviceToDevice);
    if (e != cudaSuccess) { vf_height_destroy(t); return vf_set_error(VF_E_CUDA, "height clone: %s", cudaGetErrorString(e)); }
    *out = t;
    return VF_OK;
}

extern "C" int vf_height_minmax(const vf_height* t, int clamp, float* lo, float* hi) {
    if (!t || !lo || !hi) return vf_set_error(VF_E_INVALID, "null argument");
    VF_CUDA(cudaSetDevice(t->device));
    cudaStream_t st = vf_util_stream();
    if (!st) return VF_E_CUDA;
    return vf_dem_minmax_device(t->d, t->w, t->h, (uint32_t)(t->pitch / 4), clamp, lo, hi, st);
}

extern "C" int vf_height_stats(const vf_height* t, float out4[4]) {
    if (!t || !out4) return vf_set_error(VF_E_INVALID, "null argument");
    VF_CUDA(cudaSetDevice(t->device));
    std::lock_guard<std::mutex> lk(vf_util_mutex());
    cudaStream_t st = vf_util_stream();
    if (!st) return VF_E_CUDA;
    return stats_device(t->d, t->w, t->h, (uint32_t)(t->pitch / 4), out4, st);
}

extern "C" int vf_height_normalize(vf_height* t, int mode, float eps, float lo, float hi) {
    if (!t) return vf_set_error(VF_E_INVALID, "null argument");
    if (mode != 0 && mode != 1) return vf_set_error(VF_E_INVALID, "mode must be 'minmax' or 'zscore'");
    VF_CUDA(cudaSetDevice(t->device));
    std::lock_guard<std::mutex> lk(vf_util_mutex());
    cudaStream_t st = vf_util_stream();
    if (!st) return VF_E_CUDA;
    float s4[4];
    const int rc = stats_device(t->d, t->w, t->h, (uint32_t)(t->pitch / 4), s4, st);
    if (rc) return rc;
    float a, scale;
    normalize_params(s4, mode, eps, lo, hi, a, scale);
    k_dem_normalize<<<grid_for((size_t)t->w * t->h, 256), 256, 0, st>>>(t->d, t->w, t->h, (uint32_t)(t->pitch / 4), mode, a, scale, lo);
    VF_CUDA(cudaGetLastError());
    VF_CUDA(cudaStreamSynchronize(st));
    return VF_OK;
}

// the one-shot host-to-host forms of the ingest (kept for hosts that want the scaled copy back)
extern "C" int vf_dem_ingest_f64(const double* src, size_t n, float scale, float* dst) {
    if (!src || !dst) return vf_set_error(VF_E_INVALID, "null argument");
    if (n == 0) return VF_OK;
    if (n > 0xffffffffull) return vf_set_error(VF_E_INVALID, "more than 2^32 samples");
    vf_height* t = nullptr;
    int rc = vf_height_ingest_f64(src, (uint32_t)n, 1u, scale, &t);
    if (rc) return rc;
    rc = vf_height_read_patch(t, 0, 0, (uint32_t)n, 1u, dst);
    vf_height_destroy(t);
    return rc;
}

extern "C" int vf_dem_ingest_f32(const float* src, size_t n, float scale, float* dst) {
    if (!src || !dst) return vf_set_error(VF_E_INVALID, "null argument");
    if (n == 0) return VF_OK;
    if (n > 0xffffffffull) return vf_set_error(VF_E_INVALID, "more than 2^32 samples");
    vf_height* t = nullptr;
    int rc = vf_height_ingest_f32(src, (uint32_t)n, 1u, scale, &t);
    if (rc) return rc;
    rc = vf_height_read_patch(t, 0, 0, (uint32_t)n, 1u, dst);
    vf_height_destroy(t);
    return rc;
}

// ---- allocation-free variant (no counterpart in the reference) ----------
// vf_dem_minmax for callers that must not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc): input and output are DEVICE pointers, the work is enqueued on the caller's stream, and the scratch the
// kernels need -- the accumulator block of k_dem_moments and the gathered sample of k_dem_gather / k_dem_select -- lives
// in ONE file-scope device array of fixed size: 256 B + 131 072 floats = 512.25 KiB, held by every GPU that loads the
// library until the process exits.  The sample of terrain_stats.rs:23-27 never exceeds 131 071 values (n in (65536,
// 131072) has step 1), so the array bounds the sample for every n the gather can index (n <= 2^32 - 1).
// Every call shares the array and finds the previous call's contents in it: the accumulators are re-initialised by the
// first kernel of each call, the sample is overwritten up to the m values the call reads.  Calls that overlap on the GPU
// (two streams, two host threads) race on the array: serialise them.
#define VF_NOALLOC_MAX_SAMPLES 131072u
#define VF_NOALLOC_MAX_N 0xffffffffull
struct VfNoAllocScratch {
    Moments m;                                            // 48 bytes
    unsigned char pad[256 - sizeof(Moments)];
    float vals[VF_NOALLOC_MAX_SAMPLES];
};
static __device__ __align__(256) VfNoAllocScratch g_noalloc_scratch;

// In host code `g_noalloc_scratch` is only a stand-in for the symbol, not a device address: the address is asked for once
// per device, with that device current, and cached.
static VfNoAllocScratch* g_noalloc_ptr[64] = {nullptr};
static std::mutex g_noalloc_mu;
static int noalloc_scratch(VfNoAllocScratch** out) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) { (void)cudaGetLastError(); return vf_set_error(VF_E_NO_DEVICE, "No suitable GPU adapter (no current CUDA device)"); }
    if (dev >= 64) return vf_set_error(VF_E_INVALID, "device %d: at most 64 devices are supported", dev);
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) { (void)cudaGetLastError(); return vf_set_error(VF_E_CUDA, "cudaGetSymbolAddress(scratch) on device %d: %s", dev, cudaGetErrorString(e)); }
        g_noalloc_ptr[dev] = static_cast<VfNoAllocScratch*>(p);
    }
    *out = g_noalloc_ptr[dev];
    return VF_OK;
}

__global__ void k_noalloc_begin(Moments* m) {            // what must not carry over from the previous call
    m->acc[0] = 0xffffffffu; m->acc[1] = 0u; m->acc[2] = 0u; m->acc[3] = 0u;
    m->accd[0] = 0.0; m->accd[1] = 0.0;
    m->first = ~0ull;
    m->out2[0] = 0.0f; m->out2[1] = 0.0f;
}
// (lo, hi) from the accumulators; a zero extreme takes the sign of the first zero of the array (as fix-up of stats_device)
__global__ void k_noalloc_finish(const float* __restrict__ data, const Moments* __restrict__ m, float* __restrict__ out2) {
    float lo = vf_key_float(m->acc[0]), hi = vf_key_float(m->acc[1]);
    if (m->first != ~0ull) {
        const float z = data[m->first];
        if (lo == 0.0f) lo = z;
        if (hi == 0.0f) hi = z;
    }
    out2[0] = lo; out2[1] = hi;
}

extern "C" int vf_dem_minmax_noalloc_async(const float* d_data, size_t n, int clamp, float* d_out2, void* cuda_stream) {
    if (!d_data || !d_out2) return vf_set_error(VF_E_INVALID, "null argument");
    if (n == 0) return vf_set_error(VF_E_INVALID, "heightmap slice empty");
    if ((unsigned long long)n > VF_NOALLOC_MAX_N)
        return vf_set_error(VF_E_INVALID, "vf_dem_minmax_noalloc_async: n = %zu exceeds the maximum of %llu values", n, VF_NOALLOC_MAX_N);
    VfNoAllocScratch* sc = nullptr;
    const int rc = noalloc_scratch(&sc);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    if (!clamp) {
        k_noalloc_begin<<<1, 1, 0, st>>>(&sc->m);
        k_dem_moments<<<grid_for(n, 256), 256, 0, st>>>(d_data, (uint32_t)n, 1u, (uint32_t)n, 0ull, 0.0f, sc->m.acc, sc->m.accd, &sc->m.first);
        k_noalloc_finish<<<1, 1, 0, st>>>(d_data, &sc->m, d_out2);
    } else {
        size_t step, m;
        sample_plan(n, step, m);
        if (m > VF_NOALLOC_MAX_SAMPLES)
            return vf_set_error(VF_E_INVALID, "vf_dem_minmax_noalloc_async: sample of %zu values exceeds the scratch array (%u)", m, VF_NOALLOC_MAX_SAMPLES);
        const uint32_t r1 = (uint32_t)((float)m * 0.01f), r99 = (uint32_t)((float)m * 0.99f);      // terrain_stats.rs:32-33
        k_dem_gather<<<grid_for(m, 256), 256, 0, st>>>(d_data, (uint32_t)n, (uint32_t)n, step, (uint32_t)m, sc->vals);
        k_dem_select<<<1, 1024, 0, st>>>(sc->vals, (uint32_t)m, r1, r99, d_out2);
    }
    VF_CUDA(cudaGetLastError());
    return VF_OK;
}
