// vf_context.cu -- device selection, error reporting, shared colour tables, pinned memory.
// replaces: WgpuContext (src/lib.rs:23-61) and the per-object device setup of
// src/terrain/mod.rs:277-294 / src/scene/mod.rs:63-70.
#include <math.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "vf_internal.h"

static thread_local char g_err[512] = "";

int vf_set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

extern "C" const char* vf_last_error(void) { return g_err; }

static std::mutex g_mu;
static int g_device = -1;                 // device selected by the last vf_init
static const int kMaxDevices = 64;
static float* g_thr_dev[kMaxDevices] = {nullptr};   // the sRGB threshold table, one copy per device that has been used
static float g_thr_host[256];
static bool g_thr_built = false;

// sRGB OETF + round-half-up to 8 bits, in double (same formula as the store the reference
// gets from the Rgba8UnormSrgb target, src/terrain/mod.rs:219).
static int srgb_code(float x) {
    double c = (double)x;
    if (!(c > 0.0)) c = 0.0;
    if (c > 1.0) c = 1.0;
    const double e = c <= 0.0031308 ? 12.92 * c : 1.055 * pow(c, 1.0 / 2.4) - 0.055;
    int q = (int)floor(e * 255.0 + 0.5);
    return q < 0 ? 0 : (q > 255 ? 255 : q);
}

// thr[k] = smallest non-negative float whose code is >= k (bisection over the float bit
// patterns, which are monotone for non-negative floats).
static void build_thresholds() {
    if (g_thr_built) return;
    g_thr_host[0] = 0.0f;
    for (int k = 1; k < 256; ++k) {
        uint32_t lo = 0u, hi = 0x3f800000u;            // code(0)=0 < k <= code(1)=255
        while (hi - lo > 1u) {
            const uint32_t mid = lo + (hi - lo) / 2u;
            float fm;
            memcpy(&fm, &mid, 4);
            if (srgb_code(fm) >= k) hi = mid; else lo = mid;
        }
        memcpy(&g_thr_host[k], &hi, 4);
    }
    g_thr_built = true;
}

const float* vf_srgb_thresholds_host() {
    std::lock_guard<std::mutex> lk(g_mu);
    build_thresholds();
    return g_thr_host;
}

extern "C" int vf_device_count(int* count) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        *count = 0;
        (void)cudaGetLastError();
        return vf_set_error(VF_E_NO_DEVICE, "No suitable GPU adapter (CUDA: %s)", cudaGetErrorString(e));
    }
    *count = n;
    return VF_OK;
}

extern "C" int vf_init(int device) {
    std::lock_guard<std::mutex> lk(g_mu);
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        (void)cudaGetLastError();
        return vf_set_error(VF_E_NO_DEVICE, "No suitable GPU adapter (CUDA: %s)",
                            e != cudaSuccess ? cudaGetErrorString(e) : "no devices");
    }
    if (device < 0) {
        if (g_device >= 0) device = g_device;
        else {
            // one process per GPU: honour torchrun's LOCAL_RANK when present
            const char* lr = getenv("LOCAL_RANK");
            device = lr ? atoi(lr) % n : 0;
        }
    }
    if (device >= n) return vf_set_error(VF_E_INVALID, "device %d out of range (%d devices)", device, n);
    if (device >= kMaxDevices) return vf_set_error(VF_E_INVALID, "device %d: at most %d devices are supported", device, kMaxDevices);
    VF_CUDA(cudaSetDevice(device));
    if (!g_thr_dev[device]) {                 // scenes of another device keep using their own copy
        build_thresholds();
        float* p = nullptr;
        VF_CUDA(cudaMalloc(&p, sizeof g_thr_host));
        VF_CUDA(cudaMemcpy(p, g_thr_host, sizeof g_thr_host, cudaMemcpyHostToDevice));
        g_thr_dev[device] = p;
    }
    g_device = device;
    return VF_OK;
}

int vf_require_device() {
    {
        std::lock_guard<std::mutex> lk(g_mu);
        if (g_device >= 0 && g_thr_dev[g_device]) {
            cudaError_t e = cudaSetDevice(g_device);
            if (e == cudaSuccess) return VF_OK;
        }
    }
    return vf_init(-1);
}

// table of the CURRENT device (nullptr if vf_init never ran for it)
const float* vf_srgb_thresholds_dev() {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return nullptr;
    return g_thr_dev[dev];
}
int vf_current_device() { return g_device; }

extern "C" int vf_device_info(int device, vf_device_info_t* out) {
    int n = 0;
    int rc = vf_device_count(&n);
    if (rc) return rc;
    if (device < 0 || device >= n) return vf_set_error(VF_E_INVALID, "device %d out of range", device);
    cudaDeviceProp p;
    VF_CUDA(cudaGetDeviceProperties(&p, device));
    memset(out, 0, sizeof *out);
    strncpy(out->name, p.name, sizeof out->name - 1);
    out->cc_major = p.major; out->cc_minor = p.minor; out->sm_count = p.multiProcessorCount;
    out->total_mem_bytes = (uint64_t)p.totalGlobalMem;
    out->pci_device_id = p.pciDeviceID; out->pci_bus_id = p.pciBusID;
    out->l2_bytes = p.l2CacheSize; out->max_smem_per_block = (int)p.sharedMemPerBlockOptin;
    return VF_OK;
}

extern "C" int vf_versions(int* driver, int* runtime) {
    int d = 0, r = 0;
    if (cudaDriverGetVersion(&d) != cudaSuccess) { (void)cudaGetLastError(); d = 0; }
    if (cudaRuntimeGetVersion(&r) != cudaSuccess) { (void)cudaGetLastError(); r = 0; }
    if (driver) *driver = d;
    if (runtime) *runtime = r;
    return VF_OK;
}

#define VF_STR2(x) #x
#define VF_STR(x) VF_STR2(x)
extern "C" const char* vf_build_info(void) {
    return "libvfcuda: nvcc " VF_STR(__CUDACC_VER_MAJOR__) "." VF_STR(__CUDACC_VER_MINOR__) "." VF_STR(__CUDACC_VER_BUILD__)
           ", -gencode arch=compute_100a,code=sm_100a -O3 --fmad=false -lineinfo, built " __DATE__;
}

extern "C" int vf_host_alloc(size_t bytes, void** out) {
    int rc = vf_require_device();
    if (rc) return rc;
    void* p = nullptr;
    cudaError_t e = cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        return vf_set_error(VF_E_OOM, "cudaHostAlloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
    }
    *out = p;
    return VF_OK;
}

extern "C" int vf_host_free(void* p) {
    if (!p) return VF_OK;
    VF_CUDA(cudaFreeHost(p));
    return VF_OK;
}

// ---- utility stream + scratch, per device ------------------------------------------------
static std::mutex g_util_mu;
static cudaStream_t g_util_stream[kMaxDevices] = {nullptr};
static void* g_scratch[kMaxDevices] = {nullptr};
static size_t g_scratch_bytes[kMaxDevices] = {0};
static void* g_pinned = nullptr;
static size_t g_pinned_bytes = 0;

std::mutex& vf_util_mutex() { return g_util_mu; }

static int current_device() {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) { (void)cudaGetLastError(); return -1; }
    return dev;
}

cudaStream_t vf_util_stream() {
    const int dev = current_device();
    if (dev < 0) { vf_set_error(VF_E_CUDA, "no current CUDA device"); return nullptr; }
    if (!g_util_stream[dev] && cudaStreamCreateWithFlags(&g_util_stream[dev], cudaStreamNonBlocking) != cudaSuccess) {
        vf_set_error(VF_E_CUDA, "utility stream: %s", cudaGetErrorString(cudaGetLastError()));
        g_util_stream[dev] = nullptr;
    }
    return g_util_stream[dev];
}

void* vf_scratch(size_t bytes) {
    const int dev = current_device();
    if (dev < 0) { vf_set_error(VF_E_CUDA, "no current CUDA device"); return nullptr; }
    if (bytes > g_scratch_bytes[dev]) {
        if (g_scratch[dev]) { cudaStreamSynchronize(g_util_stream[dev]); cudaFree(g_scratch[dev]); g_scratch[dev] = nullptr; g_scratch_bytes[dev] = 0; }
        const size_t want = (bytes + (1u << 20) - 1) & ~(size_t)((1u << 20) - 1);
        if (cudaMalloc(&g_scratch[dev], want) != cudaSuccess) {
            (void)cudaGetLastError();
            vf_set_error(VF_E_OOM, "scratch allocation of %zu bytes failed", want);
            return nullptr;
        }
        g_scratch_bytes[dev] = want;
    }
    return g_scratch[dev];
}

void* vf_pinned_scratch(size_t bytes) {
    if (bytes > g_pinned_bytes) {
        if (g_pinned) { cudaFreeHost(g_pinned); g_pinned = nullptr; g_pinned_bytes = 0; }
        if (cudaHostAlloc(&g_pinned, bytes, cudaHostAllocDefault) != cudaSuccess) {
            (void)cudaGetLastError();
            vf_set_error(VF_E_OOM, "pinned staging allocation of %zu bytes failed", bytes);
            return nullptr;
        }
        g_pinned_bytes = bytes;
    }
    return g_pinned;
}

// ---- palettes (bytes generated by tools/gen_lut_assets.py) -----------------------------
static const uint8_t g_palettes[3][1024] = {
#include "vf_lut_data.inc"
};

const uint8_t* vf_palette_bytes(int lut_id) {
    return (lut_id >= 0 && lut_id < 3) ? g_palettes[lut_id] : nullptr;
}

extern "C" int vf_colormap_rgba8(int lut_id, uint8_t out[1024]) {
    const uint8_t* p = vf_palette_bytes(lut_id);
    if (!p) return vf_set_error(VF_E_INVALID, "unknown colormap id %d", lut_id);
    memcpy(out, p, 1024);
    return VF_OK;
}
