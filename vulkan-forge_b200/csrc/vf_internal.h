// vf_internal.h -- host-side internals shared by the translation units of libvfcuda.so
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

#include "../../include/vf_cuda.h"

int vf_set_error(int code, const char* fmt, ...);
int vf_require_device();                 // VF_OK or VF_E_NO_DEVICE (lazy vf_init(-1))
const float* vf_srgb_thresholds_dev();   // 256 floats on the current device (thr[0] = 0)
int vf_current_device();                 // device of the last vf_init (-1: none yet)
const float* vf_srgb_thresholds_host();
const uint8_t* vf_palette_bytes(int lut_id);   // 1024 bytes or nullptr

// a device-resident R32F height image (C ABI handle); pitch in BYTES, a multiple of 16
struct vf_height {
    float* d = nullptr;
    uint32_t w = 0, h = 0;
    size_t pitch = 0;
    int device = 0;
};
// Per-device helpers for the stateless entry points (triangle, grid, DEM statistics): one non-blocking utility stream
// and one cached scratch allocation that only grows, instead of a cudaMalloc / cudaFree pair and the legacy default
// stream per call.  Callers hold vf_util_lock() for the duration of the call (the entry points are short).
#include <mutex>
std::mutex& vf_util_mutex();
cudaStream_t vf_util_stream();                 // of the current device; nullptr on failure (vf_last_error set)
void* vf_scratch(size_t bytes);                // >= bytes on the current device, 256-byte aligned; nullptr on failure
void* vf_pinned_scratch(size_t bytes);         // small pinned host staging buffer (grows); nullptr on failure

#define VF_CUDA(call)                                                                     \
    do {                                                                                  \
        cudaError_t e_ = (call);                                                          \
        if (e_ != cudaSuccess)                                                            \
            return vf_set_error(e_ == cudaErrorMemoryAllocation ? VF_E_OOM : VF_E_CUDA,   \
                                "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_),   \
                                __FILE__, __LINE__);                                      \
    } while (0)
