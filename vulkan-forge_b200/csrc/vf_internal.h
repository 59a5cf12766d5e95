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

#define VF_CUDA(call)                                                                     \
    do {                                                                                  \
        cudaError_t e_ = (call);                                                          \
        if (e_ != cudaSuccess)                                                            \
            return vf_set_error(e_ == cudaErrorMemoryAllocation ? VF_E_OOM : VF_E_CUDA,   \
                                "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_),   \
                                __FILE__, __LINE__);                                      \
    } while (0)
