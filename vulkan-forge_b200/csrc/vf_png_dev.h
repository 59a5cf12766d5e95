// vf_png_dev.h -- device-side PNG encoder (vf_png_dev.cu), used by vf_scene for render_png / render_batch_png
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

struct VfPngDev {
    uint32_t W = 0, H = 0, rows_per_band = 0, nbands = 0;
    uint32_t raw_cap = 0, band_cap = 0, file_cap = 0;     // bytes: filtered band, encoded band, whole file (worst case)
    int slots = 0;
    bool small_bands = false;                              // bands of <= 20 KiB: the <20, 4> instance of k_png_encode
    size_t smem_encode = 0, smem_pack = 0;
    uint8_t* d_bands = nullptr;                            // nbands x band_cap
    uint32_t* d_band_size = nullptr; uint32_t* d_band_adler = nullptr;
    uint8_t* d_files = nullptr;                            // slots x file_cap: finished .png files
    uint32_t* d_file_size = nullptr;                       // per slot (0xffffffff: did not fit, cannot happen with file_cap): the device's view of
    uint32_t* h_file_size = nullptr;                       // ... mapped pinned host memory: k_png_pack stores the size where the host reads it (no copy on the stream)
    uint32_t* d_tables = nullptr;                          // IHDR preamble | x^(8r) | x^(8 * 64 j) mod the CRC polynomial
};

size_t vf_png_dev_bound(int w, int h);                     // upper bound of the file size
int vf_png_dev_create(int w, int h, int slots, VfPngDev** out);
void vf_png_dev_destroy(VfPngDev* e);
size_t vf_png_dev_bytes(const VfPngDev* e);                // device memory held
// encode the RGBA8 frame at d_frame into file slot `slot` (two kernels on `st`)
int vf_png_dev_encode(VfPngDev* e, const uint8_t* d_frame, int slot, cudaStream_t st);
