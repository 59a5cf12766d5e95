// vf_png_kernels.cuh -- the two kernels of the device-side PNG encoder (host side: vf_png_dev.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define VF_PNG_THREADS 256
#define VF_PNG_PIECE 64u                      // CRC piece length in bytes
#define VF_PNG_POLY 0xEDB88320u

__device__ __forceinline__ uint32_t vf_rev(uint32_t code, uint32_t nbits) { return __brev(code) >> (32u - nbits); }

// fixed Huffman code of a literal byte: (bits, nbits), already reversed for the LSB-first stream (RFC 1951 3.2.6)
__device__ __forceinline__ uint32_t vf_lit_code(uint32_t v, uint32_t& nbits) {
    if (v < 144u) { nbits = 8u; return vf_rev(0x30u + v, 8u); }
    nbits = 9u;
    return vf_rev(0x190u + (v - 144u), 9u);
}
// match of `len` (3..258) at distance 4: length code + extra bits + distance code 3 (5 bits).  RFC 1951 3.2.5: lengths
// 3..10 are symbols 257..264; from 11 on four symbols share e = 1..5 extra bits, base length 3 + (4 + j) 2^e; 258 is 285.
__device__ __forceinline__ uint32_t vf_match_code(uint32_t len, uint32_t& nbits) {
    uint32_t sym, eb = 0u, ev = 0u;
    if (len == 258u) sym = 285u;
    else if (len <= 10u) sym = 254u + len;
    else {
        const uint32_t x = len - 3u;
        eb = 29u - (uint32_t)__clz((int)x);                     // floor(log2 x) - 2
        sym = 265u + 4u * (eb - 1u) + ((x >> eb) - 4u);
        ev = x & ((1u << eb) - 1u);
    }
    uint32_t bits, n;
    if (sym <= 279u) { n = 7u; bits = vf_rev(sym - 256u, 7u); }
    else { n = 8u; bits = vf_rev(0xC0u + (sym - 280u), 8u); }
    bits |= ev << n; n += eb;
    bits |= vf_rev(3u, 5u) << n; n += 5u;                       // distance 4 -> code 3, no extra bits
    nbits = n;
    return bits;
}

// a * b mod P in the reflected representation (x^0 is bit 31), as zlib's multmodp
__device__ __forceinline__ uint32_t vf_gf_mul(uint32_t a, uint32_t b) {
    uint32_t p = 0u;
    for (uint32_t m = 1u << 31; m; m >>= 1) {
        if (a & m) p ^= b;
        b = (b & 1u) ? (b >> 1) ^ VF_PNG_POLY : b >> 1;
    }
    return p;
}

#include "vf_png_encode.cuh"
#include "vf_png_pack.cuh"
