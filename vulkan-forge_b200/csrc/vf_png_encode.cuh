// vf_png_encode.cuh -- k_png_encode: one CTA per band of rows -> one fixed-Huffman deflate block (included by
// vf_png_kernels.cuh).
// Dynamic shared memory: raw bytes (as words) | output words | per-word token table + per-word bit offsets.
// The byte stream is handled a WORD at a time: a match at distance 4 is "this word's byte equals the previous word's
// byte", one __vcmpeq4 per word.
//   analysis (thread = a contiguous segment of words): match bits of the segment in registers; runs shorter than 3 are
//            dropped with shifts; the runs that remain (few) are cut into matches of 258 from their start, found with bit
//            scans; the result is 16 bits per word -- which of its bytes are literals, and the match that starts in it;
//   bit counts follow from that table with popcounts (a literal is 8 bits, 9 from 144 up), one block scan gives every
//            word its place in the stream;
//   emission (thread = every 256th word, so that a warp shares its literal-heavy stretches evenly): the word's <= 5
//            tokens are assembled in a register and ORed into the stream.
// A band ends on a byte boundary (empty stored block, the Z_SYNC_FLUSH idiom), so bands concatenate as bytes.
#pragma once

// SEG_WORDS: most words of a band one thread analyses (a band holds at most 256 * SEG_WORDS words).  Two instances:
// <20, 4> for bands up to 20 KiB (the default band is two rows of a 1080p frame: 64 registers, 50 KB of shared memory,
// 4 CTAs per SM) and <40, 2> for frames whose single row is longer than that (up to 8192 pixels).
// token table entry: bits 0..3 literal flags of the word's bytes | 4: a match starts here | 5..6: at which byte | 7..15: its length
template <int SEG_WORDS, int MIN_CTAS>
__global__ void __launch_bounds__(VF_PNG_THREADS, MIN_CTAS)
k_png_encode(const uint8_t* __restrict__ frame, uint32_t W, uint32_t H, uint32_t rows_per_band, uint32_t raw_cap, uint32_t out_cap,
             uint8_t* __restrict__ bands, uint32_t* __restrict__ band_size, uint32_t* __restrict__ band_adler) {
    extern __shared__ __align__(16) unsigned char png_smem[];
    uint8_t* raw = png_smem;
    const uint32_t* raww = reinterpret_cast<const uint32_t*>(png_smem);
    uint32_t* out = reinterpret_cast<uint32_t*>(png_smem + raw_cap);
    uint16_t* meta = reinterpret_cast<uint16_t*>(png_smem + raw_cap + out_cap);         // raw_cap / 4 entries
    uint16_t* woff = meta + raw_cap / 4u;                                                // bit offset of a word within its segment
    __shared__ int s_a[VF_PNG_THREADS], s_b[VF_PNG_THREADS];
    __shared__ uint32_t s_c[VF_PNG_THREADS];
    __shared__ uint32_t s_lit[256];                              // literal code | nbits << 16
    const uint32_t tid = threadIdx.x, band = blockIdx.x, nbands = gridDim.x;
    const uint32_t y0 = band * rows_per_band, rows = min(rows_per_band, H - y0);
    const uint32_t stride = 4u * W + 1u, N = rows * stride, Nw = (N + 3u) >> 2;

    { uint32_t nb; const uint32_t c = vf_lit_code(tid, nb); s_lit[tid] = c | (nb << 16); }
    for (uint32_t w = tid; w < out_cap / 4u; w += VF_PNG_THREADS) out[w] = 0u;
    if (tid == 0u) reinterpret_cast<uint32_t*>(raw)[Nw - 1u] = 0u;          // bytes past N in the last word
    __syncthreads();
    // ---- filter: type byte 1 (Sub) + pixel minus the pixel to its left, bytewise ----
    const uint32_t* px = reinterpret_cast<const uint32_t*>(frame) + (size_t)y0 * W;
    for (uint32_t r = 0; r < rows; ++r) {
        const uint32_t* prow = px + (size_t)r * W;
        uint8_t* orow = raw + r * stride + 1u;
        for (uint32_t x = tid; x < W; x += VF_PNG_THREADS) {
            const uint32_t cur = __ldg(prow + x), prev = x ? __ldg(prow + x - 1u) : 0u;
            const uint32_t d = __vsub4(cur, prev);
            uint8_t* o = orow + 4u * x;
            o[0] = (uint8_t)d; o[1] = (uint8_t)(d >> 8); o[2] = (uint8_t)(d >> 16); o[3] = (uint8_t)(d >> 24);
        }
        if (tid == 0u) raw[r * stride] = 1u;
    }
    __syncthreads();

    // ---- analysis: thread t owns words [w0, w1), bytes [s0, s1) ----
    constexpr int NW = (SEG_WORDS + 7) / 8;              // 32-bit words of match bits per segment
    const uint32_t segw = (Nw + VF_PNG_THREADS - 1u) / VF_PNG_THREADS;       // <= SEG_WORDS
    const uint32_t w0 = min(tid * segw, Nw), w1 = min(w0 + segw, Nw);
    const uint32_t s0 = min(4u * w0, N), s1 = min(4u * w1, N), nbytes_seg = s1 - s0;
    uint32_t m[NW + 1];                                          // match bit of byte s0 + k: bit k & 31 of m[k >> 5]
#pragma unroll
    for (int k = 0; k <= NW; ++k) m[k] = 0u;
    uint32_t adler_a = 0u, adler_b = 0u;                        // sum b, sum (N - i) b
    {
        uint32_t prev = w0 ? raww[w0 - 1u] : 0u;
#pragma unroll
        for (int k = 0; k < SEG_WORDS; ++k) {
            if (w0 + k < w1) {
                const uint32_t cur = raww[w0 + k];
                uint32_t bits = ((__vcmpeq4(cur, prev) & 0x01010101u) * 0x01020408u) >> 24;    // the four byte flags as bits 0..3
                if (w0 + k == 0u) bits = 0u;                    // the first four bytes have nothing 4 back
                const uint32_t i = 4u * (w0 + k);
                if (i + 4u > N) bits &= (1u << (N - i)) - 1u;   // bytes past the end of the band: breaks
                m[k >> 3] |= bits << (4 * (k & 7));
                const uint32_t b0 = cur & 255u, b1 = (cur >> 8) & 255u, b2 = (cur >> 16) & 255u, b3 = cur >> 24;
                adler_a += b0 + b1 + b2 + b3;                   // (bytes past N are zero)
                adler_b += (N - i) * (b0 + b1 + b2 + b3) - (b1 + 2u * b2 + 3u * b3);
                prev = cur;
            }
        }
    }
    adler_b %= 65521u;                                           // <= 40 words x 4 x 2^24: no overflow before this
#include "vf_png_encode_tail.inc"
}
