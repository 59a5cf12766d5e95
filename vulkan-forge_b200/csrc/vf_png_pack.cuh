// vf_png_pack.cuh -- k_png_pack: one CTA per band; the band becomes one IDAT chunk at its final offset in the file
// (included by vf_png_kernels.cuh).  length, type, data (first band: + zlib header, last band: + Adler-32 of the whole
// stream, folded from the per-band sums), CRC-32 in parallel (64-byte pieces, each shifted by x^(8 * bytes after it) mod P
// with a table of powers).  CTA 0 adds signature + IHDR, the last CTA IEND and the file size.
// pow_piece[j] = x^(8 * VF_PNG_PIECE * j) mod P, pow_byte[r] = x^(8 r) mod P (r < VF_PNG_PIECE)
#pragma once

__global__ void __launch_bounds__(VF_PNG_THREADS)
k_png_pack(const uint8_t* __restrict__ bands, uint32_t out_cap, const uint32_t* __restrict__ band_size,
           const uint32_t* __restrict__ band_adler, const uint8_t* __restrict__ preamble /* 33 bytes */,
           const uint32_t* __restrict__ pow_piece, const uint32_t* __restrict__ pow_byte,
           uint8_t* __restrict__ file, uint32_t file_cap, uint32_t* __restrict__ file_size) {
    // nothing that follows in a render stream reads this kernel's output on the device: the next frame's geometry pass,
    // when launched as a programmatic dependent, may start beside it
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    extern __shared__ __align__(16) unsigned char png_smem[];
    __shared__ uint32_t s_tab[256];
    __shared__ uint32_t s_red[VF_PNG_THREADS];
    __shared__ uint32_t s_ad[3][VF_PNG_THREADS];
    const uint32_t tid = threadIdx.x, band = blockIdx.x, nbands = gridDim.x;
    const bool first = band == 0u, last = band + 1u == nbands;
    {   // byte-wise CRC table
        uint32_t c = tid;
        for (int k = 0; k < 8; ++k) c = (c & 1u) ? (c >> 1) ^ VF_PNG_POLY : c >> 1;
        s_tab[tid] = c;
    }
    // offset of this chunk: preamble + every chunk before it (12 bytes of framing each; the first holds the zlib header)
    uint32_t part = 0u;
    for (uint32_t k = tid; k < band; k += VF_PNG_THREADS) part += band_size[k] + 12u;
    s_red[tid] = part;
    uint32_t adler = 0u;
    if (last) {
        // Adler-32 of the whole filtered stream from the per-band sums (a, b, n) = (sum of bytes, sum (n - i) b_i, bytes):
        // two consecutive pieces combine as (a1 + a2, b1 + b2 + n2 a1, n1 + n2), an associative operation, so threads
        // fold contiguous runs of bands and a tree joins them in order; the initial A = 1 adds n to B at the end
        const uint32_t per = (nbands + VF_PNG_THREADS - 1u) / VF_PNG_THREADS;
        uint32_t A = 0u, B = 0u, n = 0u;
        for (uint32_t k = tid * per; k < min(nbands, (tid + 1u) * per); ++k) {
            const uint32_t a2 = band_adler[3u * k], b2 = band_adler[3u * k + 1u], n2 = band_adler[3u * k + 2u] % 65521u;
            B = (uint32_t)((B + b2 + (unsigned long long)n2 * A) % 65521u);
            A = (A + a2) % 65521u; n = (n + n2) % 65521u;
        }
        s_ad[0][tid] = A; s_ad[1][tid] = B; s_ad[2][tid] = n;
    }
    __syncthreads();
    for (uint32_t d = 1u; d < VF_PNG_THREADS; d <<= 1) {
        if (tid % (2u * d) == 0u) {
            s_red[tid] += s_red[tid + d];
            if (last) {
                const uint32_t A1 = s_ad[0][tid], B1 = s_ad[1][tid], n1 = s_ad[2][tid];
                const uint32_t A2 = s_ad[0][tid + d], B2 = s_ad[1][tid + d], n2 = s_ad[2][tid + d];
                s_ad[0][tid] = (A1 + A2) % 65521u;
                s_ad[1][tid] = (uint32_t)((B1 + B2 + (unsigned long long)n2 * A1) % 65521u);
                s_ad[2][tid] = (n1 + n2) % 65521u;
            }
        }
        __syncthreads();
    }
    const uint32_t offset = 33u + s_red[0] + (band ? 2u : 0u);
    if (last) adler = (((s_ad[1][0] + s_ad[2][0]) % 65521u) << 16) | ((1u + s_ad[0][0]) % 65521u);
    const uint32_t bsz = band_size[band];
    const uint32_t dl = bsz + (first ? 2u : 0u) + (last ? 4u : 0u);        // chunk data length
    const uint32_t n = 4u + dl;                                  // bytes under the CRC: type + data
    __syncthreads();                                             // s_red is reused below
    // The chunk is assembled in shared memory at the same alignment (mod 4) it will have in the file, so that it goes
    // out in aligned 32-bit stores: wbuf = [length][IDAT][data][crc], wbuf[0] <-> file[offset]
    uint8_t* wbuf = png_smem + 16u + (offset & 3u);
    uint8_t* chunk = wbuf + 4u;
    if (tid == 0u) {
        wbuf[0] = (uint8_t)(dl >> 24); wbuf[1] = (uint8_t)(dl >> 16); wbuf[2] = (uint8_t)(dl >> 8); wbuf[3] = (uint8_t)dl;
        chunk[0] = 'I'; chunk[1] = 'D'; chunk[2] = 'A'; chunk[3] = 'T';
        if (first) { chunk[4] = 0x78; chunk[5] = 0x01; }
    }
    const uint32_t* src = reinterpret_cast<const uint32_t*>(bands + (size_t)band * out_cap);
    uint8_t* body = chunk + 4u + (first ? 2u : 0u);
    for (uint32_t k = tid; k < (bsz + 3u) / 4u; k += VF_PNG_THREADS) {
        const uint32_t v = __ldg(src + k);
        uint8_t* o = body + 4u * k;
        o[0] = (uint8_t)v; o[1] = (uint8_t)(v >> 8); o[2] = (uint8_t)(v >> 16); o[3] = (uint8_t)(v >> 24);   // (a few bytes past bsz: overwritten below)
    }
    __syncthreads();
    if (last && tid == 0u) { body[bsz] = (uint8_t)(adler >> 24); body[bsz + 1u] = (uint8_t)(adler >> 16); body[bsz + 2u] = (uint8_t)(adler >> 8); body[bsz + 3u] = (uint8_t)adler; }
    __syncthreads();
    // CRC-32: pieces counted from the END (piece j ends VF_PNG_PIECE * j bytes before the end), each run from a zero
    // register and multiplied by x^(8 * bytes after it); the initial all-ones register is shifted by the whole length
    uint32_t acc = 0u;
    const uint32_t npieces = (n + VF_PNG_PIECE - 1u) / VF_PNG_PIECE;
    for (uint32_t j = tid; j < npieces; j += VF_PNG_THREADS) {
        const uint32_t e = n - j * VF_PNG_PIECE, b = e > VF_PNG_PIECE ? e - VF_PNG_PIECE : 0u;
        uint32_t c = 0u;
        for (uint32_t k = b; k < e; ++k) c = s_tab[(c ^ chunk[k]) & 255u] ^ (c >> 8);
        acc ^= j ? vf_gf_mul(pow_piece[j], c) : c;
    }
    if (tid == 0u) acc ^= vf_gf_mul(vf_gf_mul(pow_piece[n / VF_PNG_PIECE], pow_byte[n % VF_PNG_PIECE]), 0xffffffffu);
    s_red[tid] = acc;
    __syncthreads();
    for (uint32_t d = VF_PNG_THREADS / 2u; d; d >>= 1) { if (tid < d) s_red[tid] ^= s_red[tid + d]; __syncthreads(); }
    const uint32_t crc = s_red[0] ^ 0xffffffffu;
    uint32_t total = 8u + n;                                     // bytes of wbuf to write
    if (tid == 0u) {
        uint8_t* c4 = chunk + n;
        c4[0] = (uint8_t)(crc >> 24); c4[1] = (uint8_t)(crc >> 16); c4[2] = (uint8_t)(crc >> 8); c4[3] = (uint8_t)crc;
        if (last) { const uint8_t iend[12] = {0, 0, 0, 0, 'I', 'E', 'N', 'D', 0xae, 0x42, 0x60, 0x82}; for (int k = 0; k < 12; ++k) c4[4 + k] = iend[k]; }
    }
    if (last) total += 12u;
    __syncthreads();
    const bool fits = offset + total <= file_cap;
    if (fits) {
        // aligned words of the file that lie entirely inside [offset, offset + total); the ragged ends byte by byte
        const uint32_t a0 = (offset + 3u) & ~3u, a1 = (offset + total) & ~3u;
        if (a0 <= a1) {                                          // wbuf[g - offset] <-> file[g], same alignment mod 4
            for (uint32_t g = a0 + 4u * tid; g < a1; g += 4u * VF_PNG_THREADS)
                *reinterpret_cast<uint32_t*>(file + g) = *reinterpret_cast<const uint32_t*>(wbuf + (g - offset));
            if (tid < a0 - offset) file[offset + tid] = wbuf[tid];
            if (tid < offset + total - a1) file[a1 + tid] = wbuf[a1 - offset + tid];
        } else if (tid < total) {
            file[offset + tid] = wbuf[tid];
        }
        if (first) for (uint32_t k = tid; k < 33u; k += VF_PNG_THREADS) file[k] = preamble[k];
    }
    if (last && tid == 0u) {                                                          // (file_size lies in mapped host memory)
        *file_size = fits ? offset + total : 0xffffffffu;                             // all ones: did not fit
        __threadfence_system();
    }
}
