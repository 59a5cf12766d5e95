// vf_png.cpp -- RGBA8 PNG writer (zlib deflate; no libpng on the build machines).
// replaces: image::RgbaImage::save, src/scene/mod.rs:331-333, src/terrain/mod.rs:487-489,
// src/lib.rs:330-332.  Byte identity with the `image` crate is neither needed nor possible;
// the reference's tests only pin "file exists / size > N / size differs".
// The image is split into horizontal bands that are filtered (Sub) and deflated on separate
// host threads as independent raw-deflate segments joined by Z_SYNC_FLUSH boundaries, so large
// frames do not serialise on a single core.
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <zlib.h>

#include <algorithm>
#include <thread>
#include <vector>

#include "vf_internal.h"

static void put_u32(std::vector<uint8_t>& v, uint32_t x) {
    v.push_back((uint8_t)(x >> 24)); v.push_back((uint8_t)(x >> 16));
    v.push_back((uint8_t)(x >> 8)); v.push_back((uint8_t)x);
}

static bool write_chunk(FILE* f, const char type[4], const uint8_t* data, size_t len) {
    uint8_t hdr[8] = {(uint8_t)(len >> 24), (uint8_t)(len >> 16), (uint8_t)(len >> 8), (uint8_t)len,
                      (uint8_t)type[0], (uint8_t)type[1], (uint8_t)type[2], (uint8_t)type[3]};
    uint32_t crc = crc32(0L, hdr + 4, 4);
    if (len) crc = crc32(crc, data, (uInt)len);
    uint8_t tail[4] = {(uint8_t)(crc >> 24), (uint8_t)(crc >> 16), (uint8_t)(crc >> 8), (uint8_t)crc};
    return fwrite(hdr, 1, 8, f) == 8 && (len == 0 || fwrite(data, 1, len, f) == len) &&
           fwrite(tail, 1, 4, f) == 4;
}

struct Band {
    int y0, y1;
    std::vector<uint8_t> out;   // raw deflate bytes
    uint32_t adler = 1;         // adler32 of the band's filtered bytes
    size_t raw_len = 0;
    bool ok = true;
};

static void deflate_band(const uint8_t* rgba, int w, Band* b, bool last) {
    const size_t stride = (size_t)w * 4;
    std::vector<uint8_t> raw((stride + 1) * (size_t)(b->y1 - b->y0));
    uint8_t* p = raw.data();
    for (int y = b->y0; y < b->y1; ++y) {
        const uint8_t* row = rgba + (size_t)y * stride;
        *p++ = 1;   // filter type Sub
        for (size_t i = 0; i < stride; ++i) p[i] = (uint8_t)(row[i] - (i >= 4 ? row[i - 4] : 0));
        p += stride;
    }
    b->raw_len = raw.size();
    b->adler = adler32(1L, raw.data(), (uInt)raw.size());
    z_stream zs;
    memset(&zs, 0, sizeof zs);
    if (deflateInit2(&zs, 3, Z_DEFLATED, -15, 8, Z_DEFAULT_STRATEGY) != Z_OK) { b->ok = false; return; }
    b->out.resize(deflateBound(&zs, (uLong)raw.size()) + 16);
    zs.next_in = raw.data(); zs.avail_in = (uInt)raw.size();
    zs.next_out = b->out.data(); zs.avail_out = (uInt)b->out.size();
    const int rc = deflate(&zs, last ? Z_FINISH : Z_FULL_FLUSH);
    b->ok = last ? rc == Z_STREAM_END : rc == Z_OK;
    b->out.resize(b->out.size() - zs.avail_out);
    deflateEnd(&zs);
}

extern "C" int vf_png_write(const char* path, const uint8_t* rgba, int w, int h) {
    if (!path || !rgba || w <= 0 || h <= 0) return vf_set_error(VF_E_INVALID, "invalid PNG arguments");
    if (((uint64_t)w * 4 + 1) * (uint64_t)h > 0x7fffffffull)
        return vf_set_error(VF_E_INVALID, "image too large for the PNG writer");
    const int hw = (int)std::max(1u, std::thread::hardware_concurrency());
    const int nb = std::max(1, std::min({hw, 16, h / 64 + 1}));
    std::vector<Band> bands(nb);
    for (int i = 0; i < nb; ++i) {
        bands[i].y0 = (int)((int64_t)h * i / nb);
        bands[i].y1 = (int)((int64_t)h * (i + 1) / nb);
    }
    std::vector<std::thread> th;
    for (int i = 1; i < nb; ++i) th.emplace_back(deflate_band, rgba, w, &bands[i], i == nb - 1);
    deflate_band(rgba, w, &bands[0], nb == 1);
    for (auto& t : th) t.join();
    std::vector<uint8_t> idat;
    idat.push_back(0x78); idat.push_back(0x5e);            // zlib header (32K window, level hint)
    uint32_t adler = 1;
    for (int i = 0; i < nb; ++i) {
        if (!bands[i].ok) return vf_set_error(VF_E_IO, "deflate failed");
        idat.insert(idat.end(), bands[i].out.begin(), bands[i].out.end());
        adler = i == 0 ? bands[i].adler : adler32_combine(adler, bands[i].adler, (z_off_t)bands[i].raw_len);
    }
    put_u32(idat, adler);
    FILE* f = fopen(path, "wb");
    if (!f) return vf_set_error(VF_E_IO, "cannot open '%s' for writing", path);
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    std::vector<uint8_t> ihdr;
    put_u32(ihdr, (uint32_t)w); put_u32(ihdr, (uint32_t)h);
    ihdr.push_back(8); ihdr.push_back(6); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
    bool ok = fwrite(sig, 1, 8, f) == 8 && write_chunk(f, "IHDR", ihdr.data(), ihdr.size()) &&
              write_chunk(f, "IDAT", idat.data(), idat.size()) && write_chunk(f, "IEND", nullptr, 0);
    ok = (fclose(f) == 0) && ok;
    if (!ok) return vf_set_error(VF_E_IO, "short write to '%s'", path);
    return VF_OK;
}
