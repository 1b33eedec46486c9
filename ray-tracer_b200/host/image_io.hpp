// Encoders for the step right after the render path: the example drivers' output formats,
// fed with the top-down 8-bit image that rt_resolve_u8 produces (sqrt-gamma, x255, min 255, truncation — the
// arithmetic of examples/book-one.rs:95-97 runs on the device; these functions only serialise bytes).
//
//   PPM  examples/book-one.rs:28-30,90-100   "P3\n<w> <h>\n255\n" then one "r g b\n" line per pixel, rows top -> bottom.
//   PNG  examples/main.rs:105-135            8-bit RGBA, alpha 255, row 0 = top (put_pixel(x, height-1-y)).  The reference
//        delegates to the `image` crate; any valid PNG stream of the same pixels is equivalent, so this writer emits
//        stored (uncompressed) deflate blocks: no zlib dependency, byte-exact pixels after any decoder.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <zlib.h>

#include "jpeg_decode.hpp"

namespace rth {

inline std::string encode_ppm_p3(const uint8_t* rgb_top_down, uint32_t width, uint32_t height) {
    std::string out;
    out.reserve((size_t)width * height * 12 + 32);
    out += "P3\n" + std::to_string(width) + " " + std::to_string(height) + "\n255\n";
    char line[16];
    const size_t n = (size_t)width * height;
    for (size_t i = 0; i < n; ++i) {
        const int len = std::snprintf(line, sizeof(line), "%u %u %u\n", (unsigned)rgb_top_down[3 * i], (unsigned)rgb_top_down[3 * i + 1],
                                      (unsigned)rgb_top_down[3 * i + 2]);
        out.append(line, (size_t)len);
    }
    return out;
}

namespace detail {
inline uint32_t crc32_update(uint32_t crc, const uint8_t* p, size_t n) {
    static uint32_t table[256];
    static bool ready = false;
    if (!ready) {
        for (uint32_t i = 0; i < 256; ++i) {
            uint32_t c = i;
            for (int k = 0; k < 8; ++k) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
            table[i] = c;
        }
        ready = true;
    }
    for (size_t i = 0; i < n; ++i) crc = table[(crc ^ p[i]) & 0xFFu] ^ (crc >> 8);
    return crc;
}
inline void put_be32(std::vector<uint8_t>& v, uint32_t x) {
    v.push_back((uint8_t)(x >> 24)); v.push_back((uint8_t)(x >> 16)); v.push_back((uint8_t)(x >> 8)); v.push_back((uint8_t)x);
}
inline void put_chunk(std::vector<uint8_t>& png, const char type[4], const std::vector<uint8_t>& data) {
    put_be32(png, (uint32_t)data.size());
    const size_t start = png.size();
    png.insert(png.end(), type, type + 4);
    png.insert(png.end(), data.begin(), data.end());
    const uint32_t crc = crc32_update(0xFFFFFFFFu, png.data() + start, png.size() - start) ^ 0xFFFFFFFFu;
    put_be32(png, crc);
}
}  // namespace detail

// RGBA8 PNG (colour type 6), alpha 255, filter 0 on every row, zlib stream of stored blocks.
inline std::vector<uint8_t> encode_png_rgba8(const uint8_t* rgb_top_down, uint32_t width, uint32_t height) {
    using namespace detail;
    std::vector<uint8_t> raw;
    raw.reserve((size_t)height * (1 + 4 * (size_t)width));
    for (uint32_t y = 0; y < height; ++y) {
        raw.push_back(0);  // filter type None
        for (uint32_t x = 0; x < width; ++x) {
            const uint8_t* p = rgb_top_down + 3 * ((size_t)y * width + x);
            raw.push_back(p[0]); raw.push_back(p[1]); raw.push_back(p[2]); raw.push_back(255);
        }
    }
    std::vector<uint8_t> z;
    z.reserve(raw.size() + raw.size() / 65535 * 5 + 16);
    z.push_back(0x78); z.push_back(0x01);  // zlib header: deflate, 32 K window, no preset dictionary
    size_t pos = 0;
    do {
        const size_t len = std::min<size_t>(65535, raw.size() - pos);
        z.push_back(pos + len == raw.size() ? 1 : 0);  // BFINAL, BTYPE = 00 (stored)
        z.push_back((uint8_t)(len & 0xFF)); z.push_back((uint8_t)(len >> 8));
        z.push_back((uint8_t)(~len & 0xFF)); z.push_back((uint8_t)((~len >> 8) & 0xFF));
        z.insert(z.end(), raw.begin() + (long)pos, raw.begin() + (long)(pos + len));
        pos += len;
    } while (pos < raw.size());
    uint32_t a = 1, b = 0;  // Adler-32 of the uncompressed data
    for (size_t i = 0; i < raw.size(); ++i) { a = (a + raw[i]) % 65521u; b = (b + a) % 65521u; }
    put_be32(z, (b << 16) | a);

    std::vector<uint8_t> png = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    std::vector<uint8_t> ihdr;
    put_be32(ihdr, width); put_be32(ihdr, height);
    ihdr.push_back(8); ihdr.push_back(6); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
    put_chunk(png, "IHDR", ihdr);
    put_chunk(png, "IDAT", z);
    put_chunk(png, "IEND", {});
    return png;
}

// ---- decoders (texture ingestion; examples/main.rs:266 opens earthmap.jpg with the `image` crate) ----------
// PNG (8-bit, colour types 0 / 2 / 3 / 4 / 6, non-interlaced; zlib inflates the IDAT stream) and PPM (P3 / P6, maxval <= 255) to
// top-down 8-bit RGB — what `image.as_rgb8()` hands the texture closure of main.rs:267-280.  JPEG is not decoded.
// Returns 0, -1 malformed (bad signature / CRC / truncated), -2 unsupported (16-bit, interlaced, maxval > 255).
namespace detail {
inline uint32_t get_be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }
inline int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}
}  // namespace detail

inline int decode_png(const uint8_t* data, size_t n, uint32_t& width, uint32_t& height, std::vector<uint8_t>& rgb) {
    using namespace detail;
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    if (n < 8 || std::memcmp(data, sig, 8) != 0) return -1;
    size_t pos = 8;
    uint32_t depth = 0, ctype = 0, interlace = 0;
    bool have_ihdr = false, have_end = false;
    std::vector<uint8_t> idat, plte;
    while (pos + 12 <= n && !have_end) {
        const uint32_t len = get_be32(data + pos);
        if (pos + 12 + (size_t)len > n) return -1;
        const uint8_t* type = data + pos + 4;
        const uint8_t* body = data + pos + 8;
        if ((crc32_update(0xFFFFFFFFu, type, 4 + (size_t)len) ^ 0xFFFFFFFFu) != get_be32(body + len)) return -1;
        if (!std::memcmp(type, "IHDR", 4)) {
            if (len != 13) return -1;
            width = get_be32(body); height = get_be32(body + 4); depth = body[8]; ctype = body[9]; interlace = body[12];
            have_ihdr = true;
        } else if (!std::memcmp(type, "PLTE", 4)) {
            plte.assign(body, body + len);
        } else if (!std::memcmp(type, "IDAT", 4)) {
            idat.insert(idat.end(), body, body + len);
        } else if (!std::memcmp(type, "IEND", 4)) {
            have_end = true;
        }
        pos += 12 + (size_t)len;
    }
    if (!have_ihdr || !have_end || width == 0 || height == 0) return -1;
    if (depth != 8 || interlace != 0) return -2;
    const uint32_t channels = ctype == 0 ? 1u : ctype == 2 ? 3u : ctype == 3 ? 1u : ctype == 4 ? 2u : ctype == 6 ? 4u : 0u;
    if (!channels) return -2;
    if (ctype == 3 && plte.empty()) return -1;
    const size_t stride = (size_t)width * channels;
    std::vector<uint8_t> raw((stride + 1) * height);
    uLongf out_len = (uLongf)raw.size();
    if (uncompress(raw.data(), &out_len, idat.data(), (uLong)idat.size()) != Z_OK || out_len != raw.size()) return -1;
    std::vector<uint8_t> prev(stride, 0), cur(stride);
    rgb.resize((size_t)width * height * 3);
    for (uint32_t y = 0; y < height; ++y) {
        const uint8_t* line = raw.data() + (stride + 1) * y;
        const uint8_t filter = line[0];
        if (filter > 4) return -1;
        for (size_t i = 0; i < stride; ++i) {
            const int a = i >= channels ? cur[i - channels] : 0, b = prev[i], c = i >= channels ? prev[i - channels] : 0;
            const int x = line[1 + i];
            const int v = filter == 0 ? x : filter == 1 ? x + a : filter == 2 ? x + b : filter == 3 ? x + ((a + b) >> 1) : x + paeth(a, b, c);
            cur[i] = (uint8_t)v;
        }
        for (uint32_t xx = 0; xx < width; ++xx) {
            uint8_t* o = rgb.data() + 3 * ((size_t)y * width + xx);
            const uint8_t* p = cur.data() + (size_t)xx * channels;
            if (ctype == 2 || ctype == 6) { o[0] = p[0]; o[1] = p[1]; o[2] = p[2]; }
            else if (ctype == 3) { if ((size_t)p[0] * 3 + 2 >= plte.size()) return -1; o[0] = plte[p[0] * 3]; o[1] = plte[p[0] * 3 + 1]; o[2] = plte[p[0] * 3 + 2]; }
            else { o[0] = o[1] = o[2] = p[0]; }
        }
        prev.swap(cur);
    }
    return 0;
}

inline int decode_ppm(const uint8_t* data, size_t n, uint32_t& width, uint32_t& height, std::vector<uint8_t>& rgb) {
    if (n < 2 || data[0] != 'P' || (data[1] != '3' && data[1] != '6')) return -1;
    const bool text = data[1] == '3';
    size_t pos = 2;
    auto token = [&](uint32_t& out) -> bool {  // next unsigned integer, skipping whitespace and # comments
        for (;;) {
            while (pos < n && (data[pos] == ' ' || data[pos] == '\t' || data[pos] == '\n' || data[pos] == '\r')) ++pos;
            if (pos < n && data[pos] == '#') { while (pos < n && data[pos] != '\n') ++pos; continue; }
            break;
        }
        if (pos >= n || data[pos] < '0' || data[pos] > '9') return false;
        uint64_t v = 0;
        while (pos < n && data[pos] >= '0' && data[pos] <= '9') { v = v * 10 + (data[pos] - '0'); if (v > 0xFFFFFFFFull) return false; ++pos; }
        out = (uint32_t)v;
        return true;
    };
    uint32_t maxval = 0;
    if (!token(width) || !token(height) || !token(maxval) || width == 0 || height == 0 || maxval == 0) return -1;
    if (maxval > 255) return -2;
    const size_t count = (size_t)width * height * 3;
    rgb.resize(count);
    if (text) {
        for (size_t i = 0; i < count; ++i) { uint32_t v; if (!token(v) || v > maxval) return -1; rgb[i] = (uint8_t)v; }
    } else {
        ++pos;  // the single whitespace byte after maxval
        if (pos + count > n) return -1;
        std::memcpy(rgb.data(), data + pos, count);
    }
    return 0;
}

inline int decode_image(const uint8_t* data, size_t n, uint32_t& width, uint32_t& height, std::vector<uint8_t>& rgb) {
    if (n >= 2 && data[0] == 'P') return decode_ppm(data, n, width, height, rgb);
    if (n >= 2 && data[0] == 0xFF && data[1] == 0xD8) return jpeg::decode(data, n, width, height, rgb);  // earthmap.jpg (examples/main.rs:266)
    return decode_png(data, n, width, height, rgb);
}

}  // namespace rth
