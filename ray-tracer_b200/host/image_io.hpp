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
#include <string>
#include <vector>

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

}  // namespace rth
