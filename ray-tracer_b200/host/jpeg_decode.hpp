// Baseline / extended-sequential JPEG decoder (ITU T.81 Huffman, 8-bit, 1 or 3 components, any sampling factors up to 2x2 per
// axis ratio, restart intervals) for texture ingestion: examples/main.rs:266 opens ./earthmap.jpg with the `image` crate and
// the texture closure of main.rs:267-280 reads it through `as_rgb8()`.  JPEG is lossy and T.81 leaves the IDCT and the chroma
// upsampling to the decoder within a tolerance; this decoder follows the de-facto reference (IJG libjpeg 6b defaults: the
// "islow" integer IDCT, triangle-filter "fancy" upsampling of 2:1 subsampled chroma, fixed-point YCbCr -> RGB), which is what
// PIL — the checker in tests/test_host_abi.py — links.  Progressive (SOF2), arithmetic-coded, 12-bit and 4-component streams are
// reported as unsupported (-2), malformed streams as -1.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>

namespace rth {
namespace jpeg {

struct Huff {  // canonical Huffman table (T.81 Annex C / F.2.2.3)
    bool present = false;
    uint8_t values[256];
    int32_t mincode[17], maxcode[18], valptr[17];
    void build(const uint8_t counts[16], const uint8_t* vals, int n) {
        present = true;
        std::memcpy(values, vals, (size_t)n);
        int code = 0, k = 0;
        for (int len = 1; len <= 16; ++len) {
            valptr[len] = k;
            mincode[len] = code;
            code += counts[len - 1];
            k += counts[len - 1];
            maxcode[len] = counts[len - 1] ? code - 1 : -1;
            code <<= 1;
        }
        maxcode[17] = 0x7FFFFFFF;
    }
};

struct Component {
    int id = 0, h = 1, v = 1, tq = 0, td = 0, ta = 0;
    int pred = 0;             // DC predictor
    int bw = 0, bh = 0;       // plane size in blocks (padded to whole MCUs)
    std::vector<uint8_t> px;  // plane, (bw * 8) x (bh * 8)
};

struct BitReader {
    const uint8_t* p;
    const uint8_t* end;
    uint32_t acc = 0;
    int nbits = 0;
    bool hit_marker = false;
    bool exhausted = false;  // ran past the end of the data without meeting a marker: a truncated stream
    void fill() {
        while (nbits <= 24) {
            uint32_t byte = 0;
            if (!hit_marker && p < end) {
                byte = *p;
                if (byte == 0xFF) {
                    if (p + 1 < end && p[1] == 0x00) p += 2;               // stuffed zero
                    else { hit_marker = true; byte = 0; }                  // a marker: feed zeros until the caller consumes it
                } else ++p;
            } else { if (!hit_marker) exhausted = true; hit_marker = true; }
            acc |= byte << (24 - nbits);
            nbits += 8;
        }
    }
    int bits(int n) {  // n in [0, 16]
        if (n == 0) return 0;
        if (nbits < n) fill();
        const int v = (int)(acc >> (32 - n));
        acc <<= n;
        nbits -= n;
        return v;
    }
    int decode(const Huff& h) {
        if (nbits < 16) fill();
        int code = 0;
        for (int len = 1; len <= 16; ++len) {
            code = (code << 1) | (int)(acc >> 31);
            acc <<= 1;
            --nbits;
            if (h.maxcode[len] >= 0 && code <= h.maxcode[len] && code >= h.mincode[len]) return h.values[h.valptr[len] + code - h.mincode[len]];
        }
        return -1;
    }
    void reset() { acc = 0; nbits = 0; hit_marker = false; }
};

inline int extend(int v, int t) { return v < (1 << (t - 1)) ? v - (1 << t) + 1 : v; }  // T.81 F.2.2.1

static const uint8_t kZigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                    41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                    30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// IJG "islow" inverse DCT (jidctint.c of libjpeg 6b: 13-bit constants, 2 extra bits between the passes), output range-limited.
inline void idct_islow(const int32_t* in, uint8_t* out, int stride) {
    constexpr int CB = 13, P1 = 2;
    constexpr int32_t F0298 = 2446, F0390 = 3196, F0541 = 4433, F0765 = 6270, F0899 = 7373, F1175 = 9633, F1501 = 12299, F1847 = 15137, F1961 = 16069,
                      F2053 = 16819, F2562 = 20995, F3072 = 25172;
    auto descale = [](int64_t x, int n) { return (int32_t)((x + ((int64_t)1 << (n - 1))) >> n); };
    int32_t ws[64];
    for (int c = 0; c < 8; ++c) {
        const int32_t* i = in + c;
        int32_t* w = ws + c;
        if (!(i[8] | i[16] | i[24] | i[32] | i[40] | i[48] | i[56])) {
            const int32_t dc = i[0] * (1 << P1);
            for (int r = 0; r < 8; ++r) w[8 * r] = dc;
            continue;
        }
        int64_t z2 = i[16], z3 = i[48];
        int64_t z1 = (z2 + z3) * F0541;
        int64_t tmp2 = z1 - z3 * F1847, tmp3 = z1 + z2 * F0765;
        z2 = i[0]; z3 = i[32];
        int64_t tmp0 = (z2 + z3) * (1 << CB), tmp1 = (z2 - z3) * (1 << CB);
        const int64_t t10 = tmp0 + tmp3, t13 = tmp0 - tmp3, t11 = tmp1 + tmp2, t12 = tmp1 - tmp2;
        tmp0 = i[56]; tmp1 = i[40]; tmp2 = i[24]; tmp3 = i[8];
        z1 = tmp0 + tmp3; z2 = tmp1 + tmp2; z3 = tmp0 + tmp2;
        int64_t z4 = tmp1 + tmp3;
        const int64_t z5 = (z3 + z4) * F1175;
        tmp0 *= F0298; tmp1 *= F2053; tmp2 *= F3072; tmp3 *= F1501;
        z1 *= -F0899; z2 *= -F2562; z3 *= -F1961; z4 *= -F0390;
        z3 += z5; z4 += z5;
        tmp0 += z1 + z3; tmp1 += z2 + z4; tmp2 += z2 + z3; tmp3 += z1 + z4;
        w[0] = descale(t10 + tmp3, CB - P1); w[56] = descale(t10 - tmp3, CB - P1);
        w[8] = descale(t11 + tmp2, CB - P1); w[48] = descale(t11 - tmp2, CB - P1);
        w[16] = descale(t12 + tmp1, CB - P1); w[40] = descale(t12 - tmp1, CB - P1);
        w[24] = descale(t13 + tmp0, CB - P1); w[32] = descale(t13 - tmp0, CB - P1);
    }
    auto limit = [](int32_t v) { v += 128; return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); };
    for (int r = 0; r < 8; ++r) {
        const int32_t* w = ws + 8 * r;
        uint8_t* o = out + (size_t)r * stride;
        int64_t z2 = w[2], z3 = w[6];
        int64_t z1 = (z2 + z3) * F0541;
        int64_t tmp2 = z1 - z3 * F1847, tmp3 = z1 + z2 * F0765;
        int64_t tmp0 = ((int64_t)w[0] + w[4]) * (1 << CB), tmp1 = ((int64_t)w[0] - w[4]) * (1 << CB);
        const int64_t t10 = tmp0 + tmp3, t13 = tmp0 - tmp3, t11 = tmp1 + tmp2, t12 = tmp1 - tmp2;
        tmp0 = w[7]; tmp1 = w[5]; tmp2 = w[3]; tmp3 = w[1];
        z1 = tmp0 + tmp3; z2 = tmp1 + tmp2; z3 = tmp0 + tmp2;
        int64_t z4 = tmp1 + tmp3;
        const int64_t z5 = (z3 + z4) * F1175;
        tmp0 *= F0298; tmp1 *= F2053; tmp2 *= F3072; tmp3 *= F1501;
        z1 *= -F0899; z2 *= -F2562; z3 *= -F1961; z4 *= -F0390;
        z3 += z5; z4 += z5;
        tmp0 += z1 + z3; tmp1 += z2 + z4; tmp2 += z2 + z3; tmp3 += z1 + z4;
        constexpr int S = CB + P1 + 3;
        o[0] = limit(descale(t10 + tmp3, S)); o[7] = limit(descale(t10 - tmp3, S));
        o[1] = limit(descale(t11 + tmp2, S)); o[6] = limit(descale(t11 - tmp2, S));
        o[2] = limit(descale(t12 + tmp1, S)); o[5] = limit(descale(t12 - tmp1, S));
        o[3] = limit(descale(t13 + tmp0, S)); o[4] = limit(descale(t13 - tmp0, S));
    }
}

// libjpeg's "fancy" (triangle filter) upsampling, jdsample.c: h2v1 and h2v2; 1:1 copies; other ratios replicate.
inline void upsample_h2v1(const uint8_t* in, int w, uint8_t* out) {  // one row, w input samples -> 2 w outputs
    if (w == 1) { out[0] = out[1] = in[0]; return; }
    out[0] = in[0];
    out[1] = (uint8_t)((in[0] * 3 + in[1] + 2) >> 2);
    for (int i = 1; i < w - 1; ++i) {
        const int v = in[i] * 3;
        out[2 * i] = (uint8_t)((v + in[i - 1] + 1) >> 2);
        out[2 * i + 1] = (uint8_t)((v + in[i + 1] + 2) >> 2);
    }
    out[2 * w - 2] = (uint8_t)((in[w - 1] * 3 + in[w - 2] + 1) >> 2);
    out[2 * w - 1] = in[w - 1];
}
inline void upsample_h2v2_row(const uint8_t* near_row, const uint8_t* far_row, int w, uint8_t* out) {  // one OUTPUT row
    if (w == 1) { const int s = near_row[0] * 3 + far_row[0]; out[0] = (uint8_t)((s * 4 + 8) >> 4); out[1] = (uint8_t)((s * 4 + 7) >> 4); return; }
    int thiscol = near_row[0] * 3 + far_row[0], nextcol = near_row[1] * 3 + far_row[1], lastcol;
    out[0] = (uint8_t)((thiscol * 4 + 8) >> 4);
    out[1] = (uint8_t)((thiscol * 3 + nextcol + 7) >> 4);
    lastcol = thiscol; thiscol = nextcol;
    for (int i = 1; i < w - 1; ++i) {
        nextcol = near_row[i + 1] * 3 + far_row[i + 1];
        out[2 * i] = (uint8_t)((thiscol * 3 + lastcol + 8) >> 4);
        out[2 * i + 1] = (uint8_t)((thiscol * 3 + nextcol + 7) >> 4);
        lastcol = thiscol; thiscol = nextcol;
    }
    out[2 * w - 2] = (uint8_t)((thiscol * 3 + lastcol + 8) >> 4);
    out[2 * w - 1] = (uint8_t)((thiscol * 4 + 7) >> 4);
}

// Returns 0 ok, -1 malformed, -2 unsupported.  rgb: top-down 8-bit RGB.
inline int decode(const uint8_t* data, size_t n, uint32_t& width, uint32_t& height, std::vector<uint8_t>& rgb) {
    if (n < 4 || data[0] != 0xFF || data[1] != 0xD8) return -1;
    uint16_t qt[4][64];
    bool qt_present[4] = {false, false, false, false};
    Huff dc[4], ac[4];
    std::vector<Component> comps;
    int restart_interval = 0, hmax = 1, vmax = 1;
    bool adobe = false;
    int adobe_transform = -1;
    bool have_frame = false, decoded = false;
    size_t pos = 2;
    auto be16 = [&](size_t at) { return (int)((data[at] << 8) | data[at + 1]); };
    while (pos + 4 <= n && !decoded) {
        if (data[pos] != 0xFF) return -1;
        while (pos < n && data[pos] == 0xFF) ++pos;  // fill bytes
        if (pos >= n) return -1;
        const int marker = data[pos++];
        if (marker == 0xD9) break;                       // EOI
        if (marker == 0x01 || (marker >= 0xD0 && marker <= 0xD7)) continue;  // stand-alone markers
        if (pos + 2 > n) return -1;
        const int len = be16(pos);
        if (len < 2 || pos + (size_t)len > n) return -1;
        const uint8_t* seg = data + pos + 2;
        const int seglen = len - 2;
        switch (marker) {
            case 0xDB: {  // DQT
                int at = 0;
                while (at < seglen) {
                    const int pq = seg[at] >> 4, tq = seg[at] & 15;
                    ++at;
                    if (tq > 3 || pq > 1 || at + 64 * (pq + 1) > seglen) return -1;
                    for (int i = 0; i < 64; ++i) { qt[tq][kZigzag[i]] = (uint16_t)(pq ? ((seg[at] << 8) | seg[at + 1]) : seg[at]); at += pq + 1; }
                    qt_present[tq] = true;
                }
            } break;
            case 0xC4: {  // DHT
                int at = 0;
                while (at < seglen) {
                    if (at + 17 > seglen) return -1;
                    const int tc = seg[at] >> 4, th = seg[at] & 15;
                    if (tc > 1 || th > 3) return -1;
                    int total = 0;
                    for (int i = 0; i < 16; ++i) total += seg[at + 1 + i];
                    if (total > 256 || at + 17 + total > seglen) return -1;
                    (tc ? ac[th] : dc[th]).build(seg + at + 1, seg + at + 17, total);
                    at += 17 + total;
                }
            } break;
            case 0xC0: case 0xC1: {  // SOF0 baseline, SOF1 extended sequential (Huffman)
                if (seglen < 6 || have_frame) return -1;
                if (seg[0] != 8) return -2;  // 12-bit
                height = (uint32_t)((seg[1] << 8) | seg[2]);
                width = (uint32_t)((seg[3] << 8) | seg[4]);
                const int nc = seg[5];
                if (width == 0 || height == 0 || seglen < 6 + 3 * nc) return -1;
                if (nc != 1 && nc != 3) return -2;  // CMYK / YCCK
                comps.resize((size_t)nc);
                for (int i = 0; i < nc; ++i) {
                    Component& c = comps[(size_t)i];
                    c.id = seg[6 + 3 * i]; c.h = seg[7 + 3 * i] >> 4; c.v = seg[7 + 3 * i] & 15; c.tq = seg[8 + 3 * i];
                    if (c.h < 1 || c.h > 4 || c.v < 1 || c.v > 4 || c.tq > 3) return -1;
                    hmax = std::max(hmax, c.h); vmax = std::max(vmax, c.v);
                }
                have_frame = true;
            } break;
            case 0xC2: case 0xC3: case 0xC5: case 0xC6: case 0xC7: case 0xC9: case 0xCA: case 0xCB: case 0xCD: case 0xCE: case 0xCF:
                return -2;  // progressive, lossless, differential, arithmetic
            case 0xDD: if (seglen < 2) return -1; restart_interval = (seg[0] << 8) | seg[1]; break;
            case 0xEE: if (seglen >= 12 && !std::memcmp(seg, "Adobe", 5)) { adobe = true; adobe_transform = seg[11]; } break;
            case 0xDA: {  // SOS: the one scan of a sequential frame (all components interleaved, or a single component)
                if (!have_frame || seglen < 1) return -1;
                const int ns = seg[0];
                if (ns != (int)comps.size() || seglen < 1 + 2 * ns + 3) return ns < (int)comps.size() ? -2 : -1;  // multi-scan sequential files: unsupported
                for (int i = 0; i < ns; ++i) {
                    Component* c = nullptr;
                    for (auto& cc : comps) if (cc.id == seg[1 + 2 * i]) c = &cc;
                    if (!c) return -1;
                    c->td = seg[2 + 2 * i] >> 4; c->ta = seg[2 + 2 * i] & 15;
                    if (c->td > 3 || c->ta > 3 || !dc[c->td].present || !ac[c->ta].present || !qt_present[c->tq]) return -1;
                }
                if (comps.size() == 1) { comps[0].h = comps[0].v = 1; hmax = vmax = 1; }  // a single-component scan is never interleaved
                const int mcux = ((int)width + 8 * hmax - 1) / (8 * hmax), mcuy = ((int)height + 8 * vmax - 1) / (8 * vmax);
                for (auto& c : comps) { c.bw = mcux * c.h; c.bh = mcuy * c.v; c.px.assign((size_t)c.bw * 8 * c.bh * 8, 0); c.pred = 0; }
                BitReader br{data + pos + (size_t)len, data + n};
                int until_restart = restart_interval, next_rst = 0;
                int32_t block[64];
                for (int my = 0; my < mcuy; ++my)
                    for (int mx = 0; mx < mcux; ++mx) {
                        if (restart_interval && until_restart == 0) {
                            // byte-align, expect RSTn
                            br.reset();
                            const uint8_t* q = br.p;
                            while (q + 1 < br.end && !(q[0] == 0xFF && q[1] >= 0xD0 && q[1] <= 0xD7)) ++q;
                            if (q + 1 >= br.end || q[1] != 0xD0 + next_rst) return -1;
                            br.p = q + 2;
                            next_rst = (next_rst + 1) & 7;
                            until_restart = restart_interval;
                            for (auto& c : comps) c.pred = 0;
                        }
                        for (auto& c : comps)
                            for (int by = 0; by < c.v; ++by)
                                for (int bx = 0; bx < c.h; ++bx) {
                                    std::memset(block, 0, sizeof(block));
                                    const int t = br.decode(dc[c.td]);
                                    if (t < 0 || t > 11) return -1;
                                    c.pred += t ? extend(br.bits(t), t) : 0;
                                    block[0] = c.pred * qt[c.tq][0];
                                    for (int k = 1; k < 64;) {
                                        const int rs = br.decode(ac[c.ta]);
                                        if (rs < 0) return -1;
                                        const int r = rs >> 4, s = rs & 15;
                                        if (s == 0) { if (r == 15) { k += 16; continue; } break; }  // ZRL / EOB
                                        k += r;
                                        if (k > 63) return -1;
                                        block[kZigzag[k]] = extend(br.bits(s), s) * qt[c.tq][kZigzag[k]];
                                        ++k;
                                    }
                                    const int px = (mx * c.h + bx) * 8, py = (my * c.v + by) * 8;
                                    idct_islow(block, c.px.data() + (size_t)py * (c.bw * 8) + px, c.bw * 8);
                                }
                        if (restart_interval) --until_restart;
                        if (br.exhausted) return -1;
                    }
                decoded = true;
            } break;
            default: break;  // APPn, COM, ...
        }
        pos += (size_t)len;
    }
    if (!decoded) return -1;

    // upsample every component to the luma grid (padded planes), then crop + colour-convert
    const int W = (int)width, H = (int)height;
    const int fullw = comps[0].bw * 8 * hmax / comps[0].h, fullh = comps[0].bh * 8 * vmax / comps[0].v;
    std::vector<std::vector<uint8_t>> full(comps.size());
    for (size_t ci = 0; ci < comps.size(); ++ci) {
        const Component& c = comps[ci];
        const int cw = c.bw * 8, ch = c.bh * 8, fx = hmax / c.h, fy = vmax / c.v;
        if (hmax % c.h || vmax % c.v) return -2;  // fractional ratios
        std::vector<uint8_t>& o = full[ci];
        if (fx == 1 && fy == 1) { o = c.px; continue; }
        o.assign((size_t)fullw * fullh, 0);
        if (fx == 2 && fy == 1) {
            for (int y = 0; y < ch; ++y) upsample_h2v1(c.px.data() + (size_t)y * cw, cw, o.data() + (size_t)y * fullw);
        } else if (fx == 2 && fy == 2) {
            for (int y = 0; y < ch; ++y) {
                const uint8_t* row = c.px.data() + (size_t)y * cw;
                const uint8_t* above = c.px.data() + (size_t)std::max(y - 1, 0) * cw;
                const uint8_t* below = c.px.data() + (size_t)std::min(y + 1, ch - 1) * cw;
                upsample_h2v2_row(row, above, cw, o.data() + (size_t)(2 * y) * fullw);
                upsample_h2v2_row(row, below, cw, o.data() + (size_t)(2 * y + 1) * fullw);
            }
        } else {  // replication (libjpeg's int_upsample)
            for (int y = 0; y < fullh; ++y)
                for (int x = 0; x < fullw; ++x) o[(size_t)y * fullw + x] = c.px[(size_t)(y / fy) * cw + x / fx];
        }
    }
    rgb.assign((size_t)W * H * 3, 0);
    const bool ycc = comps.size() == 3 && !(adobe && adobe_transform == 0);
    auto clamp8 = [](int v) { return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); };
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            uint8_t* o = rgb.data() + ((size_t)y * W + x) * 3;
            const size_t at = (size_t)y * fullw + x;
            if (comps.size() == 1) { o[0] = o[1] = o[2] = full[0][at]; continue; }
            const int Y = full[0][at], cb = full[1][at] - 128, cr = full[2][at] - 128;
            if (!ycc) { o[0] = (uint8_t)Y; o[1] = full[1][at]; o[2] = full[2][at]; continue; }
            // jdcolor.c: 16-bit fixed point, FIX(x) = (int)(x * 65536 + 0.5), arithmetic right shifts
            const int r = Y + ((91881 * cr + 32768) >> 16);
            const int g = Y + ((-22554 * cb - 46802 * cr + 32768) >> 16);
            const int b = Y + ((116130 * cb + 32768) >> 16);
            o[0] = clamp8(r); o[1] = clamp8(g); o[2] = clamp8(b);
        }
    return 0;
}

}  // namespace jpeg
}  // namespace rth
