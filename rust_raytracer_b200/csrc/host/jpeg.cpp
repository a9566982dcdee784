// jpeg.cpp — built-in JPEG decoder of the front-end, so that `image <file>.jpg` textures (texture/image.rs:21-33,
// buffer.rs:30-48: `image::open(path)?.decode()?.into_rgb32f()`) load without a host-language plug-in.
//
// The reference decodes through the `image 0.25.1` crate (a third-party dependency absent from /root/reference; parity
// unpinned — SURVEY.md 8c). What is restated here is the published algorithm itself, ITU-T T.81 sequential DCT mode with
// Huffman coding ("baseline" SOF0 and 8-bit "extended" SOF1): DQT / DHT / DRI / SOF / SOS parsing, entropy decoding with
// byte stuffing and restart markers, interleaved and non-interleaved scans, dequantisation, an 8x8 inverse DCT, chroma
// upsampling (replication, or the triangle filter most decoders apply to 2x subsampled chroma) and JFIF YCbCr -> RGB.
// Every JPEG the reference ships (scenes/resource/*.jpg) is SOF0 4:4:4 or greyscale, some with restart intervals.
// Progressive, arithmetic-coded, lossless, 12-bit and CMYK files are refused with a message (the decoder callback, when
// one is registered, still gets them). Output: RGB8, top row first — what into_rgb32f sees before its /255.
#include <cmath>
#include <cstdio>
#include <cstring>

#include "host_scene.h"

namespace rthost {

namespace {

const uint8_t kZigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                             41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                             30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

struct Huffman { // T.81 annex C / F.2.2.3: canonical codes, decoded length by length
    bool present = false;
    uint8_t vals[256];
    int mincode[17], maxcode[18], valptr[17];
    void build(const uint8_t counts[16], const uint8_t* symbols) {
        int code = 0, k = 0;
        for (int len = 1; len <= 16; len++) {
            valptr[len] = k;
            mincode[len] = code;
            code += counts[len - 1];
            k += counts[len - 1];
            maxcode[len] = counts[len - 1] ? code - 1 : -1;
            code <<= 1;
        }
        maxcode[17] = 0x7FFFFFFF;
        std::memcpy(vals, symbols, (size_t)k);
        present = true;
    }
};

struct BitReader { // entropy-coded segment: 0xFF00 is a stuffed 0xFF, any other 0xFFxx ends the segment (zeros are fed)
    const uint8_t* p = nullptr;
    size_t n = 0, pos = 0;
    uint32_t acc = 0;
    int cnt = 0;
    bool at_marker = false;
    void fill() {
        while (cnt <= 24) {
            uint32_t b = 0;
            if (!at_marker && pos < n) {
                b = p[pos];
                if (b == 0xFF) {
                    const int b2 = pos + 1 < n ? p[pos + 1] : 0xD9;
                    if (b2 == 0) pos += 2;
                    else { at_marker = true; b = 0; }
                } else pos++;
            }
            acc |= b << (24 - cnt);
            cnt += 8;
        }
    }
    int bit() {
        if (cnt < 1) fill();
        const int v = (int)(acc >> 31);
        acc <<= 1; cnt--;
        return v;
    }
    int bits(int nb) {
        if (nb == 0) return 0;
        if (cnt < nb) fill();
        const int v = (int)(acc >> (32 - nb));
        acc <<= nb; cnt -= nb;
        return v;
    }
    void reset() { acc = 0; cnt = 0; at_marker = false; }
};

struct Component {
    int id = 0, h = 1, v = 1, tq = 0, td = 0, ta = 0;
    int pred = 0;
    int blocks_w = 0, blocks_h = 0; // blocks held (padded to whole MCUs)
    int width = 0, height = 0;      // samples that carry image data: ceil(image * h / hmax)
    std::vector<uint8_t> plane;     // blocks_w * 8 x blocks_h * 8
};

struct Decoder {
    const uint8_t* p;
    size_t n;
    std::string& err;
    uint16_t qt[4][64];
    bool qt_present[4] = {false, false, false, false};
    Huffman dc[4], ac[4];
    Component comp[4];
    int n_comp = 0, width = 0, height = 0, hmax = 1, vmax = 1, restart_interval = 0;
    int adobe_transform = -1;
    bool have_frame = false;
    float idct_c[8][8];

    Decoder(const uint8_t* data, size_t size, std::string& e) : p(data), n(size), err(e) {
        for (int u = 0; u < 8; u++)
            for (int x = 0; x < 8; x++)
                idct_c[u][x] = (float)((u == 0 ? std::sqrt(0.125) : 0.5) * std::cos((2 * x + 1) * u * 3.14159265358979323846 / 16.0));
    }
    bool fail(const std::string& m) { err = "jpeg: " + m; return false; }

    int decode_symbol(BitReader& br, const Huffman& h) {
        int code = 0;
        for (int len = 1; len <= 16; len++) {
            code = (code << 1) | br.bit();
            if (h.maxcode[len] >= 0 && code <= h.maxcode[len] && code >= h.mincode[len]) return h.vals[h.valptr[len] + code - h.mincode[len]];
        }
        return -1;
    }
    static int extend(int v, int t) { return t == 0 ? 0 : (v < (1 << (t - 1)) ? v - (1 << t) + 1 : v); } // T.81 F.2.2.1

    // one 8x8 block of component c into its plane at block (bx, by)
    bool decode_block(BitReader& br, Component& c, int bx, int by) {
        float coef[64];
        for (float& f : coef) f = 0.f;
        const uint16_t* q = qt[c.tq];
        const int t = decode_symbol(br, dc[c.td]);
        if (t < 0 || t > 11) return fail("bad DC code");
        c.pred += extend(br.bits(t), t);
        coef[0] = (float)(c.pred * (int)q[0]);
        for (int k = 1; k < 64;) {
            const int rs = decode_symbol(br, ac[c.ta]);
            if (rs < 0) return fail("bad AC code");
            const int r = rs >> 4, s = rs & 15;
            if (s == 0) {
                if (r != 15) break; // end of block
                k += 16;
                continue;
            }
            k += r;
            if (k > 63) return fail("AC run past the block");
            coef[kZigzag[k]] = (float)(extend(br.bits(s), s) * (int)q[k]);
            k++;
        }
        // separable inverse DCT (T.81 A.3.3), level shift, clamp
        float tmp[64];
        for (int y = 0; y < 8; y++)
            for (int x = 0; x < 8; x++) {
                float s = 0.f;
                for (int u = 0; u < 8; u++) s += idct_c[u][x] * coef[y * 8 + u];
                tmp[y * 8 + x] = s;
            }
        const int stride = c.blocks_w * 8;
        uint8_t* out = c.plane.data() + (size_t)by * 8 * stride + (size_t)bx * 8;
        for (int x = 0; x < 8; x++)
            for (int y = 0; y < 8; y++) {
                float s = 0.f;
                for (int v = 0; v < 8; v++) s += idct_c[v][y] * tmp[v * 8 + x];
                const int val = (int)std::lrintf(s) + 128;
                out[(size_t)y * stride + x] = (uint8_t)(val < 0 ? 0 : (val > 255 ? 255 : val));
            }
        return true;
    }

    bool read_frame(size_t at, size_t len) {
        if (len < 6) return fail("short SOF");
        if (p[at] != 8) return fail("only 8-bit samples are supported");
        height = (p[at + 1] << 8) | p[at + 2];
        width = (p[at + 3] << 8) | p[at + 4];
        n_comp = p[at + 5];
        if (width <= 0 || height <= 0) return fail("empty image");
        if (n_comp != 1 && n_comp != 3) return fail("only greyscale and 3-component images are supported");
        if (len < 6 + (size_t)n_comp * 3) return fail("short SOF");
        hmax = vmax = 1;
        for (int i = 0; i < n_comp; i++) {
            Component& c = comp[i];
            c.id = p[at + 6 + i * 3];
            c.h = p[at + 7 + i * 3] >> 4; c.v = p[at + 7 + i * 3] & 15;
            c.tq = p[at + 8 + i * 3];
            if (c.h < 1 || c.h > 4 || c.v < 1 || c.v > 4 || c.tq > 3) return fail("bad component parameters");
            hmax = std::max(hmax, c.h); vmax = std::max(vmax, c.v);
        }
        const int mcus_w = (width + 8 * hmax - 1) / (8 * hmax), mcus_h = (height + 8 * vmax - 1) / (8 * vmax);
        for (int i = 0; i < n_comp; i++) {
            Component& c = comp[i];
            c.blocks_w = mcus_w * c.h; c.blocks_h = mcus_h * c.v;
            c.width = (width * c.h + hmax - 1) / hmax; c.height = (height * c.v + vmax - 1) / vmax;
            c.plane.assign((size_t)c.blocks_w * 8 * (size_t)c.blocks_h * 8, 0);
        }
        have_frame = true;
        return true;
    }

    // one scan; `at` = first byte after the SOS header; returns the position of the marker that ends it
    bool read_scan(const int* scan_comp, int ns, size_t& at) {
        BitReader br;
        br.p = p; br.n = n; br.pos = at;
        for (int i = 0; i < ns; i++) {
            Component& c = comp[scan_comp[i]];
            c.pred = 0;
            if (!qt_present[c.tq]) return fail("scan uses an undefined quantisation table");
            if (!dc[c.td].present || !ac[c.ta].present) return fail("scan uses an undefined Huffman table");
        }
        int mcus_w, mcus_h;
        if (ns == 1) { // non-interleaved: the component's own blocks that carry data (T.81 A.2.2)
            const Component& c = comp[scan_comp[0]];
            mcus_w = (c.width + 7) / 8; mcus_h = (c.height + 7) / 8;
        } else {
            mcus_w = (width + 8 * hmax - 1) / (8 * hmax); mcus_h = (height + 8 * vmax - 1) / (8 * vmax);
        }
        int until_restart = restart_interval, next_rst = 0;
        for (int my = 0; my < mcus_h; my++)
            for (int mx = 0; mx < mcus_w; mx++) {
                if (restart_interval > 0 && until_restart == 0) {
                    // byte-align, expect RSTn
                    br.reset();
                    size_t q = br.pos;
                    while (q + 1 < n && !(p[q] == 0xFF && p[q + 1] >= 0xD0 && p[q + 1] <= 0xD7)) {
                        if (p[q] == 0xFF && p[q + 1] != 0 && p[q + 1] != 0xFF) return fail("restart marker missing");
                        q++;
                    }
                    if (q + 1 >= n) return fail("restart marker missing");
                    if ((p[q + 1] & 7) != next_rst) return fail("restart markers out of order");
                    next_rst = (next_rst + 1) & 7;
                    br.pos = q + 2;
                    for (int i = 0; i < ns; i++) comp[scan_comp[i]].pred = 0;
                    until_restart = restart_interval;
                }
                if (ns == 1) {
                    if (!decode_block(br, comp[scan_comp[0]], mx, my)) return false;
                } else
                    for (int i = 0; i < ns; i++) {
                        Component& c = comp[scan_comp[i]];
                        for (int v = 0; v < c.v; v++)
                            for (int h = 0; h < c.h; h++)
                                if (!decode_block(br, c, mx * c.h + h, my * c.v + v)) return false;
                    }
                until_restart--;
            }
        // the next marker
        size_t q = br.pos;
        while (q + 1 < n && !(p[q] == 0xFF && p[q + 1] != 0 && p[q + 1] != 0xFF && !(p[q + 1] >= 0xD0 && p[q + 1] <= 0xD7))) q++;
        at = q;
        return true;
    }

    bool run() {
        if (n < 4 || p[0] != 0xFF || p[1] != 0xD8) return fail("not a JPEG file");
        size_t at = 2;
        int scanned = 0; // bit per component
        while (at + 3 < n) {
            if (p[at] != 0xFF) { at++; continue; }
            const int m = p[at + 1];
            if (m == 0xFF) { at++; continue; }            // fill byte
            if (m == 0xD8 || m == 0x01 || (m >= 0xD0 && m <= 0xD7)) { at += 2; continue; }
            if (m == 0xD9) break;                          // EOI
            const size_t len = ((size_t)p[at + 2] << 8) | p[at + 3];
            if (len < 2 || at + 2 + len > n) return fail("truncated segment");
            const size_t body = at + 4, blen = len - 2;
            if (m == 0xC0 || m == 0xC1) { if (!read_frame(body, blen)) return false; }
            else if (m == 0xC2) return fail("progressive JPEG is not supported by the built-in decoder");
            else if ((m >= 0xC3 && m <= 0xCF) && m != 0xC4 && m != 0xC8 && m != 0xCC) return fail("lossless / hierarchical / arithmetic-coded JPEG is not supported");
            else if (m == 0xCC) return fail("arithmetic-coded JPEG is not supported");
            else if (m == 0xDB) { // DQT
                size_t q = body;
                while (q < body + blen) {
                    const int pq = p[q] >> 4, tq = p[q] & 15;
                    if (tq > 3 || pq > 1) return fail("bad DQT");
                    q++;
                    if (q + (size_t)(pq ? 128 : 64) > body + blen) return fail("short DQT");
                    for (int k = 0; k < 64; k++) { qt[tq][k] = pq ? (uint16_t)((p[q] << 8) | p[q + 1]) : p[q]; q += pq ? 2 : 1; }
                    qt_present[tq] = true;
                }
            } else if (m == 0xC4) { // DHT
                size_t q = body;
                while (q < body + blen) {
                    const int tc = p[q] >> 4, th = p[q] & 15;
                    if (tc > 1 || th > 3 || q + 17 > body + blen) return fail("bad DHT");
                    int total = 0;
                    for (int k = 0; k < 16; k++) total += p[q + 1 + k];
                    if (total > 256 || q + 17 + (size_t)total > body + blen) return fail("bad DHT");
                    (tc ? ac[th] : dc[th]).build(p + q + 1, p + q + 17);
                    q += 17 + (size_t)total;
                }
            } else if (m == 0xDD) { if (blen < 2) return fail("short DRI"); restart_interval = (p[body] << 8) | p[body + 1]; }
            else if (m == 0xEE && blen >= 12 && std::memcmp(p + body, "Adobe", 5) == 0) adobe_transform = p[body + 11];
            else if (m == 0xDA) { // SOS
                if (!have_frame) return fail("SOS before SOF");
                const int ns = p[body];
                if (ns < 1 || ns > n_comp || blen < 1 + (size_t)ns * 2 + 3) return fail("bad SOS");
                int scan_comp[4];
                for (int i = 0; i < ns; i++) {
                    const int cid = p[body + 1 + i * 2];
                    int ci = -1;
                    for (int k = 0; k < n_comp; k++) if (comp[k].id == cid) ci = k;
                    if (ci < 0) return fail("SOS names an unknown component");
                    comp[ci].td = p[body + 2 + i * 2] >> 4; comp[ci].ta = p[body + 2 + i * 2] & 15;
                    if (comp[ci].td > 3 || comp[ci].ta > 3) return fail("bad table selector");
                    scan_comp[i] = ci; scanned |= 1 << ci;
                }
                size_t pos = at + 2 + len;
                if (!read_scan(scan_comp, ns, pos)) return false;
                at = pos;
                continue;
            }
            at += 2 + len;
        }
        if (!have_frame) return fail("no frame header");
        if (scanned != (1 << n_comp) - 1) return fail("not every component was coded");
        return true;
    }

    // component plane -> full resolution (width x height)
    std::vector<uint8_t> upsampled(const Component& c) const {
        const int stride = c.blocks_w * 8;
        std::vector<uint8_t> out((size_t)width * height);
        const int fh = hmax / c.h, fv = vmax / c.v;
        if (fh * c.h != hmax || fv * c.v != vmax || (fh != 1 && fh != 2) || (fv != 1 && fv != 2)) { // odd ratios: replication
            for (int y = 0; y < height; y++)
                for (int x = 0; x < width; x++) out[(size_t)y * width + x] = c.plane[(size_t)(y * c.v / vmax) * stride + (size_t)(x * c.h / hmax)];
            return out;
        }
        // 2x: the triangle filter (3/4 nearer, 1/4 farther sample), vertically then horizontally, edges replicated
        const int cw = c.width, chh = c.height;
        std::vector<int> row((size_t)cw);
        for (int y = 0; y < height; y++) {
            const int sy = fv == 2 ? y / 2 : y;
            const uint8_t* r0 = c.plane.data() + (size_t)std::min(sy, chh - 1) * stride;
            int scale = 1;
            if (fv == 2) {
                const int ny = (y & 1) ? std::min(sy + 1, chh - 1) : std::max(sy - 1, 0);
                const uint8_t* r1 = c.plane.data() + (size_t)ny * stride;
                for (int x = 0; x < cw; x++) row[(size_t)x] = 3 * r0[x] + r1[x];
                scale = 4;
            } else for (int x = 0; x < cw; x++) row[(size_t)x] = r0[x];
            uint8_t* o = out.data() + (size_t)y * width;
            if (fh == 2) {
                for (int x = 0; x < width; x++) {
                    const int sx = std::min(x / 2, cw - 1);
                    const int nx = (x & 1) ? std::min(sx + 1, cw - 1) : std::max(sx - 1, 0);
                    const int s = 3 * row[(size_t)sx] + row[(size_t)nx];
                    o[x] = (uint8_t)((s + (scale == 4 ? ((x & 1) ? 7 : 8) : ((x & 1) ? 2 : 1))) / (4 * scale));
                }
            } else
                for (int x = 0; x < width; x++) o[x] = (uint8_t)(scale == 4 ? (row[(size_t)std::min(x, cw - 1)] + ((y & 1) ? 2 : 1)) / 4 : row[(size_t)std::min(x, cw - 1)]);
        }
        return out;
    }
};

} // namespace

int decode_jpeg(const std::string& path, ImageData& out, std::string& err) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) { err = "cannot open image " + path; return RTGPU_ERR_IO; }
    std::vector<uint8_t> file;
    uint8_t buf[65536];
    size_t got;
    while ((got = std::fread(buf, 1, sizeof buf, f)) > 0) file.insert(file.end(), buf, buf + got);
    std::fclose(f);
    Decoder d(file.data(), file.size(), err);
    if (!d.run()) { err += " (" + path + ")"; return RTGPU_ERR_UNSUPPORTED; }
    out.width = d.width; out.height = d.height; out.format = RTGPU_IMG_RGB8;
    out.bytes.resize((size_t)d.width * d.height * 3);
    const size_t npx = (size_t)d.width * d.height;
    if (d.n_comp == 1) {
        const std::vector<uint8_t> y = d.upsampled(d.comp[0]);
        for (size_t i = 0; i < npx; i++) out.bytes[i * 3] = out.bytes[i * 3 + 1] = out.bytes[i * 3 + 2] = y[i];
        return 0;
    }
    const std::vector<uint8_t> c0 = d.upsampled(d.comp[0]), c1 = d.upsampled(d.comp[1]), c2 = d.upsampled(d.comp[2]);
    const bool rgb = d.adobe_transform == 0 || (d.comp[0].id == 'R' && d.comp[1].id == 'G' && d.comp[2].id == 'B');
    auto clamp8 = [](int v) { return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); };
    for (size_t i = 0; i < npx; i++) {
        if (rgb) { out.bytes[i * 3] = c0[i]; out.bytes[i * 3 + 1] = c1[i]; out.bytes[i * 3 + 2] = c2[i]; continue; }
        // JFIF: R = Y + 1.402 Cr', G = Y - 0.344136 Cb' - 0.714136 Cr', B = Y + 1.772 Cb' (16-bit fixed point, rounded)
        const int y = c0[i], cb = (int)c1[i] - 128, cr = (int)c2[i] - 128;
        out.bytes[i * 3] = clamp8(y + (int)std::floor((91881.0 * cr + 32768.0) / 65536.0));
        out.bytes[i * 3 + 1] = clamp8(y + (int)std::floor((-22554.0 * cb - 46802.0 * cr + 32768.0) / 65536.0));
        out.bytes[i * 3 + 2] = clamp8(y + (int)std::floor((116130.0 * cb + 32768.0) / 65536.0));
    }
    return 0;
}

} // namespace rthost
