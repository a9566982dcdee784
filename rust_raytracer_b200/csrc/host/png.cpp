// png.cpp — image file I/O for the front-end. The reference goes through the `image`
// crate (buffer.rs:30-48 `decode()?.into_rgb32f()`, output.rs:23-39 `img.save`). Here PNG
// is decoded/encoded with zlib directly; other formats (JPEG) go to the registered
// decoder callback. Decoded texels are kept as integers (RGB8 / RGB16); the /255 or
// /65535 conversion of into_rgb32f happens where the texel is used.
#include <zlib.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "host_scene.h"

namespace rthost {

int decode_via_callback(const std::string& path, ImageData& out, std::string& err);
bool image_decoder_registered();
int decode_jpeg(const std::string& path, ImageData& out, std::string& err); // jpeg.cpp

static uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

static int paeth(int a, int b, int c) {
    int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

static int decode_png(const std::string& path, ImageData& out, std::string& err) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) { err = "cannot open image " + path; return RTGPU_ERR_IO; }
    std::vector<uint8_t> file;
    uint8_t buf[65536];
    size_t n;
    while ((n = std::fread(buf, 1, sizeof buf, f)) > 0) file.insert(file.end(), buf, buf + n);
    std::fclose(f);
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if (file.size() < 8 || std::memcmp(file.data(), sig, 8) != 0) { err = "not a PNG: " + path; return RTGPU_ERR_PARSE; }
    uint32_t W = 0, H = 0; int depth = 0, ctype = 0, interlace = 0;
    std::vector<uint8_t> idat, palette;
    size_t pos = 8;
    while (pos + 12 <= file.size()) {
        uint32_t len = be32(&file[pos]);
        const uint8_t* type = &file[pos + 4];
        const uint8_t* data = &file[pos + 8];
        if (pos + 12 + len > file.size()) break;
        if (!std::memcmp(type, "IHDR", 4)) { W = be32(data); H = be32(data + 4); depth = data[8]; ctype = data[9]; interlace = data[12]; }
        else if (!std::memcmp(type, "PLTE", 4)) palette.assign(data, data + len);
        else if (!std::memcmp(type, "IDAT", 4)) idat.insert(idat.end(), data, data + len);
        else if (!std::memcmp(type, "IEND", 4)) break;
        pos += 12 + len;
    }
    if (W == 0 || H == 0 || interlace != 0 || (depth != 8 && depth != 16) || (ctype == 3 && depth != 8)) {
        err = "unsupported PNG layout (need non-interlaced 8/16-bit): " + path;
        return RTGPU_ERR_UNSUPPORTED;
    }
    int channels = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
    if (!channels) { err = "bad PNG colour type: " + path; return RTGPU_ERR_PARSE; }
    size_t bpp = (size_t)channels * depth / 8, stride = bpp * W;
    std::vector<uint8_t> raw((stride + 1) * H);
    uLongf raw_len = (uLongf)raw.size();
    if (uncompress(raw.data(), &raw_len, idat.data(), (uLong)idat.size()) != Z_OK || raw_len != raw.size()) {
        err = "PNG inflate failed: " + path; return RTGPU_ERR_PARSE;
    }
    std::vector<uint8_t> pix(stride * H);
    for (uint32_t y = 0; y < H; y++) {
        const uint8_t* src = &raw[(stride + 1) * y];
        uint8_t* dst = &pix[stride * y];
        const uint8_t* up = y ? &pix[stride * (y - 1)] : nullptr;
        int filter = src[0];
        for (size_t x = 0; x < stride; x++) {
            int a = x >= bpp ? dst[x - bpp] : 0, b = up ? up[x] : 0, c = (up && x >= bpp) ? up[x - bpp] : 0, v = src[x + 1];
            switch (filter) {
            case 0: break;
            case 1: v += a; break;
            case 2: v += b; break;
            case 3: v += (a + b) / 2; break;
            case 4: v += paeth(a, b, c); break;
            default: err = "bad PNG filter: " + path; return RTGPU_ERR_PARSE;
            }
            dst[x] = (uint8_t)v;
        }
    }
    out.width = (int)W; out.height = (int)H;
    size_t npx = (size_t)W * H;
    if (depth == 8) {
        out.format = RTGPU_IMG_RGB8;
        out.bytes.resize(npx * 3);
        for (size_t i = 0; i < npx; i++) {
            const uint8_t* p = &pix[i * bpp];
            uint8_t r, g, b;
            if (ctype == 3) {
                size_t k = (size_t)p[0] * 3;
                if (k + 3 > palette.size()) r = g = b = 0;
                else { r = palette[k]; g = palette[k + 1]; b = palette[k + 2]; }
            }
            else if (channels <= 2) r = g = b = p[0];
            else { r = p[0]; g = p[1]; b = p[2]; }
            out.bytes[i * 3] = r; out.bytes[i * 3 + 1] = g; out.bytes[i * 3 + 2] = b;
        }
    } else {
        out.format = RTGPU_IMG_RGB16;
        out.bytes.resize(npx * 6);
        uint16_t* o = (uint16_t*)out.bytes.data();
        for (size_t i = 0; i < npx; i++) {
            const uint8_t* p = &pix[i * bpp];
            auto rd = [&](int ch) { return (uint16_t)((p[ch * 2] << 8) | p[ch * 2 + 1]); };
            if (channels <= 2) o[i * 3] = o[i * 3 + 1] = o[i * 3 + 2] = rd(0);
            else { o[i * 3] = rd(0); o[i * 3 + 1] = rd(1); o[i * 3 + 2] = rd(2); }
        }
    }
    return 0;
}

int decode_image(const std::string& path, ImageData& out, std::string& err) {
    size_t dot = path.rfind('.');
    std::string ext = dot == std::string::npos ? "" : path.substr(dot + 1);
    for (char& c : ext) c = (char)std::tolower((unsigned char)c);
    if (ext == "png") return decode_png(path, out, err);
    { FILE* f = std::fopen(path.c_str(), "rb"); if (!f) { err = "cannot open image " + path; return RTGPU_ERR_IO; } std::fclose(f); }
    // a registered plug-in keeps every other format (the Python front-end registers PIL); without one, JPEG is decoded
    // here (jpeg.cpp: sequential Huffman JPEG — every .jpg of the reference tree), and what the built-in decoder
    // refuses (progressive, CMYK ...) is reported with its message
    if (image_decoder_registered()) return decode_via_callback(path, out, err);
    if (ext == "jpg" || ext == "jpeg") return decode_jpeg(path, out, err);
    return decode_via_callback(path, out, err);
}

// the built-in decoders alone (PNG, sequential JPEG), whatever plug-in is registered
int decode_image_builtin(const std::string& path, ImageData& out, std::string& err) {
    size_t dot = path.rfind('.');
    std::string ext = dot == std::string::npos ? "" : path.substr(dot + 1);
    for (char& c : ext) c = (char)std::tolower((unsigned char)c);
    if (ext == "png") return decode_png(path, out, err);
    if (ext == "jpg" || ext == "jpeg") return decode_jpeg(path, out, err);
    err = "no built-in decoder for " + path;
    return RTGPU_ERR_UNSUPPORTED;
}

} // namespace rthost

using namespace rthost;

// PNG writer: 8-bit RGB, filter 0, zlib default compression (output.rs:23-39 saves RGB8).
extern "C" int rtgpu_save_png(const char* path, const uint8_t* rgb, int width, int height) {
    if (!path || !rgb || width <= 0 || height <= 0) { set_error("rtgpu_save_png: bad argument"); return RTGPU_ERR_INVALID; }
    size_t stride = (size_t)width * 3;
    std::vector<uint8_t> raw((stride + 1) * height);
    for (int y = 0; y < height; y++) {
        raw[(stride + 1) * y] = 0;
        std::memcpy(&raw[(stride + 1) * y + 1], rgb + stride * y, stride);
    }
    uLongf clen = compressBound((uLong)raw.size());
    std::vector<uint8_t> comp(clen);
    if (compress2(comp.data(), &clen, raw.data(), (uLong)raw.size(), 6) != Z_OK) { set_error("PNG deflate failed"); return RTGPU_ERR_IO; }
    FILE* f = std::fopen(path, "wb");
    if (!f) { set_error(std::string("cannot write ") + path); return RTGPU_ERR_IO; }
    auto chunk = [&](const char* type, const uint8_t* data, uint32_t len) {
        uint8_t hdr[8] = {(uint8_t)(len >> 24), (uint8_t)(len >> 16), (uint8_t)(len >> 8), (uint8_t)len,
                          (uint8_t)type[0], (uint8_t)type[1], (uint8_t)type[2], (uint8_t)type[3]};
        std::fwrite(hdr, 1, 8, f);
        if (len) std::fwrite(data, 1, len, f);
        uLong crc = crc32(0L, hdr + 4, 4);
        if (len) crc = crc32(crc, data, len);
        uint8_t c[4] = {(uint8_t)(crc >> 24), (uint8_t)(crc >> 16), (uint8_t)(crc >> 8), (uint8_t)crc};
        std::fwrite(c, 1, 4, f);
    };
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    std::fwrite(sig, 1, 8, f);
    uint8_t ihdr[13] = {(uint8_t)(width >> 24), (uint8_t)(width >> 16), (uint8_t)(width >> 8), (uint8_t)width,
                        (uint8_t)(height >> 24), (uint8_t)(height >> 16), (uint8_t)(height >> 8), (uint8_t)height, 8, 2, 0, 0, 0};
    chunk("IHDR", ihdr, 13);
    chunk("IDAT", comp.data(), (uint32_t)clen);
    chunk("IEND", nullptr, 0);
    std::fclose(f);
    return 0;
}
