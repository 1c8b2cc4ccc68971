// Minimal PNG reader (zlib inflate + the five scanline filters) producing RGBA8, i.e. what the
// reference gets from stbi_load(path,&x,&y,&n,4) (src/win32_main.cpp:1728).  Non-interlaced images,
// bit depth 8 or 16, colour types 0/2/3/4/6.  Written from the PNG specification; no stb code.
#include "host_internal.h"

#include <zlib.h>

#include <cstdio>
#include <exception>
#include <cstdlib>
#include <cstring>

namespace pth {

static uint32_t be32(const uint8_t *p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

static int paeth(int a, int b, int c) {
    int p = a + b - c;
    int pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
    if (pa <= pb && pa <= pc) return a;
    if (pb <= pc) return b;
    return c;
}

bool load_png_rgba8(const char *path, std::vector<uint8_t> &rgba, int &w, int &h) {
    FILE *f = std::fopen(path, "rb");
    if (!f) { set_error(std::string("cannot open ") + path); return false; }
    std::vector<uint8_t> file;
    std::fseek(f, 0, SEEK_END);
    long sz = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    file.resize((size_t)sz);
    size_t got = std::fread(file.data(), 1, (size_t)sz, f);
    std::fclose(f);
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    if (got != (size_t)sz || sz < 33 || std::memcmp(file.data(), sig, 8) != 0) { set_error("not a PNG"); return false; }

    uint32_t width = 0, height = 0;
    int depth = 0, ctype = 0, interlace = 0;
    std::vector<uint8_t> idat, plte, trns;
    size_t pos = 8;
    bool seen_ihdr = false, seen_iend = false;
    while (pos + 12 <= file.size() && !seen_iend) {
        uint32_t len = be32(&file[pos]);
        const uint8_t *type = &file[pos + 4];
        const uint8_t *data = &file[pos + 8];
        if (pos + 12 + (size_t)len > file.size()) { set_error("truncated PNG chunk"); return false; }
        if (!std::memcmp(type, "IHDR", 4) && len >= 13) {
            width = be32(data); height = be32(data + 4);
            depth = data[8]; ctype = data[9]; interlace = data[12];
            seen_ihdr = true;
        } else if (!std::memcmp(type, "PLTE", 4)) {
            plte.assign(data, data + len);
        } else if (!std::memcmp(type, "tRNS", 4)) {
            trns.assign(data, data + len);
        } else if (!std::memcmp(type, "IDAT", 4)) {
            idat.insert(idat.end(), data, data + len);
        } else if (!std::memcmp(type, "IEND", 4)) {
            seen_iend = true;
        }
        pos += 12 + (size_t)len;
    }
    if (!seen_ihdr || width == 0 || height == 0) { set_error("PNG without IHDR"); return false; }
    if (interlace != 0) { set_error("interlaced PNG not supported"); return false; }
    if (!(depth == 8 || depth == 16) && !(ctype == 3 && (depth == 1 || depth == 2 || depth == 4 || depth == 8))) {
        set_error("unsupported PNG bit depth"); return false;
    }
    int channels;
    switch (ctype) {
        case 0: channels = 1; break;
        case 2: channels = 3; break;
        case 3: channels = 1; break;
        case 4: channels = 2; break;
        case 6: channels = 4; break;
        default: set_error("unsupported PNG colour type"); return false;
    }
    size_t bits_pp = (size_t)channels * depth;
    size_t bpp = (bits_pp + 7) / 8;                    // filter unit
    // IHDR is untrusted: bound the image by what the compressed data can possibly hold before anything is allocated
    // (deflate expands by at most 1032 : 1), and keep every size computation far from wrapping
    if (width > (1u << 16) || height > (1u << 16)) { set_error("PNG larger than 65536 pixels on a side"); return false; }
    size_t stride = ((size_t)width * bits_pp + 7) / 8;
    if ((stride + 1) * (size_t)height > idat.size() * 1032 + 1024) { set_error("PNG image data shorter than its header claims"); return false; }
    std::vector<uint8_t> raw((stride + 1) * (size_t)height);
    uLongf out_len = (uLongf)raw.size();
    int zr = uncompress(raw.data(), &out_len, idat.data(), (uLong)idat.size());
    if (zr != Z_OK || out_len != raw.size()) { set_error("PNG inflate failed"); return false; }

    // undo the scanline filters in place
    std::vector<uint8_t> prev(stride, 0);
    for (uint32_t y = 0; y < height; y++) {
        uint8_t *row = &raw[(stride + 1) * (size_t)y];
        int ft = row[0];
        uint8_t *cur = row + 1;
        for (size_t i = 0; i < stride; i++) {
            int a = i >= bpp ? cur[i - bpp] : 0;
            int b = prev[i];
            int c = i >= bpp ? prev[i - bpp] : 0;
            int v = cur[i];
            switch (ft) {
                case 0: break;
                case 1: v += a; break;
                case 2: v += b; break;
                case 3: v += (a + b) >> 1; break;
                case 4: v += paeth(a, b, c); break;
                default: set_error("bad PNG filter"); return false;
            }
            cur[i] = (uint8_t)v;
        }
        std::memcpy(prev.data(), cur, stride);
    }

    rgba.assign((size_t)width * height * 4, 255);
    for (uint32_t y = 0; y < height; y++) {
        const uint8_t *cur = &raw[(stride + 1) * (size_t)y + 1];
        uint8_t *dst = &rgba[(size_t)y * width * 4];
        for (uint32_t x = 0; x < width; x++) {
            uint8_t s[4] = {0, 0, 0, 255};
            if (ctype == 3) {
                int idx;
                if (depth == 8) idx = cur[x];
                else { int per = 8 / depth; int sh = (per - 1 - (int)(x % per)) * depth; idx = (cur[x / per] >> sh) & ((1 << depth) - 1); }
                if ((size_t)idx * 3 + 2 < plte.size()) { s[0] = plte[idx * 3]; s[1] = plte[idx * 3 + 1]; s[2] = plte[idx * 3 + 2]; }
                if ((size_t)idx < trns.size()) s[3] = trns[idx];
            } else {
                size_t step = depth / 8;                       // 16-bit samples: keep the high byte
                const uint8_t *p = cur + (size_t)x * channels * step;
                switch (ctype) {
                    case 0: s[0] = s[1] = s[2] = p[0]; break;
                    case 2: s[0] = p[0]; s[1] = p[step]; s[2] = p[2 * step]; break;
                    case 4: s[0] = s[1] = s[2] = p[0]; s[3] = p[step]; break;
                    case 6: s[0] = p[0]; s[1] = p[step]; s[2] = p[2 * step]; s[3] = p[3 * step]; break;
                }
            }
            dst[4 * x] = s[0]; dst[4 * x + 1] = s[1]; dst[4 * x + 2] = s[2]; dst[4 * x + 3] = s[3];
        }
    }
    w = (int)width; h = (int)height;
    return true;
}

}  // namespace pth

extern "C" int pt_host_load_png_rgba8(const char *path, uint8_t **rgba_out, int *width, int *height) {
    try {
        std::vector<uint8_t> px; int w, h;
        if (!pth::load_png_rgba8(path, px, w, h)) return -1;
        uint8_t *o = (uint8_t *)std::malloc(px.size());
        if (!o) { pth::set_error("out of memory"); return -1; }
        std::memcpy(o, px.data(), px.size());
        *rgba_out = o; *width = w; *height = h;
        return 0;
    } catch (const std::exception &e) {          // no exception crosses the C ABI
        pth::set_error(std::string("PNG: ") + e.what());
        return -1;
    }
}
