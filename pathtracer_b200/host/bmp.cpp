// .BMP output in the reference's exact byte layout (WriteDIBImage, src/win32_main.cpp:358-391;
// bitmap_header_t, include/ray.hpp:5-28): a 58-byte packed header -- the 54-byte BITMAPFILEHEADER +
// BITMAPINFOHEADER followed by 4 zero bytes the reference's struct carries -- BitmapOffset 58,
// info size 40, 32 bpp, positive height (bottom-up), then width*height raw 0xAARRGGBB words.
#include "host_internal.h"

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace {
#pragma pack(push, 1)
struct BmpHeader {
    uint16_t file_type;
    uint32_t file_size;
    uint16_t reserved1, reserved2;
    uint32_t bitmap_offset;
    uint32_t info_size;
    int32_t width, height;
    uint16_t planes, bits_per_pixel;
    uint32_t compression, size_image;
    int32_t x_ppm, y_ppm;
    uint32_t clr_used, clr_important;
    uint8_t trailing[4];
};
#pragma pack(pop)
static_assert(sizeof(BmpHeader) == 58, "reference header is 58 bytes");
}  // namespace

extern "C" int pt_host_write_bmp(const char *path, const uint32_t *bgra, int width, int height) {
    if (!path || !bgra || width <= 0 || height <= 0) { pth::set_error("pt_host_write_bmp: bad argument"); return -1; }
    BmpHeader h;
    std::memset(&h, 0, sizeof(h));
    uint32_t pixel_bytes = (uint32_t)width * (uint32_t)height * 4u;
    h.file_type = 0x4D42;
    h.file_size = (uint32_t)sizeof(h) + pixel_bytes;
    h.bitmap_offset = (uint32_t)sizeof(h);
    h.info_size = 40;
    h.width = width;
    h.height = height;
    h.planes = 1;
    h.bits_per_pixel = 32;
    FILE *f = std::fopen(path, "wb");
    if (!f) { pth::set_error(std::string("[ERROR] Unable to write output file,fopen error: ") + path); return -1; }
    bool ok = std::fwrite(&h, sizeof(h), 1, f) == 1 && std::fwrite(bgra, pixel_bytes, 1, f) == 1;
    std::fclose(f);
    if (!ok) { pth::set_error("short write"); return -1; }
    return 0;
}

extern "C" int pt_host_read_bmp(const char *path, uint32_t **bgra_out, int *width, int *height) {
    FILE *f = std::fopen(path, "rb");
    if (!f) { pth::set_error(std::string("cannot open ") + path); return -1; }
    uint8_t hdr[54];
    if (std::fread(hdr, 1, 54, f) != 54 || hdr[0] != 'B' || hdr[1] != 'M') { std::fclose(f); pth::set_error("not a BMP"); return -1; }
    uint32_t off; int32_t w, h; uint16_t bpp;
    std::memcpy(&off, hdr + 10, 4); std::memcpy(&w, hdr + 18, 4); std::memcpy(&h, hdr + 22, 4); std::memcpy(&bpp, hdr + 28, 2);
    if (bpp != 32 || w <= 0 || h == 0 || h == INT32_MIN) { std::fclose(f); pth::set_error("only 32 bpp BMPs are supported"); return -1; }
    int ah = h < 0 ? -h : h;
    // the header's sizes are untrusted: the pixel array must fit between `off` and the end of the file before anything is allocated
    std::fseek(f, 0, SEEK_END);
    const long file_size = std::ftell(f);
    const uint64_t need = (uint64_t)w * (uint64_t)ah * 4u;
    if (file_size < 54 || off < 54 || (uint64_t)off > (uint64_t)file_size || need > (uint64_t)file_size - off) {
        std::fclose(f); pth::set_error("truncated BMP"); return -1;
    }
    uint32_t *px = (uint32_t *)std::malloc((size_t)need);
    if (!px) { std::fclose(f); pth::set_error("out of memory"); return -1; }
    std::fseek(f, (long)off, SEEK_SET);
    size_t got = std::fread(px, 4, (size_t)w * ah, f);
    std::fclose(f);
    if (got != (size_t)w * ah) { std::free(px); pth::set_error("truncated BMP"); return -1; }
    if (h < 0) {   // top-down -> bottom-up
        for (int y = 0; y < ah / 2; y++)
            for (int x = 0; x < w; x++) { uint32_t t = px[(size_t)y * w + x]; px[(size_t)y * w + x] = px[(size_t)(ah - 1 - y) * w + x]; px[(size_t)(ah - 1 - y) * w + x] = t; }
    }
    *bgra_out = px; *width = w; *height = ah;
    return 0;
}
