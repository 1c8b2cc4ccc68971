// TEST INFRASTRUCTURE: mutation fuzzer for the host library's file readers (PNG, BMP, GLB), built with
// -fsanitize=address,undefined against the host sources (tests/test_host.py::test_file_readers_survive_mutated_files).
// Every reader must either succeed or fail with its error code on ANY bytes: no out-of-bounds access, no undefined behaviour,
// no exception escaping the C ABI, no allocation beyond what the file's size justifies.
//   usage: fuzz_host <seed.glb> <work dir> <iterations>
#include <zlib.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <vector>

#include "pt_host.h"

static std::vector<uint8_t> read_all(const char *p) {
    std::vector<uint8_t> v;
    FILE *f = std::fopen(p, "rb");
    if (!f) return v;
    std::fseek(f, 0, SEEK_END);
    long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    v.resize((size_t)n);
    if (std::fread(v.data(), 1, v.size(), f) != v.size()) v.clear();
    std::fclose(f);
    return v;
}
static void write_all(const std::string &p, const std::vector<uint8_t> &v) {
    FILE *f = std::fopen(p.c_str(), "wb");
    std::fwrite(v.data(), 1, v.size(), f);
    std::fclose(f);
}
static void be32(std::vector<uint8_t> &o, uint32_t x) { for (int s = 24; s >= 0; s -= 8) o.push_back((uint8_t)(x >> s)); }
static void chunk(std::vector<uint8_t> &o, const char *type, const std::vector<uint8_t> &d) {
    be32(o, (uint32_t)d.size());
    size_t at = o.size();
    o.insert(o.end(), type, type + 4);
    o.insert(o.end(), d.begin(), d.end());
    be32(o, (uint32_t)crc32(0, &o[at], (uInt)(4 + d.size())));
}
// a small valid PNG: w x h, colour type ct, depth 8, every filter type in turn
static std::vector<uint8_t> make_png(int w, int h, int ct) {
    const int ch = ct == 0 ? 1 : ct == 2 ? 3 : ct == 3 ? 1 : ct == 4 ? 2 : 4;
    std::vector<uint8_t> raw;
    for (int y = 0; y < h; y++) {
        raw.push_back((uint8_t)(y % 5));
        for (int x = 0; x < w * ch; x++) raw.push_back((uint8_t)(x * 7 + y * 13));
    }
    uLongf zl = compressBound((uLong)raw.size());
    std::vector<uint8_t> z(zl);
    compress(z.data(), &zl, raw.data(), (uLong)raw.size());
    z.resize(zl);
    std::vector<uint8_t> o = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A}, ihdr;
    be32(ihdr, (uint32_t)w); be32(ihdr, (uint32_t)h);
    ihdr.push_back(8); ihdr.push_back((uint8_t)ct); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
    chunk(o, "IHDR", ihdr);
    if (ct == 3) { std::vector<uint8_t> pl(48); for (size_t i = 0; i < pl.size(); i++) pl[i] = (uint8_t)(i * 5); chunk(o, "PLTE", pl); chunk(o, "tRNS", {1, 2, 3}); }
    chunk(o, "IDAT", z);
    chunk(o, "IEND", {});
    return o;
}

static std::mt19937 g(12345);
static std::vector<uint8_t> mutate(std::vector<uint8_t> v) {
    if (v.empty()) return v;
    const int kind = (int)(g() % 6);
    const int n = 1 + (int)(g() % 4);
    for (int k = 0; k < n; k++) {
        size_t i = g() % v.size();
        switch (kind) {
            case 0: v[i] = (uint8_t)g(); break;                                             // random byte
            case 1: v[i] ^= (uint8_t)(1u << (g() % 8)); break;                              // bit flip
            case 2: { static const uint8_t m[] = {0, 1, 0x7f, 0x80, 0xff, 0xfe}; v[i] = m[g() % 6]; break; }
            case 3: if (i + 4 <= v.size()) { uint32_t x = (g() % 2) ? 0xffffffffu : (uint32_t)g(); std::memcpy(&v[i], &x, 4); } break;   // a 32-bit field
            case 4: v.resize(i); if (v.empty()) return v; break;                            // truncate
            case 5: { size_t j = g() % v.size(); std::swap(v[i], v[j]); break; }
        }
    }
    return v;
}
// GLB: mutate inside the JSON chunk more often than in the BIN chunk (that is where the indices and offsets live)
static std::vector<uint8_t> mutate_glb(const std::vector<uint8_t> &seed) {
    std::vector<uint8_t> v = seed;
    if (v.size() < 28) return mutate(v);
    uint32_t jl;
    std::memcpy(&jl, &v[12], 4);
    if (g() % 4 == 0 || 20 + (size_t)jl > v.size()) return mutate(v);
    static const char *digits[] = {"-1", "0", "9", "99999999", "4294967296", "-2147483649", "1e30", "[", "]", "{", "}", "\"", ":", ","};
    for (int k = 0; k < 1 + (int)(g() % 3); k++) {
        size_t i = 20 + g() % jl;
        if (g() % 2) {                                                                     // overwrite a number with an edge value
            while (i < 20 + jl && !(v[i] >= '0' && v[i] <= '9')) i++;
            const char *d = digits[g() % 14];
            for (size_t c = 0; d[c] && i + c < 20 + (size_t)jl; c++) v[i + c] = (uint8_t)d[c];
        } else v[i] = (uint8_t)(32 + g() % 95);
    }
    return v;
}

int main(int argc, char **argv) {
    if (argc < 4) { std::printf("usage: fuzz_host seed.glb workdir iterations\n"); return 2; }
    const std::string dir = argv[2];
    const int iters = std::atoi(argv[3]);
    const std::vector<uint8_t> glb = read_all(argv[1]);
    if (glb.empty()) { std::printf("cannot read %s\n", argv[1]); return 2; }
    std::vector<std::vector<uint8_t>> pngs;
    for (int ct : {0, 2, 3, 4, 6}) pngs.push_back(make_png(9, 7, ct));
    std::vector<uint32_t> px(11 * 5);
    for (size_t i = 0; i < px.size(); i++) px[i] = (uint32_t)(i * 2654435761u);
    const std::string bmp_path = dir + "/seed.bmp";
    if (pt_host_write_bmp(bmp_path.c_str(), px.data(), 11, 5)) return 2;
    const std::vector<uint8_t> bmp = read_all(bmp_path.c_str());
    long ok[3] = {0, 0, 0}, bad[3] = {0, 0, 0};
    const std::string f = dir + "/case.bin";
    // the seeds themselves must load
    for (auto &p : pngs) { write_all(f, p); uint8_t *o = nullptr; int w, h; if (pt_host_load_png_rgba8(f.c_str(), &o, &w, &h) || w != 9 || h != 7) { std::printf("seed PNG does not load\n"); return 1; } std::free(o); }
    for (int it = 0; it < iters; it++) {
        try {
            {   write_all(f, mutate(pngs[it % pngs.size()]));
                uint8_t *o = nullptr; int w = 0, h = 0;
                if (pt_host_load_png_rgba8(f.c_str(), &o, &w, &h) == 0) { ok[0]++; volatile uint8_t s = 0; for (size_t i = 0; i < (size_t)w * h * 4; i++) s ^= o[i]; std::free(o); } else bad[0]++;
            }
            {   write_all(f, mutate(bmp));
                uint32_t *o = nullptr; int w = 0, h = 0;
                if (pt_host_read_bmp(f.c_str(), &o, &w, &h) == 0) { ok[1]++; volatile uint32_t s = 0; for (size_t i = 0; i < (size_t)w * h; i++) s ^= o[i]; std::free(o); } else bad[1]++;
            }
            if (it % 4 == 0) {
                write_all(f, mutate_glb(glb));
                float *pts = nullptr; int32_t *mats = nullptr; pt_material *nm = nullptr; uint32_t nt = 0, nn = 0;
                if (pt_host_load_glb(f.c_str(), &pts, &mats, &nt, &nm, &nn, 3) == 0) {
                    ok[2]++; volatile float s = 0; for (size_t i = 0; i < (size_t)nt * 9; i++) s += pts[i];
                    std::free(pts); std::free(mats); std::free(nm);
                } else bad[2]++;
            }
        } catch (const std::exception &e) {
            std::printf("iteration %d: exception escaped the C ABI: %s\n", it, e.what());
            return 1;
        }
    }
    std::printf("fuzz ok: png %ld loaded / %ld rejected, bmp %ld / %ld, glb %ld / %ld\n", ok[0], bad[0], ok[1], bad[1], ok[2], bad[2]);
    return 0;
}
