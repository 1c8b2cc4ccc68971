// pathtracer -- the reference program's command line on top of the B200 render core.
//
// Mirrors main() + MasterThread of the reference (src/win32_main.cpp:214-279, 879-987):
//     ParseArgs -> LoadWorld -> DefineCamera -> [render] -> WriteDIBImage("test.bmp")
// with [render] = pt_upload_scene / pt_render / pt_readback (include/pt_api.h) instead of the CPU thread pool,
// and the GDI live viewer replaced by optional progressive .BMP snapshots (--snapshots N).
// --gpus N shards one frame over N devices of this box from ONE process: a context per device, interleaved
// 32x32 tiles (gather) or sample ranges (accumulator sum) exchanged by peer copies over NVLink.
// There is no CPU fallback: without a CUDA device the program exits with the library's error.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "pt_api.h"
#include "pt_host.h"

static double now_s() {
    using namespace std::chrono;
    return duration<double>(steady_clock::now().time_since_epoch()).count();
}

static int die(const char *what, pt_context *ctx) {
    std::fprintf(stderr, "pathtracer: %s: %s\n", what, pt_last_error(ctx));
    return 1;
}

int main(int argc, char **argv) {
    pt_host_options opt;
    if (pt_host_parse_args(argc, argv, &opt)) return 0;                 // -h: help printed (main.cpp:2185)

    const int W = opt.width > 0 ? opt.width : 1280, H = opt.height > 0 ? opt.height : 720;   // main.cpp:218-219
    const int G = opt.gpus > 0 ? opt.gpus : 1;
    const int lens = opt.lens;                                          // (pt_host_scene forces it for world 4, main.cpp:2030)
    const uint32_t tex_flags = (opt.normals ? PT_SCENE_NORMAL_MAPS : 0) | (opt.metalness ? PT_SCENE_METALNESS_MAPS : 0) |
                               (opt.roughness ? PT_SCENE_ROUGHNESS_MAPS : 0);

    pt_host_world *world = nullptr;
    if (pt_host_load_world(opt.world, opt.res_dir, opt.scene_seed, &world)) {
        std::fprintf(stderr, "pathtracer: cannot build world %d: %s\n", opt.world, pt_host_last_error());
        return 1;
    }
    pt_scene scene;
    if (pt_host_scene(world, W, H, lens, tex_flags, &scene)) {
        std::fprintf(stderr, "pathtracer: %s\n", pt_host_last_error());
        return 1;
    }

    std::vector<pt_context *> ctx(G, nullptr);
    for (int g = 0; g < G; g++) {
        pt_config cfg = {sizeof(pt_config), g, opt.lbvh ? PT_BVH_DEVICE_LBVH : PT_BVH_HOST_SAH, opt.kernel, nullptr};
        if (pt_create(&cfg, &ctx[g])) return die("pt_create", nullptr);
        if (pt_upload_scene(ctx[g], &scene)) return die("pt_upload_scene", ctx[g]);
    }

    const uint32_t spp = (uint32_t)opt.pp * (uint32_t)opt.pp;
    const int chunks = opt.snapshots > 0 ? opt.snapshots : 1;
    std::vector<uint32_t> image((size_t)W * H);
    std::vector<float> hdr(opt.hdr_dump ? (size_t)W * H * 4 : 0);
    pt_stats stats;
    std::memset(&stats, 0, sizeof(stats));
    const double t0 = now_s();
    double render_ms = 0.0;

    for (int c = 0; c < chunks; c++) {
        // this chunk's slice of the p*p sample list, then each device's share of it
        const uint32_t c_lo = (uint32_t)((uint64_t)spp * c / chunks), c_hi = (uint32_t)((uint64_t)spp * (c + 1) / chunks);
        for (int g = 0; g < G; g++) {
            pt_render_params rp;
            std::memset(&rp, 0, sizeof(rp));
            rp.struct_size = sizeof(rp);
            rp.width = W; rp.height = H; rp.pp = opt.pp; rp.seed = opt.seed;
            if (opt.shard_samples) {
                rp.sample_begin = c_lo + (uint32_t)((uint64_t)(c_hi - c_lo) * g / G);
                rp.sample_end = c_lo + (uint32_t)((uint64_t)(c_hi - c_lo) * (g + 1) / G);
                // device 0 keeps the running sum of all chunks; the others start from zero every chunk
                rp.flags = (g == 0 ? c == 0 : true) ? PT_RENDER_CLEAR : 0;
            } else {
                rp.sample_begin = c_lo; rp.sample_end = c_hi;
                rp.tile_rank = (uint32_t)g; rp.tile_count = (uint32_t)G;
                rp.flags = c == 0 ? PT_RENDER_CLEAR : 0;
            }
            if (rp.sample_end == rp.sample_begin) rp.sample_begin = rp.sample_end = spp;   // empty share ([0,0) would mean "all")
            if (pt_render(ctx[g], &rp)) return die("pt_render", ctx[g]);             // asynchronous: all devices run together
        }
        for (int g = 1; g < G; g++) {
            int rc = opt.shard_samples ? pt_add_accum_from(ctx[0], ctx[g]) : pt_gather_tiles_from(ctx[0], ctx[g], (uint32_t)g, (uint32_t)G);
            if (rc) return die("multi-GPU exchange", ctx[0]);
        }
        pt_stats st;
        if (pt_readback(ctx[0], image.data(), (opt.hdr_dump && c == chunks - 1 && (G == 1 || opt.shard_samples)) ? hdr.data() : nullptr, &st))
            return die("pt_readback", ctx[0]);
        render_ms += st.render_ms;
        stats = st;
        if (chunks > 1 || c == chunks - 1) {
            if (pt_host_write_bmp(opt.out_path, image.data(), W, H)) {
                std::fprintf(stderr, "pathtracer: cannot write %s\n", opt.out_path);
                return 1;
            }
            if (chunks > 1) std::printf("snapshot %d/%d: %u of %u spp\n", c + 1, chunks, c_hi, spp);
        }
    }
    const double wall = now_s() - t0;

    if (opt.hdr_dump) {
        if (G > 1 && !opt.shard_samples) std::fprintf(stderr, "pathtracer: --hdr with tile sharding is not supported (use --shard samples)\n");
        else {
            std::string path = std::string(opt.out_path) + ".hdr.bin";
            FILE *f = std::fopen(path.c_str(), "wb");
            if (f) {
                int32_t hd[2] = {W, H};
                std::fwrite(hd, sizeof(hd), 1, f);
                std::fwrite(hdr.data(), sizeof(float), hdr.size(), f);
                std::fclose(f);
            }
        }
    }
    const double samples = (double)W * H * spp;
    std::printf("Done. Image written to %s.\n", opt.out_path);                         // main.cpp:984
    std::printf("[pt_b200] world %d, %dx%d, %u spp, %d GPU(s)%s: %.3f s wall, device 0 kernels %.2f ms, %.1f Msamples/s\n", opt.world, W, H, spp,
                G, G > 1 ? (opt.shard_samples ? " (samples)" : " (tiles)") : "", wall, render_ms, wall > 0 ? samples / wall / 1e6 : 0.0);
    for (int g = 0; g < G; g++) pt_destroy(ctx[g]);
    pt_host_free_world(world);
    return 0;
}
