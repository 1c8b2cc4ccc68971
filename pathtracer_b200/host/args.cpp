// Command line of the drop-in executable: the reference's ParseArgs/PrintHelp behaviour
// (src/win32_main.cpp:2076-2195) plus new long options that the reference scan never sees.
//
// Reference behaviour kept on purpose:
//   * every argv string that starts with '-' is scanned CHARACTER BY CHARACTER, so "-nmr" is three flags;
//   * t/p/w read atoi() of the rest of the string but do not consume it, so each digit then prints
//     "Warning: invalid program arugment -<digit>" (main.cpp:2187-2189);
//   * clamps: t in [0, min(cores,512)], p in [0,1000], w in [1,5] (main.cpp:2162-2181);
//   * defaults: p=4, world 1, pinhole, all texture kinds on, threads = physical cores (main.cpp:2111-2163).
// New (strings starting with "--" are NOT fed to the character scan):
//   --width N --height N --gpus N --shard tiles|samples --seed N --scene-seed N --out PATH --res DIR
//   --lbvh --kernel mega|wavefront
#include "host_internal.h"

#include <unistd.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <set>
#include <string>

namespace {

int physical_cores() {
    // the reference counts RelationProcessorCore entries (main.cpp:2145-2150)
    FILE *f = std::fopen("/proc/cpuinfo", "r");
    std::set<std::pair<int, int>> cores;
    if (f) {
        char line[512];
        int phys = 0, core = -1;
        while (std::fgets(line, sizeof(line), f)) {
            if (!std::strncmp(line, "physical id", 11)) { const char *c = std::strchr(line, ':'); if (c) phys = std::atoi(c + 1); }
            else if (!std::strncmp(line, "core id", 7)) { const char *c = std::strchr(line, ':'); if (c) core = std::atoi(c + 1); }
            else if (line[0] == '\n') { if (core >= 0) cores.insert({phys, core}); core = -1; }
        }
        if (core >= 0) cores.insert({phys, core});
        std::fclose(f);
    }
    int n = (int)cores.size();
    if (n == 0) n = (int)sysconf(_SC_NPROCESSORS_ONLN);
    return n > 0 ? n : 1;
}

}  // namespace

extern "C" void pt_host_print_help(void) {
    std::printf("usage: Pathtracer.exe [options]\n");
    std::printf("\nPhysically-based Path Tracer capable of rendering various geometrical shapes, including triangles.\n");
    std::printf("Written by Noah J. Cabral.\n\n");
    std::printf("optional arguments:\n");
    std::printf("\tt<int>                        - Set the number of threads to use.\n");
    std::printf("\tp<int>                        - Set the rays to shoot per pixel.\n");
    std::printf("\tw<int>                        - Set the world number to load. Possible options:\n"
                "\t\t1:\tDefault scene.\n"
                "\t\t2:\tMetal-roughness test.\n"
                "\t\t3:\tCornell box.\n"
                "\t\t4:\tRay Tracing in One Weekend book cover.\n"
                "\t\t5:\tMario N64 model.\n");
    std::printf("\td                             - Enable depth of field via thin-lens approximation.\n");
    std::printf("\tn                             - Disable loading normal map textures.\n");
    std::printf("\tm                             - Disable loading metalness material textures.\n");
    std::printf("\tr                             - Disable loading roughness material textures.\n");
    std::printf("\th                             - Print this help menu.\n");
    std::printf("\nB200 render core additions (long options):\n");
    std::printf("\t--width N --height N          - Frame size (reference: fixed 1280x720 window).\n");
    std::printf("\t--gpus N                      - GPUs of this box to shard the frame over.\n");
    std::printf("\t--shard tiles|samples         - Interleaved 32x32 tiles (peer-copy gather) or sample ranges (peer copy + add of the float sums).\n");
    std::printf("\t--seed N                      - Philox key of the render.\n");
    std::printf("\t--scene-seed N                - mt19937 seed of world 4's sphere layout.\n");
    std::printf("\t--out PATH                    - Output .BMP (default test.bmp).\n");
    std::printf("\t--res DIR                     - Directory holding mario.glb and the rusty-metal textures.\n");
    std::printf("\t--lbvh                        - Build the BVH on the device (LBVH) instead of host SAH.\n");
    std::printf("\t--kernel mega|wavefront       - Force a render kernel.\n");
    std::printf("\t--snapshots N                 - Write the image after each of N sample chunks (progressive view).\n");
    std::printf("\t--hdr                         - Also dump the float accumulator to <out>.hdr.bin (for image_compare).\n");
}

extern "C" int pt_host_parse_args(int argc, char **argv, pt_host_options *o) {
    std::memset(o, 0, sizeof(*o));
    o->pp = 4;
    o->normals = o->metalness = o->roughness = 1;
    o->world = 1;
    o->width = 1280;
    o->height = 720;
    o->gpus = 1;
    o->seed = 1;
    o->scene_seed = 7;
    std::snprintf(o->out_path, sizeof(o->out_path), "test.bmp");
    std::snprintf(o->res_dir, sizeof(o->res_dir), "res");

    int systemcores = physical_cores();
    int maxthreads = std::min(systemcores, 512);
    o->system_cores = systemcores;
    o->threads = maxthreads;

    for (int i = 0; i < argc && argv[i]; i++) {
        const char *a = argv[i];
        if (a[0] != '-') continue;
        if (a[1] == '-' && a[2] != '\0') {
            std::string key(a + 2), val;
            size_t eq = key.find('=');
            bool has_val = false;
            if (eq != std::string::npos) { val = key.substr(eq + 1); key = key.substr(0, eq); has_val = true; }
            auto need = [&]() -> const char * {
                if (has_val) return val.c_str();
                if (i + 1 < argc && argv[i + 1]) return argv[++i];
                return "";
            };
            if (key == "width") o->width = std::max(1, std::atoi(need()));
            else if (key == "height") o->height = std::max(1, std::atoi(need()));
            else if (key == "gpus") o->gpus = std::max(1, std::atoi(need()));
            else if (key == "shard") { std::string s = need(); o->shard_samples = (s == "samples"); }
            else if (key == "seed") o->seed = std::strtoull(need(), nullptr, 10);
            else if (key == "scene-seed") o->scene_seed = (uint32_t)std::strtoul(need(), nullptr, 10);
            else if (key == "out") std::snprintf(o->out_path, sizeof(o->out_path), "%s", need());
            else if (key == "res") std::snprintf(o->res_dir, sizeof(o->res_dir), "%s", need());
            else if (key == "lbvh") o->lbvh = 1;
            else if (key == "snapshots") o->snapshots = std::max(1, std::atoi(need()));
            else if (key == "hdr") o->hdr_dump = 1;
            else if (key == "kernel") { std::string s = need(); o->kernel = s == "mega" ? PT_KERNEL_MEGA : s == "wavefront" ? PT_KERNEL_WAVEFRONT : PT_KERNEL_AUTO; }
            else std::printf("Warning: invalid program arugment --%s\n", key.c_str());
            continue;
        }
        for (const char *p = a; *p;) {
            char c = *p++;
            switch (c) {
                case 't': o->threads = std::max(0, std::min(maxthreads, std::atoi(p))); break;
                case 'p': o->pp = std::max(0, std::min(1000, std::atoi(p))); break;
                case 'n': o->normals = 0; break;
                case 'm': o->metalness = 0; break;
                case 'r': o->roughness = 0; break;
                case 'h': pt_host_print_help(); o->help = 1; return 1;
                case 'w': o->world = std::max(0, std::min(4, std::atoi(p) - 1)) + 1; break;
                case 'd': o->lens = 1; break;
                case '-': break;
                default: std::printf("Warning: invalid program arugment -%c\n", c); break;
            }
        }
    }
    std::printf("System has %d core(s).\n", systemcores);
    std::printf("Using %d thread(s).\n\n", o->threads);
    return 0;
}
