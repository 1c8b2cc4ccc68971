// Internal C++ types of libpt_host.so (the C surface is include/pt_host.h).
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "pt_host.h"

namespace pth {

struct HostError : std::runtime_error {
    explicit HostError(const std::string &m) : std::runtime_error(m) {}
};

struct Texture {
    int w = 0, h = 0;
    std::vector<uint8_t> rgba;
};

struct World {
    int kind = 0;                       // 1..5
    bool force_lens = false;            // world 4 (src/win32_main.cpp:2030)
    std::vector<pt_material> materials;
    std::vector<pt_sphere> spheres;
    std::vector<pt_plane> planes;
    std::vector<pt_quad> quads;
    std::vector<pt_aabb> aabbs;
    std::vector<float> tri_points;      // 9 floats per triangle
    std::vector<int32_t> tri_mats;      // 1 per triangle
    Texture textures[4];
    pt_camera camera;                   // user parameters as LoadWorld leaves them
};

void build_world(World &w, int world, const std::string &res_dir, uint32_t scene_seed);
void define_camera(pt_camera &c, int width, int height);

bool load_png_rgba8(const char *path, std::vector<uint8_t> &rgba, int &w, int &h);
bool load_glb(const char *path, std::vector<float> &tri_points, std::vector<int32_t> &tri_mats,
              std::vector<pt_material> &materials);

void set_error(const std::string &m);
const std::string &last_error();

}  // namespace pth

struct pt_host_world {
    pth::World w;
};
