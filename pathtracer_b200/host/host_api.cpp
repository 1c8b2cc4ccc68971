// C surface of libpt_host.so (include/pt_host.h).
#include "host_internal.h"

#include <exception>

#include <cstdlib>
#include <cstring>

namespace pth {
static thread_local std::string g_error;
void set_error(const std::string &m) { g_error = m; }
const std::string &last_error() { return g_error; }
}  // namespace pth

extern "C" const char *pt_host_last_error(void) { return pth::last_error().c_str(); }

extern "C" int pt_host_load_world(int world, const char *res_dir, uint32_t scene_seed, pt_host_world **out) {
    if (!out) return -1;
    *out = nullptr;
    pt_host_world *hw = new pt_host_world();
    try {
        pth::build_world(hw->w, world, res_dir ? res_dir : "res", scene_seed);
    } catch (const std::exception &e) {
        pth::set_error(e.what());
        delete hw;
        return -1;
    }
    *out = hw;
    return 0;
}

extern "C" void pt_host_free_world(pt_host_world *w) { delete w; }

extern "C" void pt_host_define_camera(pt_camera *cam, int width, int height) { pth::define_camera(*cam, width, height); }

extern "C" int pt_host_scene(pt_host_world *hw, int width, int height, int lens, uint32_t tex_flags, pt_scene *s) {
    if (!hw || !s || width <= 0 || height <= 0) { pth::set_error("pt_host_scene: bad argument"); return -1; }
    pth::World &w = hw->w;
    std::memset(s, 0, sizeof(*s));
    s->struct_size = (uint32_t)sizeof(pt_scene);
    s->flags = tex_flags & (PT_SCENE_NORMAL_MAPS | PT_SCENE_METALNESS_MAPS | PT_SCENE_ROUGHNESS_MAPS);
    if (w.kind == 4) s->flags |= PT_SCENE_JUST_COSINE;          // main.cpp:654-655
    s->n_materials = (uint32_t)w.materials.size();
    s->n_spheres = (uint32_t)w.spheres.size();
    s->n_planes = (uint32_t)w.planes.size();
    s->n_quads = (uint32_t)w.quads.size();
    s->n_aabbs = (uint32_t)w.aabbs.size();
    s->n_triangles = (uint32_t)w.tri_mats.size();
    s->materials = w.materials.data();
    s->spheres = w.spheres.data();
    s->planes = w.planes.data();
    s->quads = w.quads.data();
    s->aabbs = w.aabbs.data();
    s->tri_points = w.tri_points.data();
    s->tri_materials = w.tri_mats.data();
    for (int i = 0; i < 4; i++) {
        s->textures[i].width = w.textures[i].w;
        s->textures[i].height = w.textures[i].h;
        s->textures[i].rgba8 = w.textures[i].rgba.empty() ? nullptr : w.textures[i].rgba.data();
    }
    s->camera = w.camera;
    // LoadWorld: c->use_pinhole = g_use_pinhole (main.cpp:1801); world 4 forces the lens (main.cpp:2030)
    s->camera.use_pinhole = (w.force_lens || lens) ? 0 : 1;
    pth::define_camera(s->camera, width, height);
    return 0;
}

extern "C" int pt_host_load_glb(const char *path, float **tri_points_out, int32_t **tri_mats_out, uint32_t *n_tris,
                                pt_material **new_materials_out, uint32_t *n_new_materials, uint32_t first_new_index) {
    try {
        std::vector<float> pts;
        std::vector<int32_t> mats;
        std::vector<pt_material> materials(first_new_index);
        if (!pth::load_glb(path, pts, mats, materials)) return -1;
        // +4: never malloc(0); an empty vector's data() may be NULL, which memcpy must not see
        auto dup = [](const void *src, size_t bytes) -> void * {
            void *p = std::malloc(bytes + 4);
            if (p && bytes) std::memcpy(p, src, bytes);
            return p;
        };
        *n_tris = (uint32_t)mats.size();
        *n_new_materials = (uint32_t)materials.size() - first_new_index;
        *tri_points_out = (float *)dup(pts.data(), pts.size() * sizeof(float));
        *tri_mats_out = (int32_t *)dup(mats.data(), mats.size() * sizeof(int32_t));
        *new_materials_out = (pt_material *)dup(materials.data() + first_new_index, (*n_new_materials) * sizeof(pt_material));
        if (!*tri_points_out || !*tri_mats_out || !*new_materials_out) {
            std::free(*tri_points_out); std::free(*tri_mats_out); std::free(*new_materials_out);
            pth::set_error("out of memory");
            return -1;
        }
        return 0;
    } catch (const std::exception &e) {          // no exception crosses the C ABI
        pth::set_error(std::string("GLB: ") + e.what());
        return -1;
    }
}
