// Host-side scene construction: the five built-in worlds, restated from the reference's LoadWorld
// (src/win32_main.cpp:1788-2046) with AddSky (2048-2051), AddSunDirectionalLight (2053-2067) and
// MakeGroundPlane (2069-2074).  Output is plain pt_* PODs (include/pt_api.h); nothing here renders.
#include "host_internal.h"

#include <cmath>
#include <cstring>
#include <random>

namespace pth {

static pt_v3 v3(float x, float y, float z) { pt_v3 r = {x, y, z}; return r; }

static pt_material default_material() {
    // material_t's member initialisers, include/ray.hpp:63-78
    pt_material m;
    std::memset(&m, 0, sizeof(m));
    m.alpha = 1.f;
    m.ior = 1.f;
    m.roughness = 1.f;
    return m;
}

static uint32_t push_material(World &w, const pt_material &m) {
    w.materials.push_back(m);
    return (uint32_t)w.materials.size() - 1;
}

static void add_sky(World &w, pt_v3 color) {
    pt_material m = default_material();
    m.emit = color;
    push_material(w, m);
}

// the "sun": an emissive (15,15,15) sphere far away; it becomes spheres[0], the light the
// diffuse estimator importance-samples (main.cpp:683)
static void add_sun(World &w) {
    pt_material m = default_material();
    m.emit = v3(15.f, 15.f, 15.f);
    uint32_t light = push_material(w, m);
    pt_sphere s = {v3(2000, 2000, 2000), 1000.f, light};
    w.spheres.push_back(s);
}

static pt_plane ground_plane(uint32_t mat) {
    pt_plane p = {v3(0, 0, 1), 0.f, mat};
    return p;
}

static void camera_defaults(pt_camera &c) {
    // main.cpp:1801-1806
    std::memset(&c, 0, sizeof(c));
    c.use_pinhole = 1;
    c.pos = v3(0, -10, 1);
    c.fov = 45.f;
    c.focal_distance = 5.f;
    c.aperture_radius = 0.035f;
    c.target = v3(0, 0, 0);
}

static const pt_v3 kSkyBlue = {65 / 255.f, 108 / 255.f, 162 / 255.f};

static void world_default(World &w, const std::string &res_dir) {
    // main.cpp:1809-1843
    add_sky(w, kSkyBlue);
    add_sun(w);

    pt_material ground = default_material();
    ground.albedo_tex = 1;
    ground.metalness_tex = 2;
    ground.metal_color = v3(0.562f, 0.565f, 0.578f);
    ground.roughness_tex = 3;
    ground.normal_tex = 4;
    uint32_t gm = push_material(w, ground);
    w.spheres.push_back(pt_sphere{v3(0, 0, -1000), 1000.f, gm});

    static const char *names[4] = {"rusty-metal_albedo.png", "rusty-metal_metallic.png",
                                   "rusty-metal_roughness.png", "rusty-metal_normal-ogl.png"};
    for (int i = 0; i < 4; i++) {
        Texture t;
        std::string path = res_dir + "/" + names[i];
        if (!load_png_rgba8(path.c_str(), t.rgba, t.w, t.h))
            throw HostError("cannot load texture " + path + ": " + last_error());
        w.textures[i] = std::move(t);
    }

    pt_material m = default_material();
    m.albedo = v3(0.7f, 0.25f, 0.3f);
    m.roughness = 0.f;
    w.spheres.push_back(pt_sphere{v3(0, 0, 0), 1.0f, push_material(w, m)});

    m = default_material();
    m.albedo = v3(0.0f, 0.8f, 0.0f);
    m.metalness = 0.8f;
    m.metal_color = v3(0.562f, 0.565f, 0.578f);
    m.roughness = 0.f;
    w.spheres.push_back(pt_sphere{v3(-2, 0, 2), 1.0f, push_material(w, m)});

    m = default_material();
    m.albedo = v3(0.3f, 0.25f, 0.7f);
    m.roughness = 0.f;
    w.spheres.push_back(pt_sphere{v3(-1, -5, 0), 1.0f, push_material(w, m)});

    w.camera.fov = 30.f;
}

static void world_brdf_test(World &w) {
    // main.cpp:1903-1929: 11x11 spheres, metalness along i, roughness along j
    add_sky(w, kSkyBlue);
    add_sun(w);
    pt_material g = default_material();
    g.albedo = v3(0.5f, 0.5f, 0.5f);
    w.planes.push_back(ground_plane(push_material(w, g)));
    for (int i = 0; i < 11; i++)
        for (int j = 0; j < 11; j++) {
            pt_v3 color = v3(1.0f, 0.782f, 0.344f);
            pt_material m = default_material();
            m.albedo = color;
            m.metalness = i / 10.f;
            m.metal_color = color;
            m.roughness = j / 10.f;
            uint32_t id = push_material(w, m);
            // V3(i/2.f, 11/2.f-j/2.f, 0.2): the literal 0.2 is a double narrowed to float
            w.spheres.push_back(pt_sphere{v3(i / 2.f, 11 / 2.f - j / 2.f, (float)0.2), 0.2f, id});
        }
    w.camera.target = v3(2.5f, 2.5f, 0);
    w.camera.pos = v3(2.5f, 7, 2);
    w.camera.fov = 50.f;
    w.camera.focal_distance = 10.f;
}

static void world_cornell(World &w) {
    // main.cpp:1844-1901: 800 x 555 x 555 box of five quads, emissive sphere in the middle
    add_sky(w, v3(0, 0, 0));
    const int left = 0, right = 800, bottom = 0, top = 555, front = 0, back = 555;
    pt_material m = default_material();
    m.albedo = v3((float).65, (float).05, (float).05);
    uint32_t red = push_material(w, m);
    m.albedo = v3((float).73, (float).73, (float).73);
    uint32_t white = push_material(w, m);
    m.albedo = v3((float).12, (float).45, (float).15);
    uint32_t green = push_material(w, m);
    m = default_material();
    m.emit = v3(15.f, 15.f, 15.f);
    uint32_t light = push_material(w, m);

    auto F = [](int a) { return (float)a; };
    w.quads.push_back(pt_quad{v3(F(right), F(bottom), F(front)), v3(0, 0, F(top - bottom)), v3(0, F(back - front), 0), green});
    w.quads.push_back(pt_quad{v3(F(left), F(bottom), F(front)), v3(0, F(back - front), 0), v3(0, 0, F(top - bottom)), red});
    w.spheres.push_back(pt_sphere{v3((right - left) / 2.f, (back - front) / 2.f, (top - bottom) / 2.f), 65.f, light});
    w.quads.push_back(pt_quad{v3(F(left), F(front), F(top)), v3(0, F(back - front), 0), v3(F(right - left), 0, 0), white});
    w.quads.push_back(pt_quad{v3(F(left), F(back), F(bottom)), v3(F(right - left), 0, 0), v3(0, 0, F(top - bottom)), white});
    w.quads.push_back(pt_quad{v3(F(left), F(bottom), F(front)), v3(F(right - left), 0, 0), v3(0, F(back - front), 0), white});

    w.camera.fov = 40;
    w.camera.pos = v3((right - left) / 2.f, F(front - 800), (top - bottom) / 2.f);
    w.camera.target = v3((right - left) / 2.f, F(front), (top - bottom) / 2.f);
}

// std::mt19937 + uniform_real_distribution<float>(0,1): the reference's RandomUnilateral
// (include/ray_math.hpp:245-259).  Same libstdc++ => same numbers for the same seed.
struct Mt {
    std::mt19937 gen;
    std::uniform_real_distribution<float> dis{0.0, 1.0};
    explicit Mt(uint32_t seed) : gen(seed) {}
    float u() { return dis(gen); }
    // RandomV3() is V3(R(),R(),R()) (ray_math.hpp:267-269); g++ evaluates the three call
    // arguments last-to-first, so z is drawn first.  Checked against the oracle's scene dump.
    pt_v3 v3_args_rtl() { float z = u(); float y = u(); float x = u(); return pt_v3{x, y, z}; }
};

static void world_rtow(World &w, uint32_t scene_seed) {
    // main.cpp:1960-2036.  The layout consumes the global RNG; draw order follows what g++ does with
    // the reference's expressions (function arguments right-to-left, braced initialisers left-to-right).
    add_sky(w, v3(1.f, 1.f, 1.f));
    pt_material ground = default_material();
    ground.albedo = v3(0.5f, 0.5f, 0.5f);
    w.spheres.push_back(pt_sphere{v3(0, 0, -1000), 1000.f, push_material(w, ground)});

    Mt rng(scene_seed);
    pt_material material = ground;   // the reference reuses one `material` variable (main.cpp:1796, 1994)
    for (int a = -11; a < 11; a++)
        for (int b = -11; b < 11; b++) {
            float choose_mat = rng.u();
            // V3(a + 0.9*R(), b + 0.9*R(), 0.2): arguments right-to-left => y's draw precedes x's
            float ry = rng.u();
            float rx = rng.u();
            pt_v3 center = v3((float)(a + 0.9 * rx), (float)(b + 0.9 * ry), (float)0.2);
            float dx = center.x - 4.f, dy = center.y - 0.f, dz = center.z - (float)0.2;
            float mag = (float)std::sqrt(dx * dx + dy * dy + dz * dz);
            if (mag > 0.9) {
                if (choose_mat < 0.8) {
                    // Hadamard(RandomV3(), RandomV3()): second argument first
                    pt_v3 q = rng.v3_args_rtl();
                    pt_v3 p = rng.v3_args_rtl();
                    material = default_material();
                    material.albedo = v3(p.x * q.x, p.y * q.y, p.z * q.z);
                } else {
                    // { .metalness = R(), .metalColor = 0.5f*RandomV3()+V3(.5,.5,.5), .roughness = 1.f - material.metalness }
                    // braced list: left to right; `material.metalness` is still the PREVIOUS material's value
                    float prev_metalness = material.metalness;
                    pt_material nm = default_material();
                    nm.metalness = rng.u();
                    pt_v3 r = rng.v3_args_rtl();
                    nm.metal_color = v3(0.5f * r.x + 0.5f, 0.5f * r.y + 0.5f, 0.5f * r.z + 0.5f);
                    nm.roughness = 1.f - prev_metalness;
                    material = nm;
                }
                uint32_t id = push_material(w, material);
                w.spheres.push_back(pt_sphere{center, 0.2f, id});
            }
        }

    material = default_material();
    material.albedo = v3((float)0.4, (float)0.2, (float)0.1);
    w.spheres.push_back(pt_sphere{v3(-4, 0, 1), 1.f, push_material(w, material)});

    material = default_material();
    material.metalness = 1.f;
    material.metal_color = v3((float)0.7, (float)0.6, (float)0.5);
    material.roughness = 0.f;
    w.spheres.push_back(pt_sphere{v3(4, 0, 1), 1.f, push_material(w, material)});

    w.force_lens = true;                    // main.cpp:2030
    w.camera.target = v3(0, 0, 0);
    w.camera.pos = v3(13, 3, 2);
    w.camera.fov = 20.f;
    w.camera.focal_distance = 10.f;
}

static void world_mario(World &w, const std::string &res_dir) {
    // main.cpp:1930-1959
    add_sky(w, kSkyBlue);
    add_sun(w);
    pt_material g = default_material();
    g.albedo = v3(0.5f, 0.5f, 0.5f);
    w.planes.push_back(ground_plane(push_material(w, g)));

    std::string path = res_dir + "/mario.glb";
    if (!load_glb(path.c_str(), w.tri_points, w.tri_mats, w.materials))
        throw HostError("cannot load " + path + ": " + last_error());

    w.camera.target = v3(0, 0, 1);
    w.camera.pos = v3(-5, -5, 1);
    w.camera.fov = 30.f;
}

void build_world(World &w, int world, const std::string &res_dir, uint32_t scene_seed) {
    camera_defaults(w.camera);
    w.kind = world;
    switch (world) {
        case 1: world_default(w, res_dir); break;
        case 2: world_brdf_test(w); break;
        case 3: world_cornell(w); break;
        case 4: world_rtow(w, scene_seed); break;
        case 5: world_mario(w, res_dir); break;
        default: throw HostError("world must be 1..5");
    }
}

}  // namespace pth
