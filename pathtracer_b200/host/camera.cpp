// DefineCamera restated (src/win32_main.cpp:2197-2231).  Expression shapes and evaluation order follow
// the reference so the derived camera is bit-identical when built without FMA contraction.
#include "host_internal.h"

#include <cmath>

namespace pth {

static const float kPi = 3.14159265358979323846264338327f;   // ray_math.hpp:7
static const float kFixedFocalLength = 0.098f;               // main.cpp:94

static pt_v3 sub(pt_v3 a, pt_v3 b) { return pt_v3{a.x - b.x, a.y - b.y, a.z - b.z}; }
static pt_v3 cross(pt_v3 a, pt_v3 b) {
    return pt_v3{a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y};
}
static pt_v3 normalize(pt_v3 a) {
    float m = a.x * a.x + a.y * a.y + a.z * a.z;
    m = std::sqrt(m);
    a.x /= m; a.y /= m; a.z /= m;
    return a;
}

void define_camera(pt_camera &c, int width, int height) {
    c.axis_z = normalize(sub(c.pos, c.target));
    c.axis_x = normalize(cross(pt_v3{0, 0, 1}, c.axis_z));
    c.axis_y = normalize(cross(c.axis_z, c.axis_x));

    if (!c.use_pinhole)
        c.focal_length = 1.f / (1.f / kFixedFocalLength - 1.f / c.focal_distance);
    else
        c.focal_length = kFixedFocalLength;

    // full opening angle is 2*fov: tan(fov), not tan(fov/2) (main.cpp:2211)
    c.film_w = std::tan((kPi / 180.f) * c.fov) * 2.f * c.focal_length;
    c.film_h = c.film_w;
    if (width > height)
        c.film_h = c.film_w * (float)height / (float)width;
    else if (height > width)
        c.film_w = c.film_h * (float)width / (float)height;

    c.half_film_w = c.film_w / 2.0f;
    c.half_film_h = c.film_h / 2.0f;

    c.frustum_center = pt_v3{c.pos.x - c.focal_length * c.axis_z.x, c.pos.y - c.focal_length * c.axis_z.y,
                             c.pos.z - c.focal_length * c.axis_z.z};

    // film space is [-1,1], so half a pixel is 1/W (main.cpp:2228-2231)
    c.half_pixel_w = 1.0f / (float)width;
    c.half_pixel_h = 1.0f / (float)height;
}

}  // namespace pth
