/*
 * pt_oracle.c -- plain-C, scalar restatement of the reference's render hot path.
 * TEST INFRASTRUCTURE ONLY (see pt_oracle.h).  Citations are into /root/reference:
 *   main.cpp = src/win32_main.cpp, math.hpp = include/ray_math.hpp, ray.hpp = include/ray.hpp.
 *
 * Written to be bit-exact with the reference when fed the same mt19937 stream: every float
 * expression keeps the reference's association order, libm calls are the float overloads the
 * C++ reference resolves to, and the build disables FMA contraction (oracle/Makefile).
 */
#define _GNU_SOURCE
#include "pt_oracle.h"

#include <float.h>
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define PI_F 3.14159265358979323846264338327f /* math.hpp:7 */
#define TOLERANCE ((float)1e-9)              /* math.hpp:10 */
#define MIN_HIT_DISTANCE ((float)1e-4)       /* main.cpp:90 */
#define MAX_BOUNCE_COUNT 4                   /* main.cpp:86 */
#define N_AIR 1.003f                         /* main.cpp:93 */
#define FIXED_FOCAL_LENGTH 0.098f            /* main.cpp:94 */
#define MIN_ROUGHNESS ((float)(0.01f))       /* main.cpp:95 */

typedef pt_v3 v3;

/* ---------------------------------------------------------------- vectors (math.hpp:73-241) */
static inline v3 V3(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 add(v3 a, v3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 sub(v3 a, v3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 scl(float a, v3 b) { return V3(a * b.x, a * b.y, a * b.z); }
static inline v3 neg(v3 a) { return V3(-a.x, -a.y, -a.z); }
static inline float dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline v3 cross(v3 a, v3 b) {
    return V3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
static inline v3 hadamard(v3 a, v3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline float magnitude(v3 a) { return sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); }
static inline float magnitude_sq(v3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
static inline v3 normalize(v3 a) {
    float m = magnitude(a);
    a.x /= m; a.y /= m; a.z /= m;
    return a;
}
static inline v3 lerp3(v3 a, v3 b, float p) { return add(scl(1.0f - p, a), scl(p, b)); }   /* math.hpp:306-308 */
static inline int v3_eq(v3 a, v3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
static inline int v3_isnan(v3 a) { return a.x != a.x || a.y != a.y || a.z != a.z; }           /* math.hpp:501-504 */
static inline float maxf_macro(float a, float b) { return (a > b) ? a : b; }                /* Win32 max() */
static inline float minf_macro(float a, float b) { return (a < b) ? a : b; }                /* Win32 min() */

/* ---------------------------------------------------------------- RNG */
/* mt19937 (Matsumoto & Nishimura) as std::mt19937; float mapping as libstdc++'s
   uniform_real_distribution<float>(0,1) -> generate_canonical<float,24>: one 32-bit draw,
   float(x) / 2^32, and the >= 1 result replaced by nextafter(1,0). */
typedef struct { uint32_t mt[624]; int idx; } mt_state;
static mt_state g_mt;
static int g_mt_seeded = 0;

static void mt_seed(mt_state *m, uint32_t seed) {
    m->mt[0] = seed;
    for (int i = 1; i < 624; i++) m->mt[i] = 1812433253u * (m->mt[i - 1] ^ (m->mt[i - 1] >> 30)) + (uint32_t)i;
    m->idx = 624;
}
static uint32_t mt_next(mt_state *m) {
    if (m->idx >= 624) {
        for (int i = 0; i < 624; i++) {
            uint32_t y = (m->mt[i] & 0x80000000u) | (m->mt[(i + 1) % 624] & 0x7fffffffu);
            m->mt[i] = m->mt[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        }
        m->idx = 0;
    }
    uint32_t y = m->mt[m->idx++];
    y ^= y >> 11;
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= y >> 18;
    return y;
}
static inline float u32_to_unit_float(uint32_t x) {
    float r = (float)x / 4294967296.0f;
    if (r >= 1.0f) r = 0x1.fffffep-1f;
    return r;
}

/* Philox4x32-10 (Salmon et al., SC'11).  Stream layout shared with the CUDA kernels:
   key = (seed lo, seed hi); counter = (block, a, b, tag); uniform k = word k%4 of block k/4. */
static void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
}

typedef struct {
    int mode;
    /* philox */
    uint32_t k0, k1, a, b, tag, count;
    uint32_t buf[4];
} rng_t;

static void rng_philox(rng_t *r, uint64_t seed, uint32_t a, uint32_t b, uint32_t tag) {
    r->mode = PTO_RNG_PHILOX;
    r->k0 = (uint32_t)seed; r->k1 = (uint32_t)(seed >> 32);
    r->a = a; r->b = b; r->tag = tag; r->count = 0;
}
static void rng_mt(rng_t *r) { memset(r, 0, sizeof(*r)); r->mode = PTO_RNG_MT19937; }

static float rnd(rng_t *r) {   /* RandomUnilateral, math.hpp:250-259 */
    if (r->mode == PTO_RNG_MT19937) {
        if (!g_mt_seeded) { mt_seed(&g_mt, 5489u); g_mt_seeded = 1; }
        return u32_to_unit_float(mt_next(&g_mt));
    }
    uint32_t w = r->count & 3u;
    if (w == 0) {
        uint32_t c[4] = {r->count >> 2, r->a, r->b, r->tag};
        philox4x32_10(c, r->k0, r->k1);
        memcpy(r->buf, c, sizeof(c));
    }
    r->count++;
    return u32_to_unit_float(r->buf[w]);
}

/* PHILOX mode only: skip to the next 4-uniform block.  Every iteration of RayCast's sampling loop starts
   on a fresh block, so the CUDA kernels need exactly one Philox evaluation per path vertex. */
static void rng_align_block(rng_t *r) {
    if (r->mode == PTO_RNG_PHILOX) r->count = (r->count + 3u) & ~3u;
}

void pto_seed_mt(uint32_t seed) { mt_seed(&g_mt, seed); g_mt_seeded = 1; }
void pto_random_mt(uint32_t n, float *out) { rng_t r; rng_mt(&r); for (uint32_t i = 0; i < n; i++) out[i] = rnd(&r); }
float pto_philox_uniform(uint64_t seed, uint32_t a, uint32_t b, uint32_t tag, uint32_t k) {
    uint32_t c[4] = {k >> 2, a, b, tag};
    philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
    return u32_to_unit_float(c[k & 3u]);
}

/* ---------------------------------------------------------------- primitive tests */
/* RaySphereIntersect, main.cpp:2355-2379: near root only, root term must exceed 1e-9 */
static float ray_sphere(v3 o, v3 d, float min_hit, const pt_sphere *s, v3 *normal) {
    v3 rel = sub(o, s->p);
    float a = dot(d, d);
    float b = 2.0f * dot(rel, d);
    float c = dot(rel, rel) - s->r * s->r;
    float denom = 2.0f * a;
    float disc = b * b - 4.0f * a * c;
    if (disc < 0.f) return min_hit;
    float root = sqrtf(disc);
    if (root > TOLERANCE) {
        float t = (-b - root) / denom;
        if (t > min_hit) {
            *normal = normalize(add(scl(t, d), rel));
            return t;
        }
    }
    return min_hit;
}

/* RayIntersectPlane, math.hpp:334-341 */
static float ray_plane(v3 o, v3 d, v3 N, float dd, float min_hit) {
    float denom = dot(N, d);
    if ((denom < -TOLERANCE) || (denom > TOLERANCE)) return (dd - dot(N, o)) / denom;
    return min_hit;
}

/* RayIntersectPlanarShape<>, math.hpp:357-381 */
static float ray_planar(int quad, v3 o, v3 d, float min_hit, v3 A, v3 u, v3 v) {
    v3 N = cross(u, v);
    v3 Nunit = normalize(N);
    float dd = dot(A, Nunit);
    float t = ray_plane(o, d, Nunit, dd, min_hit);
    if (t != min_hit) {
        v3 p = sub(add(o, scl(t, d)), A);
        v3 w = scl(1.f / dot(N, N), N);
        float alpha = dot(w, cross(p, v));
        float beta = dot(w, cross(u, p));
        if (quad) { if (alpha >= 0.f && alpha <= 1.f && beta >= 0.f && beta <= 1.f) return t; }
        else      { if (alpha >= 0.f && beta >= 0.f && (alpha + beta) <= 1.f) return t; }
    }
    return min_hit;
}

/* RayIntersectWithAABB2, math.hpp:398-482: faces tested in the order front(-z) back(+z) left(-x)
   right(+x) top(+y) bottom(-y); first face with t >= 0 whose hit point lies inside the box wins */
static float ray_aabb(v3 o, v3 d, float min_hit, const pt_aabb *b, int *face) {
    for (int j = 0; j < 6; j++) {
        float plane, t, x, y, z;
        switch (j) {
            case 0: case 1:
                plane = (j == 0) ? b->min.z : b->max.z;
                if (d.z == 0.f) continue;
                t = (plane - o.z) / d.z;
                if (t < 0.f) continue;
                y = o.y + d.y * t; x = o.x + d.x * t; z = plane;
                break;
            case 2: case 3:
                plane = (j == 2) ? b->min.x : b->max.x;
                if (d.x == 0.f) continue;
                t = (plane - o.x) / d.x;
                if (t < 0.f) continue;
                y = o.y + d.y * t; z = o.z + d.z * t; x = plane;
                break;
            default:
                plane = (j == 4) ? b->max.y : b->min.y;
                if (d.y == 0.f) continue;
                t = (plane - o.y) / d.y;
                if (t < 0.f) continue;
                x = o.x + d.x * t; z = o.z + d.z * t; y = plane;
                break;
        }
        if (x >= b->min.x && x <= b->max.x && y >= b->min.y && y <= b->max.y && z >= b->min.z && z <= b->max.z) {
            *face = j;
            return t;
        }
    }
    return min_hit;
}

static const v3 k_face_normals[6] = {{0.f, 0.f, -1.f}, {0.f, 0.f, 1.f}, {-1.f, 0.f, 0.f},
                                     {1.f, 0.f, 0.f}, {0.f, 1.f, 0.f}, {0.f, -1.f, 0.f}};   /* main.cpp:540-543 */

/* RayCastIntersect, main.cpp:406-556: spheres, quads (min t 0.02), planes, triangles, boxes;
   strict '<' keeps the first of equal hits.  The reference walks an octree over the triangles
   (main.cpp:476-526) whose nearest hit equals this brute-force loop. */
static pt_hit closest_hit(const pt_scene *s, v3 o, v3 d) {
    const float min_hit = MIN_HIT_DISTANCE;
    pt_hit h;
    h.t = FLT_MAX; h.kind = PT_HIT_MISS; h.index = -1; h.mat = 0; h.normal = V3(0, 0, 0);
    for (uint32_t i = 0; i < s->n_spheres; i++) {
        v3 N; float t = ray_sphere(o, d, min_hit, &s->spheres[i], &N);
        if (t > min_hit && t < h.t) { h.t = t; h.mat = s->spheres[i].mat; h.normal = N; h.kind = PT_HIT_SPHERE; h.index = (int32_t)i; }
    }
    for (uint32_t i = 0; i < s->n_quads; i++) {
        const pt_quad *q = &s->quads[i];
        v3 N = normalize(cross(q->u, q->v));
        float mh = 0.02;   /* double literal narrowed to float, main.cpp:446 */
        float t = ray_planar(1, o, d, mh, q->point, q->u, q->v);
        if (t > mh && t < h.t) { h.t = t; h.mat = q->mat; h.normal = N; h.kind = PT_HIT_QUAD; h.index = (int32_t)i; }
    }
    for (uint32_t i = 0; i < s->n_planes; i++) {
        const pt_plane *p = &s->planes[i];
        float t = ray_plane(o, d, p->n, p->d, min_hit);
        if (t > min_hit && t < h.t) { h.t = t; h.mat = p->mat; h.normal = p->n; h.kind = PT_HIT_PLANE; h.index = (int32_t)i; }
    }
    for (uint32_t i = 0; i < s->n_triangles; i++) {
        const float *P = s->tri_points + 9 * (size_t)i;
        v3 A = V3(P[0], P[1], P[2]), B = V3(P[3], P[4], P[5]), C = V3(P[6], P[7], P[8]);
        v3 u = sub(B, A), v = sub(C, A);
        float t = ray_planar(0, o, d, min_hit, A, u, v);
        if (t > min_hit && t < h.t) {
            h.t = t; h.mat = (uint32_t)s->tri_materials[i]; h.normal = normalize(cross(u, v));
            h.kind = PT_HIT_TRIANGLE; h.index = (int32_t)i;
        }
    }
    for (uint32_t i = 0; i < s->n_aabbs; i++) {
        int face = 0;
        float t = ray_aabb(o, d, min_hit, &s->aabbs[i], &face);
        if (t > min_hit && t < h.t) { h.t = t; h.mat = s->aabbs[i].mat; h.normal = k_face_normals[face]; h.kind = PT_HIT_AABB; h.index = (int32_t)i; }
    }
    return h;
}

void pto_intersect(const pt_scene *s, uint32_t n, const float *o, const float *d, pt_hit *hits) {
    for (uint32_t i = 0; i < n; i++)
        hits[i] = closest_hit(s, V3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), V3(d[3 * i], d[3 * i + 1], d[3 * i + 2]));
}

/* ---------------------------------------------------------------- textures */
/* SampleTexture, main.cpp:1680-1709.  uv is in texels; wrap by integer modulo; bilinear on
   float3 texels = byte/255.f (LoadTexture, main.cpp:1737-1739) */
static v3 texel(const pt_texture *t, int x, int y) {
    const uint8_t *p = t->rgba8 + 4 * ((size_t)y * t->width + x);
    return V3((float)p[0] / 255.f, (float)p[1] / 255.f, (float)p[2] / 255.f);
}
static v3 sample_texture(const pt_texture *tex, float u, float v) {
    int x1, x2, y1, y2;
    float s, t;
    u = fabsf(u); v = fabsf(v);
    x1 = (int)u;
    y1 = (int)v;
    s = u - (float)x1;
    t = v - (float)y1;
    s = minf_macro(1.f, maxf_macro(s, 0.f));
    t = minf_macro(1.f, maxf_macro(t, 0.f));
    x1 = x1 % tex->width;
    x2 = (x1 + 1) % tex->width;
    y1 = y1 % tex->height;
    y2 = (y1 + 1) % tex->height;
    v3 top = lerp3(texel(tex, x1, y1), texel(tex, x2, y1), s);
    v3 bottom = lerp3(texel(tex, x1, y2), texel(tex, x2, y2), s);
    return lerp3(top, bottom, t);
}
/* BespokeSampleTexture, main.cpp:1675-1678: planar mapping uv = hit.xy, half a texture per world unit */
static v3 bespoke_sample(const pt_texture *tex, float u, float v) {
    return sample_texture(tex, u * (float)tex->width * 0.5f, v * (float)tex->height * 0.5f);
}
void pto_sample_texture(const pt_scene *s, int idx, float u, float v, float *rgb) {
    v3 c = bespoke_sample(&s->textures[idx], u, v);
    rgb[0] = c.x; rgb[1] = c.y; rgb[2] = c.z;
}

/* ---------------------------------------------------------------- BRDF / sampling helpers */
static float burley(float r) { return r * r * r * r; }                                  /* main.cpp:2252-2254 */
static int effectively_smooth(float r) { return r < MIN_ROUGHNESS; }                    /* main.cpp:1783-1786 */

static v3 schlick_metal(float F0, float cos_theta, float metalness, v3 surface) {       /* main.cpp:1752-1756 */
    v3 vF0 = lerp3(V3(F0, F0, F0), surface, metalness);
    return add(vF0, scl(powf(1.f - cos_theta, 5.f), sub(V3(1.f, 1.f, 1.f), vF0)));
}

static float hammon(v3 N, v3 L, v3 V, float roughness) {                                /* main.cpp:1773-1781 */
    float a2 = burley(roughness);
    float NdotV = dot(N, V);
    float NdotL = dot(N, L);
    float numerator = 2.f * NdotL * NdotV;
    float denom = NdotV * sqrtf(a2 + (1.f - a2) * NdotL * NdotL) + NdotL * sqrtf(a2 + (1 - a2) * NdotV * NdotV);
    return numerator / denom;
}

static v3 brdf_specular(v3 N, v3 L, v3 V, v3 H, float roughness) {                      /* main.cpp:1610-1620 */
    float s = hammon(N, L, V, roughness) * (fabsf(dot(H, L)) / fabsf(dot(N, L)) / fabsf(dot(H, N)));
    return scl(s, V3(1.f, 1.f, 1.f));
}

static v3 brdf_diff(const pt_scene *s, const pt_material *m, v3 p) {                    /* main.cpp:1595-1608 */
    float pi_term = 1.f / PI_F;
    if (m->albedo_tex == 0) return scl(pi_term, m->albedo);
    return scl(pi_term, bespoke_sample(&s->textures[m->albedo_tex - 1], p.x, p.y));
}

static void build_onb(v3 w, v3 *a, v3 *b, v3 *c) {                                       /* main.cpp:2288-2305 */
    v3 unit_w = normalize(w);
    int is_x = fabsf(unit_w.x) > 0.9;     /* float |x| compared against the double 0.9 */
    v3 A = is_x ? V3(0, 1, 0) : V3(1, 0, 0);
    v3 v = normalize(cross(unit_w, A));
    v3 u = cross(unit_w, v);
    *a = u; *b = v; *c = unit_w;
}

static v3 random_cosine_hemisphere(rng_t *r) {                                           /* main.cpp:2258-2268 */
    float r1 = rnd(r);
    float r2 = rnd(r);
    float phi = 2.f * PI_F * r1;
    float x = cosf(phi) * sqrtf(r2);
    float y = sinf(phi) * sqrtf(r2);
    float z = sqrtf(1.f - r2);
    return V3(x, y, z);
}

static v3 random_half_vector_ggx(rng_t *r, float roughness) {                            /* main.cpp:2271-2285 */
    float a2 = burley(roughness);
    float z1 = rnd(r);
    float z2 = rnd(r);
    float phi = 2.f * PI_F * z1;
    float theta = acosf(sqrtf((1.f - z2) / (1.f + z2 * (a2 - 1.f))));
    float x = cosf(phi) * sinf(theta);
    float y = sinf(phi) * sinf(theta);
    float z = cosf(theta);
    return V3(x, y, z);
}

static v3 random_to_sphere(rng_t *r, const pt_sphere *s, v3 from) {                      /* main.cpp:2332-2353 */
    float ds = magnitude_sq(sub(from, s->p));
    float term1 = 1.f - s->r * s->r / ds;
    if (term1 < 0.f) return V3(0.f, 0.f, 0.f);       /* inside the light: no draws consumed */
    float r1 = rnd(r);
    float r2 = rnd(r);
    float z = 1.f + r2 * (sqrtf(term1) - 1.f);
    float term2 = 1 - z * z;
    float phi = 2 * PI_F * r1;
    float x = cosf(phi) * sqrtf(term2);
    float y = sinf(phi) * sqrtf(term2);
    return V3(x, y, z);
}

static float pdf_cosine(v3 dir) {                                                        /* main.cpp:290-299 */
    return maxf_macro(0.f, dot(V3(0, 0, 1), dir) / PI_F);
}
static float pdf_to_sphere(v3 dir, const pt_sphere *s, v3 from) {                        /* main.cpp:324-340 */
    v3 N;
    if (ray_sphere(from, dir, MIN_HIT_DISTANCE, s, &N) <= MIN_HIT_DISTANCE) return 0.f;
    float cos_theta_max = sqrtf(1.f - s->r * s->r / magnitude_sq(sub(from, s->p)));
    float solid_angle = 2.f * PI_F * (1.f - cos_theta_max);
    return 1.f / solid_angle;
}

/* ---------------------------------------------------------------- RayCast, main.cpp:558-823 */
static v3 ray_cast(const pt_scene *s, rng_t *rng, v3 o, v3 d, int depth) {
    v3 radiance = V3(0, 0, 0);
    if (depth >= MAX_BOUNCE_COUNT) return radiance;

    v3 ray_o = o, ray_d = d;
    pt_hit p = closest_hit(s, ray_o, ray_d);
    const pt_material mat = s->materials[p.mat];
    v3 N = p.normal;
    v3 next_normal = p.normal;

    int hit_sky = (p.mat == 0);
    int hit_light = !v3_eq(mat.emit, V3(0, 0, 0));

    do {
        if (hit_sky || hit_light) break;

        float cos_theta, F0, metalness, roughness, NdotV, NdotL, px = 0.f;
        v3 H = V3(0, 0, 0), L = V3(0, 0, 0), V, pure_bounce, brdf, ks, kd, tx, ty, tz;

        cos_theta = dot(next_normal, ray_d);
        cos_theta = (cos_theta > 0.f) ? dot(scl(-1.f, next_normal), ray_d) : cos_theta;
        F0 = ((N_AIR - mat.ior) / (N_AIR + mat.ior)) * ((N_AIR - mat.ior) / (N_AIR + mat.ior));

        ray_o = add(ray_o, scl(p.t, ray_d));
        pure_bounce = sub(ray_d, scl(2.0f * cos_theta, next_normal));
        V = neg(ray_d);

        float u = ray_o.x, v = ray_o.y;                 /* planar "uv = hit.xy" mapping, main.cpp:613 */
        if (mat.metalness_tex == 0 || !(s->flags & PT_SCENE_METALNESS_MAPS)) metalness = mat.metalness;
        else metalness = bespoke_sample(&s->textures[mat.metalness_tex - 1], u, v).x;
        if (mat.roughness_tex == 0 || !(s->flags & PT_SCENE_ROUGHNESS_MAPS)) roughness = mat.roughness;
        else roughness = bespoke_sample(&s->textures[mat.roughness_tex - 1], u, v).x;
        if ((s->flags & PT_SCENE_NORMAL_MAPS) && mat.normal_tex != 0) {
            N = bespoke_sample(&s->textures[mat.normal_tex - 1], u, v);
            N = normalize(sub(scl(2.f, N), V3(1.f, 1.f, 1.f)));      /* used as a world-space vector */
        }

        NdotV = dot(N, V);
        if (NdotV <= 0.f) break;                         /* back-face: emission only, main.cpp:647-648 */

        build_onb(N, &tx, &ty, &tz);
        const int just_cosine = (s->flags & PT_SCENE_JUST_COSINE) != 0;

        for (;;) {                                       /* do { ... } while(true), main.cpp:660-792 */
            rng_align_block(rng);
            int specular = rnd(rng) > 0.5f;

            if (specular && effectively_smooth(roughness)) {
                L = pure_bounce;
                px = 1.f;
            } else if (!specular) {
                int sample_cosine = rnd(rng) > 0.5f;
                const pt_sphere *light = &s->spheres[0];
                v3 rdir = V3(0, 0, 0), lobe;
                if (sample_cosine || just_cosine) {
                    rdir = random_cosine_hemisphere(rng);
                } else {
                    v3 direction = sub(light->p, ray_o);
                    rdir = random_to_sphere(rng, light, ray_o);
                    build_onb(direction, &tx, &ty, &tz);   /* basis stays re-aimed for any retry */
                }
                if (v3_eq(rdir, V3(0, 0, 0))) continue;
                lobe = normalize(add(add(scl(rdir.x, tx), scl(rdir.y, ty)), scl(rdir.z, tz)));
                L = lobe;
                H = scl(1.f / magnitude(add(L, V)), add(L, V));
                if (!just_cosine)
                    px = 0.5f * pdf_cosine(normalize(rdir)) + 0.5f * pdf_to_sphere(lobe, light, ray_o);
                else
                    px = pdf_cosine(normalize(rdir));
                if (px == 0.f) continue;
            } else {
                v3 rdir = random_half_vector_ggx(rng, roughness);
                H = normalize(add(add(scl(rdir.x, tx), scl(rdir.y, ty)), scl(rdir.z, N)));
                L = sub(scl(2.f * dot(V, H), H), V);
                px = 1.f;
            }

            if ((NdotL = dot(N, L)) > 0.f) {
                if (effectively_smooth(roughness)) {
                    ks = schlick_metal(F0, NdotL, metalness, mat.metal_color);
                } else if ((dot(H, V) > 0.f) && (cos_theta = dot(H, L)) > 0.f) {
                    ks = schlick_metal(F0, cos_theta, metalness, mat.metal_color);
                } else {
                    break;
                }
                kd = sub(V3(1.f, 1.f, 1.f), ks);
                kd = lerp3(kd, V3(0, 0, 0), metalness);

                if (specular && effectively_smooth(roughness)) brdf = ks;
                else if (specular) brdf = hadamard(ks, brdf_specular(N, L, V, H, roughness));
                else brdf = scl(NdotL, hadamard(kd, brdf_diff(s, &mat, ray_o)));

                v3 child = ray_cast(s, rng, ray_o, L, depth + 1);
                radiance = add(radiance, scl(2.f * (1.f / px), hadamard(child, brdf)));
            }
            break;
        }
    } while (0);

    radiance = add(radiance, mat.emit);
    return radiance;
}

void pto_raycast(const pt_scene *s, int rng_mode, uint64_t seed, uint32_t n, const float *o, const float *d, float *rgb) {
    rng_t r;
    rng_mt(&r);
    for (uint32_t i = 0; i < n; i++) {
        if (rng_mode == PTO_RNG_PHILOX) rng_philox(&r, seed, i, 0, 2);
        v3 c = ray_cast(s, &r, V3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), V3(d[3 * i], d[3 * i + 1], d[3 * i + 2]), 0);
        rgb[3 * i] = c.x; rgb[3 * i + 1] = c.y; rgb[3 * i + 2] = c.z;
    }
}

/* ---------------------------------------------------------------- resolve */
/* TonemapPass (main.cpp:1664-1673; note e = 0.4), LinearToSRGB (math.hpp:320-332),
   255*s truncated, BGRAPack4x8 with A = 255 (math.hpp:196-202, main.cpp:1175-1182) */
static float tonemap1(float x) {
    const float a = 2.51f, b = 0.03f, c = 2.43f, d = 0.59f, e = 0.4f;
    float q = (x * (a * x + b)) / (e + x * (c * x + d));
    float m = (1.0f < q) ? 1.0f : q;           /* std::min(v, 1) */
    return (0.0f < m) ? m : 0.0f;              /* std::max(0, m) */
}
static float linear_to_srgb(float L) {
    if (L < 0.0f) L = 0.0f;
    if (L > 1.0f) L = 1.0f;
    float S = L * 12.92f;
    if (L > 0.0031308f) S = 1.055f * powf(L, 1.0f / 2.4f) - 0.055f;
    return S;
}
static uint32_t resolve_pixel(v3 color) {
    float r = 255.0f * linear_to_srgb(tonemap1(color.x));
    float g = 255.0f * linear_to_srgb(tonemap1(color.y));
    float b = 255.0f * linear_to_srgb(tonemap1(color.z));
    return (((uint32_t)255.0f) << 24) | (((uint32_t)r) << 16) | (((uint32_t)g) << 8) | ((uint32_t)b);
}
void pto_resolve(uint32_t n, const float *rgb, uint32_t *bgra) {
    for (uint32_t i = 0; i < n; i++) bgra[i] = resolve_pixel(V3(rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]));
}

/* ---------------------------------------------------------------- camera + per-pixel sampling */
void pto_primary_rays(const pt_scene *s, int width, int height, float *o, float *d) {
    const pt_camera *c = &s->camera;
    for (int y = 0; y < height; y++)
        for (int x = 0; x < width; x++) {
            const float fy = -1.0f + 2.0f * (float)y / (float)height;
            const float fx = -1.0f + 2.0f * (float)x / (float)width;
            v3 P = add(add(c->frustum_center, scl(fx * c->half_film_w, c->axis_x)), scl(fy * c->half_film_h, c->axis_y));
            v3 dir = normalize(sub(P, c->pos));
            size_t i = (size_t)y * width + x;
            o[3 * i] = c->pos.x; o[3 * i + 1] = c->pos.y; o[3 * i + 2] = c->pos.z;
            d[3 * i] = dir.x; d[3 * i + 1] = dir.y; d[3 * i + 2] = dir.z;
        }
}

static const float k_poisson_disk[12][2] = {          /* main.cpp:1097-1110 */
    {0.0f, 0.0f}, {-0.94201624f, -0.39906216f}, {0.94558609f, -0.76890725f}, {-0.094184101f, -0.92938870f},
    {0.34495938f, 0.29387760f}, {-0.91588581f, 0.45771432f}, {-0.81544232f, -0.87912464f},
    {-0.38277543f, 0.27676845f}, {0.97484398f, 0.75648379f}, {0.44323325f, -0.97511554f},
    {0.53742981f, -0.47373420f}, {-0.26496911f, -0.41893023f}};

#define PTO_MAX_NAN_RETRIES 8   /* PHILOX mode only: the reference retries without bound (main.cpp:1068, 1166) */

/* RNG streams in PHILOX mode (shared with the CUDA kernels; DESIGN.md "RNG streams"):
     path stream      (seed; pixel, sample, tag = 0 | retry << 8): pinhole jitter = uniforms 0,1, then RayCast's draws
     lens film stream (seed; pixel, rayIndex, tag = 1): the two film-jitter uniforms of outer sample rayIndex */
static void render_pixel(const pt_scene *s, int width, int height, int pp, int rng_mode, uint64_t seed,
                         int x, int y, uint32_t s_begin, uint32_t s_end,
                         double *sum, double *sumsq, v3 *mean_out) {
    const pt_camera *c = &s->camera;
    const float fy = -1.0f + 2.0f * (float)y / (float)height;      /* main.cpp:1003-1006: row 0 = bottom */
    const float fx = -1.0f + 2.0f * (float)x / (float)width;
    const uint32_t pixel = (uint32_t)y * (uint32_t)width + (uint32_t)x;
    const float contrib = 1.0f / (float)pp / (float)pp;
    v3 color = V3(0, 0, 0);
    double acc[3] = {0, 0, 0}, acc2[3] = {0, 0, 0};
    rng_t rng;
    rng_mt(&rng);

    for (uint32_t sidx = s_begin; sidx < s_end; sidx++) {
        const int i = (int)(sidx / (uint32_t)pp), j = (int)(sidx % (uint32_t)pp);
        v3 L;
        int retry = 0;
        if (c->use_pinhole) {
            for (;;) {                                            /* main.cpp:1041-1069 */
                if (rng_mode == PTO_RNG_PHILOX) rng_philox(&rng, seed, pixel, sidx, 0u | ((uint32_t)retry << 8));
                float llx = fx - 1.f * c->half_pixel_w;
                float lly = fy - 1.f * c->half_pixel_h;
                float stepX = 1.f / pp * c->half_pixel_w * 2.f;
                float stepY = 1.f / pp * c->half_pixel_h * 2.f;
                float xs = llx + (float)i / pp * c->half_pixel_w + stepX * 0.5f;
                float ys = lly + (float)j / pp * c->half_pixel_h + stepY * 0.5f;
                xs += (rnd(&rng) - 0.5f) * stepX;
                ys += (rnd(&rng) - 0.5f) * stepY;
                v3 P = add(add(c->frustum_center, scl(xs * c->half_film_w, c->axis_x)), scl(ys * c->half_film_h, c->axis_y));
                v3 dir = normalize(sub(P, c->pos));
                L = ray_cast(s, &rng, c->pos, dir, 0);
                if (!v3_isnan(L)) break;
                if (rng_mode == PTO_RNG_PHILOX && ++retry >= PTO_MAX_NAN_RETRIES) { L = V3(0, 0, 0); break; }
            }
        } else {
            /* main.cpp:1114-1167: outer film sample i (= rayIndex), inner lens sample j (= rayIndex2) */
            float u0, u1;
            if (rng_mode == PTO_RNG_PHILOX) {
                u0 = pto_philox_uniform(seed, pixel, (uint32_t)i, 1, 0);
                u1 = pto_philox_uniform(seed, pixel, (uint32_t)i, 1, 1);
            } else {
                /* sequential stream: the film jitter is drawn once per outer sample, when j == 0 */
                static _Thread_local float su0, su1;
                if (j == 0) { su0 = rnd(&rng); su1 = rnd(&rng); }
                u0 = su0; u1 = su1;
            }
            float offX = fx + ((2.0f * u0 - 1.0f) * c->half_pixel_w);
            float offY = fy + ((2.0f * u1 - 1.0f) * c->half_pixel_h);
            v3 P = add(add(c->frustum_center, scl(offX * c->half_film_w, c->axis_x)), scl(offY * c->half_film_h, c->axis_y));
            v3 dir = normalize(sub(P, c->pos));
            float fpd = 1.f / (1.f / FIXED_FOCAL_LENGTH - 1.f / c->focal_length);
            v3 N = neg(c->axis_z);
            v3 plane_point = add(add(c->pos, c->axis_x), scl(fpd, N));
            float dd = dot(N, plane_point);
            float t = ray_plane(c->pos, dir, N, dd, MIN_HIT_DISTANCE);
            v3 focal = add(c->pos, scl(t, dir));
            const float *ds = k_poisson_disk[((uint32_t)j * (uint32_t)i) % 12u];
            v3 ro = add(add(c->pos, scl(ds[0] * c->aperture_radius, c->axis_x)), scl(ds[1] * c->aperture_radius, c->axis_y));
            v3 rd = normalize(sub(focal, ro));
            for (;;) {
                if (rng_mode == PTO_RNG_PHILOX) rng_philox(&rng, seed, pixel, sidx, 0u | ((uint32_t)retry << 8));
                L = ray_cast(s, &rng, ro, rd, 0);
                if (!v3_isnan(L)) break;
                if (rng_mode == PTO_RNG_PHILOX && ++retry >= PTO_MAX_NAN_RETRIES) { L = V3(0, 0, 0); break; }
            }
        }
        color = add(color, scl(contrib, L));
        acc[0] += L.x; acc[1] += L.y; acc[2] += L.z;
        acc2[0] += (double)L.x * L.x; acc2[1] += (double)L.y * L.y; acc2[2] += (double)L.z * L.z;
    }
    if (sum) { sum[0] = acc[0]; sum[1] = acc[1]; sum[2] = acc[2]; }
    if (sumsq) { sumsq[0] = acc2[0]; sumsq[1] = acc2[1]; sumsq[2] = acc2[2]; }
    *mean_out = color;
}

void pto_render(const pt_scene *s, int width, int height, int pp, int rng_mode, uint64_t seed,
                int x0, int y0, int w, int h, uint32_t s_begin, uint32_t s_end,
                double *sum, double *sumsq, float *mean_f, uint32_t *ldr) {
    if (s_end == 0) s_end = (uint32_t)pp * (uint32_t)pp;
    for (int yy = 0; yy < h; yy++)
        for (int xx = 0; xx < w; xx++) {
            size_t idx = (size_t)yy * w + xx;
            v3 mean;
            render_pixel(s, width, height, pp, rng_mode, seed, x0 + xx, y0 + yy, s_begin, s_end,
                         sum ? sum + 3 * idx : NULL, sumsq ? sumsq + 3 * idx : NULL, &mean);
            if (mean_f) { mean_f[3 * idx] = mean.x; mean_f[3 * idx + 1] = mean.y; mean_f[3 * idx + 2] = mean.z; }
            if (ldr) ldr[idx] = resolve_pixel(mean);
        }
}

/* ---------------------------------------------------------------- threaded whole-frame render (PHILOX) */
typedef struct {
    const pt_scene *s;
    int width, height, pp, y_next_chunk;
    uint64_t seed;
    float *mean_f;
    uint32_t *ldr;
    int *next_row;
    pthread_mutex_t *lock;
} job_t;

static void *worker(void *arg) {
    job_t *j = (job_t *)arg;
    for (;;) {
        pthread_mutex_lock(j->lock);
        int y = *j->next_row;
        *j->next_row = y + 4;
        pthread_mutex_unlock(j->lock);
        if (y >= j->height) break;
        int rows = j->height - y < 4 ? j->height - y : 4;
        pto_render(j->s, j->width, j->height, j->pp, PTO_RNG_PHILOX, j->seed, 0, y, j->width, rows, 0, 0, NULL, NULL,
                   j->mean_f ? j->mean_f + 3 * (size_t)y * j->width : NULL, j->ldr ? j->ldr + (size_t)y * j->width : NULL);
    }
    return NULL;
}

double pto_render_threads(const pt_scene *s, int width, int height, int pp, uint64_t seed, int threads,
                          float *mean_f, uint32_t *ldr) {
    if (threads < 1) threads = 1;
    if (threads > 1024) threads = 1024;
    pthread_t *tid = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)threads);
    pthread_mutex_t lock;
    pthread_mutex_init(&lock, NULL);
    int next_row = 0;
    job_t job = {s, width, height, pp, 0, seed, mean_f, ldr, &next_row, &lock};
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int i = 0; i < threads; i++) pthread_create(&tid[i], NULL, worker, &job);
    for (int i = 0; i < threads; i++) pthread_join(tid[i], NULL);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    pthread_mutex_destroy(&lock);
    free(tid);
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}
