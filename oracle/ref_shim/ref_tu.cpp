/*
 * oracle/_ref harness: the reference renderer itself, compiled UNMODIFIED from
 * /root/reference/src/win32_main.cpp (included below where it lies; no reference
 * source is copied into this repository), plus exports that let the tests drive
 * the reference's own statics (RayCastIntersect, RayCast, RenderTexel,
 * MasterThread, LoadWorld, DefineCamera).
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing here is linked into libpt_b200.so; only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load what this file builds (oracle/_ref/libref_harness.so,
 * oracle/_ref/pathtracer_ref).
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <float.h>
#include <assert.h>
#include <stdint.h>
#include <stddef.h>
#include <stdarg.h>
#include <limits.h>
#include <sys/types.h>
#include <sys/stat.h>
#include <unistd.h>
#include <fcntl.h>
#include <malloc.h>
#include <algorithm>
#include <random>
#include <memory>
#include <atomic>
#include <vector>
#include <map>

/* glibc's <sys/types.h> already owns quad_t, <signal.h> owns stack_t */
#define quad_t  ref_quad_t
#define stack_t ref_stack_t
#define main    ref_main
#include "win32_main.cpp"      /* found via -I/root/reference/src : the unmodified reference TU */
#undef main

int    __argc = 0;
char **__argv = nullptr;

extern "C" {
volatile int64_t pt_shim_first_worker_ns = 0;
volatile int64_t pt_shim_last_join_ns = 0;
int pt_shim_core_count = 0;
}

/* ------------------------------------------------------------------------- */
/* headless Inf-Forge: the window is a size, the blit is a no-op              */
/* (reference: src/inf_forge_win.c:127-219, 279-345)                          */
/* ------------------------------------------------------------------------- */
static int g_shim_client_w = 1280, g_shim_client_h = 720;
extern "C" {
IF_window_handle_t IF_create_window_ex(IF_create_window_info_t *info) {
    IF_window_handle_t h; h.opaque_handle = (void *)1;
    const char *e = getenv("PT_REF_SIZE");           /* "WxH"; default = what main() asks for */
    int w = info->width, hh = info->height;
    if (e) sscanf(e, "%dx%d", &w, &hh);
    g_shim_client_w = w; g_shim_client_h = hh;
    return h;
}
IF_rect_t IF_get_window_clientarea(IF_window_handle_t) {
    IF_rect_t r; r.width = g_shim_client_w; r.height = g_shim_client_h; r.x = 0; r.y = 0; return r;
}
IF_BOOL IF_win_poll_message(IF_window_handle_t, IF_winmsg_t *msg) { msg->isvalid = 0; usleep(1000); return 1; }
void IF_blit_to_window_surface(IF_window_handle_t, void *, IF_texture_info_t *) {}
void IF_close_window(IF_window_handle_t) {}
}

/* ------------------------------------------------------------------------- */
/* harness state: one loaded copy of each world per process                   */
/* (GenerateAccelerationStructure keeps a function-static scratch pointer it  */
/*  frees, main.cpp:1208/1444, so world 5 can be built only once)             */
/* ------------------------------------------------------------------------- */
struct saved_world {
    bool valid = false;
    unsigned scene_seed = 0;
    world_t world;
    camera_t cam_user;            /* as LoadWorld left it (before DefineCamera) */
    mipchain_t tex[4];
};
static saved_world g_saved[WORLD_KIND_COUNT];

static void silence_stdout(int *saved_fd) {
    fflush(stdout);
    *saved_fd = dup(1);
    int nul = open("/dev/null", O_WRONLY);
    dup2(nul, 1);
    close(nul);
}
static void restore_stdout(int saved_fd) {
    fflush(stdout);
    dup2(saved_fd, 1);
    close(saved_fd);
}

extern "C" {

void ref_seed(unsigned s) { gen.seed(s); dis.reset(); }

void ref_set_core_count(int c) { pt_shim_core_count = c; }

/* flags: bit0 = -d (thin lens), bit1 = -n, bit2 = -m, bit3 = -r            (main.cpp:2173-2184) */
int ref_load(const char *run_dir, int world1, int W, int H, int pp, int flags, unsigned scene_seed)
{
    if (world1 < 1 || world1 > (int)WORLD_KIND_COUNT) return -1;
    /* GenerateAccelerationStructure never zeroes its 21 MB scratch grid (main.cpp:1208) and only works
       when malloc hands back fresh zero pages; pin the mmap threshold so big blocks always are. */
    static bool s_mallopt = false;
    if (!s_mallopt) { mallopt(M_MMAP_THRESHOLD, 1 << 20); s_mallopt = true; }
    if (run_dir && run_dir[0] && chdir(run_dir) != 0) return -2;
    world_kind_t kind = (world_kind_t)(world1 - 1);

    g_sc = 8;
    g_pp = pp;
    g_tc = 1;
    g_use_pinhole = !(flags & 1);
    g_bNormals    = !(flags & 2);
    g_bMetalness  = !(flags & 4);
    g_bRoughness  = !(flags & 8);
    g_worldKind   = kind;

    if (g_image.pixelPointer) free(g_image.pixelPointer);
    g_image = AllocateImage(W, H);
    memset(g_image.pixelPointer, 0, (size_t)W * H * 4);

    saved_world &s = g_saved[kind];
    if (s.valid && s.scene_seed != scene_seed && kind == WORLD_RAYTRACING_ONE_WEEKEND) s.valid = false;
    if (!s.valid) {
        g_materials = nullptr; g_lights = nullptr; g_planes = nullptr; g_spheres = nullptr;
        g_quads = nullptr; g_meshes = nullptr; g_aabbs = nullptr;
        memset(g_textures, 0, sizeof(g_textures));
        g_world = {};
        g_camera = {};
        ref_seed(scene_seed);
        LoadWorld(kind, &g_camera);
        s.valid = true;
        s.scene_seed = scene_seed;
        s.world = g_world;
        s.cam_user = g_camera;
        memcpy(s.tex, g_textures, sizeof(g_textures));
    }
    g_world = s.world;
    g_materials = g_world.materials; g_lights = g_world.lights; g_planes = g_world.planes;
    g_spheres = g_world.spheres; g_quads = g_world.quads; g_meshes = g_world.meshes; g_aabbs = g_world.aabbs;
    memcpy(g_textures, s.tex, sizeof(g_textures));
    g_camera = s.cam_user;
    /* LoadWorld: c->use_pinhole = g_use_pinhole (main.cpp:1801), world 4 forces the lens (main.cpp:2030) */
    g_camera.use_pinhole = (kind == WORLD_RAYTRACING_ONE_WEEKEND) ? false : g_use_pinhole;

    int fd; silence_stdout(&fd);
    DefineCamera(&g_camera);
    restore_stdout(fd);
    return 0;
}

/* counts: [0] materials [1] spheres [2] planes [3] quads [4] aabbs [5] mesh points [6] textures loaded */
void ref_get_counts(int *c)
{
    c[0] = nc_sbcount(g_world.materials);
    c[1] = nc_sbcount(g_world.spheres);
    c[2] = nc_sbcount(g_world.planes);
    c[3] = nc_sbcount(g_world.quads);
    c[4] = nc_sbcount(g_world.aabbs);
    c[5] = nc_sbcount(g_world.meshes) ? (int)g_world.meshes[0].pointCount : 0;
    int nt = 0;
    for (int i = 0; i < 4; i++) if (g_textures[i].mips) nt++;
    c[6] = nt;
    c[7] = (int)sizeof(camera_t);
}

int ref_sizeof(int what)
{
    switch (what) {
        case 0: return (int)sizeof(material_t);
        case 1: return (int)sizeof(sphere_t);
        case 2: return (int)sizeof(plane_t);
        case 3: return (int)sizeof(ref_quad_t);
        case 4: return (int)sizeof(aabb_t);
        case 5: return (int)sizeof(camera_t);
    }
    return -1;
}

/* raw copies of the reference's own PODs */
void ref_get_scene(void *materials, void *spheres, void *planes, void *quads, void *aabbs,
                   float *points, int *point_mats, void *camera)
{
    int c[8]; ref_get_counts(c);
    if (materials) memcpy(materials, g_world.materials, (size_t)c[0] * sizeof(material_t));
    if (spheres)   memcpy(spheres,   g_world.spheres,   (size_t)c[1] * sizeof(sphere_t));
    if (planes)    memcpy(planes,    g_world.planes,    (size_t)c[2] * sizeof(plane_t));
    if (quads)     memcpy(quads,     g_world.quads,     (size_t)c[3] * sizeof(ref_quad_t));
    if (aabbs)     memcpy(aabbs,     g_world.aabbs,     (size_t)c[4] * sizeof(aabb_t));
    if (c[5]) {
        if (points)     memcpy(points, g_world.meshes[0].points, (size_t)c[5] * sizeof(v3));
        if (point_mats) memcpy(point_mats, g_world.meshes[0].matIndices, (size_t)c[5] * sizeof(int));
    }
    if (camera) memcpy(camera, &g_camera, sizeof(camera_t));
}

/* texture idx 0..3 mip 0 (float3 texels as the reference holds them, main.cpp:1733-1740) */
int ref_get_texture(int idx, int *w, int *h, const float **data)
{
    if (idx < 0 || idx > 3 || !g_textures[idx].mips) return -1;
    texture_t t = g_textures[idx].mips[0];
    *w = t.width; *h = t.height; *data = (const float *)t.data;
    return 0;
}

/* ---- closest hit through the reference's RayCastIntersect; primitive id recovered by
        re-running the reference's own per-primitive functions in its order (main.cpp:418-553) ---- */
/* kind: 0 miss, 1 sphere, 2 quad, 3 plane, 4 triangle, 5 aabb */
void ref_intersect(int n, const float *o, const float *d, float *t_out, unsigned *mat_out,
                   float *normal_out, int *kind_out, int *index_out)
{
    const float minHit = MIN_HIT_DISTANCE;
    for (int r = 0; r < n; r++) {
        v3 ro = V3(o[3*r], o[3*r+1], o[3*r+2]);
        v3 rd = V3(d[3*r], d[3*r+1], d[3*r+2]);
        ray_payload_t p = RayCastIntersect(&g_world, ro, rd);
        if (t_out) t_out[r] = p.hitDistance;
        if (mat_out) mat_out[r] = p.hitMatIndex;
        if (normal_out) { normal_out[3*r] = p.normal.x; normal_out[3*r+1] = p.normal.y; normal_out[3*r+2] = p.normal.z; }
        if (!kind_out) continue;

        float best = FLT_MAX; int kind = 0, index = -1;
        for (unsigned i = 0; i < (unsigned)nc_sbcount(g_world.spheres); i++) {
            v3 N; float t = RaySphereIntersect(ro, rd, minHit, g_world.spheres[i], &N);
            if (t > minHit && t < best) { best = t; kind = 1; index = (int)i; }
        }
        for (unsigned i = 0; i < (unsigned)nc_sbcount(g_world.quads); i++) {
            ref_quad_t q = g_world.quads[i];
            float mh = 0.02;
            float t = RayIntersectPlanarShape<PLANAR_QUAD>(ro, rd, mh, q.point, q.u, q.v);
            if (t > mh && t < best) { best = t; kind = 2; index = (int)i; }
        }
        for (unsigned i = 0; i < (unsigned)nc_sbcount(g_world.planes); i++) {
            plane_t pl = g_world.planes[i];
            float t = RayIntersectPlane(ro, rd, pl.n, pl.d, minHit);
            if (t > minHit && t < best) { best = t; kind = 3; index = (int)i; }
        }
        if (nc_sbcount(g_world.meshes)) {
            mesh_t &mesh = g_world.meshes[0];
            for (unsigned tri = 0; tri * 3 < mesh.pointCount; tri++) {
                v3 A = mesh.points[tri*3], B = mesh.points[tri*3+1], C = mesh.points[tri*3+2];
                float t = RayIntersectPlanarShape<PLANAR_TRIANGLE>(ro, rd, minHit, A, B - A, C - A);
                if (t > minHit && t < best) { best = t; kind = 4; index = (int)tri; }
            }
        }
        for (unsigned i = 0; i < (unsigned)nc_sbcount(g_world.aabbs); i++) {
            bool early; int face;
            float t = RayIntersectWithAABB2(ro, rd, minHit, g_world.aabbs[i], &early, &face);
            if (t > minHit && t < best) { best = t; kind = 5; index = (int)i; }
        }
        /* the recovered nearest t must be the one RayCastIntersect reported (octree == brute force) */
        if (best != p.hitDistance) kind = -kind - 100;
        kind_out[r] = kind;
        index_out[r] = index;
    }
}

/* radiance of arbitrary rays through the reference's RayCast, consuming the global RNG in order */
void ref_raycast(int n, const float *o, const float *d, float *rgb)
{
    for (int r = 0; r < n; r++) {
        v3 c = RayCast(&g_world, V3(o[3*r], o[3*r+1], o[3*r+2]), V3(d[3*r], d[3*r+1], d[3*r+2]), 0);
        rgb[3*r] = c.x; rgb[3*r+1] = c.y; rgb[3*r+2] = c.z;
    }
}

/* jitter-free camera ray through the lower-left corner of every pixel
   (film mapping of main.cpp:1003-1006 and 1059-1064 with the random offsets set to zero) */
void ref_primary_rays(float *o, float *d)
{
    const camera_t &c = g_camera;
    for (unsigned y = 0; y < g_image.height; y++)
    for (unsigned x = 0; x < g_image.width; x++) {
        const float fy = -1.0f + 2.0f * (float)y / (float)g_image.height;
        const float fx = -1.0f + 2.0f * (float)x / (float)g_image.width;
        v3 P = c.frustrumCenter + (fx * c.halfFilmWidth * c.axisX) + (fy * c.halfFilmHeight * c.axisY);
        v3 dir = Normalize(P - c.pos);
        size_t i = (size_t)y * g_image.width + x;
        o[3*i] = c.pos.x; o[3*i+1] = c.pos.y; o[3*i+2] = c.pos.z;
        d[3*i] = dir.x;   d[3*i+1] = dir.y;   d[3*i+2] = dir.z;
    }
}

/* the reference's own per-tile entry point (the drop-in seam, main.cpp:45, 990) */
void ref_render_texel(int x0, int y0, int w, int h, unsigned *out)
{
    texel_t t; t.width = w; t.height = h; t.xPos = x0; t.yPos = y0;
    unsigned *base = g_image.pixelPointer + (size_t)y0 * g_image.width + x0;
    RenderTexel(base, t);
    for (int y = 0; y < h; y++)
        memcpy(out + (size_t)y * w, g_image.pixelPointer + (size_t)(y0 + y) * g_image.width + x0, (size_t)w * 4);
}

/*
 * HDR statistics around the reference's RayCast.  The two sampling loops are the harness's
 * restatement of main.cpp:1041-1074 (pinhole) and 1114-1169 (thin lens); with the same seed they
 * draw the same numbers in the same order, so `ldr` equals what ref_render_texel writes (tested).
 * sum / sumsq: per pixel, per channel, in double, over the spp = pp*pp accepted samples.
 * mean_f: the float accumulation the reference itself performs (color += contrib*radiance).
 */
void ref_render_hdr(int x0, int y0, int w, int h, double *sum, double *sumsq, float *mean_f, unsigned *ldr)
{
    static const float disk[12][2] = {
        {0.0f, 0.0f}, {-0.94201624f, -0.39906216f}, {0.94558609f, -0.76890725f}, {-0.094184101f, -0.92938870f},
        {0.34495938f, 0.29387760f}, {-0.91588581f, 0.45771432f}, {-0.81544232f, -0.87912464f},
        {-0.38277543f, 0.27676845f}, {0.97484398f, 0.75648379f}, {0.44323325f, -0.97511554f},
        {0.53742981f, -0.47373420f}, {-0.26496911f, -0.41893023f}};
    const camera_t &c = g_camera;
    const int pp = g_pp;
    for (int yy = 0; yy < h; yy++)
    for (int xx = 0; xx < w; xx++) {
        unsigned x = x0 + xx, y = y0 + yy;
        const float fy = -1.0f + 2.0f * (float)y / (float)g_image.height;
        const float fx = -1.0f + 2.0f * (float)x / (float)g_image.width;
        double s[3] = {0, 0, 0}, s2[3] = {0, 0, 0};
        v3 color = {};
        float contrib = 1.0f / (float)pp / (float)pp;
        if (c.use_pinhole) {
            for (int i = 0; i < pp; i++)
            for (int j = 0; j < pp; j++) {
                float llx = fx - 1.f * c.halfFilmPixelW;
                float lly = fy - 1.f * c.halfFilmPixelH;
                float stepX = 1.f / pp * c.halfFilmPixelW * 2.f;
                float stepY = 1.f / pp * c.halfFilmPixelH * 2.f;
                float xs = llx + float(i) / pp * c.halfFilmPixelW + stepX * 0.5f;
                float ys = lly + float(j) / pp * c.halfFilmPixelH + stepY * 0.5f;
                xs += (RandomUnilateral() - 0.5f) * stepX;
                ys += (RandomUnilateral() - 0.5f) * stepY;
                v3 P = c.frustrumCenter + (xs * c.halfFilmWidth * c.axisX) + (ys * c.halfFilmHeight * c.axisY);
                v3 dir = Normalize(P - c.pos);
                v3 L = RayCast(&g_world, c.pos, dir, 0);
                if (IsNaN(L)) { j--; continue; }
                color = color + contrib * L;
                for (int k = 0; k < 3; k++) { s[k] += L.E[k]; s2[k] += (double)L.E[k] * L.E[k]; }
            }
        } else {
            for (unsigned ri = 0; ri < (unsigned)pp; ri++) {
                float offX = fx + (RandomBilateral() * c.halfFilmPixelW);
                float offY = fy + (RandomBilateral() * c.halfFilmPixelH);
                v3 P = c.frustrumCenter + (offX * c.halfFilmWidth * c.axisX) + (offY * c.halfFilmHeight * c.axisY);
                v3 dir = Normalize(P - c.pos);
                float fpd = 1.f / (1.f / FIXED_FOCAL_LENGTH - 1.f / c.focalLength);
                v3 N = -c.axisZ;
                v3 planePoint = c.pos + c.axisX + fpd * N;
                float dd = Dot(N, planePoint);
                float t = RayIntersectPlane(c.pos, dir, N, dd, MIN_HIT_DISTANCE);
                v3 focal = c.pos + t * dir;
                for (unsigned r2 = 0; r2 < (unsigned)pp; r2++) {
                    const float *ds = disk[(r2 * ri) % 12];
                    v3 ro = c.pos + ds[0] * c.aperatureRadius * c.axisX + ds[1] * c.aperatureRadius * c.axisY;
                    v3 rdir = Normalize(focal - ro);
                    v3 L = RayCast(&g_world, ro, rdir, 0);
                    if (IsNaN(L)) { r2--; continue; }
                    color = color + contrib * L;
                    for (int k = 0; k < 3; k++) { s[k] += L.E[k]; s2[k] += (double)L.E[k] * L.E[k]; }
                }
            }
        }
        size_t idx = (size_t)yy * w + xx;
        for (int k = 0; k < 3; k++) {
            if (sum) sum[3*idx+k] = s[k];
            if (sumsq) sumsq[3*idx+k] = s2[k];
            if (mean_f) mean_f[3*idx+k] = color.E[k];
        }
        if (ldr) {
            v3 tm = TonemapPass(color);
            v4 px = {255.0f * LinearToSRGB(tm.r), 255.0f * LinearToSRGB(tm.g), 255.0f * LinearToSRGB(tm.b), 255.0f};
            ldr[idx] = BGRAPack4x8(px);
        }
    }
}

/* resolve of one HDR colour through the reference's own TonemapPass / LinearToSRGB / BGRAPack4x8 */
unsigned ref_resolve(float r, float g, float b)
{
    v3 tm = TonemapPass(V3(r, g, b));
    v4 px = {255.0f * LinearToSRGB(tm.r), 255.0f * LinearToSRGB(tm.g), 255.0f * LinearToSRGB(tm.b), 255.0f};
    return BGRAPack4x8(px);
}

/*
 * Whole-frame render through the reference's own scheduler (MasterThread -> RenderThread ->
 * RenderTexel, main.cpp:851-987) on `threads` workers.  Returns seconds from the first worker's
 * creation to the last worker's join; pixels = g_image (row 0 = bottom).
 * Writes test.bmp into the run directory exactly as the reference does (main.cpp:984).
 */
double ref_render_threads(int threads, unsigned *pixels)
{
    g_tc = threads < 1 ? 1 : (threads > MAX_THREAD_COUNT ? MAX_THREAD_COUNT : threads);
    pt_shim_first_worker_ns = 0;
    pt_shim_last_join_ns = 0;
    int fd; silence_stdout(&fd);
    HANDLE m = CreateThread(nullptr, 0, MasterThread, nullptr, 0, nullptr);
    WaitForSingleObject(m, INFINITE);
    restore_stdout(fd);
    if (pixels) memcpy(pixels, g_image.pixelPointer, (size_t)g_image.width * g_image.height * 4);
    return (double)(pt_shim_last_join_ns - pt_shim_first_worker_ns) * 1e-9;
}

/* raw mt19937 draws through the reference's RandomUnilateral (math.hpp:250-259) */
void ref_random(int n, float *out) { for (int i = 0; i < n; i++) out[i] = RandomUnilateral(); }

/* small helpers on the hot path, exposed one by one for unit-level parity checks */
void ref_sample_texture(int idx, float u, float v, float *rgb)
{
    v3 c = BespokeSampleTexture(g_textures[idx].mips[0], v2{u, v});
    rgb[0] = c.x; rgb[1] = c.y; rgb[2] = c.z;
}

/* the reference's box primitive (RayIntersectWithAABB2, ray_math.hpp:398-482, boxes from MakeAABB :484-492): none of
   the five worlds holds one, so it is tapped directly.  Per ray: nearest accepted box as RayCastIntersect's loop does
   (main.cpp:529-553): t > minHitDistance && t < best, in box order; outputs t (FLT_MAX = miss), box index, face. */
void ref_aabb_test(int n_rays, const float *o, const float *d, int n_boxes, const float *origins, const float *half_dims,
                   float *t_out, int *index_out, int *face_out)
{
    const float minHit = MIN_HIT_DISTANCE;
    for (int r = 0; r < n_rays; r++) {
        v3 ro = V3(o[3*r], o[3*r+1], o[3*r+2]), rd = V3(d[3*r], d[3*r+1], d[3*r+2]);
        float best = FLT_MAX; int bi = -1, bf = -1;
        for (int b = 0; b < n_boxes; b++) {
            aabb_t box = MakeAABB(V3(origins[3*b], origins[3*b+1], origins[3*b+2]), V3(half_dims[3*b], half_dims[3*b+1], half_dims[3*b+2]));
            bool early; int face = -1;
            float t = RayIntersectWithAABB2(ro, rd, minHit, box, &early, &face);
            if ((t > minHit) && (t < best)) { best = t; bi = b; bf = face; }
        }
        t_out[r] = best; index_out[r] = bi; face_out[r] = bf;
    }
}

} /* extern "C" */

#ifdef PT_REF_B200
/* ------------------------------------------------------------------------- */
/* pathtracer_ref_b200: THE DROP-IN, demonstrated.  The reference program -- its ParseArgs, LoadWorld,     */
/* GLB/texture loaders, DefineCamera and WriteDIBImage, compiled unmodified above -- with ONE routine      */
/* substituted: the thread shim hands CreateThread(MasterThread) (main.cpp:239-246) this function instead, */
/* which is the patch of INTEGRATION.md section 1 verbatim: g_world/g_camera/g_textures -> pt_upload_scene  */
/* -> pt_render -> pt_readback into g_image.pixelPointer, then the reference's own tail (main.cpp:983-986). */
/* Extra knob: PT_SEED=n (the reference seeds from random_device).                                         */
/* ------------------------------------------------------------------------- */
#include "pt_api.h"

static DWORD WINAPI MasterThread_b200(LPVOID)
{
    static_assert(sizeof(material_t) == sizeof(pt_material) && sizeof(sphere_t) == sizeof(pt_sphere) &&
                  sizeof(plane_t) == sizeof(pt_plane) && sizeof(ref_quad_t) == sizeof(pt_quad) &&
                  sizeof(aabb_t) == sizeof(pt_aabb) && sizeof(camera_t) == sizeof(pt_camera), "pt_api.h mirrors ray.hpp");

    pt_config cfg = { sizeof(pt_config), /*device*/0, PT_BVH_HOST_SAH, PT_KERNEL_AUTO, /*stream*/nullptr };
    pt_context *pt = nullptr;
    if (pt_create(&cfg, &pt) != PT_OK) { fprintf(stderr, "pt: %s\n", pt_last_error(nullptr)); exit(1); }

    pt_scene sc = {};
    sc.struct_size = sizeof(pt_scene);
    sc.flags = (g_bNormals ? PT_SCENE_NORMAL_MAPS : 0) | (g_bMetalness ? PT_SCENE_METALNESS_MAPS : 0) |
               (g_bRoughness ? PT_SCENE_ROUGHNESS_MAPS : 0) |
               (g_worldKind == WORLD_RAYTRACING_ONE_WEEKEND ? PT_SCENE_JUST_COSINE : 0);      /* main.cpp:654-655 */
    sc.n_materials = nc_sbcount(g_world.materials);  sc.materials = (const pt_material *)g_world.materials;
    sc.n_spheres   = nc_sbcount(g_world.spheres);    sc.spheres   = (const pt_sphere *)g_world.spheres;
    sc.n_planes    = nc_sbcount(g_world.planes);     sc.planes    = (const pt_plane *)g_world.planes;
    sc.n_quads     = nc_sbcount(g_world.quads);      sc.quads     = (const pt_quad *)g_world.quads;
    sc.n_aabbs     = nc_sbcount(g_world.aabbs);      sc.aabbs     = (const pt_aabb *)g_world.aabbs;

    /* meshes[0] only (main.cpp:483): 3 consecutive v3 per triangle, material = that of its first vertex (main.cpp:520) */
    int *triMat = nullptr;
    if (nc_sbcount(g_world.meshes)) {
        mesh_t &m = g_world.meshes[0];
        sc.n_triangles = m.pointCount / 3;
        sc.tri_points  = (const float *)m.points;
        triMat = (int *)malloc(sizeof(int) * (sc.n_triangles ? sc.n_triangles : 1));
        for (unsigned t = 0; t < sc.n_triangles; t++) triMat[t] = m.matIndices[3 * t];
        sc.tri_materials = triMat;
    }

    /* g_textures[i].mips[0] holds float3 = byte/255.f (main.cpp:1737-1739); the library wants the bytes back */
    unsigned char *texBytes[4] = {};
    for (int i = 0; i < 4; i++) if (g_textures[i].mips && g_textures[i].mips[0].data) {
        texture_t t = g_textures[i].mips[0];
        size_t n = (size_t)t.width * t.height;
        texBytes[i] = (unsigned char *)malloc(n * 4);
        for (size_t p = 0; p < n; p++) {
            texBytes[i][4*p+0] = (unsigned char)(t.data[p].x * 255.f + 0.5f);
            texBytes[i][4*p+1] = (unsigned char)(t.data[p].y * 255.f + 0.5f);
            texBytes[i][4*p+2] = (unsigned char)(t.data[p].z * 255.f + 0.5f);
            texBytes[i][4*p+3] = 255;
        }
        sc.textures[i].width = t.width; sc.textures[i].height = t.height;
        sc.textures[i].rgba8 = texBytes[i];
    }
    memcpy(&sc.camera, &g_camera, sizeof(pt_camera));             /* camera_t, ray.hpp:176-186: same field order */

    int rc = pt_upload_scene(pt, &sc);                            /* copies: host pointers may go */
    for (int i = 0; i < 4; i++) free(texBytes[i]);
    free(triMat);

    pt_render_params rp = {};
    rp.struct_size = sizeof(rp);
    rp.width = g_image.width; rp.height = g_image.height;
    rp.pp = g_pp;                                                  /* main.cpp:113 */
    rp.seed = getenv("PT_SEED") ? strtoull(getenv("PT_SEED"), nullptr, 10) : (uint64_t)std::random_device{}();
    rp.flags = PT_RENDER_CLEAR;
    if (rc == PT_OK) rc = pt_render(pt, &rp);
    if (rc == PT_OK) rc = pt_readback(pt, (uint32_t *)g_image.pixelPointer, nullptr, nullptr);   /* row 0 = bottom */
    if (rc != PT_OK) { fprintf(stderr, "pt: %s\n", pt_last_error(pt)); exit(1); }
    pt_destroy(pt);

    WriteDIBImage(g_image, "test.bmp");                            /* main.cpp:983-986 */
    printf("Done. Image written to test.bmp.\n");
    ExitThread(0);
}

static LPTHREAD_START_ROUTINE b200_thread_hook(LPTHREAD_START_ROUTINE fn) { return fn == MasterThread ? MasterThread_b200 : fn; }
#endif

#ifdef PT_REF_MAIN
/* pathtracer_ref: the reference program, CLI and all (ParseArgs -> ... -> test.bmp).
   Extra environment knobs of the shim only: PT_REF_SIZE=WxH, PT_REF_SEED=n, PT_REF_CORES=n. */
int main(int argc, char **argv)
{
    __argc = argc;
    __argv = argv;
    mallopt(M_MMAP_THRESHOLD, 1 << 20);   /* see ref_load */
    if (const char *s = getenv("PT_REF_SEED")) ref_seed((unsigned)strtoul(s, nullptr, 10));
    if (const char *s = getenv("PT_REF_CORES")) pt_shim_core_count = atoi(s);
#ifdef PT_REF_B200
    pt_shim_thread_hook = b200_thread_hook;
#endif
    int rc = ref_main();
    if (pt_shim_first_worker_ns && pt_shim_last_join_ns) {
        double sec = (double)(pt_shim_last_join_ns - pt_shim_first_worker_ns) * 1e-9;
        double samples = (double)g_image.width * g_image.height * g_pp * g_pp;
        fprintf(stderr, "[pt_ref] %ux%u pp=%d threads=%d render=%.6f s  %.4f Msamples/s\n",
                g_image.width, g_image.height, g_pp, g_tc, sec, samples / sec * 1e-6);
    }
    return rc;
}
#endif
