// Device-side scene layout and small math helpers of the B200 render core (sm_100a only).
// No tensor cores anywhere: nothing on this path is a dense contraction (BASELINE.json north_star).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "pt_api.h"

#define PT_MAX_BOUNCE_COUNT 4          // src/win32_main.cpp:86
#define PT_MIN_HIT 1e-4f               // src/win32_main.cpp:90
#define PT_QUAD_MIN_HIT 0.02f          // src/win32_main.cpp:446
#define PT_TOLERANCE 1e-9f             // include/ray_math.hpp:10
#define PT_PI 3.14159265358979323846f
#define PT_MAX_NAN_RETRIES 8           // the reference redraws NaN samples without bound (main.cpp:1068, 1166)
#define PT_TILE 32                     // THREAD_GROUP_SIZE, src/win32_main.cpp:88
#define PT_STACK_DEPTH 48

// primitive id = kind << 28 | index ; kinds ordered as the reference tests them (main.cpp:418-553),
// so "smaller id wins an exact tie" reproduces the strict '<' of the reference loops
#define PT_ID(kind, index) (((uint32_t)(kind) << 28) | (uint32_t)(index))
#define PT_ID_MISS 0xFFFFFFFFu

// ---- BVH2 node, 64 bytes = four 16-byte vector loads -------------------------------------------
//   b0 = (L.min.x, L.min.y, L.min.z, L.max.x)   b1 = (L.max.y, L.max.z, R.min.x, R.min.y)
//   b2 = (R.min.z, R.max.x, R.max.y, R.max.z)   link = (left, right, 0, 0)
//   child >= 0: inner node index; child < 0: leaf, ~child = count << 24 | first  (count <= 127)
struct __align__(16) PtBvhNode {
    float4 b0, b1, b2;
    int4 link;
};

// Triangle / quad record, 48 bytes = three 16-byte loads.  With N = u x v, w = N/(N.N) the reference's
// alpha = w.(p x v), beta = w.(u x p) (ray_math.hpp:369-372) are the triple products p.(v x w), p.(w x u):
//   r0 = (Nunit, d = A.Nunit)   r1 = (v x w, -A.(v x w))   r2 = (w x u, -A.(w x u))
struct __align__(16) PtPlanar {
    float4 r0, r1, r2;
};

struct PtBvh {
    const PtBvhNode *nodes;   // nodes[0] is the root (always an inner node)
    uint32_t n_nodes;
    uint32_t n_prims;
};

struct PtTexture {
    const uchar4 *texels;     // RGBA8, row-major
    const float *unit255;     // 256 floats: (float)i / 255.f computed by IEEE division on the host (main.cpp:1737-1739)
    int w, h;
    int wmask, hmask;         // w-1 / h-1 when a power of two (x % w == x & mask), else 0
};

// everything the kernels read; passed by value as a __grid_constant__ kernel parameter
struct PtDeviceScene {
    // linear ("analytic") lists, tested one by one: all planes, quads, boxes, and the spheres that are
    // either few or too large to be useful in a BVH (the r=1000 ground and sun spheres)
    const float4 *lin_spheres;      // (c.xyz, r)
    const uint32_t *lin_sphere_ids; // original sphere index
    uint32_t n_lin_spheres;
    const float4 *planes;           // (n.xyz, d)
    uint32_t n_planes;
    const PtPlanar *quads;
    uint32_t n_quads;
    const float4 *aabbs;            // 2 per box: (min.xyz, 0), (max.xyz, 0)
    uint32_t n_aabbs;

    // BVH over the remaining spheres; leaves index bvh_spheres (ids map back to original indices)
    PtBvh sphere_bvh;
    const float4 *bvh_spheres;
    const uint32_t *bvh_sphere_ids;

    // BVH over all triangles; leaves index tris (ids map back to original triangle indices)
    PtBvh tri_bvh;
    const PtPlanar *tris;
    const uint32_t *tri_ids;
    const float4 *tri_normals;      // unit normal by ORIGINAL triangle index (main.cpp:508)

    // material table + per-primitive material index
    const pt_material *materials;
    uint32_t n_materials;
    const uint32_t *sphere_mat, *plane_mat, *quad_mat, *aabb_mat, *tri_mat;   // indexed by ORIGINAL index
    const float4 *all_spheres;      // every sphere by original index (normals, light sampling)
    uint32_t n_spheres;

    PtTexture tex[4];
    uint32_t flags;                 // PT_SCENE_*
    float4 light;                   // spheres[0] (c.xyz, r): the importance-sampled light (main.cpp:683)
    pt_camera cam;

    // cuts of the two BVHs into <= 32 clusters for the warp-cooperative closest hit (pt_cluster.h; PT_F_COOP
    // variants only; NULL when the tree has no such cut).  [l] = (min.xyz, first | count << 24), [32 + l] = (max.xyz, 0).
    // Kept at the end: the other members' parameter offsets -- and the SASS of the default variants -- do not move.
    const float4 *tri_clusters, *sph_clusters;
    uint32_t n_tri_clusters, n_sph_clusters;
};

// ---- tiles -> ranks ------------------------------------------------------------------------------
// Tile (tx, ty) of the row-major 32x32 tile grid belongs to rank (tx + skew * ty) % count: a DIAGONAL interleave.  (The plain
// "tile id % count" of round 1 degenerates to vertical stripes whenever the tile-row length is a multiple of count -- 1280 and
// 3840 pixels at 8 ranks -- and the per-rank render times then spread by 6-10 %, profiles/r02e_scale_*.)  skew is coprime with
// count, so every row and every run of `count` rows deals its tiles evenly.  A rank's tiles are enumerated by SLOT:
// slot j = ty * slots + m with slots = ceil(tiles_x / count), tx = first(ty) + count * m; a slot whose tx falls outside the
// grid is empty (the kernels' own x < width test skips it).  Every rank has tiles_y * slots slots: packs are equally long.
struct PtTileMap {
    uint32_t tiles_x, tiles_y, count, skew, slots;      // slots per tile row; n_slots = tiles_y * slots
};
__host__ __device__ inline uint32_t pt_tile_skew(uint32_t count) {
    if (count <= 1) return 0;
    uint32_t k = count / 2 + 1;
    for (;; k++) {                                       // smallest k >= count/2 + 1 coprime with count
        uint32_t a = k, b = count;
        while (b) { uint32_t t = a % b; a = b; b = t; }
        if (a == 1) return k % count;
    }
}
__host__ __device__ inline PtTileMap pt_tile_map(int width, int height, uint32_t count) {
    PtTileMap m;
    m.tiles_x = (uint32_t)(width + PT_TILE - 1) / PT_TILE;
    m.tiles_y = (uint32_t)(height + PT_TILE - 1) / PT_TILE;
    m.count = count ? count : 1;
    m.skew = pt_tile_skew(m.count);
    m.slots = (m.tiles_x + m.count - 1) / m.count;
    return m;
}
// slot j of `rank` -> tile coordinates; tx >= tiles_x (or ty >= tiles_y) marks an empty slot
__host__ __device__ inline void pt_slot_tile(const PtTileMap &m, uint32_t rank, uint32_t j, uint32_t &tx, uint32_t &ty) {
    ty = j / m.slots;
    const uint32_t k = j - ty * m.slots;
    const uint32_t first = (rank + m.count - (m.skew * ty) % m.count) % m.count;
    tx = first + m.count * k;
}
// tile coordinates -> owning rank and its slot there
__host__ __device__ inline void pt_tile_slot(const PtTileMap &m, uint32_t tx, uint32_t ty, uint32_t &rank, uint32_t &j) {
    rank = (tx + m.skew * ty) % m.count;
    j = ty * m.slots + tx / m.count;
}

struct PtHit {
    float t;
    uint32_t id;     // PT_ID(kind,index) or PT_ID_MISS
};

// ---- float3 helpers -----------------------------------------------------------------------------
__device__ __forceinline__ float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
__device__ __forceinline__ float3 f3(const pt_v3 &v) { return make_float3(v.x, v.y, v.z); }
__device__ __forceinline__ float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ float3 operator*(float s, float3 a) { return f3(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ float dot3(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ float dot3(float4 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ float3 cross3(float3 a, float3 b) {
    return f3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
// Fast divide and reciprocal square root: the bare MUFU.RCP / MUFU.RSQ.  The CUDA intrinsics (__fdividef, rsqrtf) wrap the
// same MUFU in range handling (|y| > 2^126 for the divide, denormal arguments for the rsqrt) -- cases this path never
// produces -- at ~5 instructions per call, 80 instructions of k_render<624>.  Same values on the normal range: every output
// array bit-identical, +0.6 ... +2.2 % per world (tools/ab_identity.py, profiles/r02a_ab_compare_old_accl_mufu.json).
__device__ __forceinline__ float pt_fdiv(float x, float y) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(y));
    return x * r;
}
__device__ __forceinline__ float pt_rsqrt(float x) {
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float3 normalize3(float3 a) {
    float inv = pt_rsqrt(dot3(a, a));
    return inv * a;
}
__device__ __forceinline__ bool isnan3(float3 a) { return (a.x != a.x) || (a.y != a.y) || (a.z != a.z); }

// ---- Philox4x32-10 counter RNG (Salmon et al., SC'11) -----------------------------------------
// Stream layout (DESIGN.md "RNG streams", shared with oracle/pt_oracle.c):
//   key = (seed lo, seed hi); counter = (block, a, b, tag); uniform k of a stream = word k%4 of block k/4
//   path stream      a = pixel index, b = sample index, tag = 0 | retry << 8
//   lens film stream a = pixel index, b = rayIndex,     tag = 1
//   pt_trace_paths   a = ray index,   b = 0,            tag = 2
__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        c0 = hi1 ^ c1 ^ k0;
        c1 = lo1;
        c2 = hi0 ^ c3 ^ k1;
        c3 = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}

// float(x)/2^32 with the >= 1 result clamped to nextafter(1,0): libstdc++'s generate_canonical<float,24>
// over one 32-bit draw, i.e. what the reference's RandomUnilateral returns (ray_math.hpp:245-259)
__device__ __forceinline__ float u32_to_unit(uint32_t x) {
    return fminf(__uint2float_rn(x) * 2.3283064365386963e-10f, 0.99999994f);
}

// One Philox block (4 uniforms) per path vertex: the camera uses block 0 (words 0,1), and every iteration of
// RayCast's sampling loop (main.cpp:660) starts on the next block -- coin, [coin, r1, r2] or [z1, z2].
struct PtRng {
    uint32_t k0, k1, a, b, tag, blk;
    __device__ __forceinline__ void init(uint64_t seed, uint32_t a_, uint32_t b_, uint32_t tag_) {
        k0 = (uint32_t)seed; k1 = (uint32_t)(seed >> 32);
        a = a_; b = b_; tag = tag_; blk = 0;
    }
    // one block of an arbitrary stream of the same seed (the thin-lens film stream)
    __device__ __forceinline__ uint4 block_of(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) const {
        return philox4x32_10(c0, c1, c2, c3, k0, k1);
    }
    __device__ __forceinline__ uint4 next_block() {
        uint4 r = block_of(blk, a, b, tag);
        blk++;
        return r;
    }
};

// ---- compile-time scene features: one k_render instantiation per combination that occurs ---------------
// (smaller code -> fewer instruction-cache misses, fewer registers -> more resident warps)
enum : uint32_t {
    PT_F_TEX = 1u,        // some material references a texture
    PT_F_QUADS = 2u,
    PT_F_AABBS = 4u,
    PT_F_SPH_BVH = 8u,
    PT_F_TRI_BVH = 16u,
    PT_F_LIGHT = 32u,     // importance-sample spheres[0] (everything but PT_SCENE_JUST_COSINE)
    PT_F_ALL = 63u | 128u,
    PT_F_DEFER = 64u,     // kernel policy, not a scene feature: park lanes that need a BRDF sample (pt_kernels.cuh)
    PT_F_COOP = 512u,     // kernel policy: sparse BVH walks (few lanes of the warp) are traced one ray at a time by all 32 lanes
    PT_F_LENS = 128u      // thin-lens camera (-d, always in world 4); pinhole-only variants drop that code
};
