// Kernels of the render core.  One persistent-per-warp "megakernel" with path regeneration does camera
// rays -> closest hit -> shading -> per-pixel accumulation entirely in registers; resolve, parity/debug
// and tile pack/unpack kernels are small.  (The wavefront variant lives in pt_wavefront.cuh.)
#pragma once
#include "pt_device.cuh"
#include "pt_shade.cuh"
#include "pt_trace.cuh"

#ifndef PT_RENDER_THREADS
#define PT_RENDER_THREADS 128
#endif
// CTAs per SM the register allocation aims at, per variant (measured, profiles/r01d_sweep_occupancy.log, r02a_ab_compare_mb7_on_final.json):
// triangles 8 (64 regs, a few spills, +6 %; 7 loses 1.1 %), everything else 7 (72 regs; textures: +2.3 % over 6).
// -DPT_RENDER_MIN_BLOCKS=n overrides it for all variants (tools/build_variants.sh).
__host__ __device__ constexpr int pt_render_min_blocks(uint32_t F) {
#ifdef PT_RENDER_MIN_BLOCKS
    return PT_RENDER_MIN_BLOCKS;
#else
    return (F & PT_F_TRI_BVH) ? 8 : 7;
#endif
}

// k_render / k_render_wf instantiations, from specialised to general: the first whose mask covers the scene runs
#define PT_RENDER_VARIANTS(X) \
    X(PT_F_TEX | PT_F_LIGHT | PT_F_DEFER) X(PT_F_SPH_BVH | PT_F_LIGHT | PT_F_DEFER) X(PT_F_QUADS | PT_F_LIGHT) X(PT_F_SPH_BVH | PT_F_LENS) \
    X(PT_F_TRI_BVH | PT_F_LIGHT | PT_F_DEFER) /* the same families with -d */ X(PT_F_TEX | PT_F_LIGHT | PT_F_DEFER | PT_F_LENS) \
    X(PT_F_SPH_BVH | PT_F_LIGHT | PT_F_LENS | PT_F_DEFER) X(PT_F_QUADS | PT_F_LIGHT | PT_F_LENS) X(PT_F_TRI_BVH | PT_F_LIGHT | PT_F_DEFER | PT_F_LENS) \
    X(PT_F_ALL | PT_F_DEFER)
// megakernel only: the same families with PT_F_COOP (sparse walks traced warp-cooperatively); launched when the scene's
// one tree has a cluster cut and the threshold is > 0 (default: triangle trees, 12 lanes; pt_tuning.coop_max overrides)
#define PT_RENDER_COOP_VARIANTS(X) \
    X(PT_F_TRI_BVH | PT_F_LIGHT | PT_F_DEFER | PT_F_COOP) X(PT_F_SPH_BVH | PT_F_LIGHT | PT_F_DEFER | PT_F_COOP) \
    X(PT_F_SPH_BVH | PT_F_LENS | PT_F_COOP) X(PT_F_TRI_BVH | PT_F_LIGHT | PT_F_DEFER | PT_F_LENS | PT_F_COOP)

#define PT_N_COUNTERS 4             // [0] rays [1] NaN retries [2] samples dropped after PT_MAX_NAN_RETRIES [3] -

struct PtRenderArgs {
    int width, height, pp;
    uint32_t pp_magic;                  // floor(2^32/pp)+1: s / pp == __umulhi(s, pp_magic) for s < pp*pp <= 10^6
    uint32_t s_begin, s_end;            // sample range of the pp*pp list
    uint32_t tile_rank;                 // this launch renders the tiles that tiles.{count, skew} deal to tile_rank (pt_device.cuh)
    PtTileMap tiles;
    uint32_t n_my_tiles;                // = tiles.tiles_y * tiles.slots: the rank's tile SLOTS (some may be empty)
    uint32_t log2_sample_lanes;         // S = lanes of a warp that share one pixel; P = 32/S pixels per warp
    uint32_t log2_pool_pixels;          // wavefront kernel: pixels per warp pool
    uint32_t shade_batch;               // megakernel, PT_F_DEFER variants: parked lanes before the BRDF code runs
    uint64_t seed;
    float4 *accum;                      // width*height: sum r,g,b + number of samples
    unsigned long long *counters;       // [0] rays, [1] NaN retries, [2] samples
    uint32_t coop_max;                  // PT_F_COOP variants: walks entered by <= coop_max lanes are traced cooperatively
};

// ---- camera (RenderTexel's two sampling loops, src/win32_main.cpp:1032-1170) ----------------------
__constant__ float2 c_poisson_disk[12] = {   // main.cpp:1097-1110
    {0.0f, 0.0f}, {-0.94201624f, -0.39906216f}, {0.94558609f, -0.76890725f}, {-0.094184101f, -0.92938870f},
    {0.34495938f, 0.29387760f}, {-0.91588581f, 0.45771432f}, {-0.81544232f, -0.87912464f},
    {-0.38277543f, 0.27676845f}, {0.97484398f, 0.75648379f}, {0.44323325f, -0.97511554f},
    {0.53742981f, -0.47373420f}, {-0.26496911f, -0.41893023f}};

// What a camera ray needs from its pixel, computed once per pixel instead of once per sample (two IEEE divisions
// and 1/pp were ~8 % of a sample's instructions in k_render): the film position of the pixel's lower-left corner.
struct PtPixelCam {
    float fx, fy, inv_pp;
    uint32_t pixel;
};
__device__ __forceinline__ PtPixelCam pt_pixel_cam(int W, int H, int pp, int x, int y) {
    PtPixelCam p;
    p.fy = -1.0f + 2.0f * (float)y / (float)H;      // row 0 = bottom; pixel -> its lower-left corner
    p.fx = -1.0f + 2.0f * (float)x / (float)W;
    p.inv_pp = 1.f / (float)pp;
    p.pixel = (uint32_t)y * (uint32_t)W + (uint32_t)x;
    return p;
}

// Camera ray of sample s of a pixel; leaves rng positioned after the camera's own draws.
template <uint32_t F>
__device__ __forceinline__ void pt_camera_ray(const pt_camera &c, const PtPixelCam &px, int pp, uint32_t pp_magic, uint32_t s,
                                              uint32_t retry, uint64_t seed, PtRng &rng, float3 &o, float3 &d) {
    const float fy = px.fy, fx = px.fx;
    const uint32_t pixel = px.pixel;
    const float3 ax = f3(c.axis_x), ay = f3(c.axis_y), fc = f3(c.frustum_center), pos = f3(c.pos);
    rng.init(seed, pixel, s, retry << 8);
    if (!(F & PT_F_LENS) || c.use_pinhole) {
        // stratum (i,j): origins advance by hpw/pp, each stratum is 2*hpw/pp wide (main.cpp:1044-1057)
        const uint32_t i = __umulhi(s, pp_magic), j = s - i * (uint32_t)pp;
        const float inv_pp = px.inv_pp;
        const float stepX = inv_pp * c.half_pixel_w * 2.f, stepY = inv_pp * c.half_pixel_h * 2.f;
        float xs = (fx - c.half_pixel_w) + (float)i * inv_pp * c.half_pixel_w + stepX * 0.5f;
        float ys = (fy - c.half_pixel_h) + (float)j * inv_pp * c.half_pixel_h + stepY * 0.5f;
        const uint4 u = rng.next_block();                     // block 0 of the path stream: the two jitters
        xs += (u32_to_unit(u.x) - 0.5f) * stepX;
        ys += (u32_to_unit(u.y) - 0.5f) * stepY;
        float3 P = fc + (xs * c.half_film_w) * ax + (ys * c.half_film_h) * ay;
        o = pos;
        d = normalize3(P - pos);
    } else {
        // outer film sample ri jittered +-1 pixel, focal point on the focal plane, lens point from the
        // Poisson table at (ri*r2) % 12 (main.cpp:1114-1161)
        const uint32_t ri = __umulhi(s, pp_magic), r2 = s - ri * (uint32_t)pp;
        uint4 fj = rng.block_of(0u, pixel, ri, 1u);          // rng.init above set the same seed
        float offX = fx + (2.0f * u32_to_unit(fj.x) - 1.0f) * c.half_pixel_w;
        float offY = fy + (2.0f * u32_to_unit(fj.y) - 1.0f) * c.half_pixel_h;
        float3 P = fc + (offX * c.half_film_w) * ax + (offY * c.half_film_h) * ay;
        float3 dir = normalize3(P - pos);
        const float fpd = 1.f / (1.f / 0.098f - 1.f / c.focal_length);       // == focal_distance
        const float3 N = -f3(c.axis_z);
        const float3 pp0 = pos + ax + fpd * N;
        float t = pt_fdiv(dot3(N, pp0) - dot3(N, pos), dot3(N, dir));
        float3 focal = pos + t * dir;
        float2 ds = c_poisson_disk[(r2 * ri) % 12u];
        o = pos + (ds.x * c.aperture_radius) * ax + (ds.y * c.aperture_radius) * ay;
        d = normalize3(focal - o);
    }
}

// ---- the render megakernel -----------------------------------------------------------------------
// A warp owns P = 32/S adjacent pixels; the S lanes of a pixel stride through its sample list.  Every
// loop iteration each live lane advances its path by ONE segment (trace + shade); a lane whose path
// ended starts its next sample in the same iteration (path regeneration), so lanes at different
// depths share the instruction stream instead of idling behind the longest path.
template <uint32_t F>
__global__ void __launch_bounds__(PT_RENDER_THREADS, pt_render_min_blocks(F))
k_render(const __grid_constant__ PtDeviceScene S, const __grid_constant__ PtRenderArgs A) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t lS = A.log2_sample_lanes, lP = 5u - lS;
    const uint32_t Sn = 1u << lS;
    const uint32_t pw_log2 = (lP + 1u) >> 1, ph_log2 = lP - pw_log2;     // pixel block of a warp: pw x ph
    const uint32_t groups_x_log2 = 5u - pw_log2;                          // 32/pw groups per tile row
    const uint32_t groups_per_tile_log2 = 10u - lP;

    const uint32_t tile_j = warp >> groups_per_tile_log2;
    const uint32_t grp = warp & ((1u << groups_per_tile_log2) - 1u);
    bool active = tile_j < A.n_my_tiles;
    uint32_t tx, ty;
    pt_slot_tile(A.tiles, A.tile_rank, tile_j, tx, ty);      // an empty slot lands outside the image: inactive below
    tx *= PT_TILE;
    ty *= PT_TILE;
    const uint32_t gx = grp & ((1u << groups_x_log2) - 1u), gy = grp >> groups_x_log2;
    const uint32_t pig = lane >> lS, sl = lane & (Sn - 1u);
    const int x = (int)(tx + (gx << pw_log2) + (pig & ((1u << pw_log2) - 1u)));
    const int y = (int)(ty + (gy << ph_log2) + (pig >> pw_log2));
    active = active && x < A.width && y < A.height;

    const PtPixelCam pxc = pt_pixel_cam(A.width, A.height, A.pp, x, y);
    uint32_t retry = 0, n_acc = 0, n_rays = 0, n_nan = 0;
    // Per-pixel sums in a CANONICAL order, whatever the launch shape: sample s belongs to logical lane
    // (s - s_begin) % 32; a physical lane owns 32/S logical lanes and keeps one accumulator for each (local
    // memory, touched once per finished sample); at the end they are added in the order of a 32-lane xor
    // butterfly.  So S = 8 (full frames) and S = 32 (small multi-GPU shares) give bit-identical pixels.
    // (Measured, profiles/r02a_ab_compare_old_accl.json: walking the logical lanes one after the other with the running
    // sum in three registers -- no load + store per sample -- is bit-identical and 3-5 % SLOWER on every world: four more
    // live registers at the 64-register cap cost more than the local-memory traffic they remove.)
    const uint32_t n_logical = 32u >> lS;
    uint32_t s = active ? A.s_begin + sl : A.s_end;
    float accl[32][3];
#pragma unroll 1
    for (uint32_t k = 0; k < n_logical; k++) accl[k][0] = accl[k][1] = accl[k][2] = 0.f;
    float3 o = f3(0, 0, 0), d = f3(0, 0, 1), beta = f3(1, 1, 1), L = f3(0, 0, 0);
    int depth = 0;
    PtHit h;
    h.t = 3.402823466e+38f;
    h.id = PT_ID_MISS;
    PtRng rng;
    rng.init(A.seed, 0, 0, 0);

    // Lane states.  Where only ~30 % of the segments continue with a BRDF sample (w1, w5), shading right
    // after the trace runs the longest straight-line code of the kernel at a third of the lanes.  With
    // PT_F_DEFER a lane that needs a BRDF sample PARKS in NEED_SHADE while the others keep regenerating
    // camera rays, until `shade_batch` lanes are parked (or nobody can regenerate); then the sampling code
    // runs once for all of them.  Measured (profiles/r01d_sweep_defer.txt, r01e): w1 +18 %, w5 +8 %, w2 +7 %; it
    // loses 2 % on w3/w4, whose variants therefore shade at once.  The same deferral for the BVH walk lost everywhere.
    // Scheduling never changes a sample's value: the Philox stream is keyed by (pixel, sample) and every
    // lane adds its own samples in order.
    enum { NO_RAY = 0, NEW_RAY = 1, NEED_SHADE = 2 };
    uint32_t state = NO_RAY;

    constexpr bool kCoop = (F & PT_F_COOP) && (((F & PT_F_SPH_BVH) != 0) != ((F & PT_F_TRI_BVH) != 0));
    // (Resumable per-lane walks that yield to shading / regeneration once some lanes hold a finished walk were built for both
    // kinds of tree, verified bit-identical and measured slower at every threshold: profiles/r02_rejected_experiments.txt #3.)

    while (true) {
        if (state == NO_RAY && s < A.s_end) {
            // thin-lens variants keep the per-sample evaluation: holding fx, fy across the loop cost w4 1.5 % (profiles/r01h_ab_*)
            if (F & PT_F_LENS) pt_camera_ray<F>(S.cam, pt_pixel_cam(A.width, A.height, A.pp, x, y), A.pp, A.pp_magic, s, retry, A.seed, rng, o, d);
            else pt_camera_ray<F>(S.cam, pxc, A.pp, A.pp_magic, s, retry, A.seed, rng, o, d);
            beta = f3(1, 1, 1);
            L = f3(0, 0, 0);
            depth = 0;
            state = NEW_RAY;
        }
        const uint32_t n_live = __popc(__ballot_sync(0xffffffffu, state != NO_RAY));
        if (n_live == 0) break;

        // PT_F_COOP: every new ray does the linear lists and the root step itself; then the warp votes.  When few lanes
        // enter the tree (bounce rays that cross the mesh's box while their neighbours see sky or ground -- 2/3 of the
        // traversal's warp instructions, profiles/r01e_bvh_walk_occupancy_w5.txt) their rays are traced one after the
        // other by all 32 lanes over the cluster cut (pt_coop_trace); otherwise each lane walks its own ray as before.
        if (kCoop) {
            constexpr int LEAF = (F & PT_F_TRI_BVH) ? PT_LEAF_TRIS : PT_LEAF_SPHERES;
            const PtBvh &bvh = (F & PT_F_TRI_BVH) ? S.tri_bvh : S.sphere_bvh;
            int node0 = PT_STACK_SENTINEL, far0 = PT_STACK_SENTINEL;
            PtRayInv rinv;
            float a = 0.f;
            if (state == NEW_RAY) {
                h.t = 3.402823466e+38f;
                h.id = PT_ID_MISS;
                a = pt_dot_exact(d);
                pt_hit_analytic<F>(S, o, d, a, h);
                rinv = pt_ray_inv(o, d);
                pt_bvh_root_step(bvh, rinv, h.t, node0, far0);
            }
            __syncwarp();
            uint32_t m = __ballot_sync(0xffffffffu, node0 != PT_STACK_SENTINEL);
            if (m != 0u) {
                if ((uint32_t)__popc(m) <= A.coop_max) {
                    const float4 *cl = (F & PT_F_TRI_BVH) ? S.tri_clusters : S.sph_clusters;
                    const uint32_t ncl = (F & PT_F_TRI_BVH) ? S.n_tri_clusters : S.n_sph_clusters;
#pragma unroll 1
                    while (m != 0u) {
                        const int src = __ffs((int)m) - 1;
                        m &= m - 1u;
                        pt_coop_trace<LEAF>(S, cl, ncl, src, lane, o, d, h);
                    }
                } else if (node0 != PT_STACK_SENTINEL) {
                    pt_bvh_walk_from<LEAF>(S, bvh, rinv, o, d, a, h, node0, far0);
                }
            }
            __syncwarp();
        }

        // ---- closest hit, emission, termination rules of RayCast (main.cpp:579-590, 799) ----
        bool done = false;
        if (state == NEW_RAY) {
            if (!kCoop) h = pt_closest_hit<F>(S, o, d);
            n_rays++;
            float3 n;
            uint32_t mat;
            pt_hit_surface<F>(S, h, o, d, n, mat);
            const float3 e = pt_load_emit(S, mat);
            L = L + beta * e;
            const bool emissive = (e.x != 0.f) || (e.y != 0.f) || (e.z != 0.f);
            done = (mat == 0) || emissive || (depth == PT_MAX_BOUNCE_COUNT - 1);
            if (!(F & PT_F_DEFER)) {
                // shade at once, normal and material still in registers (best where most segments continue)
                if (!done) done = !pt_shade_vertex<F>(S, rng, mat, n, h.t, o, d, beta);
                depth++;
            } else {
                state = NEED_SHADE;
            }
        }

        // ---- BRDF sample for the parked lanes (RayCast's sampling loop, main.cpp:657-781) ----
        if (F & PT_F_DEFER) {
            __syncwarp();
            const uint32_t nS = __popc(__ballot_sync(0xffffffffu, state == NEED_SHADE && !done));
            const uint32_t nD = __popc(__ballot_sync(0xffffffffu, done && s + Sn < A.s_end));   // lanes that will regenerate
            if (nS >= A.shade_batch || nD == 0) {
                if (state == NEED_SHADE && !done) {
                    float3 n;
                    uint32_t mat;
                    pt_hit_surface<F>(S, h, o, d, n, mat);
                    done = !pt_shade_vertex<F>(S, rng, mat, n, h.t, o, d, beta);
                    depth++;
                    state = NEW_RAY;
                }
            }
            __syncwarp();
        }
        if (done) {
            state = NO_RAY;
            if (isnan3(L) && retry + 1 < PT_MAX_NAN_RETRIES) {
                retry++;                                                    // redraw this sample, main.cpp:1068
                n_nan++;
            } else {
                if (isnan3(L)) {                                            // still NaN after the redraws: black, and reported (pt_stats.nan_dropped)
                    L = f3(0, 0, 0);
                    if (A.counters) atomicAdd(A.counters + 2, 1ull);
                }
                const uint32_t kk = ((s - A.s_begin) >> lS) & (n_logical - 1u);
                accl[kk][0] += L.x;
                accl[kk][1] += L.y;
                accl[kk][2] += L.z;
                n_acc++;
                s += Sn;
                retry = 0;
            }
        }
    }

    // the butterfly levels that pair logical lanes of the same physical lane (offsets 16 .. S) ...
#pragma unroll 1
    for (uint32_t ks = 16u >> lS; ks >= 1u; ks >>= 1)
#pragma unroll 1
        for (uint32_t k = 0; k < ks; k++) {
            accl[k][0] += accl[k + ks][0];
            accl[k][1] += accl[k + ks][1];
            accl[k][2] += accl[k + ks][2];
        }
    float3 acc = f3(accl[0][0], accl[0][1], accl[0][2]);
    // ... then the levels across the S lanes of each pixel (offsets S/2 .. 1)
    for (uint32_t off = Sn >> 1; off > 0; off >>= 1) {
        acc.x += __shfl_xor_sync(0xffffffffu, acc.x, off);
        acc.y += __shfl_xor_sync(0xffffffffu, acc.y, off);
        acc.z += __shfl_xor_sync(0xffffffffu, acc.z, off);
        n_acc += __shfl_xor_sync(0xffffffffu, n_acc, off);
    }
    if (active && sl == 0) {
        float4 *p = A.accum + (size_t)y * A.width + x;
        float4 v = *p;
        v.x += acc.x; v.y += acc.y; v.z += acc.z; v.w += (float)n_acc;
        *p = v;
    }
    // statistics: one atomic per warp
    for (int off = 16; off > 0; off >>= 1) {
        n_rays += __shfl_xor_sync(0xffffffffu, n_rays, off);
        n_nan += __shfl_xor_sync(0xffffffffu, n_nan, off);
    }
    if (lane == 0 && A.counters) {
        atomicAdd(A.counters + 0, (unsigned long long)n_rays);
        if (n_nan) atomicAdd(A.counters + 1, (unsigned long long)n_nan);
    }
}

// ---- resolve: mean -> ACES (e = 0.4) -> sRGB -> truncating BGRA pack (main.cpp:1172-1182) -------
// Arithmetic kept in the reference's order with explicit round-to-nearest ops (no FMA contraction) and
// powf evaluated in double, so that the bytes equal the CPU's for the same HDR input.
__device__ __forceinline__ float pt_tonemap1(float x) {
    float num = __fmul_rn(x, __fadd_rn(__fmul_rn(2.51f, x), 0.03f));
    float den = __fadd_rn(0.4f, __fmul_rn(x, __fadd_rn(__fmul_rn(2.43f, x), 0.59f)));
    float q = __fdiv_rn(num, den);
    float m = (1.0f < q) ? 1.0f : q;            // std::min(v,1): NaN stays NaN ...
    return (0.0f < m) ? m : 0.0f;               // ... std::max(0,NaN) = 0   (ray_math.hpp:298-304)
}
__device__ __forceinline__ float pt_linear_to_srgb(float Lc) {                // ray_math.hpp:320-332
    if (Lc < 0.0f) Lc = 0.0f;
    if (Lc > 1.0f) Lc = 1.0f;
    float Sv = __fmul_rn(Lc, 12.92f);
    if (Lc > 0.0031308f) {
        float p = (float)pow((double)Lc, (double)(1.0f / 2.4f));
        Sv = __fsub_rn(__fmul_rn(1.055f, p), 0.055f);
    }
    return Sv;
}
__device__ __forceinline__ uint32_t pt_resolve_pixel(float r, float g, float b) {
    uint32_t R = (uint32_t)__fmul_rn(255.0f, pt_linear_to_srgb(pt_tonemap1(r)));
    uint32_t G = (uint32_t)__fmul_rn(255.0f, pt_linear_to_srgb(pt_tonemap1(g)));
    uint32_t B = (uint32_t)__fmul_rn(255.0f, pt_linear_to_srgb(pt_tonemap1(b)));
    return (255u << 24) | (R << 16) | (G << 8) | B;                          // BGRAPack4x8, ray_math.hpp:196-202
}

__global__ void k_resolve(const float4 *__restrict__ accum, uint32_t *__restrict__ bgra, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 a = accum[i];
    float inv = a.w > 0.f ? __fdiv_rn(1.0f, a.w) : 0.f;
    bgra[i] = pt_resolve_pixel(__fmul_rn(a.x, inv), __fmul_rn(a.y, inv), __fmul_rn(a.z, inv));
}

__global__ void k_resolve_colors(const float *__restrict__ rgb, uint32_t *__restrict__ out, uint32_t n) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = pt_resolve_pixel(rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]);
}

// ---- parity / debug kernels ----------------------------------------------------------------------
__device__ __forceinline__ void pt_store_hit(const PtDeviceScene &S, const PtHit &h, float3 o, float3 d, pt_hit *out) {
    float3 n;
    uint32_t mat;
    pt_hit_surface<PT_F_ALL>(S, h, o, d, n, mat);
    pt_hit r;
    r.t = h.t;
    r.kind = (h.id == PT_ID_MISS) ? PT_HIT_MISS : (int32_t)(h.id >> 28);
    r.index = (h.id == PT_ID_MISS) ? -1 : (int32_t)(h.id & 0x0FFFFFFFu);
    r.mat = mat;
    r.normal.x = n.x; r.normal.y = n.y; r.normal.z = n.z;
    *out = r;
}

__global__ void k_trace_rays(const __grid_constant__ PtDeviceScene S, const float *__restrict__ o, const float *__restrict__ d,
                             pt_hit *__restrict__ out, uint32_t n) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float3 ro = f3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), rd = f3(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
    PtHit h = pt_closest_hit<PT_F_ALL>(S, ro, rd);
    pt_store_hit(S, h, ro, rd, out + i);
}

// jitter-free ray through the lower-left corner of each pixel, direction computed with IEEE
// mul/add/sqrt/div in the reference's order so the ray itself is bit-identical to the CPU's
__global__ void k_primary_hits(const __grid_constant__ PtDeviceScene S, int W, int H, pt_hit *__restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= W * H) return;
    int x = i % W, y = i / W;
    const pt_camera &c = S.cam;
    float fy = __fadd_rn(-1.0f, __fdiv_rn(__fmul_rn(2.0f, (float)y), (float)H));
    float fx = __fadd_rn(-1.0f, __fdiv_rn(__fmul_rn(2.0f, (float)x), (float)W));
    float sx = __fmul_rn(fx, c.half_film_w), sy = __fmul_rn(fy, c.half_film_h);
    float px = __fadd_rn(__fadd_rn(c.frustum_center.x, __fmul_rn(sx, c.axis_x.x)), __fmul_rn(sy, c.axis_y.x));
    float py = __fadd_rn(__fadd_rn(c.frustum_center.y, __fmul_rn(sx, c.axis_x.y)), __fmul_rn(sy, c.axis_y.y));
    float pz = __fadd_rn(__fadd_rn(c.frustum_center.z, __fmul_rn(sx, c.axis_x.z)), __fmul_rn(sy, c.axis_y.z));
    float dx = __fsub_rn(px, c.pos.x), dy = __fsub_rn(py, c.pos.y), dz = __fsub_rn(pz, c.pos.z);
    float m = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
    float3 rd = f3(__fdiv_rn(dx, m), __fdiv_rn(dy, m), __fdiv_rn(dz, m));
    float3 ro = f3(c.pos);
    PtHit h = pt_closest_hit<PT_F_ALL>(S, ro, rd);
    pt_store_hit(S, h, ro, rd, out + i);
}

__global__ void k_trace_paths(const __grid_constant__ PtDeviceScene S, const float *__restrict__ o, const float *__restrict__ d,
                              uint64_t seed, float *__restrict__ rgb, uint32_t n) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    PtRng rng;
    rng.init(seed, i, 0u, 2u);
    float3 L = pt_trace_path<PT_F_ALL>(S, rng, f3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), f3(d[3 * i], d[3 * i + 1], d[3 * i + 2]));
    rgb[3 * i] = L.x; rgb[3 * i + 1] = L.y; rgb[3 * i + 2] = L.z;
}

// ---- tile pack / unpack for the multi-GPU gather --------------------------------------------------
// packed layout: tile slot j of the rank (pt_slot_tile) occupies 1024 uint32 at offset j*1024, row-major inside the tile;
// pixels outside the image -- and whole empty slots -- are zero
__global__ void k_pack_tiles(const uint32_t *__restrict__ bgra, uint32_t *__restrict__ packed, int W, int H,
                             const PtTileMap tm, uint32_t rank, uint32_t n_my_tiles) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_my_tiles * 1024u) return;
    uint32_t j = i >> 10, p = i & 1023u;
    uint32_t tx, ty;
    pt_slot_tile(tm, rank, j, tx, ty);
    int x = (int)(tx * PT_TILE + (p & 31u)), y = (int)(ty * PT_TILE + (p >> 5));
    packed[i] = (x < W && y < H) ? bgra[(size_t)y * W + x] : 0u;
}
__global__ void k_unpack_tiles(uint32_t *__restrict__ bgra, const uint32_t *__restrict__ packed, int W, int H,
                               const PtTileMap tm, uint32_t rank, uint32_t n_my_tiles) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_my_tiles * 1024u) return;
    uint32_t j = i >> 10, p = i & 1023u;
    uint32_t tx, ty;
    pt_slot_tile(tm, rank, j, tx, ty);
    int x = (int)(tx * PT_TILE + (p & 31u)), y = (int)(ty * PT_TILE + (p >> 5));
    if (x < W && y < H) bgra[(size_t)y * W + x] = packed[i];
}

// resolve + pack in one pass: own tile j -> packed[j*1024 ..], resolved from the accumulator (no full-frame resolve first)
__global__ void k_resolve_pack_tiles(const float4 *__restrict__ accum, uint32_t *__restrict__ packed, int W, int H,
                                     const PtTileMap tm, uint32_t rank, uint32_t n_my_tiles) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_my_tiles * 1024u) return;
    uint32_t j = i >> 10, p = i & 1023u;
    uint32_t tx, ty;
    pt_slot_tile(tm, rank, j, tx, ty);
    int x = (int)(tx * PT_TILE + (p & 31u)), y = (int)(ty * PT_TILE + (p >> 5));
    uint32_t v = 0u;
    if (x < W && y < H) {
        float4 a = accum[(size_t)y * W + x];
        float inv = a.w > 0.f ? __fdiv_rn(1.0f, a.w) : 0.f;
        v = pt_resolve_pixel(__fmul_rn(a.x, inv), __fmul_rn(a.y, inv), __fmul_rn(a.z, inv));
    }
    packed[i] = v;
}
// scatter the packs of ALL ranks (gathered back to back, `stride` words per rank) into the frame: one thread per pixel
__global__ void k_unpack_gathered(uint32_t *__restrict__ bgra, const uint32_t *__restrict__ gathered, int W, int H,
                                  const PtTileMap tm, uint32_t stride) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (uint32_t)W * (uint32_t)H) return;
    uint32_t x = i % (uint32_t)W, y = i / (uint32_t)W;
    uint32_t r, j;
    pt_tile_slot(tm, x / PT_TILE, y / PT_TILE, r, j);
    bgra[i] = gathered[(size_t)r * stride + j * 1024u + (y % PT_TILE) * PT_TILE + (x % PT_TILE)];
}
// resolve n pixels from any accumulator to any BGRA buffer (a rank's share after a reduce-scatter)
__global__ void k_resolve_range(const float4 *__restrict__ accum, uint32_t *__restrict__ bgra, uint32_t n) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 a = accum[i];
    float inv = a.w > 0.f ? __fdiv_rn(1.0f, a.w) : 0.f;
    bgra[i] = pt_resolve_pixel(__fmul_rn(a.x, inv), __fmul_rn(a.y, inv), __fmul_rn(a.z, inv));
}

// resolve this rank's tiles straight into a BGRA frame that may live on another GPU (mapped peer memory): thread i
// handles pixel i & 1023 of own tile i >> 10, so a warp stores one 128-byte row segment of a tile
__global__ void k_resolve_tiles_to(const float4 *__restrict__ accum, uint32_t *__restrict__ dst, int W, int H,
                                   const PtTileMap tm, uint32_t rank, uint32_t n_my_tiles) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_my_tiles * 1024u) return;
    uint32_t j = i >> 10, p = i & 1023u;
    uint32_t tx, ty;
    pt_slot_tile(tm, rank, j, tx, ty);
    int x = (int)(tx * PT_TILE + (p & 31u)), y = (int)(ty * PT_TILE + (p >> 5));
    if (x >= W || y >= H) return;
    float4 a = accum[(size_t)y * W + x];
    float inv = a.w > 0.f ? __fdiv_rn(1.0f, a.w) : 0.f;
    dst[(size_t)y * W + x] = pt_resolve_pixel(__fmul_rn(a.x, inv), __fmul_rn(a.y, inv), __fmul_rn(a.z, inv));
}

// ---- FP32 roofline probe: 8 independent FMA chains per thread ------------------------------------------
__global__ void k_fma_peak(float *out, int iters) {
    float a[8];
    const float b = 1.0000001f, c = 1e-7f;
#pragma unroll
    for (int k = 0; k < 8; k++) a[k] = (float)(threadIdx.x + k) * 1e-3f;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
#pragma unroll
            for (int k = 0; k < 8; k++) a[k] = fmaf(a[k], b, c);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; k++) s += a[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
