// Wavefront-in-SM variant of the render loop (PT_KERNEL_WAVEFRONT).
//
// The megakernel (pt_kernels.cuh) keeps a path in the registers of one lane, so a warp instruction only
// serves the lanes whose paths happen to be in the same stage: measured 12 of 32 threads per instruction
// on w5 (profiles/r01b).  Here a warp owns a POOL of paths (2..4 x 32) whose state lives in shared
// memory, plus one index list per stage:
//
//     END   : the path ended (or the slot is fresh) -> accumulate, start the next sample (camera ray)
//     SHADE : a surface was hit and the path continues -> BRDF sample, next ray
//     BVH   : the new ray enters the root of a BVH    -> walk it; a lane whose ray is finished takes the
//             next ray of the list at once (persistent refill), so the walk runs with full lanes until
//             the list is empty instead of idling behind the longest ray
//     HIT   : a walked ray has its final hit          -> emission, termination rules -> END or SHADE
//
// Every iteration the warp runs the stage with the MOST pending paths, up to 32 of them, one per lane.
// END and SHADE finish with the cheap analytic part of the closest hit and the BVH-root test of their
// new ray, which decides the next list.  Only __syncwarp() is needed: no CTA barriers, no atomics.
// The arithmetic is pt_trace.cuh / pt_shade.cuh unchanged, and the Philox streams are keyed by
// (pixel, sample), so the image is the same estimator as the megakernel's, sample for sample.
#pragma once
#include <string>

#include "pt_kernels.cuh"

#ifndef PT_WF_POOL
#define PT_WF_POOL 64            // paths per warp (multiple of 32)
#endif
#ifndef PT_WF_WARPS
#define PT_WF_WARPS 4            // warps per CTA
#endif
#ifndef PT_WF_MIN_BLOCKS
#define PT_WF_MIN_BLOCKS 6
#endif
#ifndef PT_WF_REFILL
#define PT_WF_REFILL 0           // 1: persistent BVH stage (idle lanes take the next ray of the list).  Correct (bit-identical
#endif                           //    images) but measured slower with 64-path pools: 11.7 vs 13.0 Gsamples/s on w5 (profiles/r01e)

enum { WF_OX, WF_OY, WF_OZ, WF_DX, WF_DY, WF_DZ, WF_BX, WF_BY, WF_BZ, WF_LX, WF_LY, WF_LZ,
       WF_AX, WF_AY, WF_AZ, WF_T, WF_ID, WF_S, WF_MISC, WF_FIELDS };
enum { WF_BVH = 0, WF_END = 1, WF_SHADE = 2, WF_HIT = 3, WF_NONE = 4 };
#define WF_FRESH (1u << 24)      // misc: depth | retry << 8 | philox block << 16 | WF_FRESH

static_assert(PT_WF_POOL >= 32 && PT_WF_POOL <= 256 && (PT_WF_POOL & (PT_WF_POOL - 1)) == 0, "PT_WF_POOL: power of two in [32, 256] (8-bit list entries)");

struct PtWfWarp {                                  // shared memory of one warp: SoA, field-major
    float f[WF_FIELDS][PT_WF_POOL];                // path p of field k at bank (p % 32): <= POOL/32-way conflicts
    uint8_t list[4][PT_WF_POOL];
};

__device__ __forceinline__ float3 wf_ld3(const PtWfWarp &W, int k, uint32_t p) { return f3(W.f[k][p], W.f[k + 1][p], W.f[k + 2][p]); }
__device__ __forceinline__ void wf_st3(PtWfWarp &W, int k, uint32_t p, float3 v) { W.f[k][p] = v.x; W.f[k + 1][p] = v.y; W.f[k + 2][p] = v.z; }

template <uint32_t F>
__global__ void __launch_bounds__(PT_WF_WARPS * 32, PT_WF_MIN_BLOCKS)
k_render_wf(const __grid_constant__ PtDeviceScene S, const __grid_constant__ PtRenderArgs A) {
    __shared__ PtWfWarp sm[PT_WF_WARPS];
    PtWfWarp &W = sm[threadIdx.x >> 5];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;

    // the warp's pixel group: P = 2^lP pixels (pw x ph block of a 32x32 tile), Sn = POOL/P sample slots each
    const uint32_t lP = A.log2_pool_pixels;
    const uint32_t Sn = (uint32_t)PT_WF_POOL >> lP;
    const uint32_t pw_log2 = (lP + 1u) >> 1, ph_log2 = lP - pw_log2;
    const uint32_t groups_x_log2 = 5u - pw_log2, groups_per_tile_log2 = 10u - lP;
    const uint32_t tile_j = warp >> groups_per_tile_log2;
    const uint32_t grp = warp & ((1u << groups_per_tile_log2) - 1u);
    if (tile_j >= A.n_my_tiles) return;                                   // whole warp
    uint32_t ttx, tty;
    pt_slot_tile(A.tiles, A.tile_rank, tile_j, ttx, tty);
    if (ttx >= A.tiles.tiles_x) return;                                   // an empty slot: whole warp
    const uint32_t gx = grp & ((1u << groups_x_log2) - 1u), gy = grp >> groups_x_log2;
    const uint32_t x0 = ttx * PT_TILE + (gx << pw_log2), y0 = tty * PT_TILE + (gy << ph_log2);
    // pool slot p: pixel-in-group q = p % P, sample slot sl = p / P
#define WF_PIXEL_X(p) ((int)(x0 + (((p) & ((1u << lP) - 1u)) & ((1u << pw_log2) - 1u))))
#define WF_PIXEL_Y(p) ((int)(y0 + (((p) & ((1u << lP) - 1u)) >> pw_log2)))

    uint32_t nB = 0, nE = 0, nS = 0, nH = 0;  // list lengths (warp-uniform)
    constexpr bool kOneBvh = PT_WF_REFILL && (((F & PT_F_SPH_BVH) != 0) != ((F & PT_F_TRI_BVH) != 0));   // refill needs one tree type
    uint32_t n_rays = 0, n_nan = 0;           // per-lane statistics

    // every slot starts "fresh" on the END list: END skips the accumulate and shoots the first camera ray
    for (uint32_t p = lane; p < (uint32_t)PT_WF_POOL; p += 32u) {
        const uint32_t s0 = A.s_begin + (p >> lP);
        const bool ok = WF_PIXEL_X(p) < A.width && WF_PIXEL_Y(p) < A.height && s0 < A.s_end;
        W.f[WF_S][p] = __uint_as_float(s0);
        W.f[WF_MISC][p] = __uint_as_float(WF_FRESH);
        wf_st3(W, WF_AX, p, f3(0, 0, 0));
        const uint32_t m = __ballot_sync(0xffffffffu, ok);
        if (ok) W.list[WF_END][nE + __popc(m & lt_mask)] = (uint8_t)p;
        nE += __popc(m);
    }

    while (true) {
        __syncwarp();
        // the fullest list runs next
        uint32_t st, cnt;
        if (nB >= nE && nB >= nS && nB >= nH) { st = WF_BVH; cnt = nB; }
        else if (nH >= nE && nH >= nS) { st = WF_HIT; cnt = nH; }
        else if (nS >= nE) { st = WF_SHADE; cnt = nS; }
        else { st = WF_END; cnt = nE; }
        if (cnt == 0) break;
        const uint32_t take = cnt < 32u ? cnt : 32u;
        bool have = lane < take;
        uint32_t p = have ? (uint32_t)W.list[st][cnt - 1u - lane] : 0u;
        if (st == WF_BVH) nB -= take; else if (st == WF_SHADE) nS -= take; else if (st == WF_HIT) nH -= take; else nE -= take;

        uint32_t dest = WF_NONE, misc = 0;
        bool has_ray = false, has_hit = false;
        float3 o = f3(0, 0, 0), d = f3(0, 0, 1);
        PtHit h;
        h.t = 3.402823466e+38f;
        h.id = PT_ID_MISS;

        if (st == WF_END) {
            if (have) {
                misc = __float_as_uint(W.f[WF_MISC][p]);
                uint32_t s = __float_as_uint(W.f[WF_S][p]);
                uint32_t retry = (misc >> 8) & 0xffu;
                if (!(misc & WF_FRESH)) {
                    float3 L = wf_ld3(W, WF_LX, p);
                    const bool bad = isnan3(L);
                    if (bad && retry + 1 < PT_MAX_NAN_RETRIES) {
                        retry++;                                             // redraw this sample, main.cpp:1068
                        n_nan++;
                    } else {
                        if (bad) L = f3(0, 0, 0);
                        wf_st3(W, WF_AX, p, wf_ld3(W, WF_AX, p) + L);
                        s += Sn;
                        retry = 0;
                        W.f[WF_S][p] = __uint_as_float(s);
                    }
                }
                if (s < A.s_end) {
                    PtRng rng;
                    pt_camera_ray<F>(S.cam, pt_pixel_cam(A.width, A.height, A.pp, WF_PIXEL_X(p), WF_PIXEL_Y(p)), A.pp, A.pp_magic, s, retry, A.seed, rng, o, d);
                    wf_st3(W, WF_BX, p, f3(1, 1, 1));
                    wf_st3(W, WF_LX, p, f3(0, 0, 0));
                    misc = (retry << 8) | (rng.blk << 16);                   // depth 0
                    has_ray = true;
                }
            }
        } else if (st == WF_SHADE) {
            if (have) {
                misc = __float_as_uint(W.f[WF_MISC][p]);
                o = wf_ld3(W, WF_OX, p);
                d = wf_ld3(W, WF_DX, p);
                h.t = W.f[WF_T][p];
                h.id = __float_as_uint(W.f[WF_ID][p]);
                float3 beta = wf_ld3(W, WF_BX, p);
                float3 n;
                uint32_t mat;
                pt_hit_surface<F>(S, h, o, d, n, mat);
                PtRng rng;
                const uint32_t pixel = (uint32_t)WF_PIXEL_Y(p) * (uint32_t)A.width + (uint32_t)WF_PIXEL_X(p);
                rng.init(A.seed, pixel, __float_as_uint(W.f[WF_S][p]), misc & 0xff00u);
                rng.blk = (misc >> 16) & 0xffu;
                if (pt_shade_vertex<F>(S, rng, mat, n, h.t, o, d, beta)) {
                    wf_st3(W, WF_BX, p, beta);
                    misc = ((misc & 0xffffu) + 1u) | (rng.blk << 16);        // depth + 1
                    has_ray = true;
                } else {
                    dest = WF_END;                                           // no bounce: radiance so far stands
                }
            }
        } else if (st == WF_HIT) {
            if (have) {
                misc = __float_as_uint(W.f[WF_MISC][p]);
                o = wf_ld3(W, WF_OX, p);
                d = wf_ld3(W, WF_DX, p);
                h.t = W.f[WF_T][p];
                h.id = __float_as_uint(W.f[WF_ID][p]);
                has_hit = true;
            }
        } else if (kOneBvh) {
            // persistent walk: finished lanes hand their hit to the HIT list and take the next ray at once
            constexpr int LEAF = (F & PT_F_TRI_BVH) ? PT_LEAF_TRIS : PT_LEAF_SPHERES;
            const PtBvh &bvh = (F & PT_F_TRI_BVH) ? S.tri_bvh : S.sphere_bvh;
            PtWalker<LEAF> wk;
            bool walking = have;
            if (walking) {
                h.t = W.f[WF_T][p];
                h.id = __float_as_uint(W.f[WF_ID][p]);
                wk.begin(wf_ld3(W, WF_OX, p), wf_ld3(W, WF_DX, p), h);
            } else {
                wk.node = PT_STACK_SENTINEL;
                wk.sp = 1;
            }
            while (true) {
                const uint32_t idle = __ballot_sync(0xffffffffu, !walking);
                if (idle != 0u && nB != 0u) {
                    const uint32_t rank = __popc(idle & lt_mask);
                    if (!walking && rank < nB) {
                        p = (uint32_t)W.list[WF_BVH][nB - 1u - rank];
                        h.t = W.f[WF_T][p];
                        h.id = __float_as_uint(W.f[WF_ID][p]);
                        wk.begin(wf_ld3(W, WF_OX, p), wf_ld3(W, WF_DX, p), h);
                        walking = true;
                    }
                    const uint32_t n_idle = __popc(idle);
                    nB -= n_idle < nB ? n_idle : nB;
                }
                if (!__any_sync(0xffffffffu, walking)) break;
                if (walking) wk.advance(S, bvh);
                const bool fin = walking && wk.done();
                const uint32_t mF = __ballot_sync(0xffffffffu, fin);
                if (mF != 0u) {
                    if (fin) {
                        W.f[WF_T][p] = wk.h.t;
                        W.f[WF_ID][p] = __uint_as_float(wk.h.id);
                        W.list[WF_HIT][nH + __popc(mF & lt_mask)] = (uint8_t)p;
                        walking = false;
                    }
                    nH += __popc(mF);
                }
            }
            have = false;                    // nothing left for the common tail
        } else {
            if (have) {
                misc = __float_as_uint(W.f[WF_MISC][p]);
                o = wf_ld3(W, WF_OX, p);
                d = wf_ld3(W, WF_DX, p);
                h.t = W.f[WF_T][p];
                h.id = __float_as_uint(W.f[WF_ID][p]);
                const float a = pt_dot_exact(d);
                if ((F & PT_F_SPH_BVH) && S.sphere_bvh.n_prims) pt_bvh_walk<PT_LEAF_SPHERES>(S, S.sphere_bvh, o, d, a, h);
                if ((F & PT_F_TRI_BVH) && S.tri_bvh.n_prims) pt_bvh_walk<PT_LEAF_TRIS>(S, S.tri_bvh, o, d, a, h);
                has_hit = true;
            }
        }

        if (has_ray) {
            // front end of the closest hit: the linear lists, then "does it enter a BVH at all?"
            h.t = 3.402823466e+38f;
            h.id = PT_ID_MISS;
            pt_hit_analytic<F>(S, o, d, pt_dot_exact(d), h);
            bool deep = false;
            if (F & (PT_F_SPH_BVH | PT_F_TRI_BVH)) {
                const PtRayInv r = pt_ray_inv(o, d);
                if ((F & PT_F_SPH_BVH) && S.sphere_bvh.n_prims) deep = deep || pt_bvh_root_hit(S.sphere_bvh, r, h.t);
                if ((F & PT_F_TRI_BVH) && S.tri_bvh.n_prims) deep = deep || pt_bvh_root_hit(S.tri_bvh, r, h.t);
            }
            wf_st3(W, WF_OX, p, o);
            wf_st3(W, WF_DX, p, d);
            if (deep) {
                W.f[WF_T][p] = h.t;
                W.f[WF_ID][p] = __uint_as_float(h.id);
                dest = WF_BVH;
            } else {
                has_hit = true;
            }
        }
        if (has_hit) {
            // the segment's hit is final: emission and the termination rules of RayCast (main.cpp:579-590, 799)
            n_rays++;
            float3 n;
            uint32_t mat;
            pt_hit_surface<F>(S, h, o, d, n, mat);
            const float3 e = pt_load_emit(S, mat);
            const bool emissive = (e.x != 0.f) || (e.y != 0.f) || (e.z != 0.f);
            if (mat == 0 || emissive || (misc & 0xffu) == PT_MAX_BOUNCE_COUNT - 1) {
                wf_st3(W, WF_LX, p, wf_ld3(W, WF_LX, p) + wf_ld3(W, WF_BX, p) * e);
                dest = WF_END;
            } else {
                W.f[WF_T][p] = h.t;
                W.f[WF_ID][p] = __uint_as_float(h.id);
                dest = WF_SHADE;
            }
        }
        if (have) W.f[WF_MISC][p] = __uint_as_float(misc);

        __syncwarp();                        // every lane has popped before anyone pushes
        const uint32_t mB = __ballot_sync(0xffffffffu, dest == WF_BVH);
        const uint32_t mE = __ballot_sync(0xffffffffu, dest == WF_END);
        const uint32_t mS = __ballot_sync(0xffffffffu, dest == WF_SHADE);
        if (dest == WF_BVH) W.list[WF_BVH][nB + __popc(mB & lt_mask)] = (uint8_t)p;
        if (dest == WF_END) W.list[WF_END][nE + __popc(mE & lt_mask)] = (uint8_t)p;
        if (dest == WF_SHADE) W.list[WF_SHADE][nS + __popc(mS & lt_mask)] = (uint8_t)p;
        nB += __popc(mB);
        nE += __popc(mE);
        nS += __popc(mS);
    }

    // per pixel: add the sample slots in slot order (fixed for a given launch shape), one read-modify-write
    __syncwarp();
    for (uint32_t q = lane; q < (1u << lP); q += 32u) {
        const int x = WF_PIXEL_X(q), y = WF_PIXEL_Y(q);
        if (x < A.width && y < A.height) {
            float3 acc = f3(0, 0, 0);
            for (uint32_t sl = 0; sl < Sn; sl++) acc = acc + wf_ld3(W, WF_AX, q + (sl << lP));
            float4 *px = A.accum + (size_t)y * A.width + x;
            float4 v = *px;
            v.x += acc.x; v.y += acc.y; v.z += acc.z; v.w += (float)(A.s_end - A.s_begin);
            *px = v;
        }
    }
#undef WF_PIXEL_X
#undef WF_PIXEL_Y
    for (int off = 16; off > 0; off >>= 1) {
        n_rays += __shfl_xor_sync(0xffffffffu, n_rays, off);
        n_nan += __shfl_xor_sync(0xffffffffu, n_nan, off);
    }
    if (lane == 0 && A.counters) {
        atomicAdd(A.counters + 0, (unsigned long long)n_rays);
        if (n_nan) atomicAdd(A.counters + 1, (unsigned long long)n_nan);
    }
}

struct PtWavefront {
    int pixels_log2 = -1;        // PT_WF_PIXELS_LOG2: pixels per warp pool (-1 = auto)
    void release() {}

    int render(cudaStream_t st, const PtDeviceScene &S, PtRenderArgs A, uint32_t features, uint32_t *launches, std::string &err) {
        const uint32_t n_samples = A.s_end - A.s_begin;
        const uint32_t pool_log2 = 31u - (uint32_t)__builtin_clz((unsigned)PT_WF_POOL);
        // default: 16 pixels x POOL/16 sample slots; more pixels when there are fewer samples than slots
        uint32_t lP = pixels_log2 >= 0 ? (uint32_t)pixels_log2 : 4u;
        while (lP < pool_log2 && ((uint32_t)PT_WF_POOL >> lP) > n_samples) lP++;
        if (lP > pool_log2) lP = pool_log2;
        A.log2_pool_pixels = lP;
        const uint64_t warps = (uint64_t)A.n_my_tiles * (1024u >> lP);
        const uint64_t blocks = (warps + PT_WF_WARPS - 1) / PT_WF_WARPS;
        // the kernel derives its warp index from blockIdx.x * blockDim.x in 32 bits
        if (blocks > 0xffffffffull / (PT_WF_WARPS * 32) - 1) { err = "pt_render: frame too large for one launch"; return PT_ERR_UNSUPPORTED; }
#define X(M) if ((features & ~(uint32_t)(M)) == 0) { k_render_wf<(M)><<<(unsigned)blocks, PT_WF_WARPS * 32, 0, st>>>(S, A); *launches = 1; return PT_OK; }
        PT_RENDER_VARIANTS(X)
#undef X
        err = "pt_render: no kernel variant covers this scene";
        return PT_ERR_UNSUPPORTED;
    }
};
