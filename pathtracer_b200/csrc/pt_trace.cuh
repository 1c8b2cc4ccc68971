// Closest hit: the reference's RayCastIntersect (src/win32_main.cpp:406-556) re-organised for the GPU.
// Analytic primitives come from short linear lists; triangles and small spheres from flattened BVH2s
// with 64-byte nodes (4 x LDG.128) and 48-byte planar records (3 x LDG.128).  Nearest-hit semantics
// are the reference's: accept t > 1e-4 (0.02 for quads), strict '<', and on an exact tie the primitive
// the reference would have tested first (spheres, quads, planes, triangles, boxes; then lower index).
#pragma once
#include "pt_device.cuh"

// RaySphereIntersect, main.cpp:2355-2379: near root only; sqrt(disc) must exceed 1e-9.  Returns t or -1.
// Sphere silhouettes decide most primary-hit ids in worlds 2 and 4 (hundreds of small spheres), so this
// test keeps the reference's exact operation order with explicit round-to-nearest ops (no FMA
// contraction, IEEE sqrt and divide): t is bit-identical to the CPU's for the same ray.
__device__ __forceinline__ float pt_dot_rn(float ax, float ay, float az, float bx, float by, float bz) {
    return __fadd_rn(__fadd_rn(__fmul_rn(ax, bx), __fmul_rn(ay, by)), __fmul_rn(az, bz));
}
__device__ __forceinline__ float pt_dot_exact(float3 d) { return pt_dot_rn(d.x, d.y, d.z, d.x, d.y, d.z); }
__device__ __forceinline__ float pt_sphere_t(float3 o, float3 d, float a, float4 s) {
    float rx = __fsub_rn(o.x, s.x), ry = __fsub_rn(o.y, s.y), rz = __fsub_rn(o.z, s.z);
    float b = __fmul_rn(2.0f, pt_dot_rn(rx, ry, rz, d.x, d.y, d.z));
    float c = __fsub_rn(pt_dot_rn(rx, ry, rz, rx, ry, rz), __fmul_rn(s.w, s.w));
    float disc = __fsub_rn(__fmul_rn(b, b), __fmul_rn(__fmul_rn(4.0f, a), c));
    if (!(disc >= 0.f)) return -1.f;
    float root = __fsqrt_rn(disc);
    if (!(root > PT_TOLERANCE)) return -1.f;
    float t = __fdiv_rn(__fsub_rn(-b, root), __fmul_rn(2.0f, a));
    return (t > PT_MIN_HIT) ? t : -1.f;
}

// RayIntersectPlane, ray_math.hpp:334-341.  Returns t (any sign) or NaN when parallel.
__device__ __forceinline__ float pt_plane_t(float3 o, float3 d, float4 p) {
    float denom = dot3(p, d);
    if (!((denom < -PT_TOLERANCE) || (denom > PT_TOLERANCE))) return __int_as_float(0x7fc00000);
    return pt_fdiv(p.w - dot3(p, o), denom);
}

// RayIntersectPlanarShape<TRIANGLE|QUAD>, ray_math.hpp:357-381, on a precomputed PtPlanar record
template <bool QUAD>
__device__ __forceinline__ float pt_planar_t(float3 o, float3 d, const float4 r0, const float4 r1, const float4 r2, float min_hit) {
    float t = pt_plane_t(o, d, r0);
    if (!(t > min_hit)) return -1.f;
    float3 P = f3(fmaf(t, d.x, o.x), fmaf(t, d.y, o.y), fmaf(t, d.z, o.z));
    float alpha = dot3(r1, P) + r1.w;
    float beta = dot3(r2, P) + r2.w;
    bool in = QUAD ? (alpha >= 0.f && alpha <= 1.f && beta >= 0.f && beta <= 1.f)
                   : (alpha >= 0.f && beta >= 0.f && (alpha + beta) <= 1.f);
    return in ? t : -1.f;
}

// RayIntersectWithAABB2, ray_math.hpp:398-482: six faces in the order -z +z -x +x +y -y; the first face
// with t >= 0 whose hit point lies inside the box wins.  Returns t or -1; *face = that face's index.
__device__ __forceinline__ float pt_aabb_t(float3 o, float3 d, float4 mn, float4 mx, int *face) {
#pragma unroll
    for (int j = 0; j < 6; j++) {
        float plane, oc, dc;
        if (j < 2) { plane = (j == 0) ? mn.z : mx.z; oc = o.z; dc = d.z; }
        else if (j < 4) { plane = (j == 2) ? mn.x : mx.x; oc = o.x; dc = d.x; }
        else { plane = (j == 4) ? mx.y : mn.y; oc = o.y; dc = d.y; }
        if (dc == 0.f) continue;
        float t = (plane - oc) / dc;
        if (t < 0.f) continue;
        float x = (j == 2 || j == 3) ? plane : o.x + d.x * t;
        float y = (j >= 4) ? plane : o.y + d.y * t;
        float z = (j < 2) ? plane : o.z + d.z * t;
        if (x >= mn.x && x <= mx.x && y >= mn.y && y <= mx.y && z >= mn.z && z <= mx.z) {
            *face = j;
            return t;
        }
    }
    return -1.f;
}

__device__ __forceinline__ void pt_accept(PtHit &h, float t, uint32_t id) {
    if (t < h.t || (t == h.t && id < h.id)) { h.t = t; h.id = id; }
}

enum { PT_LEAF_SPHERES = 0, PT_LEAF_TRIS = 1 };
#define PT_STACK_SENTINEL ((int)0x80000000)

// The per-lane traversal stack: a thread-local array (local memory: L1-resident, stores write through to L2).
#define PT_DECLARE_STACK int stack[PT_STACK_DEPTH]
#define PT_STK(i) stack[(i)]
#define PT_WALK_STACK_LIMIT PT_STACK_DEPTH

__device__ __forceinline__ float pt_rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

struct PtRayInv {          // per-ray constants of the slab test
    float3 inv, ood;
};
template <bool BARE_RCP = true>
__device__ __forceinline__ PtRayInv pt_ray_inv(float3 o, float3 d) {
    // reciprocal direction with zero components nudged so that no inf*0 appears in the slab test
    float3 dd = f3(fabsf(d.x) < 1e-20f ? copysignf(1e-20f, d.x) : d.x,
                   fabsf(d.y) < 1e-20f ? copysignf(1e-20f, d.y) : d.y,
                   fabsf(d.z) < 1e-20f ? copysignf(1e-20f, d.z) : d.z);
    PtRayInv r;
    // one MUFU.RCP each (<= 1 ulp): the slab test below is conservative by 4e-7 relative and the boxes are padded.
    // (__fdividef(1.0f, x) returns the same value for 1e-20 <= |x| <= 1 but wraps the MUFU in range handling: 8 instructions.)
    // Measured (profiles/r01h_ab_*): the bare form gains everywhere (w5 with the hoisted camera constants +4.4 %, w2 +1.4 %) except
    // in the thin-lens sphere family (w4 -1.4 %, although the kernel is 16 instructions shorter), which keeps __fdividef.
    if (BARE_RCP) r.inv = f3(pt_rcp_approx(dd.x), pt_rcp_approx(dd.y), pt_rcp_approx(dd.z));
    else r.inv = f3(__fdividef(1.0f, dd.x), __fdividef(1.0f, dd.y), __fdividef(1.0f, dd.z));
    r.ood = f3(o.x * r.inv.x, o.y * r.inv.y, o.z * r.inv.z);
    return r;
}

// slab test of the two child boxes of one node against [0, tmax]; returns near/far entry distances
__device__ __forceinline__ void pt_node_slabs(const float4 b0, const float4 b1, const float4 b2, const PtRayInv &r, float tmax,
                                              bool &hl, bool &hr, float &ln, float &rn) {
    float lx0 = fmaf(b0.x, r.inv.x, -r.ood.x), lx1 = fmaf(b0.w, r.inv.x, -r.ood.x);
    float ly0 = fmaf(b0.y, r.inv.y, -r.ood.y), ly1 = fmaf(b1.x, r.inv.y, -r.ood.y);
    float lz0 = fmaf(b0.z, r.inv.z, -r.ood.z), lz1 = fmaf(b1.y, r.inv.z, -r.ood.z);
    ln = fmaxf(fmaxf(fminf(lx0, lx1), fminf(ly0, ly1)), fmaxf(fminf(lz0, lz1), 0.f));
    float lf = fminf(fminf(fmaxf(lx0, lx1), fmaxf(ly0, ly1)), fminf(fmaxf(lz0, lz1), tmax)) * 1.0000004f;
    float rx0 = fmaf(b1.z, r.inv.x, -r.ood.x), rx1 = fmaf(b2.y, r.inv.x, -r.ood.x);
    float ry0 = fmaf(b1.w, r.inv.y, -r.ood.y), ry1 = fmaf(b2.z, r.inv.y, -r.ood.y);
    float rz0 = fmaf(b2.x, r.inv.z, -r.ood.z), rz1 = fmaf(b2.w, r.inv.z, -r.ood.z);
    rn = fmaxf(fmaxf(fminf(rx0, rx1), fminf(ry0, ry1)), fmaxf(fminf(rz0, rz1), 0.f));
    float rf = fminf(fminf(fmaxf(rx0, rx1), fmaxf(ry0, ry1)), fminf(fmaxf(rz0, rz1), tmax)) * 1.0000004f;
    hl = ln <= lf;
    hr = rn <= rf;
}

// does the ray enter either child box of the root within [0, tmax]?  (the "needs deep traversal" test)
__device__ __forceinline__ bool pt_bvh_root_hit(const PtBvh &bvh, const PtRayInv &r, float tmax) {
    const float4 *nodes = reinterpret_cast<const float4 *>(bvh.nodes);
    bool hl, hr;
    float ln, rn;
    pt_node_slabs(__ldg(nodes), __ldg(nodes + 1), __ldg(nodes + 2), r, tmax, hl, hr, ln, rn);
    return hl || hr;
}

template <int LEAF>
__device__ __forceinline__ void pt_leaf_test(const PtDeviceScene &S, int node, float3 o, float3 d, float a, PtHit &h) {
    uint32_t enc = (uint32_t)(~node);
    uint32_t first = enc & 0xFFFFFFu, count = enc >> 24;
#pragma unroll 1
    for (uint32_t i = 0; i < count; i++) {
        uint32_t pos = first + i;
        if (LEAF == PT_LEAF_TRIS) {
            const float4 *rec = reinterpret_cast<const float4 *>(S.tris + pos);
            float t = pt_planar_t<false>(o, d, __ldg(rec), __ldg(rec + 1), __ldg(rec + 2), PT_MIN_HIT);
            if (t > 0.f && t <= h.t) pt_accept(h, t, PT_ID(PT_HIT_TRIANGLE, __ldg(S.tri_ids + pos)));
        } else {
            float t = pt_sphere_t(o, d, a, __ldg(S.bvh_spheres + pos));
            if (t > 0.f && t <= h.t) pt_accept(h, t, PT_ID(PT_HIT_SPHERE, __ldg(S.bvh_sphere_ids + pos)));
        }
    }
}

// Stack-based BVH2 walk in "while-while" form: every lane first descends inner nodes until it holds a
// leaf (nearer child first, culled by the best t so far), then the lanes test their leaves together.
template <int LEAF, bool BARE_RCP = true>
__device__ __forceinline__ void pt_bvh_walk(const PtDeviceScene &S, const PtBvh &bvh, float3 o, float3 d, float a, PtHit &h) {
    const PtRayInv r = pt_ray_inv<BARE_RCP>(o, d);
    PT_DECLARE_STACK;
    PT_STK(0) = PT_STACK_SENTINEL;
    int sp = 1;
    int node = 0;
    const float4 *nodes = reinterpret_cast<const float4 *>(bvh.nodes);
    while (node != PT_STACK_SENTINEL) {
        while (node >= 0) {
            const float4 b0 = __ldg(nodes + 4 * node), b1 = __ldg(nodes + 4 * node + 1), b2 = __ldg(nodes + 4 * node + 2);
            const int4 link = __ldg(reinterpret_cast<const int4 *>(nodes + 4 * node + 3));
            bool hl, hr;
            float ln, rn;
            pt_node_slabs(b0, b1, b2, r, h.t, hl, hr, ln, rn);
            if (hl && hr) {
                bool left_first = ln <= rn;
                PT_STK(sp++) = left_first ? link.y : link.x;
                node = left_first ? link.x : link.y;
            } else if (hl) node = link.x;
            else if (hr) node = link.y;
            else node = PT_STK(--sp);
        }
        if (node != PT_STACK_SENTINEL) {
            pt_leaf_test<LEAF>(S, node, o, d, a, h);
            node = PT_STK(--sp);
        }
    }
}

// The same walk as a resumable object: advance() descends to the next leaf, tests it and pops.  A lane can
// therefore finish one ray and begin() another while its neighbours are still walking (the wavefront
// kernel's BVH stage refills idle lanes from its ray list this way).
template <int LEAF>
struct PtWalker {
    PtRayInv r;
    float3 o, d;
    float a;
    PtHit h;
    int node, sp;
    int stack[PT_STACK_DEPTH];
    __device__ __forceinline__ void begin(float3 o_, float3 d_, const PtHit &h0) {
        o = o_; d = d_; h = h0;
        a = pt_dot_exact(d_);
        r = pt_ray_inv(o_, d_);
        stack[0] = PT_STACK_SENTINEL;
        sp = 1;
        node = 0;
    }
    __device__ __forceinline__ bool done() const { return node == PT_STACK_SENTINEL; }
    __device__ __forceinline__ void advance(const PtDeviceScene &S, const PtBvh &bvh) {
        const float4 *nodes = reinterpret_cast<const float4 *>(bvh.nodes);
        while (node >= 0) {
            const float4 b0 = __ldg(nodes + 4 * node), b1 = __ldg(nodes + 4 * node + 1), b2 = __ldg(nodes + 4 * node + 2);
            const int4 link = __ldg(reinterpret_cast<const int4 *>(nodes + 4 * node + 3));
            bool hl, hr;
            float ln, rn;
            pt_node_slabs(b0, b1, b2, r, h.t, hl, hr, ln, rn);
            if (hl && hr) {
                bool left_first = ln <= rn;
                stack[sp++] = left_first ? link.y : link.x;
                node = left_first ? link.x : link.y;
            } else if (hl) node = link.x;
            else if (hr) node = link.y;
            else node = stack[--sp];
        }
        if (node != PT_STACK_SENTINEL) {
            pt_leaf_test<LEAF>(S, node, o, d, a, h);
            node = stack[--sp];
        }
    }
};

// ---- sparse walks: the root step on its own, and the walk resumed below it ---------------------------
// (PT_F_COOP variants only.)  The kernel first asks every new ray whether it enters the tree at all; the warp then
// votes on HOW the entering rays are traced, so the root step is split off the walk and its outcome handed on:
// `node` = the child to descend first (PT_STACK_SENTINEL: the ray misses both), `far` = the other child when both are hit.
__device__ __forceinline__ void pt_bvh_root_step(const PtBvh &bvh, const PtRayInv &r, float tmax, int &node, int &far) {
    const float4 *nodes = reinterpret_cast<const float4 *>(bvh.nodes);
    const int4 link = __ldg(reinterpret_cast<const int4 *>(nodes + 3));
    bool hl, hr;
    float ln, rn;
    pt_node_slabs(__ldg(nodes), __ldg(nodes + 1), __ldg(nodes + 2), r, tmax, hl, hr, ln, rn);
    node = far = PT_STACK_SENTINEL;
    if (hl && hr) {
        const bool left_first = ln <= rn;
        node = left_first ? link.x : link.y;
        far = left_first ? link.y : link.x;
    } else if (hl) node = link.x;
    else if (hr) node = link.y;
}

template <int LEAF>
__device__ __forceinline__ void pt_bvh_walk_from(const PtDeviceScene &S, const PtBvh &bvh, const PtRayInv &r, float3 o, float3 d, float a,
                                                 PtHit &h, int node, int far) {
    PT_DECLARE_STACK;
    PT_STK(0) = PT_STACK_SENTINEL;
    int sp = 1;
    if (far != PT_STACK_SENTINEL) PT_STK(sp++) = far;
    const float4 *nodes = reinterpret_cast<const float4 *>(bvh.nodes);
    while (node != PT_STACK_SENTINEL) {
        while (node >= 0) {
            const float4 b0 = __ldg(nodes + 4 * node), b1 = __ldg(nodes + 4 * node + 1), b2 = __ldg(nodes + 4 * node + 2);
            const int4 link = __ldg(reinterpret_cast<const int4 *>(nodes + 4 * node + 3));
            bool hl, hr;
            float ln, rn;
            pt_node_slabs(b0, b1, b2, r, h.t, hl, hr, ln, rn);
            if (hl && hr) {
                bool left_first = ln <= rn;
                PT_STK(sp++) = left_first ? link.y : link.x;
                node = left_first ? link.x : link.y;
            } else if (hl) node = link.x;
            else if (hr) node = link.y;
            else node = PT_STK(--sp);
        }
        if (node != PT_STACK_SENTINEL) {
            pt_leaf_test<LEAF>(S, node, o, d, a, h);
            node = PT_STK(--sp);
        }
    }
}

// ---- warp-cooperative closest hit of ONE ray (the ray of lane `src`), executed by all 32 lanes ----------
// `cl` is the cut of the tree into n <= 32 clusters (pt_cluster.h).  Lane l tests cluster l's box; the clusters the
// ray enters are then taken nearest first (REDUX.MIN over the entry distances) until the next one lies behind the
// best hit; the lanes test a cluster's primitives 32 at a time with the same per-primitive functions as the walk,
// and two REDUX.MINs take the lexicographic minimum of (t, id) -- t > 0, so its bit pattern orders like the float.
// pt_accept keeps exactly that minimum over the same candidates, so the hit is bit-identical to the walk's.
// Must be called by the whole, converged warp.  ~100 warp instructions for a ray that crosses the mesh's box and
// misses; a BVH2 walk costs the warp 15-30 node steps of ~55 instructions however few lanes take part.
template <int LEAF>
__device__ __forceinline__ void pt_coop_trace(const PtDeviceScene &S, const float4 *__restrict__ cl, uint32_t n_clusters, int src,
                                              uint32_t lane, float3 o, float3 d, PtHit &h) {
    const uint32_t FULL = 0xffffffffu;
    const float3 O = f3(__shfl_sync(FULL, o.x, src), __shfl_sync(FULL, o.y, src), __shfl_sync(FULL, o.z, src));
    const float3 D = f3(__shfl_sync(FULL, d.x, src), __shfl_sync(FULL, d.y, src), __shfl_sync(FULL, d.z, src));
    float bt = __shfl_sync(FULL, h.t, src);
    uint32_t bid = __shfl_sync(FULL, h.id, src);
    const PtRayInv r = pt_ray_inv(O, D);
    const float a = (LEAF == PT_LEAF_SPHERES) ? pt_dot_exact(D) : 0.f;
    const float4 mn = __ldg(cl + lane), mx = __ldg(cl + 32 + lane);
    const float x0 = fmaf(mn.x, r.inv.x, -r.ood.x), x1 = fmaf(mx.x, r.inv.x, -r.ood.x);
    const float y0 = fmaf(mn.y, r.inv.y, -r.ood.y), y1 = fmaf(mx.y, r.inv.y, -r.ood.y);
    const float z0 = fmaf(mn.z, r.inv.z, -r.ood.z), z1 = fmaf(mx.z, r.inv.z, -r.ood.z);
    const float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), 0.f));
    const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), bt)) * 1.0000004f;
    bool pend = lane < n_clusters && tn <= tf;
    const uint32_t tnb = __float_as_uint(tn);                  // tn >= 0: ordered like the float
#pragma unroll 1
    while (true) {
        const uint32_t nb = __reduce_min_sync(FULL, pend ? tnb : 0xFFFFFFFFu);
        if (nb == 0xFFFFFFFFu || __uint_as_float(nb) > bt * 1.0000004f) break;     // none left, or all behind the best hit
        const int c = __ffs((int)__ballot_sync(FULL, pend && tnb == nb)) - 1;
        if ((int)lane == c) pend = false;
        const uint32_t fc = __shfl_sync(FULL, __float_as_uint(mn.w), c);           // first | count << 24
        const uint32_t first = fc & 0xFFFFFFu, count = fc >> 24;
#pragma unroll 1
        for (uint32_t base = 0; base < count; base += 32u) {
            const uint32_t pos = first + base + lane;
            float t = -1.f;
            if (base + lane < count) {
                if (LEAF == PT_LEAF_TRIS) {
                    const float4 *rec = reinterpret_cast<const float4 *>(S.tris + pos);
                    t = pt_planar_t<false>(O, D, __ldg(rec), __ldg(rec + 1), __ldg(rec + 2), PT_MIN_HIT);
                } else {
                    t = pt_sphere_t(O, D, a, __ldg(S.bvh_spheres + pos));
                }
            }
            const bool ok = t > 0.f && t <= bt;
            if (__any_sync(FULL, ok)) {
                const uint32_t tb = ok ? __float_as_uint(t) : 0x7f7fffffu;
                const uint32_t tmin = __reduce_min_sync(FULL, tb);
                uint32_t ii = PT_ID_MISS;
                if (ok && tb == tmin)
                    ii = (LEAF == PT_LEAF_TRIS) ? PT_ID(PT_HIT_TRIANGLE, __ldg(S.tri_ids + pos)) : PT_ID(PT_HIT_SPHERE, __ldg(S.bvh_sphere_ids + pos));
                const uint32_t imin = __reduce_min_sync(FULL, ii);
                const float tt = __uint_as_float(tmin);
                if (tt < bt || (tt == bt && imin < bid)) { bt = tt; bid = imin; }
            }
        }
    }
    if ((int)lane == src) { h.t = bt; h.id = bid; }
}

// the analytic (linear-list) part of the closest hit: spheres, quads, planes, boxes
template <uint32_t F>
__device__ __forceinline__ void pt_hit_analytic(const PtDeviceScene &S, float3 o, float3 d, float a, PtHit &h) {
#pragma unroll 1
    for (uint32_t i = 0; i < S.n_lin_spheres; i++) {
        float t = pt_sphere_t(o, d, a, __ldg(S.lin_spheres + i));
        if (t > 0.f && t <= h.t) pt_accept(h, t, PT_ID(PT_HIT_SPHERE, __ldg(S.lin_sphere_ids + i)));
    }
    if (F & PT_F_QUADS) {
#pragma unroll 1
        for (uint32_t i = 0; i < S.n_quads; i++) {
            const float4 *rec = reinterpret_cast<const float4 *>(S.quads + i);
            float t = pt_planar_t<true>(o, d, __ldg(rec), __ldg(rec + 1), __ldg(rec + 2), PT_QUAD_MIN_HIT);
            if (t > 0.f && t <= h.t) pt_accept(h, t, PT_ID(PT_HIT_QUAD, i));
        }
    }
#pragma unroll 1
    for (uint32_t i = 0; i < S.n_planes; i++) {
        float t = pt_plane_t(o, d, __ldg(S.planes + i));
        if (t > PT_MIN_HIT && t <= h.t) pt_accept(h, t, PT_ID(PT_HIT_PLANE, i));
    }
    if (F & PT_F_AABBS) {
#pragma unroll 1
        for (uint32_t i = 0; i < S.n_aabbs; i++) {
            int face;
            float t = pt_aabb_t(o, d, __ldg(S.aabbs + 2 * i), __ldg(S.aabbs + 2 * i + 1), &face);
            if (t > PT_MIN_HIT && t <= h.t) pt_accept(h, t, PT_ID(PT_HIT_AABB, i));
        }
    }
}

// closest hit over the whole scene; h.id == PT_ID_MISS and h.t == FLT_MAX on a miss (main.cpp:411-412)
template <uint32_t F>
__device__ __forceinline__ PtHit pt_closest_hit(const PtDeviceScene &S, float3 o, float3 d) {
    PtHit h;
    h.t = 3.402823466e+38f;
    h.id = PT_ID_MISS;
    const float a = pt_dot_exact(d);
    pt_hit_analytic<F>(S, o, d, a, h);
    if ((F & PT_F_SPH_BVH) && S.sphere_bvh.n_prims) pt_bvh_walk<PT_LEAF_SPHERES, !(F & PT_F_LENS)>(S, S.sphere_bvh, o, d, a, h);
    if ((F & PT_F_TRI_BVH) && S.tri_bvh.n_prims) pt_bvh_walk<PT_LEAF_TRIS>(S, S.tri_bvh, o, d, a, h);
    return h;
}

// geometric normal and material of a hit (ray_payload_t, include/ray.hpp:137-141)
template <uint32_t F>
__device__ __forceinline__ void pt_hit_surface(const PtDeviceScene &S, const PtHit &h, float3 o, float3 d,
                                               float3 &normal, uint32_t &mat) {
    uint32_t kind = h.id >> 28, idx = h.id & 0x0FFFFFFFu;
    normal = f3(0, 0, 0);
    mat = 0;
    if (h.id == PT_ID_MISS) return;
    if (kind == PT_HIT_SPHERE) {
        float4 s = __ldg(S.all_spheres + idx);
        // Normalize(t*dir + (origin - centre)), main.cpp:2374
        normal = normalize3(f3(fmaf(h.t, d.x, o.x - s.x), fmaf(h.t, d.y, o.y - s.y), fmaf(h.t, d.z, o.z - s.z)));
        mat = __ldg(S.sphere_mat + idx);
    } else if (kind == PT_HIT_PLANE) {
        float4 p = __ldg(S.planes + idx);
        normal = f3(p.x, p.y, p.z);
        mat = __ldg(S.plane_mat + idx);
    } else if ((F & PT_F_TRI_BVH) && kind == PT_HIT_TRIANGLE) {
        float4 n = __ldg(S.tri_normals + idx);
        normal = f3(n.x, n.y, n.z);
        mat = __ldg(S.tri_mat + idx);
    } else if ((F & PT_F_QUADS) && kind == PT_HIT_QUAD) {
        float4 r0 = __ldg(reinterpret_cast<const float4 *>(S.quads + idx));
        normal = f3(r0.x, r0.y, r0.z);
        mat = __ldg(S.quad_mat + idx);
    } else if ((F & PT_F_AABBS) && kind == PT_HIT_AABB) {
        int face = 0;
        pt_aabb_t(o, d, __ldg(S.aabbs + 2 * idx), __ldg(S.aabbs + 2 * idx + 1), &face);
        // face normal table, main.cpp:540-543: -z +z -x +x +y -y
        float sgn = (face == 0 || face == 2 || face == 5) ? -1.f : 1.f;
        normal = f3((face == 2 || face == 3) ? sgn : 0.f, (face >= 4) ? sgn : 0.f, (face < 2) ? sgn : 0.f);
        mat = __ldg(S.aabb_mat + idx);
    }
}
