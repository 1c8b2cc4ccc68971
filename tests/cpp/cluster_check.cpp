// CPU check of the cluster cut behind the warp-cooperative closest hit (pathtracer_b200/csrc/pt_cluster.h) and a
// lane-by-lane emulation of pt_coop_trace (pt_trace.cuh) against the BVH2 walk and brute force.  Test infrastructure:
// built by tests/test_clusters.py with g++; includes the product's host-side headers, none of its CUDA code.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <random>
#include <vector>

#include "pt_bvh.h"
#include "pt_cluster.h"

using namespace ptbvh;

namespace {
struct Hit { float t; uint32_t id; };
const float kMinHit = 1e-4f;

// Moeller-Trumbore; only has to be ONE deterministic function shared by the three tracers
float tri_t(const float *P, const float *o, const float *d) {
    float e1[3], e2[3], pv[3], tv[3], qv[3];
    for (int k = 0; k < 3; k++) { e1[k] = P[3 + k] - P[k]; e2[k] = P[6 + k] - P[k]; }
    pv[0] = d[1] * e2[2] - d[2] * e2[1]; pv[1] = d[2] * e2[0] - d[0] * e2[2]; pv[2] = d[0] * e2[1] - d[1] * e2[0];
    float det = e1[0] * pv[0] + e1[1] * pv[1] + e1[2] * pv[2];
    if (std::fabs(det) < 1e-12f) return -1.f;
    float inv = 1.f / det;
    for (int k = 0; k < 3; k++) tv[k] = o[k] - P[k];
    float u = (tv[0] * pv[0] + tv[1] * pv[1] + tv[2] * pv[2]) * inv;
    if (u < 0.f || u > 1.f) return -1.f;
    qv[0] = tv[1] * e1[2] - tv[2] * e1[1]; qv[1] = tv[2] * e1[0] - tv[0] * e1[2]; qv[2] = tv[0] * e1[1] - tv[1] * e1[0];
    float v = (d[0] * qv[0] + d[1] * qv[1] + d[2] * qv[2]) * inv;
    if (v < 0.f || u + v > 1.f) return -1.f;
    float t = (e2[0] * qv[0] + e2[1] * qv[1] + e2[2] * qv[2]) * inv;
    return t > kMinHit ? t : -1.f;
}
void accept(Hit &h, float t, uint32_t id) { if (t < h.t || (t == h.t && id < h.id)) { h.t = t; h.id = id; } }

struct Ray { float o[3], d[3], inv[3], ood[3]; };
Ray make_ray(const float *o, const float *d) {
    Ray r;
    for (int k = 0; k < 3; k++) {
        r.o[k] = o[k]; r.d[k] = d[k];
        float dd = std::fabs(d[k]) < 1e-20f ? std::copysign(1e-20f, d[k]) : d[k];
        r.inv[k] = 1.f / dd;
        r.ood[k] = o[k] * r.inv[k];
    }
    return r;
}
// pt_node_slabs for one box
bool slab(const float *mn, const float *mx, const Ray &r, float tmax, float &tn) {
    float a[3], b[3];
    for (int k = 0; k < 3; k++) { a[k] = mn[k] * r.inv[k] - r.ood[k]; b[k] = mx[k] * r.inv[k] - r.ood[k]; }
    tn = std::fmax(std::fmax(std::fmin(a[0], b[0]), std::fmin(a[1], b[1])), std::fmax(std::fmin(a[2], b[2]), 0.f));
    float tf = std::fmin(std::fmin(std::fmax(a[0], b[0]), std::fmax(a[1], b[1])), std::fmin(std::fmax(a[2], b[2]), tmax)) * 1.0000004f;
    return tn <= tf;
}

Hit brute(const float *tris, uint32_t n, const Ray &r) {
    Hit h{3.402823466e+38f, 0xFFFFFFFFu};
    for (uint32_t i = 0; i < n; i++) { float t = tri_t(tris + 9 * i, r.o, r.d); if (t > 0.f && t <= h.t) accept(h, t, i); }
    return h;
}
Hit walk(const Result &R, const float *tris, const Ray &r, uint64_t &steps) {
    Hit h{3.402823466e+38f, 0xFFFFFFFFu};
    std::vector<int32_t> st{0};
    while (!st.empty()) {
        int32_t node = st.back(); st.pop_back();
        if (node < 0) {
            uint32_t enc = (uint32_t)~node, first = enc & 0xFFFFFFu, count = enc >> 24;
            for (uint32_t i = 0; i < count; i++) { uint32_t id = R.order[first + i]; float t = tri_t(tris + 9 * id, r.o, r.d); if (t > 0.f && t <= h.t) accept(h, t, id); }
            continue;
        }
        steps++;
        const Node &n = R.nodes[node];
        float ln, rn;
        bool hl = slab(n.b, n.b + 3, r, h.t, ln), hr = slab(n.b + 6, n.b + 9, r, h.t, rn);
        if (hl && hr) { bool lf = ln <= rn; st.push_back(lf ? n.link[1] : n.link[0]); st.push_back(lf ? n.link[0] : n.link[1]); }
        else if (hl) st.push_back(n.link[0]);
        else if (hr) st.push_back(n.link[1]);
    }
    return h;
}
// pt_coop_trace, 32 "lanes" in a loop; image = ClusterCut::image()
Hit coop(const std::vector<float> &img, uint32_t ncl, const Result &R, const float *tris, const Ray &r, uint64_t &rounds, uint64_t &visited) {
    float bt = 3.402823466e+38f; uint32_t bid = 0xFFFFFFFFu;
    bool pend[32]; float tn[32];
    for (uint32_t l = 0; l < 32; l++) pend[l] = slab(&img[4 * l], &img[4 * (32 + l)], r, bt, tn[l]) && l < ncl;
    while (true) {
        uint32_t nb = 0xFFFFFFFFu;
        for (int l = 0; l < 32; l++) if (pend[l]) { uint32_t b; std::memcpy(&b, &tn[l], 4); nb = std::min(nb, b); }
        float nf; std::memcpy(&nf, &nb, 4);
        if (nb == 0xFFFFFFFFu || nf > bt * 1.0000004f) break;
        int c = -1;
        for (int l = 0; l < 32 && c < 0; l++) { uint32_t b; std::memcpy(&b, &tn[l], 4); if (pend[l] && b == nb) c = l; }
        pend[c] = false;
        visited++;
        uint32_t fc; std::memcpy(&fc, &img[4 * c + 3], 4);
        uint32_t first = fc & 0xFFFFFFu, count = fc >> 24;
        for (uint32_t base = 0; base < count; base += 32) {
            rounds++;
            uint32_t tmin = 0x7f7fffffu; bool any = false;
            float t[32]; bool ok[32];
            for (uint32_t l = 0; l < 32; l++) {
                t[l] = -1.f;
                if (base + l < count) t[l] = tri_t(tris + 9 * R.order[first + base + l], r.o, r.d);
                ok[l] = t[l] > 0.f && t[l] <= bt;
                if (ok[l]) { any = true; uint32_t b; std::memcpy(&b, &t[l], 4); tmin = std::min(tmin, b); }
            }
            if (!any) continue;
            uint32_t imin = 0xFFFFFFFFu;
            for (uint32_t l = 0; l < 32; l++) { uint32_t b; std::memcpy(&b, &t[l], 4); if (ok[l] && b == tmin) imin = std::min(imin, R.order[first + base + l]); }
            float tt; std::memcpy(&tt, &tmin, 4);
            if (tt < bt || (tt == bt && imin < bid)) { bt = tt; bid = imin; }
        }
    }
    return Hit{bt, bid};
}
}  // namespace

struct ClusterStats {
    uint32_t n_clusters, max_count, min_count, partition_ok, contain_ok;
    uint32_t n_rays, mismatch_walk, mismatch_coop, n_hits;
    double walk_steps_per_ray, coop_rounds_per_ray, coop_clusters_per_ray;
    float firsts[32], counts[32];
};

// boxes: n x 6 floats (min, max).  tris: n x 9 floats or NULL (then only the structure is checked).
// Rays: origins on a sphere of radius 2*extent around the centre, aimed at random points of the bounding box
// blown up by `spread` (so spread > 1 gives rays that cross the root box and miss, like the sparse walks).
extern "C" int cluster_check(const float *boxes, const float *tris, uint32_t n, int max_leaf, uint32_t n_rays, uint32_t seed, float spread,
                             ClusterStats *out) {
    std::vector<Box> pb(n);
    for (uint32_t i = 0; i < n; i++) for (int k = 0; k < 3; k++) { pb[i].mn[k] = boxes[6 * i + k]; pb[i].mx[k] = boxes[6 * i + 3 + k]; }
    Result R = Builder(pb, max_leaf).build();
    ClusterCut cut = build_cut(R);
    memset(out, 0, sizeof(*out));
    out->n_clusters = (uint32_t)cut.clusters.size();
    if (cut.clusters.empty()) return 1;
    // structure: clusters tile [0, n) and their boxes contain their primitives
    uint32_t next = 0; bool part = true, contain = true;
    out->min_count = 0xFFFFFFFFu;
    for (size_t c = 0; c < cut.clusters.size(); c++) {
        const Cluster &cl = cut.clusters[c];
        if (cl.first != next || cl.count == 0) part = false;
        next = cl.first + cl.count;
        out->max_count = std::max(out->max_count, cl.count); out->min_count = std::min(out->min_count, cl.count);
        out->firsts[c] = (float)cl.first; out->counts[c] = (float)cl.count;
        for (uint32_t i = 0; i < cl.count && cl.first + i < n; i++) {
            const Box &b = pb[R.order[cl.first + i]];
            for (int k = 0; k < 3; k++) if (b.mn[k] < cl.box.mn[k] || b.mx[k] > cl.box.mx[k]) contain = false;
        }
    }
    if (next != n) part = false;
    out->partition_ok = part; out->contain_ok = contain;
    if (!tris || !part) return 0;
    // the image round trip
    std::vector<float> img = cut.image();
    Box all; all.reset();
    for (auto &b : pb) all.grow(b);
    float ctr[3], ext = 0.f;
    for (int k = 0; k < 3; k++) { ctr[k] = 0.5f * (all.mn[k] + all.mx[k]); ext = std::max(ext, all.mx[k] - all.mn[k]); }
    std::mt19937 gen(seed);
    std::uniform_real_distribution<float> U(0.f, 1.f);
    uint64_t steps = 0, rounds = 0, visited = 0;
    for (uint32_t i = 0; i < n_rays; i++) {
        float o[3], tgt[3], d[3], len = 0.f;
        float z = 2.f * U(gen) - 1.f, ph = 6.2831853f * U(gen), s = std::sqrt(std::max(0.f, 1.f - z * z));
        o[0] = ctr[0] + 2.f * ext * s * std::cos(ph); o[1] = ctr[1] + 2.f * ext * s * std::sin(ph); o[2] = ctr[2] + 2.f * ext * z;
        for (int k = 0; k < 3; k++) { tgt[k] = ctr[k] + (U(gen) - 0.5f) * spread * (all.mx[k] - all.mn[k]); d[k] = tgt[k] - o[k]; len += d[k] * d[k]; }
        len = std::sqrt(len);
        for (int k = 0; k < 3; k++) d[k] /= len;
        Ray r = make_ray(o, d);
        Hit hb = brute(tris, n, r), hw = walk(R, tris, r, steps), hc = coop(img, out->n_clusters, R, tris, r, rounds, visited);
        out->n_rays++;
        if (hb.id != 0xFFFFFFFFu) out->n_hits++;
        if (hw.t != hb.t || hw.id != hb.id) out->mismatch_walk++;
        if (hc.t != hb.t || hc.id != hb.id) out->mismatch_coop++;
    }
    out->walk_steps_per_ray = (double)steps / n_rays;
    out->coop_rounds_per_ray = (double)rounds / n_rays;
    out->coop_clusters_per_ray = (double)visited / n_rays;
    return 0;
}

// per-ray work of the BVH2 walk and of the cooperative trace for caller-supplied rays (tools/walk_model.py):
// node steps, leaf primitive tests, clusters visited, 32-wide primitive rounds
extern "C" int walk_profile(const float *boxes, const float *tris, uint32_t n, int max_leaf, uint32_t n_rays, const float *o, const float *d,
                            uint32_t *steps, uint32_t *prim_tests, uint32_t *clusters, uint32_t *rounds, uint32_t *hit) {
    std::vector<Box> pb(n);
    for (uint32_t i = 0; i < n; i++) for (int k = 0; k < 3; k++) { pb[i].mn[k] = boxes[6 * i + k]; pb[i].mx[k] = boxes[6 * i + 3 + k]; }
    Result R = Builder(pb, max_leaf).build();
    ClusterCut cut = build_cut(R);
    if (cut.clusters.empty()) return 1;
    std::vector<float> img = cut.image();
    for (uint32_t i = 0; i < n_rays; i++) {
        Ray r = make_ray(o + 3 * i, d + 3 * i);
        // the walk, counting primitive tests as well
        Hit h{3.402823466e+38f, 0xFFFFFFFFu};
        std::vector<int32_t> st{0};
        uint32_t ns = 0, np = 0;
        while (!st.empty()) {
            int32_t node = st.back(); st.pop_back();
            if (node < 0) {
                uint32_t enc = (uint32_t)~node, first = enc & 0xFFFFFFu, count = enc >> 24;
                for (uint32_t k = 0; k < count; k++) { np++; uint32_t id = R.order[first + k]; float t = tri_t(tris + 9 * id, r.o, r.d); if (t > 0.f && t <= h.t) accept(h, t, id); }
                continue;
            }
            ns++;
            const Node &nd = R.nodes[node];
            float ln, rn;
            bool hl = slab(nd.b, nd.b + 3, r, h.t, ln), hr = slab(nd.b + 6, nd.b + 9, r, h.t, rn);
            if (hl && hr) { bool lf = ln <= rn; st.push_back(lf ? nd.link[1] : nd.link[0]); st.push_back(lf ? nd.link[0] : nd.link[1]); }
            else if (hl) st.push_back(nd.link[0]);
            else if (hr) st.push_back(nd.link[1]);
        }
        uint64_t rd = 0, vis = 0;
        coop(img, (uint32_t)cut.clusters.size(), R, tris, r, rd, vis);
        steps[i] = ns; prim_tests[i] = np; clusters[i] = (uint32_t)vis; rounds[i] = (uint32_t)rd; hit[i] = h.id != 0xFFFFFFFFu;
    }
    return 0;
}
