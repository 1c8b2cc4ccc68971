// Host side of the warp-cooperative closest hit (pt_trace.cuh, pt_coop_trace): a CUT of the BVH2 into at most
// 32 subtrees ("clusters"), one per lane of a warp.  A cluster is the subtree's (already padded) box plus the
// contiguous range of leaf-order primitive positions below it, so a warp can trace ONE ray with 32 lanes: every
// lane tests one cluster box, then the lanes test the primitives of each cluster the ray enters, 32 at a time.
// It serves the sparse walks of the megakernel -- the few bounce rays of a warp that cross the mesh's box while
// the other lanes have nothing to traverse (profiles/r01e_bvh_walk_occupancy_w5.txt).
// Replaces nothing in the reference: like the BVH it is an acceleration structure over RayCastIntersect's
// triangle / sphere loops (src/win32_main.cpp:418-526) and keeps their nearest-hit semantics.
#pragma once
#include <cstring>

#include "pt_bvh.h"

namespace ptbvh {

constexpr int kClusterLanes = 32;

struct Cluster {
    Box box;
    uint32_t first = 0, count = 0;
};

struct ClusterCut {
    std::vector<Cluster> clusters;      // <= 32; empty when the tree cannot be cut (see build_cut)
    // device image: 64 x 16 bytes; [l] = (min.xyz, bits: first | count << 24), [32 + l] = (max.xyz, 0)
    std::vector<float> image() const {
        std::vector<float> v(2 * kClusterLanes * 4, 0.f);
        for (int l = 0; l < kClusterLanes; l++) {
            float *mn = &v[4 * l], *mx = &v[4 * (kClusterLanes + l)];
            uint32_t first = 0, count = 0;
            if (l < (int)clusters.size()) {
                const Cluster &c = clusters[l];
                for (int k = 0; k < 3; k++) { mn[k] = c.box.mn[k]; mx[k] = c.box.mx[k]; }
                first = c.first; count = c.count;
            } else {
                for (int k = 0; k < 3; k++) mn[k] = mx[k] = 1e30f;      // never counted: lanes >= n are masked off
            }
            const uint32_t packed = first | (count << 24);          // count <= 4 * 32, first < 2^24 (as in the leaf links)
            std::memcpy(mn + 3, &packed, 4);
        }
        return v;
    }
};

namespace detail {
struct CutEntry {
    Box box;
    int32_t link;
    uint32_t first, count;
};

inline Box child_box(const Node &n, int side) {
    Box b;
    const float *p = n.b + 6 * side;
    for (int k = 0; k < 3; k++) { b.mn[k] = p[k]; b.mx[k] = p[3 + k]; }
    return b;
}

// Primitive ranges of every inner node, computed once (post-order, iterative): ranges[i] = positions below nodes[i];
// ok = false when some subtree is not one contiguous run of leaf-order positions (then the tree has no cut).
struct Ranges {
    std::vector<uint32_t> first, count;
    bool ok = true;
    bool of_link(int32_t link, uint32_t &f, uint32_t &c) const {
        if (link < 0) {
            uint32_t enc = (uint32_t)(~link);
            f = enc & 0xFFFFFFu; c = enc >> 24;
            return true;
        }
        if ((size_t)link >= first.size()) return false;
        f = first[link]; c = count[link];
        return true;
    }
};

inline Ranges node_ranges(const Result &r) {
    Ranges R;
    const size_t n = r.nodes.size();
    R.first.assign(n, 0); R.count.assign(n, 0);
    std::vector<uint8_t> state(n, 0);                 // 0 = unseen, 1 = children pushed, 2 = done
    std::vector<int32_t> stack;
    if (n) stack.push_back(0);
    while (!stack.empty() && R.ok) {
        const int32_t i = stack.back();
        if (state[i] == 2) { stack.pop_back(); continue; }
        if (state[i] == 0) {
            state[i] = 1;
            for (int side = 0; side < 2; side++) {
                const int32_t l = r.nodes[i].link[side];
                if (l >= 0) {
                    if ((size_t)l >= n || state[l] != 0) { R.ok = false; break; }      // not a tree
                    stack.push_back(l);
                }
            }
            continue;
        }
        uint32_t f0, c0, f1, c1;
        if (!R.of_link(r.nodes[i].link[0], f0, c0) || !R.of_link(r.nodes[i].link[1], f1, c1)) { R.ok = false; break; }
        if (c0 == 0) { R.first[i] = f1; R.count[i] = c1; }
        else if (c1 == 0) { R.first[i] = f0; R.count[i] = c0; }
        else if (f0 + c0 == f1) { R.first[i] = f0; R.count[i] = c0 + c1; }
        else if (f1 + c1 == f0) { R.first[i] = f1; R.count[i] = c0 + c1; }
        else R.ok = false;
        state[i] = 2;
        stack.pop_back();
    }
    return R;
}

inline float rounds(uint32_t count) { return (float)((count + kClusterLanes - 1) / kClusterLanes); }
}  // namespace detail

// Greedy cut: start from the root's two children and keep opening the inner node whose replacement by its two
// children lowers the expected number of 32-wide primitive rounds the most, cost(x) = area(x) * ceil(count(x)/32)
// (the surface-area heuristic with "one round" as the unit of work), until 32 clusters exist or nothing gains.
inline ClusterCut build_cut(const Result &r, uint32_t max_rounds_per_cluster = 4) {
    using namespace detail;
    ClusterCut out;
    if (r.nodes.empty()) return out;
    const Ranges R = node_ranges(r);
    if (!R.ok) return out;
    std::vector<CutEntry> cut;
    auto push = [&](const Node &n, int side) -> bool {
        CutEntry e;
        e.box = child_box(n, side);
        e.link = n.link[side];
        if (!R.of_link(e.link, e.first, e.count)) return false;
        if (e.count) cut.push_back(e);
        return true;
    };
    if (!push(r.nodes[0], 0) || !push(r.nodes[0], 1)) return out;
    while ((int)cut.size() < kClusterLanes) {
        int best = -1;
        float best_gain = 0.f;
        for (size_t i = 0; i < cut.size(); i++) {
            const CutEntry &e = cut[i];
            if (e.link < 0) continue;
            const Node &n = r.nodes[e.link];
            uint32_t f, c0 = 0, c1 = 0;
            if (!R.of_link(n.link[0], f, c0) || !R.of_link(n.link[1], f, c1)) return out;
            float gain = e.box.area() * rounds(e.count) - child_box(n, 0).area() * rounds(c0) - child_box(n, 1).area() * rounds(c1);
            if (e.count > max_rounds_per_cluster * kClusterLanes) gain = 1e30f + (float)e.count;      // must be opened
            if (gain > best_gain) { best_gain = gain; best = (int)i; }
        }
        if (best < 0) break;
        const Node n = r.nodes[cut[best].link];
        cut.erase(cut.begin() + best);
        if (!push(n, 0) || !push(n, 1)) return out;
    }
    for (const CutEntry &e : cut)
        if (e.count > max_rounds_per_cluster * kClusterLanes) return out;      // too big a mesh for one level of 32 clusters
    std::sort(cut.begin(), cut.end(), [](const CutEntry &a, const CutEntry &b) { return a.first < b.first; });
    for (const CutEntry &e : cut) {
        Cluster c;
        c.box = e.box; c.first = e.first; c.count = e.count;
        out.clusters.push_back(c);
    }
    return out;
}

}  // namespace ptbvh
