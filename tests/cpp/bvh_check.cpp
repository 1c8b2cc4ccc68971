// TEST INFRASTRUCTURE: the host SAH builder (pathtracer_b200/csrc/pt_bvh.h: radix-sorted keys, the parent's axis not re-sorted,
// SSE boxes) against the plain formulation of the same algorithm -- std::stable_sort of the range along every axis at every node,
// scalar boxes -- written out here.  The two must produce the IDENTICAL tree: same nodes (boxes bit for bit, links) and the
// same primitive order.  extern "C" int bvh_check(boxes, n, max_leaf) returns 0 when identical, else a code saying what differs.
#include <algorithm>
#include <cfloat>
#include <cstring>
#include <vector>

#include "pt_bvh.h"

namespace plain {
using ptbvh::Box;
using ptbvh::Node;

struct Builder {
    const std::vector<Box> &boxes;
    std::vector<float> cent;
    int max_leaf;
    std::vector<Node> nodes;
    std::vector<uint32_t> order;
    int max_depth = 0;

    Builder(const std::vector<Box> &b, int ml) : boxes(b), max_leaf(ml) {
        cent.resize(b.size() * 3);
        for (size_t i = 0; i < b.size(); i++)
            for (int k = 0; k < 3; k++) cent[3 * i + k] = 0.5f * (b[i].mn[k] + b[i].mx[k]);
    }
    static void pad(Box &b) {
        for (int k = 0; k < 3; k++) {
            float e = 1e-5f * std::max(1.f, std::max(std::fabs(b.mn[k]), std::fabs(b.mx[k])));
            b.mn[k] -= e; b.mx[k] += e;
        }
    }
    uint32_t partition(uint32_t first, uint32_t last) {
        const uint32_t n = last - first;
        float best_cost = FLT_MAX; int best_axis = -1; uint32_t best_k = 0;
        std::vector<uint32_t> idx(n), best_order;
        std::vector<float> right_area(n);
        for (int ax = 0; ax < 3; ax++) {
            for (uint32_t i = 0; i < n; i++) idx[i] = order[first + i];
            std::stable_sort(idx.begin(), idx.end(), [&](uint32_t a, uint32_t b) { return cent[3 * a + ax] < cent[3 * b + ax]; });
            Box acc; acc.reset();
            for (uint32_t i = n; i-- > 1;) { acc.grow(boxes[idx[i]]); right_area[i] = acc.area(); }
            acc.reset();
            for (uint32_t k = 1; k < n; k++) {
                acc.grow(boxes[idx[k - 1]]);
                float cost = acc.area() * (float)k + right_area[k] * (float)(n - k);
                if (cost < best_cost) { best_cost = cost; best_axis = ax; best_k = k; best_order = idx; }
            }
        }
        if (best_axis < 0) return first + n / 2;
        for (uint32_t i = 0; i < n; i++) order[first + i] = best_order[i];
        return first + best_k;
    }
    int32_t child(uint32_t first, uint32_t last, int depth, Box &box) {
        box.reset();
        for (uint32_t i = first; i < last; i++) box.grow(boxes[order[i]]);
        pad(box);
        uint32_t n = last - first;
        if (n <= (uint32_t)max_leaf) { max_depth = std::max(max_depth, depth); return ptbvh::encode_leaf(first, n); }
        uint32_t idx = (uint32_t)nodes.size();
        nodes.emplace_back();
        split(idx, first, last, depth + 1);
        return (int32_t)idx;
    }
    void split(uint32_t node, uint32_t first, uint32_t last, int depth) {
        uint32_t mid = partition(first, last);
        Box lb, rb;
        int32_t l = child(first, mid, depth, lb), r = child(mid, last, depth, rb);
        Node &n = nodes[node];
        ptbvh::store_child_box(n, 0, lb);
        ptbvh::store_child_box(n, 1, rb);
        n.link[0] = l; n.link[1] = r; n.link[2] = 0; n.link[3] = 0;
    }
    void build() {
        const uint32_t n = (uint32_t)boxes.size();
        order.resize(n);
        for (uint32_t i = 0; i < n; i++) order[i] = i;
        nodes.reserve(2 * n + 2);
        nodes.emplace_back();
        split(0, 0, n, 1);
    }
};
}  // namespace plain

extern "C" int bvh_check(const float *boxes6, uint32_t n, int max_leaf) {
    std::vector<ptbvh::Box> b(n);
    for (uint32_t i = 0; i < n; i++) { std::memcpy(b[i].mn, boxes6 + 6 * i, 12); std::memcpy(b[i].mx, boxes6 + 6 * i + 3, 12); }
    if (n <= (uint32_t)max_leaf) return 0;                  // the one-leaf root: nothing to compare
    ptbvh::Result r = ptbvh::Builder(b, max_leaf, true).build();
    plain::Builder p(b, max_leaf);
    p.build();
    if (r.order != p.order) return 1;
    if (r.nodes.size() != p.nodes.size()) return 2;
    if (std::memcmp(r.nodes.data(), p.nodes.data(), r.nodes.size() * sizeof(ptbvh::Node)) != 0) return 3;
    if (r.max_depth != p.max_depth) return 4;
    return 0;
}
