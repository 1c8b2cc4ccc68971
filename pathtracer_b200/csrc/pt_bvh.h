// Host-side binned-SAH BVH2 builder producing the flattened 64-byte node layout of pt_device.cuh.
// Replaces the reference's 64^3 voxel-grid octree (GenerateAccelerationStructure,
// src/win32_main.cpp:1188-1447); only its nearest-hit semantics are contract.
#pragma once
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <vector>

namespace ptbvh {

struct Box {
    float mn[3], mx[3];
    void reset() { for (int k = 0; k < 3; k++) { mn[k] = FLT_MAX; mx[k] = -FLT_MAX; } }
    void grow(const Box &b) { for (int k = 0; k < 3; k++) { mn[k] = std::min(mn[k], b.mn[k]); mx[k] = std::max(mx[k], b.mx[k]); } }
    void grow(const float *p) { for (int k = 0; k < 3; k++) { mn[k] = std::min(mn[k], p[k]); mx[k] = std::max(mx[k], p[k]); } }
    float area() const {
        float dx = mx[0] - mn[0], dy = mx[1] - mn[1], dz = mx[2] - mn[2];
        if (dx < 0 || dy < 0 || dz < 0) return 0.f;
        return 2.f * (dx * dy + dy * dz + dz * dx);
    }
};

struct Node {                 // same 64 bytes as PtBvhNode
    float b[12];              // L.min.xyz L.max.x | L.max.yz R.min.xy | R.min.z R.max.xyz
    int32_t link[4];          // left, right, 0, 0
};

struct Result {
    std::vector<Node> nodes;          // nodes[0] = root (inner)
    std::vector<uint32_t> order;      // position -> original primitive index
    int max_depth = 0;
};

inline int32_t encode_leaf(uint32_t first, uint32_t count) { return ~(int32_t)((count << 24) | first); }

inline void store_child_box(Node &n, int side, const Box &b) {
    if (side == 0) { n.b[0] = b.mn[0]; n.b[1] = b.mn[1]; n.b[2] = b.mn[2]; n.b[3] = b.mx[0]; n.b[4] = b.mx[1]; n.b[5] = b.mx[2]; }
    else { n.b[6] = b.mn[0]; n.b[7] = b.mn[1]; n.b[8] = b.mn[2]; n.b[9] = b.mx[0]; n.b[10] = b.mx[1]; n.b[11] = b.mx[2]; }
}

class Builder {
public:
    Builder(const std::vector<Box> &prim_boxes, int max_leaf, bool sweep = true) : boxes_(prim_boxes), max_leaf_(max_leaf), sweep_(sweep) {
        cent_.resize(boxes_.size() * 3);
        for (size_t i = 0; i < boxes_.size(); i++)
            for (int k = 0; k < 3; k++) cent_[3 * i + k] = 0.5f * (boxes_[i].mn[k] + boxes_[i].mx[k]);
    }

    Result build() {
        Result r;
        const uint32_t n = (uint32_t)boxes_.size();
        r.order.resize(n);
        for (uint32_t i = 0; i < n; i++) r.order[i] = i;
        out_ = &r;
        r.nodes.reserve(2 * n + 2);
        if (n <= (uint32_t)max_leaf_) {
            // root must be an inner node: all primitives on the left, an empty leaf on the right
            Node root{};
            Box all; all.reset();
            for (uint32_t i = 0; i < n; i++) all.grow(boxes_[i]);
            pad(all);
            Box empty; for (int k = 0; k < 3; k++) { empty.mn[k] = 1e30f; empty.mx[k] = -1e30f; }
            store_child_box(root, 0, all);
            store_child_box(root, 1, empty);
            root.link[0] = encode_leaf(0, n);
            root.link[1] = encode_leaf(0, 0);
            r.nodes.push_back(root);
            r.max_depth = 1;
            return r;
        }
        r.nodes.emplace_back();
        split(0, 0, n, 1);
        return r;
    }

private:
    static void pad(Box &b) {
        // conservative: widen by a few ulps so the slab test never culls a primitive lying on the boundary
        for (int k = 0; k < 3; k++) {
            float e = 1e-5f * std::max(1.f, std::max(std::fabs(b.mn[k]), std::fabs(b.mx[k])));
            b.mn[k] -= e; b.mx[k] += e;
        }
    }

    Box range_box(uint32_t first, uint32_t last) const {
        Box b; b.reset();
        for (uint32_t i = first; i < last; i++) b.grow(boxes_[out_->order[i]]);
        return b;
    }

    // exact sweep SAH for small ranges: sort by centroid on each axis and evaluate every split position
    uint32_t partition_sweep(uint32_t first, uint32_t last) {
        const uint32_t n = last - first;
        float best_cost = FLT_MAX; int best_axis = -1; uint32_t best_k = 0;
        std::vector<uint32_t> idx(n), best_order;
        std::vector<float> right_area(n);
        for (int ax = 0; ax < 3; ax++) {
            for (uint32_t i = 0; i < n; i++) idx[i] = out_->order[first + i];
            std::stable_sort(idx.begin(), idx.end(), [&](uint32_t a, uint32_t b) { return cent_[3 * a + ax] < cent_[3 * b + ax]; });
            Box acc; acc.reset();
            for (uint32_t i = n; i-- > 1;) { acc.grow(boxes_[idx[i]]); right_area[i] = acc.area(); }
            acc.reset();
            for (uint32_t k = 1; k < n; k++) {                 // left = idx[0..k), right = idx[k..n)
                acc.grow(boxes_[idx[k - 1]]);
                float cost = acc.area() * (float)k + right_area[k] * (float)(n - k);
                if (cost < best_cost) { best_cost = cost; best_axis = ax; best_k = k; best_order = idx; }
            }
        }
        if (best_axis < 0) return first + n / 2;
        for (uint32_t i = 0; i < n; i++) out_->order[first + i] = best_order[i];
        return first + best_k;
    }

    // chooses the split of [first,last) and returns mid; falls back to a median split when SAH finds nothing
    uint32_t partition(uint32_t first, uint32_t last) {
        if (sweep_ && last - first <= 4096) return partition_sweep(first, last);
        const int NB = 16;
        Box cb; cb.reset();
        for (uint32_t i = first; i < last; i++) cb.grow(&cent_[3 * out_->order[i]]);
        float best_cost = FLT_MAX; int best_axis = -1, best_bin = -1;
        for (int ax = 0; ax < 3; ax++) {
            float lo = cb.mn[ax], hi = cb.mx[ax];
            if (!(hi - lo > 1e-12f)) continue;
            Box bb[NB]; uint32_t cnt[NB];
            for (int b = 0; b < NB; b++) { bb[b].reset(); cnt[b] = 0; }
            float scale = NB / (hi - lo);
            for (uint32_t i = first; i < last; i++) {
                uint32_t p = out_->order[i];
                int b = std::min(NB - 1, std::max(0, (int)((cent_[3 * p + ax] - lo) * scale)));
                bb[b].grow(boxes_[p]); cnt[b]++;
            }
            float la[NB], ra[NB]; uint32_t lc[NB], rc[NB];
            Box acc; acc.reset(); uint32_t c = 0;
            for (int b = 0; b < NB; b++) { acc.grow(bb[b]); c += cnt[b]; la[b] = acc.area(); lc[b] = c; }
            acc.reset(); c = 0;
            for (int b = NB - 1; b >= 0; b--) { acc.grow(bb[b]); c += cnt[b]; ra[b] = acc.area(); rc[b] = c; }
            for (int b = 0; b < NB - 1; b++) {
                if (lc[b] == 0 || rc[b + 1] == 0) continue;
                float cost = la[b] * lc[b] + ra[b + 1] * rc[b + 1];
                if (cost < best_cost) { best_cost = cost; best_axis = ax; best_bin = b; }
            }
        }
        uint32_t mid;
        if (best_axis >= 0) {
            float lo = cb.mn[best_axis], hi = cb.mx[best_axis], scale = NB / (hi - lo);
            auto it = std::partition(out_->order.begin() + first, out_->order.begin() + last, [&](uint32_t p) {
                int b = std::min(NB - 1, std::max(0, (int)((cent_[3 * p + best_axis] - lo) * scale)));
                return b <= best_bin;
            });
            mid = (uint32_t)(it - out_->order.begin());
            if (mid > first && mid < last) return mid;
        }
        // median split along the widest centroid axis (or plain halves if all centroids coincide)
        int ax = 0;
        for (int k = 1; k < 3; k++) if (cb.mx[k] - cb.mn[k] > cb.mx[ax] - cb.mn[ax]) ax = k;
        mid = first + (last - first) / 2;
        std::nth_element(out_->order.begin() + first, out_->order.begin() + mid, out_->order.begin() + last,
                         [&](uint32_t a, uint32_t b) { return cent_[3 * a + ax] < cent_[3 * b + ax]; });
        return mid;
    }

    int32_t make_child(uint32_t first, uint32_t last, int depth, Box &box_out) {
        box_out = range_box(first, last);
        pad(box_out);
        uint32_t n = last - first;
        if (n <= (uint32_t)max_leaf_) {
            out_->max_depth = std::max(out_->max_depth, depth);
            return encode_leaf(first, n);
        }
        uint32_t idx = (uint32_t)out_->nodes.size();
        out_->nodes.emplace_back();
        split(idx, first, last, depth + 1);
        return (int32_t)idx;
    }

    void split(uint32_t node, uint32_t first, uint32_t last, int depth) {
        uint32_t mid = partition(first, last);
        Box lb, rb;
        int32_t l = make_child(first, mid, depth, lb);
        int32_t r = make_child(mid, last, depth, rb);
        Node &n = out_->nodes[node];
        store_child_box(n, 0, lb);
        store_child_box(n, 1, rb);
        n.link[0] = l; n.link[1] = r; n.link[2] = 0; n.link[3] = 0;
    }

    const std::vector<Box> &boxes_;
    std::vector<float> cent_;
    int max_leaf_;
    bool sweep_ = true;       // exact SAH below 4096 primitives per range, 16 bins above
    Result *out_ = nullptr;
};

}  // namespace ptbvh
