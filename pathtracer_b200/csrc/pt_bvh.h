// Host-side binned-SAH BVH2 builder producing the flattened 64-byte node layout of pt_device.cuh.
// Replaces the reference's 64^3 voxel-grid octree (GenerateAccelerationStructure,
// src/win32_main.cpp:1188-1447); only its nearest-hit semantics are contract.
#pragma once
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <xmmintrin.h>
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
        box4_.resize(boxes_.size());
        for (size_t i = 0; i < boxes_.size(); i++) {
            box4_[i].mn = _mm_set_ps(0.f, boxes_[i].mn[2], boxes_[i].mn[1], boxes_[i].mn[0]);
            box4_[i].mx = _mm_set_ps(0.f, boxes_[i].mx[2], boxes_[i].mx[1], boxes_[i].mx[0]);
        }
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
        split(0, 0, n, 1, -1);
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

    // exact sweep SAH for small ranges: sort by centroid on each axis and evaluate every split position.
    // Host time matters here: the BVH build is most of pt_upload_scene, i.e. of the end-to-end overhead per frame at 8 GPUs.
    // Same splits and the same tree as the plain formulation (a std::stable_sort of the range per axis, scalar boxes), but
    //   * the stable sort is a stable LSD radix sort of the ordered float bits (4 passes of 8 bits; a comparison sort of 752 keys
    //     is ~7 000 badly predicted branches, and it ran three times per node), insertion sort below 48 elements;
    //   * the axis the PARENT split on needs no sort: the incoming order is that sort, ties included;
    //   * boxes are swept as two SSE registers (min, max) -- a grow is two instructions.
    struct Box4 {
        __m128 mn, mx;
        void reset() { mn = _mm_set1_ps(FLT_MAX); mx = _mm_set1_ps(-FLT_MAX); }
        void grow(const Box4 &b) { mn = _mm_min_ps(mn, b.mn); mx = _mm_max_ps(mx, b.mx); }
        float area() const {                               // == Box::area(): 2 (dx dy + dy dz + dz dx), 0 for an empty box
            const __m128 d = _mm_sub_ps(mx, mn);
            alignas(16) float e[4];
            _mm_store_ps(e, d);
            if (e[0] < 0 || e[1] < 0 || e[2] < 0) return 0.f;
            return 2.f * (e[0] * e[1] + e[1] * e[2] + e[2] * e[0]);
        }
    };

    // idx <- order[first, first + n) stably sorted by the centroid along ax (ties keep the incoming order, as std::stable_sort would)
    void sort_by_centroid(uint32_t first, uint32_t n, int ax, std::vector<uint32_t> &idx) {
        keys_.resize(2 * (size_t)n);
        uint64_t *a = keys_.data(), *b = a + n;               // (ordered float bits << 32) | primitive
        for (uint32_t i = 0; i < n; i++) {
            const uint32_t p = out_->order[first + i];
            uint32_t bits;
            const float c = cent_[3 * p + ax];
            std::memcpy(&bits, &c, 4);
            if (bits == 0x80000000u) bits = 0;                                   // -0 == +0, as operator< has it
            bits = (bits & 0x80000000u) ? ~bits : (bits | 0x80000000u);          // the floats' order; centroids are never NaN
            a[i] = ((uint64_t)bits << 32) | p;
        }
        if (n < 48) {
            for (uint32_t i = 1; i < n; i++) {                                   // stable insertion sort on the float bits only
                const uint64_t v = a[i];
                uint32_t j = i;
                while (j > 0 && (a[j - 1] >> 32) > (v >> 32)) { a[j] = a[j - 1]; j--; }
                a[j] = v;
            }
        } else {
            for (int pass = 0; pass < 4; pass++) {
                const int shift = 32 + 8 * pass;
                uint32_t count[257] = {0};
                for (uint32_t i = 0; i < n; i++) count[((a[i] >> shift) & 255u) + 1]++;
                if (count[((a[0] >> shift) & 255u) + 1] == n) continue;          // every key has this digit: the pass is the identity
                for (int d = 0; d < 256; d++) count[d + 1] += count[d];
                for (uint32_t i = 0; i < n; i++) b[count[(a[i] >> shift) & 255u]++] = a[i];
                std::swap(a, b);
            }
        }
        for (uint32_t i = 0; i < n; i++) idx[i] = (uint32_t)a[i];
    }

    uint32_t partition_sweep(uint32_t first, uint32_t last, int sorted_axis, int *axis_out) {
        const uint32_t n = last - first;
        float best_cost = FLT_MAX; int best_axis = -1; uint32_t best_k = 0;
        if (right_area_.size() < n) right_area_.resize(n);
        for (int ax = 0; ax < 3; ax++) {
            std::vector<uint32_t> &idx = idx_[ax];
            idx.resize(n);
            if (ax == sorted_axis) {
                for (uint32_t i = 0; i < n; i++) idx[i] = out_->order[first + i];
            } else {
                sort_by_centroid(first, n, ax, idx);
            }
            Box4 acc; acc.reset();
            for (uint32_t i = n; i-- > 1;) { acc.grow(box4_[idx[i]]); right_area_[i] = acc.area(); }
            acc.reset();
            for (uint32_t k = 1; k < n; k++) {                 // left = idx[0..k), right = idx[k..n)
                acc.grow(box4_[idx[k - 1]]);
                float cost = acc.area() * (float)k + right_area_[k] * (float)(n - k);
                if (cost < best_cost) { best_cost = cost; best_axis = ax; best_k = k; }
            }
        }
        *axis_out = best_axis;
        if (best_axis < 0) return first + n / 2;
        std::copy(idx_[best_axis].begin(), idx_[best_axis].end(), out_->order.begin() + first);
        return first + best_k;
    }

    // chooses the split of [first,last) and returns mid; falls back to a median split when SAH finds nothing
    // sorted_axis: the axis along which [first,last) is already sorted (the parent's sweep split), -1 = none; *axis_out = the
    // axis this split leaves the two halves sorted along (-1 after a binned or median split)
    uint32_t partition(uint32_t first, uint32_t last, int sorted_axis, int *axis_out) {
        if (sweep_ && last - first <= 4096) return partition_sweep(first, last, sorted_axis, axis_out);
        *axis_out = -1;
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

    int32_t make_child(uint32_t first, uint32_t last, int depth, Box &box_out, int sorted_axis) {
        box_out = range_box(first, last);
        pad(box_out);
        uint32_t n = last - first;
        if (n <= (uint32_t)max_leaf_) {
            out_->max_depth = std::max(out_->max_depth, depth);
            return encode_leaf(first, n);
        }
        uint32_t idx = (uint32_t)out_->nodes.size();
        out_->nodes.emplace_back();
        split(idx, first, last, depth + 1, sorted_axis);
        return (int32_t)idx;
    }

    void split(uint32_t node, uint32_t first, uint32_t last, int depth, int sorted_axis) {
        int axis = -1;
        uint32_t mid = partition(first, last, sorted_axis, &axis);
        Box lb, rb;
        int32_t l = make_child(first, mid, depth, lb, axis);
        int32_t r = make_child(mid, last, depth, rb, axis);
        Node &n = out_->nodes[node];
        store_child_box(n, 0, lb);
        store_child_box(n, 1, rb);
        n.link[0] = l; n.link[1] = r; n.link[2] = 0; n.link[3] = 0;
    }

    const std::vector<Box> &boxes_;
    std::vector<float> cent_;
    std::vector<Box4> box4_;          // the same boxes as (min, max) SSE pairs for the sweeps
    std::vector<uint32_t> idx_[3];    // partition_sweep scratch: the range sorted along each axis
    std::vector<float> right_area_;
    std::vector<uint64_t> keys_;
    int max_leaf_;
    bool sweep_ = true;       // exact SAH below 4096 primitives per range, 16 bins above
    Result *out_ = nullptr;
};

}  // namespace ptbvh
