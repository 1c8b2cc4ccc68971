// Device LBVH build (PT_BVH_DEVICE_LBVH): Morton codes -> radix sort -> Karras radix tree -> bottom-up refit,
// emitting the same 64-byte two-children-per-node layout the host SAH builder (pt_bvh.h) produces.
// Replaces the reference's 64^3 voxel-grid octree builder (GenerateAccelerationStructure,
// src/win32_main.cpp:1188-1452) on the device; only nearest-hit semantics are contract.
//
//   k_lbvh_bounds     centroid bounds of all primitive boxes (warp shuffles + ordered-int atomics)
//   k_lbvh_keys       key = 30-bit Morton code of the centroid << 32 | primitive index  (unique keys)
//   cub radix sort    64-bit keys (the one library call: plumbing, not the render path)
//   k_lbvh_hierarchy  T. Karras, "Maximizing Parallelism in the Construction of BVHs...", HPG 2012: one thread
//                     per internal node finds its key range and split from common-prefix lengths
//   k_lbvh_refit      one thread per leaf climbs; the second arrival at a node owns it (atomic counter),
//                     unions its children's boxes and writes the node: child boxes padded like the host
//                     builder's, subtrees of <= max_leaf primitives collapsed into one leaf link
// The host then renumbers the reachable nodes depth-first (root = 0) and measures the depth.
#pragma once
#include <string>
#include <vector>

#include <cub/device/device_radix_sort.cuh>

#include "pt_api.h"
#include "pt_bvh.h"

struct PtLbvhBox { float mn[3], mx[3]; };

__device__ __forceinline__ int lbvh_f2o(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }   // order-preserving
__device__ __forceinline__ float lbvh_o2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

__global__ void k_lbvh_bounds(const PtLbvhBox *__restrict__ boxes, uint32_t n, int *__restrict__ bounds /* 3 min, 3 max (ordered ints) */) {
    float mn[3] = {3.4e38f, 3.4e38f, 3.4e38f}, mx[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        for (int k = 0; k < 3; k++) {
            float c = 0.5f * (boxes[i].mn[k] + boxes[i].mx[k]);
            mn[k] = fminf(mn[k], c);
            mx[k] = fmaxf(mx[k], c);
        }
    for (int k = 0; k < 3; k++) {
        for (int off = 16; off > 0; off >>= 1) {
            mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], off));
            mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], off));
        }
        if ((threadIdx.x & 31) == 0) {
            atomicMin(bounds + k, lbvh_f2o(mn[k]));
            atomicMax(bounds + 3 + k, lbvh_f2o(mx[k]));
        }
    }
}

__device__ __forceinline__ uint32_t lbvh_spread10(uint32_t v) {      // 10 bits -> every third bit
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

__global__ void k_lbvh_keys(const PtLbvhBox *__restrict__ boxes, uint32_t n, const int *__restrict__ bounds,
                            unsigned long long *__restrict__ keys, uint32_t *__restrict__ vals) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t q[3];
    for (int k = 0; k < 3; k++) {
        float lo = lbvh_o2f(bounds[k]), hi = lbvh_o2f(bounds[3 + k]);
        float c = 0.5f * (boxes[i].mn[k] + boxes[i].mx[k]);
        float u = hi > lo ? (c - lo) / (hi - lo) : 0.f;
        q[k] = (uint32_t)fminf(fmaxf(u * 1024.f, 0.f), 1023.f);
    }
    uint32_t code = (lbvh_spread10(q[0]) << 2) | (lbvh_spread10(q[1]) << 1) | lbvh_spread10(q[2]);
    keys[i] = ((unsigned long long)code << 32) | i;
    vals[i] = i;
}

// common-prefix length of keys i and j; -1 outside [0,n)
__device__ __forceinline__ int lbvh_delta(const unsigned long long *keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    return __clzll((long long)(keys[i] ^ keys[j]));
}

// internal node i of the radix tree: children ids use [0,n-1) = internal, n-1+k = leaf k
__global__ void k_lbvh_hierarchy(const unsigned long long *__restrict__ keys, int n, int *__restrict__ child /* 2 per node */,
                                 int *__restrict__ parent /* 2n-1 */, int2 *__restrict__ range /* first,last per internal */) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = (lbvh_delta(keys, n, i, i + 1) - lbvh_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = lbvh_delta(keys, n, i, i - d);
    int lmax = 2;
    while (lbvh_delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (lbvh_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = lbvh_delta(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) / 2;; t = (t + 1) / 2) {
        if (lbvh_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t == 1) break;
    }
    const int gamma = i + s * d + min(d, 0);
    const int first = min(i, j), last = max(i, j);
    const int left = (first == gamma) ? (n - 1 + gamma) : gamma;
    const int right = (last == gamma + 1) ? (n - 1 + gamma + 1) : gamma + 1;
    child[2 * i] = left;
    child[2 * i + 1] = right;
    parent[left] = i;
    parent[right] = i;
    range[i] = make_int2(first, last);
    if (i == 0) parent[0] = -1;
}

__device__ __forceinline__ void lbvh_pad(float *mn, float *mx) {       // as ptbvh::Builder::pad
    for (int k = 0; k < 3; k++) {
        float e = 1e-5f * fmaxf(1.f, fmaxf(fabsf(mn[k]), fabsf(mx[k])));
        mn[k] -= e;
        mx[k] += e;
    }
}

__global__ void k_lbvh_refit(const PtLbvhBox *__restrict__ boxes, const uint32_t *__restrict__ sorted, int n, int max_leaf,
                             const int *__restrict__ child, const int *__restrict__ parent, const int2 *__restrict__ range,
                             PtLbvhBox *__restrict__ nbox /* 2n-1 */, int *__restrict__ visits, ptbvh::Node *__restrict__ nodes) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    nbox[n - 1 + k] = boxes[sorted[k]];
    __threadfence();
    int node = parent[n - 1 + k];
    while (node >= 0) {
        if (atomicAdd(visits + node, 1) == 0) return;       // the sibling subtree is not finished: it will do this node
        __threadfence();
        ptbvh::Node out;
        PtLbvhBox u;
        for (int a = 0; a < 3; a++) { u.mn[a] = 3.4e38f; u.mx[a] = -3.4e38f; }
        for (int side = 0; side < 2; side++) {
            const int c = child[2 * node + side];
            PtLbvhBox b;                                     // written by another SM: read through L2
            for (int a = 0; a < 3; a++) { b.mn[a] = __ldcg(&nbox[c].mn[a]); b.mx[a] = __ldcg(&nbox[c].mx[a]); }
            for (int a = 0; a < 3; a++) { u.mn[a] = fminf(u.mn[a], b.mn[a]); u.mx[a] = fmaxf(u.mx[a], b.mx[a]); }
            lbvh_pad(b.mn, b.mx);
            float *dst = out.b + 6 * side;
            dst[0] = b.mn[0]; dst[1] = b.mn[1]; dst[2] = b.mn[2]; dst[3] = b.mx[0]; dst[4] = b.mx[1]; dst[5] = b.mx[2];
            int link;
            if (c >= n - 1) link = ~(int)((1u << 24) | (uint32_t)(c - (n - 1)));                      // single primitive
            else {
                const int2 r = range[c];
                const int cnt = r.y - r.x + 1;
                link = cnt <= max_leaf ? ~(int)(((uint32_t)cnt << 24) | (uint32_t)r.x) : c;           // collapsed subtree | inner
            }
            out.link[side] = link;
        }
        out.link[2] = out.link[3] = 0;
        nodes[node] = out;
        nbox[node] = u;
        __threadfence();
        node = parent[node];
    }
}

#define LBVH_CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { err = std::string("pt_lbvh_build: ") + cudaGetErrorString(e_); cleanup(); return PT_ERR_CUDA; } } while (0)

static inline int pt_lbvh_build(cudaStream_t st, const std::vector<ptbvh::Box> &prim_boxes, int max_leaf, ptbvh::Result &r, std::string &err) {
    const int n = (int)prim_boxes.size();
    if (n <= max_leaf || n < 2) {               // a root with one leaf child: nothing to build on the device
        r = ptbvh::Builder(prim_boxes, max_leaf).build();
        return PT_OK;
    }
    static_assert(sizeof(PtLbvhBox) == sizeof(ptbvh::Box), "box layouts must match");
    std::vector<void *> allocs;
    auto cleanup = [&]() { for (void *p : allocs) cudaFree(p); allocs.clear(); };
    auto dalloc = [&](size_t bytes) -> void * { void *p = nullptr; if (cudaMalloc(&p, bytes) != cudaSuccess) return nullptr; allocs.push_back(p); return p; };

    PtLbvhBox *d_boxes = (PtLbvhBox *)dalloc(sizeof(PtLbvhBox) * n);
    int *d_bounds = (int *)dalloc(6 * sizeof(int));
    unsigned long long *d_keys = (unsigned long long *)dalloc(sizeof(unsigned long long) * n), *d_keys2 = (unsigned long long *)dalloc(sizeof(unsigned long long) * n);
    uint32_t *d_vals = (uint32_t *)dalloc(sizeof(uint32_t) * n), *d_vals2 = (uint32_t *)dalloc(sizeof(uint32_t) * n);
    int *d_child = (int *)dalloc(sizeof(int) * 2 * (n - 1)), *d_parent = (int *)dalloc(sizeof(int) * (2 * n - 1)), *d_visits = (int *)dalloc(sizeof(int) * (n - 1));
    int2 *d_range = (int2 *)dalloc(sizeof(int2) * (n - 1));
    PtLbvhBox *d_nbox = (PtLbvhBox *)dalloc(sizeof(PtLbvhBox) * (2 * n - 1));
    ptbvh::Node *d_nodes = (ptbvh::Node *)dalloc(sizeof(ptbvh::Node) * (n - 1));
    if (!d_boxes || !d_bounds || !d_keys || !d_keys2 || !d_vals || !d_vals2 || !d_child || !d_parent || !d_visits || !d_range || !d_nbox || !d_nodes) {
        err = "pt_lbvh_build: out of device memory";
        cleanup();
        return PT_ERR_OOM;
    }
    LBVH_CK(cudaMemcpyAsync(d_boxes, prim_boxes.data(), sizeof(PtLbvhBox) * n, cudaMemcpyHostToDevice, st));
    const int init_bounds[6] = {0x7f7fffff, 0x7f7fffff, 0x7f7fffff, (int)0x80800000, (int)0x80800000, (int)0x80800000};   // +FLT_MAX / -FLT_MAX, ordered
    LBVH_CK(cudaMemcpyAsync(d_bounds, init_bounds, sizeof(init_bounds), cudaMemcpyHostToDevice, st));
    LBVH_CK(cudaMemsetAsync(d_visits, 0, sizeof(int) * (n - 1), st));
    const int T = 256, B = (n + T - 1) / T;
    k_lbvh_bounds<<<std::min(B, 1024), T, 0, st>>>(d_boxes, (uint32_t)n, d_bounds);
    k_lbvh_keys<<<B, T, 0, st>>>(d_boxes, (uint32_t)n, d_bounds, d_keys, d_vals);
    size_t tmp_bytes = 0;
    LBVH_CK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, d_keys, d_keys2, d_vals, d_vals2, n, 0, 64, st));
    void *d_tmp = dalloc(tmp_bytes ? tmp_bytes : 16);
    if (!d_tmp) { err = "pt_lbvh_build: out of device memory"; cleanup(); return PT_ERR_OOM; }
    LBVH_CK(cub::DeviceRadixSort::SortPairs(d_tmp, tmp_bytes, d_keys, d_keys2, d_vals, d_vals2, n, 0, 64, st));
    k_lbvh_hierarchy<<<B, T, 0, st>>>(d_keys2, n, d_child, d_parent, d_range);
    k_lbvh_refit<<<B, T, 0, st>>>(d_boxes, d_vals2, n, max_leaf, d_child, d_parent, d_range, d_nbox, d_visits, d_nodes);
    LBVH_CK(cudaGetLastError());

    std::vector<ptbvh::Node> raw(n - 1);
    r.order.resize(n);
    LBVH_CK(cudaMemcpyAsync(raw.data(), d_nodes, sizeof(ptbvh::Node) * (n - 1), cudaMemcpyDeviceToHost, st));
    LBVH_CK(cudaMemcpyAsync(r.order.data(), d_vals2, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost, st));
    LBVH_CK(cudaStreamSynchronize(st));
    cleanup();

    // keep the reachable inner nodes only, depth-first from the root (= Karras node 0), and measure the depth
    r.nodes.clear();
    r.nodes.reserve(n);
    r.max_depth = 0;
    struct Item { int src, dst, depth; };
    std::vector<Item> stack;
    r.nodes.push_back(raw[0]);
    stack.push_back({0, 0, 1});
    while (!stack.empty()) {
        Item it = stack.back();
        stack.pop_back();
        r.max_depth = std::max(r.max_depth, it.depth);
        for (int side = 1; side >= 0; side--) {                       // left child gets the next index
            int link = raw[it.src].link[side];
            if (link < 0) continue;
            int dst = (int)r.nodes.size();
            r.nodes.push_back(raw[link]);
            r.nodes[it.dst].link[side] = dst;
            stack.push_back({link, dst, it.depth + 1});
        }
    }
    return PT_OK;
}
#undef LBVH_CK
