// build_gpu.cuh — acceleration-structure build ON THE GPU. Replaces, like bvh.cpp, the reference's
// per-mesh octree build (src/object/mesh/octree.rs:31-210) and object BVH (src/object/bvh.rs:32-81) by one hierarchy,
// and emits the SAME 80-byte 8-wide quantised nodes (scene_types.h: BvhNode) the traversal kernels walk:
//
//   1 k_morton           63-bit Morton code of every primitive's box centre in the cube around the scene
//   2 cub::DeviceRadixSort  (code, primitive) pairs
//   3 k_radix_tree       Karras 2012: one thread per internal node finds its range and split from the common prefixes of
//                        the sorted codes (ties broken by position); children / parent links, leaf ranges
//   4 k_refit_dp         bottom-up (the second thread to arrive at a node continues upward): boxes, and the SAH-optimal
//                        wide-collapse tables of Ylitie, Karras, Laine 2017 section 3.1 — the dynamic programme bvh.cpp
//                        runs on the host, C(n, i) = cheapest representation of the subtree of n as at most i slots
//   5 k_collapse         level by level from the root: one thread per wide node follows the DP's decisions to its <= 8
//                        children, assigns octant-ordered slots, quantises the child boxes conservatively (8 bits per
//                        plane), reserves its internal children and its leaf primitives with one atomic each
//
// Primitives of a radix-tree subtree are a contiguous range of the sorted order, so a subtree of <= 3 primitives is a
// leaf slot without any gathering. The binary tree is an LBVH (spatial median splits), not a SAH tree: the build is
// ~100x faster than the host's, the tree traverses somewhat slower (measured: DESIGN.md section 7).
#pragma once
#include <cfloat>
#include <cub/device/device_radix_sort.cuh>

#include "bvh.h"

namespace rt {
namespace gpubuild {

constexpr int MAX_LEAF = 3; // a wide node's leaf slot holds at most 3 primitives (meta byte encoding)

struct Box6 { float lo[3], hi[3]; };

__device__ __forceinline__ uint64_t spread21(uint64_t x) { // 21 bits -> every third bit
    x &= 0x1FFFFFull;
    x = (x | (x << 32)) & 0x1F00000000FFFFull;
    x = (x | (x << 16)) & 0x1F0000FF0000FFull;
    x = (x | (x << 8)) & 0x100F00F00F00F00Full;
    x = (x | (x << 4)) & 0x10C30C30C30C30C3ull;
    x = (x | (x << 2)) & 0x1249249249249249ull;
    return x;
}

__global__ void k_morton(const PrimBox* boxes, uint32_t n, float ox, float oy, float oz, float inv_extent, uint64_t* keys, uint32_t* idx) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const PrimBox b = boxes[i];
    const float cx = (0.5f * (b.lo[0] + b.hi[0]) - ox) * inv_extent, cy = (0.5f * (b.lo[1] + b.hi[1]) - oy) * inv_extent,
                cz = (0.5f * (b.lo[2] + b.hi[2]) - oz) * inv_extent;
    const float s = 2097151.f; // 2^21 - 1
    const uint64_t x = (uint64_t)fminf(fmaxf(cx * s, 0.f), s), y = (uint64_t)fminf(fmaxf(cy * s, 0.f), s), z = (uint64_t)fminf(fmaxf(cz * s, 0.f), s);
    keys[i] = (spread21(x) << 2) | (spread21(y) << 1) | spread21(z);
    idx[i] = i;
}

// binary radix tree over n leaves: internal nodes [0, n - 1), leaf l is node (n - 1) + l; root = internal node 0
struct Tree2 {
    Box6* box;          // [2n - 1]
    int32_t* left;      // [n - 1]
    int32_t* right;     // [n - 1]
    int32_t* parent;    // [2n - 1]
    uint32_t* first;    // [2n - 1] first sorted position of the subtree
    uint32_t* count;    // [2n - 1] primitives of the subtree
    uint32_t* visits;   // [n - 1] arrival counter of the bottom-up pass
    uint32_t n;
};

// length of the common prefix of the keys at sorted positions i and j (-1 outside the array); equal keys are told
// apart by their positions (Karras 2012, section 4)
__device__ __forceinline__ int delta(const uint64_t* keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const uint64_t a = keys[i], b = keys[j];
    if (a != b) return __clzll((long long)(a ^ b));
    return 64 + __clz((unsigned)i ^ (unsigned)j);
}

__global__ void k_radix_tree(const uint64_t* keys, Tree2 t) {
    const int n = (int)t.n;
    const int i = (int)(blockIdx.x * blockDim.x + threadIdx.x);
    if (i >= n - 1) return;
    const int d = delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int s = lmax / 2; s >= 1; s /= 2)
        if (delta(keys, n, i, i + (l + s) * d) > dmin) l += s;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, j);
    int s = 0;
    int step = l;
    do { // binary search of the split position
        step = (step + 1) >> 1;
        if (delta(keys, n, i, i + (s + step) * d) > dnode) s += step;
    } while (step > 1);
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const int lc = lo == gamma ? (n - 1) + gamma : gamma, rc = hi == gamma + 1 ? (n - 1) + gamma + 1 : gamma + 1;
    t.left[i] = lc; t.right[i] = rc;
    t.parent[lc] = i; t.parent[rc] = i;
    t.first[i] = (uint32_t)lo; t.count[i] = (uint32_t)(hi - lo + 1);
    if (i == 0) t.parent[0] = -1;
}

// SAH-optimal collapse tables (bvh.cpp: WideBuilder::Dp): c[i - 1] = C(n, i), split[j - 2] = k of C_dist(n, j)
struct Dp { float c[7]; uint8_t split[7]; uint8_t leaf; };

__device__ __forceinline__ float box_area(const Box6& b) {
    const float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
    return 2.f * (dx * dy + dy * dz + dz * dx);
}

__device__ __forceinline__ void dp_leaf(Dp& d, float area, float c_prim) {
    const float c = area * c_prim;
#pragma unroll
    for (int i = 0; i < 7; i++) { d.c[i] = c; d.split[i] = 0; }
    d.leaf = 1;
}

__device__ __forceinline__ void dp_inner(Dp& d, const Dp& L, const Dp& R, float area, uint32_t count, float c_prim, float c_node) {
    float dist[9];
    for (int j = 2; j <= 8; j++) {
        float best = FLT_MAX; int bk = 1;
        for (int k = 1; k < j; k++) {
            if (k > 7 || j - k > 7) continue;
            const float v = L.c[k - 1] + R.c[j - k - 1];
            if (v < best) { best = v; bk = k; }
        }
        dist[j] = best; d.split[j - 2] = (uint8_t)bk;
    }
    const float c_leaf = count <= (uint32_t)MAX_LEAF ? area * (float)count * c_prim : FLT_MAX;
    const float c_int = dist[8] + area * c_node;
    d.leaf = c_leaf <= c_int ? 1 : 0;
    d.c[0] = fminf(c_leaf, c_int);
    for (int i = 2; i <= 7; i++) d.c[i - 1] = fminf(dist[i], d.c[i - 2]);
}

template <typename T> __device__ __forceinline__ void load_cg(const T* p, T& out) { // ld.global.cg, word by word
    static_assert(sizeof(T) % 4 == 0, "word-sized records");
    const uint32_t* q = reinterpret_cast<const uint32_t*>(p);
    uint32_t* o = reinterpret_cast<uint32_t*>(&out);
#pragma unroll
    for (size_t i = 0; i < sizeof(T) / 4; i++) o[i] = __ldcg(q + i);
}

// one thread per leaf: its box and table, then upward; at every internal node the first thread to arrive stops, the
// second (whose sibling's results are then complete and visible) computes the node
__global__ void k_refit_dp(const PrimBox* boxes, const uint32_t* sorted_idx, Tree2 t, Dp* dp, float c_prim, float c_node) {
    const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= t.n) return;
    const uint32_t leaf = (t.n - 1) + l;
    const PrimBox pb = boxes[sorted_idx[l]];
    Box6 b;
    for (int a = 0; a < 3; a++) { b.lo[a] = pb.lo[a]; b.hi[a] = pb.hi[a]; }
    t.box[leaf] = b; t.first[leaf] = l; t.count[leaf] = 1;
    dp_leaf(dp[leaf], box_area(b), c_prim);
    int node = t.parent[leaf];
    while (node >= 0) {
        __threadfence();
        if (atomicAdd(&t.visits[node], 1u) == 0u) return;
        __threadfence();
        const int lc = t.left[node], rc = t.right[node];
        // the children's results were written by other threads, possibly into cache lines this SM already holds for their
        // neighbours: read them past L1
        Box6 bl, br;
        Dp dl, dr;
        load_cg(&t.box[lc], bl); load_cg(&t.box[rc], br);
        load_cg(&dp[lc], dl); load_cg(&dp[rc], dr);
        Box6 nb;
        for (int a = 0; a < 3; a++) { nb.lo[a] = fminf(bl.lo[a], br.lo[a]); nb.hi[a] = fmaxf(bl.hi[a], br.hi[a]); }
        t.box[node] = nb;
        Dp d;
        dp_inner(d, dl, dr, box_area(nb), t.count[node], c_prim, c_node);
        dp[node] = d;
        node = t.parent[node];
    }
}

} // namespace gpubuild
} // namespace rt

#include "build_gpu_collapse.cuh"
