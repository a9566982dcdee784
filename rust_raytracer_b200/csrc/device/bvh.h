// bvh.h — host builder of the 8-wide quantised BVH the traversal kernel walks.
// Replaces, on the GPU path, the reference's three acceleration structures: the per-mesh
// octree (src/object/mesh/octree.rs), the object BVH (src/object/bvh.rs) and the linear
// ObjectList scan (src/object/list.rs) — all primitives of the flattened world go into
// ONE hierarchy.
#pragma once
#include <cstdint>
#include <vector>

#include "scene_types.h"

namespace rt {

struct PrimBox { float lo[3], hi[3]; };

struct WideBvh {
    std::vector<BvhNode> nodes;
    std::vector<uint32_t> prim_order; // leaf order -> input primitive index
    int max_depth = 0;
    double sah_cost = 0;
};

// Binned-SAH binary build, greedy collapse to 8-wide, octant-ordered child slots,
// conservative 8-bit quantisation of child boxes. `boxes` must already be padded.
void build_wide_bvh(const std::vector<PrimBox>& boxes, WideBvh& out);

// The builder flatten_scene uses: build_wide_bvh unless the CUDA side installs the GPU build (api.cu: RTGPU_BUILD=gpu,
// build_gpu.cuh). Host-only users of flatten.cpp (tests/hostsim) always get the host builder.
typedef void (*WideBvhBuilderFn)(const std::vector<PrimBox>& boxes, WideBvh& out);
extern WideBvhBuilderFn g_wide_bvh_builder;

// exact dequantisation used by builder checks, host simulation and the kernels alike
RT_HD float bvh_exp2(uint8_t e) {
    union { uint32_t u; float f; } c;
    c.u = (uint32_t)e << 23;
    return c.f;
}

} // namespace rt
