// flatten.h — RtgpuSceneDesc (object tree) -> flat primitive tables + wide BVH (host side).
#pragma once
#include <memory>
#include <string>
#include <utility>
#include <vector>

#include "bvh.h"
#include "rtgpu.h"
#include "scene_types.h"

namespace rt {

// std::vector whose resize() leaves trivially-constructible elements UNINITIALISED: the per-primitive arrays (48 + 96 bytes
// x 1.0 M primitives, twice: input order and leaf order) are written completely by worker threads right after they are
// sized, and value-initialising them first was a single-threaded 310 MB memset + first-touch page fault pass (~60 ms of
// the "Ready" phase of the dragon scene).
template <class T> struct DefaultInitAllocator : std::allocator<T> {
    template <class U> struct rebind { using other = DefaultInitAllocator<U>; };
    template <class U, class... A> void construct(U* p, A&&... a) {
        if constexpr (sizeof...(A) == 0) ::new ((void*)p) U; else ::new ((void*)p) U(std::forward<A>(a)...);
    }
};
template <class T> using PodVector = std::vector<T, DefaultInitAllocator<T>>;

struct HostImage { int width, height, format; std::vector<uint8_t> texels; }; // 4 channels per texel

struct HostMesh { // f64 copies for the exact path (reference layout)
    std::vector<double> positions, normals, uvs;
    std::vector<int32_t> tri_pos, tri_nrm, tri_uv;
};

struct HostScene {
    WideBvh bvh;
    PodVector<Prim> prims;          // leaf order
    PodVector<ShadeTri> shade_tris; // leaf order (analytic entries zeroed)
    std::vector<APrim> aprims;
    std::vector<Xform> xforms;
    std::vector<int32_t> chain;
    std::vector<GateBox> gates;
    std::vector<Instance> instances;
    std::vector<HostMesh> meshes;
    std::vector<Volume> volumes;
    std::vector<APrim> vol_aprims;
    std::vector<int32_t> vol_chain; // pairs (first, len) per vol_aprim
    std::vector<Material> materials;
    std::vector<TexProgram> tex_programs;
    std::vector<TexOp> tex_ops;
    std::vector<HostImage> images;
    std::vector<DevNoise> noises;
    std::vector<Light> lights;
    std::vector<APrim> light_aprims;
    std::vector<int32_t> infinite;
    std::vector<BroadcastPrim> bcast;
    uint64_t n_triangles = 0;
    uint32_t class_mask = 0;          // material classes (shade queues) that can receive paths
    int32_t n_global = 0, global_first = 0; // broadcast list (see DeviceScene)
    float bvh_lo[3] = {0, 0, 0}, bvh_hi[3] = {0, 0, 0};
    bool trace_cold = false;          // non-similarity spheres / gated quads / media present (k_trace_fast<COLD>)
    double ms_build = 0;
};

// returns 0 or a negative RtgpuStatus with `err` set
int flatten_scene(const RtgpuSceneDesc& d, HostScene& out, std::string& err);

} // namespace rt
