// flatten.h — RtgpuSceneDesc (object tree) -> flat primitive tables + wide BVH (host side).
#pragma once
#include <string>
#include <vector>

#include "bvh.h"
#include "rtgpu.h"
#include "scene_types.h"

namespace rt {

struct HostImage { int width, height, format; std::vector<uint8_t> texels; }; // 4 channels per texel

struct HostMesh { // f64 copies for the exact path (reference layout)
    std::vector<double> positions, normals, uvs;
    std::vector<int32_t> tri_pos, tri_nrm, tri_uv;
};

struct HostScene {
    WideBvh bvh;
    std::vector<Prim> prims;          // leaf order
    std::vector<ShadeTri> shade_tris; // leaf order (analytic entries zeroed)
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
