// hostsim.cpp — TEST INFRASTRUCTURE. Compiles the device headers (traverse.cuh, shade.cuh)
// for the HOST so that the traversal / shading logic can be unit-tested on a machine without
// a GPU. It is never linked into librtgpu.so and is not a fallback: the product's entry
// points only launch CUDA kernels.
#include <cstring>
#include <string>
#include <vector>

#include "batch_plan.h"
#include "flatten.h"
#include "traverse.cuh"
#include "hooks.cuh"

using namespace rt;

static std::string g_err;

struct SimScene {
    HostScene hs;
    std::vector<ExactMesh> meshes;
    std::vector<DevImage> images;
    DeviceScene ds;
};

static void assemble(SimScene& s) {
    HostScene& h = s.hs;
    std::memset(&s.ds, 0, sizeof s.ds);
    for (auto& m : h.meshes) {
        ExactMesh e{};
        e.positions = m.positions.data(); e.normals = m.normals.data(); e.uvs = m.uvs.empty() ? nullptr : m.uvs.data();
        e.tri_pos = m.tri_pos.data(); e.tri_nrm = m.tri_nrm.data(); e.tri_uv = m.tri_uv.empty() ? nullptr : m.tri_uv.data();
        e.n_triangles = (int)(m.tri_pos.size() / 3);
        s.meshes.push_back(e);
    }
    for (auto& im : h.images) { DevImage d{}; d.data = im.texels.data(); d.width = im.width; d.height = im.height; d.format = im.format; s.images.push_back(d); }
    DeviceScene& d = s.ds;
    d.nodes = h.bvh.nodes.data(); d.prims = h.prims.data(); d.shade_tris = h.shade_tris.data();
    d.aprims = h.aprims.data(); d.xforms = h.xforms.data(); d.chain = h.chain.data(); d.gates = h.gates.data();
    d.instances = h.instances.data(); d.meshes = s.meshes.data(); d.volumes = h.volumes.data();
    d.vol_aprims = h.vol_aprims.data(); d.vol_chain_first = h.vol_chain.data();
    d.materials = h.materials.data(); d.tex_programs = h.tex_programs.data(); d.tex_ops = h.tex_ops.data();
    d.images = s.images.data(); d.noises = h.noises.data(); d.lights = h.lights.data(); d.light_aprims = h.light_aprims.data();
    d.infinite = h.infinite.data(); d.bcast = h.bcast.data();
    d.n_nodes = (int)h.bvh.nodes.size(); d.n_prims = (int)h.prims.size(); d.n_aprims = (int)h.aprims.size();
    d.n_instances = (int)h.instances.size(); d.n_volumes = (int)h.volumes.size(); d.n_lights = (int)h.lights.size();
    d.n_infinite = (int)h.infinite.size(); d.n_materials = (int)h.materials.size();
    d.has_bvh = h.bvh.nodes.empty() ? 0 : 1;
    d.n_global = h.n_global; d.global_first = h.global_first;
    for (int a = 0; a < 3; a++) { d.bvh_lo[a] = h.bvh_lo[a]; d.bvh_hi[a] = h.bvh_hi[a]; }
}

extern "C" {

const char* hostsim_last_error() { return g_err.c_str(); }

int hostsim_scene_create(const RtgpuSceneDesc* desc, SimScene** out) {
    auto* s = new SimScene();
    int rc = flatten_scene(*desc, s->hs, g_err);
    if (rc) { delete s; return rc; }
    assemble(*s);
    *out = s;
    return 0;
}
int hostsim_scene_destroy(SimScene* s) { delete s; return 0; }

int hostsim_bvh_info(SimScene* s, int64_t out[4]) {
    out[0] = (int64_t)s->hs.bvh.nodes.size(); out[1] = (int64_t)s->hs.prims.size(); out[2] = s->hs.bvh.max_depth; out[3] = (int64_t)s->hs.ms_build;
    return 0;
}

// node fill statistics: out[k] (k = 0..8) nodes with k occupied slots; out[9 + j] (j = 1..3) leaf slots with j
// primitives; out[13] internal child slots; out[14] summed SAH-style area of ... (unused)
int hostsim_bvh_histogram(SimScene* s, int64_t out[16]) {
    for (int i = 0; i < 16; i++) out[i] = 0;
    for (const BvhNode& n : s->hs.bvh.nodes) {
        int used = 0;
        for (int k = 0; k < 8; k++) {
            const uint8_t m = n.meta[k];
            if (!m) continue;
            used++;
            if ((m & 0x18u) == 0x18u && (m >> 5) == 1u) out[13]++;
            else { const int c = (m >> 5) == 1 ? 1 : ((m >> 5) == 3 ? 2 : 3); out[9 + c]++; }
        }
        out[used]++;
    }
    return 0;
}

// FNV-1a over the node array and the leaf order: two builds are the same tree iff the hashes agree
uint64_t hostsim_bvh_hash(SimScene* s) {
    uint64_t h = 1469598103934665603ull;
    auto mix = [&](const void* p, size_t n) { const unsigned char* b = (const unsigned char*)p; for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 1099511628211ull; } };
    mix(s->hs.bvh.nodes.data(), s->hs.bvh.nodes.size() * sizeof(BvhNode));
    mix(s->hs.bvh.prim_order.data(), s->hs.bvh.prim_order.size() * sizeof(uint32_t));
    mix(s->hs.prims.data(), s->hs.prims.size() * sizeof(Prim));
    return h;
}

// closest hit ids + t (instance, primitive, t) for n rays
int hostsim_closest(SimScene* s, const double* rays, int64_t n, int exact, int32_t* out_inst, int32_t* out_prim, double* out_t) {
    for (int64_t i = 0; i < n; i++) {
        RayD r; r.o = x3(rays[i * 6], rays[i * 6 + 1], rays[i * 6 + 2]); r.d = x3(rays[i * 6 + 3], rays[i * 6 + 4], rays[i * 6 + 5]);
        Hit h; U2 stack[RT_STACK_SIZE];
        bool ok = exact ? closest_hit<true>(s->ds, r, 0.001, INFINITY, h, stack) : closest_hit<false>(s->ds, r, 0.001, INFINITY, h, stack);
        if (!ok) { out_inst[i] = -1; out_prim[i] = 0; out_t[i] = INFINITY; continue; }
        const Prim& p = s->ds.prims[h.prim];
        out_inst[i] = (int32_t)f2u(p.a.w);
        out_prim[i] = (f2u(p.c.w) & PRIM_ANALYTIC) ? 0 : (int32_t)f2u(p.b.w);
        out_t[i] = h.t;
    }
    return 0;
}

// bounce_hook_one (hooks.cuh) — the function k_bounce_hook runs per ray — on the host
int hostsim_bounce(SimScene* s, double light_bias, const double* rays, const double* xi, int n_xi, int64_t n, BounceRecord* out) {
    for (int64_t i = 0; i < n; i++) bounce_hook_one(s->ds, (float)light_bias, rays + i * 6, xi + i * n_xi, n_xi, out[i]);
    return 0;
}

// the broadcast list as flatten.cpp made it: per entry the flags (incl. BCAST_EMIT_CONST / BCAST_INWARD) and the emission in the .w lanes
int hostsim_bcast_info(SimScene* s, int index, uint32_t* flags, float* emission3) {
    if (index < 0 || index >= (int)s->hs.bcast.size()) return -1;
    const BroadcastPrim& b = s->hs.bcast[(size_t)index];
    std::memcpy(flags, &b.q.w, 4);
    emission3[0] = b.n.w; emission3[1] = b.w1.w; emission3[2] = b.w2.w;
    return (int)s->hs.bcast.size();
}

// plan_batches (batch_plan.h: what api.cu: enqueue_render cuts a render into) with a linear memory model of `bytes_per_path` per lane
int hostsim_plan_batches(uint64_t total_samples, uint64_t npix, uint64_t batch_paths, int lanes, double bytes_per_path, double budget,
                         uint64_t* out3) {
    const BatchPlan p = plan_batches(total_samples, npix, batch_paths, (size_t)lanes, budget, [&](uint64_t paths) { return (double)paths * bytes_per_path; });
    out3[0] = p.samples_per_batch; out3[1] = p.n_batches; out3[2] = (uint64_t)p.n_lanes;
    return p.ok ? 0 : -1;
}

}
