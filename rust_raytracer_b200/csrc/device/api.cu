// api.cu — the CUDA side of the C ABI in include/rtgpu.h: device selection, scene upload,
// the wavefront render loop, the parity hooks. No entry point here has a CPU fallback: without
// a usable CUDA device they fail with RTGPU_ERR_NO_DEVICE.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "flatten.h"
#include "batch_plan.h"
#include "kernels.cuh"
#include "build_gpu.cuh"

namespace rthost { void set_error(const std::string& msg); }

using namespace rt;

namespace {

std::vector<int> g_devices; // selected by rtgpu_init
std::mutex g_mutex, g_build_mutex;

struct CudaError { std::string msg; int code = RTGPU_ERR_CUDA; };
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) throw CudaError{std::string(#call) + ": " + cudaGetErrorString(e_)}; } while (0)

struct DeviceCtx {
    int device = -1;
    int sm_count = 0;
    std::vector<void*> allocs;
    DeviceScene ds{};
    DeviceScene* d_ds = nullptr; // device-resident copy of the descriptor (cold paths of k_trace_fast)
    // Batches of one render are spread round-robin over a few "lanes" (stream + path pool): the
    // small late bounces of one batch (each bounded below by the latency of its longest ray) overlap
    // with the large early bounces of the others instead of leaving the SMs idle.
    struct Lane {
        cudaStream_t stream = nullptr;
        cudaEvent_t done = nullptr;
        PathPool pool{};
        PathPool host_copy{};      // what d_pool holds
        PathPool* d_pool = nullptr;
        // per-batch arrays (rays[2], class queues, counters), the hit records, the tail-merging carry arrays:
        // three allocation groups, each freed and re-made when it has to grow
        std::vector<void*> base_allocs, carry_allocs;
        void* hits_alloc = nullptr;
        uint32_t capacity = 0, waste_max = 0, class_mask = 0;
        int bounces = 0;
        uint64_t hits_slots = 0, carry_t_cap = 0, carry_b_cap = 0, carry_q_cap = 0;
        size_t base_bytes = 0, hits_bytes = 0, carry_bytes = 0;
        size_t bytes() const { return base_bytes + hits_bytes + carry_bytes; }
    };
    std::vector<Lane> lanes;
    bool has_normal_maps = false;     // any material of the scene has a normal map (launch_shade)
    bool has_spheres = true;          // analytic spheres among the primitives or lights (else: shade kernels without their code, SPH = false)
    uint32_t const_tex_classes = 0;   // bit c: every texture a path of material class c can evaluate is one constant -> k_shade<c> without the texture interpreter (CTEX)
    uint32_t* d_pool_error = nullptr; // PathPool::error of every lane (kernels.cuh: pool_fail)
    // rays per camera path and bounce measured by the pilot sample of a render (enqueue_render): what the
    // tail-merging arrays are sized from. Cached per (camera, depth, light bias).
    struct Survival { uint64_t key = 0; bool valid = false; std::vector<double> all, traced; };
    Survival survival;
    bool no_merge = false; // set while rtgpu_render repeats a render whose tail-merging arrays overflowed
    char* geom_arena = nullptr; // nodes | prims
    size_t geom_nodes_bytes = 0, geom_bytes = 0, l2_persist_max = 0, l2_window_max = 0;
    cudaEvent_t fork = nullptr;
    cudaStream_t stream = nullptr;
    float4* accum = nullptr; // owned accumulation buffer (rtgpu_render)
    size_t accum_pixels = 0;
    // multi-GPU renders: the accumulator comes from a memory pool of this device that the reducing device may read
    // (cudaMemPoolSetAccess) — peer access per allocation, not cudaDeviceEnablePeerAccess, which maps EVERY allocation
    // of the peer (27 GB of path pools per GPU: ~25 ms per peer, measured) into the reducing device's address space
    cudaMemPool_t shared_pool = nullptr;
    int shared_reader = -1;   // device the pool is readable from
    bool accum_shared = false, accum_from_pool = false;
    unsigned long long* d_rays_per_bounce = nullptr;
    uint64_t launches = 0, trace_launches = 0;
    // RTGPU_FLAG_PROFILE: event pairs per kernel family (0 raygen, 1 trace, 2 shade)
    std::vector<cudaEvent_t> ev_pool;
    size_t ev_used = 0;
    std::vector<std::pair<size_t, int>> ev_spans; // (index of start event, family); stop = start + 1
    double ms_family[3] = {0, 0, 0};

    size_t span_begin(int family, cudaStream_t on) {
        if (ev_used + 2 > ev_pool.size()) {
            for (int i = 0; i < 2; i++) { cudaEvent_t e; CK(cudaEventCreate(&e)); ev_pool.push_back(e); }
        }
        const size_t at = ev_used;
        ev_used += 2;
        CK(cudaEventRecord(ev_pool[at], on));
        ev_spans.push_back({at, family});
        return at;
    }
    void span_end(size_t at, cudaStream_t on) { CK(cudaEventRecord(ev_pool[at + 1], on)); }
    void collect_spans() { // call after the stream has been synchronised
        for (auto& sp : ev_spans) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, ev_pool[sp.first], ev_pool[sp.first + 1]) == cudaSuccess) ms_family[sp.second] += ms;
        }
        ev_spans.clear(); ev_used = 0;
    }

    template <typename T, typename A> T* upload(const std::vector<T, A>& v) {
        if (v.empty()) return nullptr;
        void* p = nullptr;
        CK(cudaMalloc(&p, v.size() * sizeof(T)));
        allocs.push_back(p);
        CK(cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
        return (T*)p;
    }
    template <typename T> T* alloc(size_t n) {
        void* p = nullptr;
        CK(cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T)));
        allocs.push_back(p);
        return (T*)p;
    }
    void release() {
        if (device < 0) return;
        cudaSetDevice(device);
        for (void* p : allocs) cudaFree(p);
        allocs.clear();
        for (cudaEvent_t e : ev_pool) cudaEventDestroy(e);
        ev_pool.clear();
        if (accum_from_pool && accum) cudaFree(accum);
        accum = nullptr; accum_pixels = 0; accum_from_pool = accum_shared = false;
        if (shared_pool) cudaMemPoolDestroy(shared_pool);
        shared_pool = nullptr;
        for (Lane& l : lanes) {
            if (l.stream) cudaStreamDestroy(l.stream);
            if (l.done) cudaEventDestroy(l.done);
            for (void* p : l.base_allocs) cudaFree(p);
            for (void* p : l.carry_allocs) cudaFree(p);
            if (l.hits_alloc) cudaFree(l.hits_alloc);
        }
        lanes.clear();
        if (fork) cudaEventDestroy(fork);
        fork = nullptr;
        if (stream) cudaStreamDestroy(stream);
        stream = nullptr;
    }
};

#include "build_gpu_host.cuh"

} // namespace

struct RtgpuScene {
    double stats_gpu_build_ms = 0;
    HostScene hs;
    std::vector<DeviceCtx> devs;
    RtgpuStats stats{};
};

namespace {

int env_int(const char* name, int dflt);

void upload_scene(const HostScene& hs, DeviceCtx& c) {
    CK(cudaSetDevice(c.device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, c.device));
    c.sm_count = prop.multiProcessorCount;
    CK(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
    DeviceScene& d = c.ds;
    std::memset(&d, 0, sizeof d);
    // nodes and leaf primitives share one arena so that ONE L2 access-policy window can cover the
    // whole traversal working set (set_l2_window)
    {
        const size_t nb = (hs.bvh.nodes.size() * sizeof(BvhNode) + 255) & ~(size_t)255, pb = hs.prims.size() * sizeof(Prim);
        char* arena = c.alloc<char>(nb + pb + 256);
        if (!hs.bvh.nodes.empty()) CK(cudaMemcpy(arena, hs.bvh.nodes.data(), hs.bvh.nodes.size() * sizeof(BvhNode), cudaMemcpyHostToDevice));
        if (!hs.prims.empty()) CK(cudaMemcpy(arena + nb, hs.prims.data(), pb, cudaMemcpyHostToDevice));
        d.nodes = hs.bvh.nodes.empty() ? nullptr : (const BvhNode*)arena;
        d.prims = hs.prims.empty() ? nullptr : (const Prim*)(arena + nb);
        c.geom_arena = arena; c.geom_nodes_bytes = nb; c.geom_bytes = nb + pb;
        c.l2_persist_max = (size_t)prop.persistingL2CacheMaxSize; c.l2_window_max = (size_t)prop.accessPolicyMaxWindowSize;
    }
    d.shade_tris = c.upload(hs.shade_tris);
    d.aprims = c.upload(hs.aprims);
    d.xforms = c.upload(hs.xforms);
    d.chain = c.upload(hs.chain);
    d.gates = c.upload(hs.gates);
    d.instances = c.upload(hs.instances);
    std::vector<ExactMesh> em;
    for (const HostMesh& m : hs.meshes) {
        ExactMesh e{};
        e.positions = c.upload(m.positions); e.normals = c.upload(m.normals); e.uvs = c.upload(m.uvs);
        e.tri_pos = c.upload(m.tri_pos); e.tri_nrm = c.upload(m.tri_nrm); e.tri_uv = c.upload(m.tri_uv);
        e.n_triangles = (int32_t)(m.tri_pos.size() / 3);
        em.push_back(e);
    }
    d.meshes = c.upload(em);
    d.volumes = c.upload(hs.volumes);
    d.vol_aprims = c.upload(hs.vol_aprims);
    d.vol_chain_first = c.upload(hs.vol_chain);
    d.materials = c.upload(hs.materials);
    for (const Material& m : hs.materials) if (m.normal_map >= 0) c.has_normal_maps = true;
    c.has_spheres = env_int("RTGPU_CONST_TEX_KERNELS", 1) == 0;
    for (const APrim& a : hs.aprims) if (a.kind == APRIM_SPHERE) c.has_spheres = true;
    for (const APrim& a : hs.light_aprims) if (a.kind == APRIM_SPHERE) c.has_spheres = true;
    for (const Light& l : hs.lights) if (l.kind == LIGHT_SPHERE) c.has_spheres = true;
    c.const_tex_classes = env_int("RTGPU_CONST_TEX_KERNELS", 1) != 0 ? (1u << N_CLASSES) - 1u : 0u;
    {
        auto is_const = [&](int tex) {
            if (tex < 0) return true;
            const TexProgram& tp = hs.tex_programs[(size_t)tex];
            return tp.count == 1 && hs.tex_ops[(size_t)tp.first].op == TOP_CONST;
        };
        for (const Material& m : hs.materials)
            if (!is_const(m.albedo) || !is_const(m.roughness) || !is_const(m.normal_map)) c.const_tex_classes &= ~(1u << material_class_of_kind(m.kind));
        for (const Instance& in : hs.instances) // sky / sun: a private Emissive over `texture`, shaded by the emissive kernel
            if ((in.kind == INST_SKY || in.kind == INST_SUN) && !is_const(in.texture)) c.const_tex_classes &= ~(1u << CLASS_EMISSIVE);
    }
    d.tex_programs = c.upload(hs.tex_programs);
    d.tex_ops = c.upload(hs.tex_ops);
    std::vector<DevImage> imgs;
    for (const HostImage& im : hs.images) {
        DevImage di{};
        di.data = c.upload(im.texels); di.width = im.width; di.height = im.height; di.format = im.format;
        imgs.push_back(di);
    }
    d.images = c.upload(imgs);
    d.noises = c.upload(hs.noises);
    d.lights = c.upload(hs.lights);
    d.light_aprims = c.upload(hs.light_aprims);
    d.infinite = c.upload(hs.infinite);
    d.bcast = c.upload(hs.bcast);
    d.n_nodes = (int32_t)hs.bvh.nodes.size(); d.n_prims = (int32_t)hs.prims.size(); d.n_aprims = (int32_t)hs.aprims.size();
    d.n_instances = (int32_t)hs.instances.size(); d.n_volumes = (int32_t)hs.volumes.size(); d.n_lights = (int32_t)hs.lights.size();
    d.n_infinite = (int32_t)hs.infinite.size(); d.n_materials = (int32_t)hs.materials.size();
    d.has_bvh = hs.bvh.nodes.empty() ? 0 : 1;
    d.n_global = hs.n_global; d.global_first = hs.global_first;
    for (int a = 0; a < 3; a++) { d.bvh_lo[a] = hs.bvh_lo[a]; d.bvh_hi[a] = hs.bvh_hi[a]; }
    c.d_ds = c.alloc<DeviceScene>(1);
    CK(cudaMemcpy(c.d_ds, &d, sizeof d, cudaMemcpyHostToDevice));
    CK(cudaEventCreateWithFlags(&c.fork, cudaEventDisableTiming));
    c.d_rays_per_bounce = c.alloc<unsigned long long>(128); // [0,64) all rays, [64,128) rays that walked the BVH
    CK(cudaMemset(c.d_rays_per_bounce, 0, 128 * sizeof(unsigned long long)));
    c.d_pool_error = c.alloc<uint32_t>(1);
    CK(cudaMemset(c.d_pool_error, 0, sizeof(uint32_t)));
    // every copy / memset above went through the legacy default stream, which may return before the DMA has landed and
    // does not order with the non-blocking streams the kernels run on
    CK(cudaDeviceSynchronize());
}

int env_int(const char* name, int dflt);

// RTGPU_L2_PERSIST: 0 = off, 1 = BVH nodes, 2 = nodes + leaf primitives marked persisting in L2 for the
// kernels of this stream (the path pools stream through the same L2 and would otherwise evict them)
void set_l2_window(DeviceCtx& c, cudaStream_t stream) {
    const int mode = env_int("RTGPU_L2_PERSIST", 0);
    if (mode <= 0 || !c.geom_arena || c.l2_persist_max == 0) return;
    size_t bytes = mode == 1 ? c.geom_nodes_bytes : c.geom_bytes;
    bytes = std::min(bytes, c.l2_window_max);
    const size_t carve = std::min(c.l2_persist_max, bytes);
    CK(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, carve));
    cudaStreamAttrValue attr{};
    attr.accessPolicyWindow.base_ptr = c.geom_arena;
    attr.accessPolicyWindow.num_bytes = bytes;
    attr.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)carve / (double)bytes);
    attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    CK(cudaStreamSetAttribute(stream, cudaStreamAttributeAccessPolicyWindow, &attr));
}

// ---- path pools ------------------------------------------------------------------------------------
// A lane's memory, per path of batch capacity: 2 x 64 B ray records x (1 + waste), 32 B hit record, 2 x 4 B queue entry per
// material class present. Round 1 sized every array for "all paths alive plus every chunk half empty" (304 + 216 x group
// bytes per path: 94 GB on the headline job); now the chunk waste is bounded from the actual grids (waste_max) and the
// tail-merging arrays from the survival the pilot sample measured, with every append bounds-checked on the device.
constexpr double RT_SURVIVAL_MARGIN = 1.25;

template <typename T> T* lane_alloc(std::vector<void*>& group, size_t n, size_t& bytes) {
    void* p = nullptr;
    const size_t b = std::max<size_t>(n, 1) * sizeof(T);
    CK(cudaMalloc(&p, b));
    group.push_back(p);
    bytes += b;
    return (T*)p;
}

uint32_t ray_slots_for(uint32_t capacity, uint32_t waste_max) { return capacity + std::min(capacity, waste_max) + 64u; }
int popcount_u32(uint32_t v) { int n = 0; for (; v; v &= v - 1) n++; return n; }
size_t base_bytes_for(uint32_t capacity, uint32_t waste_max, uint32_t class_mask, int bounces) {
    return 2 * (size_t)ray_slots_for(capacity, waste_max) * sizeof(PathRay) + (size_t)ray_slots_for(capacity, waste_max) * sizeof(PathHit) +
           2 * (size_t)popcount_u32(class_mask) * capacity * sizeof(uint32_t) + (size_t)(bounces + 2) * RT_COUNTERS_PER_BOUNCE * sizeof(uint32_t);
}

DeviceCtx::Lane& lane_at(DeviceCtx& c, size_t index) {
    while (c.lanes.size() <= index) {
        DeviceCtx::Lane l;
        CK(cudaStreamCreateWithFlags(&l.stream, cudaStreamNonBlocking));
        set_l2_window(c, l.stream);
        CK(cudaEventCreateWithFlags(&l.done, cudaEventDisableTiming));
        l.d_pool = c.alloc<PathPool>(1);
        l.pool.carry_bounce = 0xFFFFFFFFu;
        c.lanes.push_back(l);
    }
    return c.lanes[index];
}

void free_group(std::vector<void*>& group, size_t& bytes) {
    for (void* p : group) CK(cudaFree(p));
    group.clear();
    bytes = 0;
}

void ensure_hits(DeviceCtx::Lane& l, uint64_t slots) {
    if (l.hits_slots >= slots) return;
    CK(cudaStreamSynchronize(l.stream));
    if (l.hits_alloc) CK(cudaFree(l.hits_alloc));
    l.hits_alloc = nullptr; l.hits_bytes = 0; l.hits_slots = 0;
    CK(cudaMalloc(&l.hits_alloc, slots * sizeof(PathHit)));
    l.hits_bytes = slots * sizeof(PathHit); l.hits_slots = slots;
    l.pool.hits = (PathHit*)l.hits_alloc; // indexed like the trace region of the bounce's ray array
}

// the per-batch arrays of lane `index` for batches of `capacity` paths
void ensure_lane_base(DeviceCtx& c, size_t index, uint32_t capacity, int bounces, uint32_t waste_max, uint32_t class_mask) {
    DeviceCtx::Lane& l = lane_at(c, index);
    PathPool& p = l.pool;
    const bool fits = l.capacity >= capacity && l.bounces >= bounces && l.waste_max >= waste_max && (class_mask & ~l.class_mask) == 0;
    if (!fits) {
        CK(cudaStreamSynchronize(l.stream)); // no kernel of this lane may still use the old arrays
        capacity = std::max(capacity, l.capacity); bounces = std::max(bounces, l.bounces);
        waste_max = std::max(waste_max, l.waste_max); class_mask |= l.class_mask;
        free_group(l.base_allocs, l.base_bytes); // superseded arrays are released, not kept until the scene dies
        p.capacity = capacity;
        p.waste_max = waste_max;
        p.ray_slots = ray_slots_for(capacity, waste_max);
        p.rays[0] = lane_alloc<PathRay>(l.base_allocs, p.ray_slots, l.base_bytes);
        p.rays[1] = lane_alloc<PathRay>(l.base_allocs, p.ray_slots, l.base_bytes);
        p.queue_cap = capacity;
        for (int h = 0; h < 2; h++)
            for (int k = 0; k < N_CLASSES; k++)
                p.class_queue[h][k] = (class_mask >> k) & 1u ? lane_alloc<uint32_t>(l.base_allocs, capacity, l.base_bytes) : nullptr;
        p.counters = lane_alloc<uint32_t>(l.base_allocs, (size_t)(bounces + 2) * RT_COUNTERS_PER_BOUNCE, l.base_bytes);
        p.error = c.d_pool_error;
        l.capacity = capacity; l.bounces = bounces; l.waste_max = waste_max; l.class_mask = class_mask;
    }
    ensure_hits(l, std::max<uint64_t>(p.ray_slots, l.carry_t_cap));
}

// the tail-merging arrays of lane `index`: trace region, bypass region (records) and class-queue entries
void ensure_lane_carry(DeviceCtx& c, size_t index, uint64_t t_cap, uint64_t b_cap, uint64_t q_cap) {
    DeviceCtx::Lane& l = lane_at(c, index);
    PathPool& p = l.pool;
    if (l.carry_t_cap < t_cap || l.carry_b_cap < b_cap || l.carry_q_cap < q_cap) {
        CK(cudaStreamSynchronize(l.stream));
        t_cap = std::max(t_cap, l.carry_t_cap); b_cap = std::max(b_cap, l.carry_b_cap); q_cap = std::max(q_cap, l.carry_q_cap);
        if (t_cap + b_cap > 0xFFFFFF00ull || q_cap > 0xFFFFFF00ull) throw CudaError{"tail-merging group too large for 32-bit record indices"};
        free_group(l.carry_allocs, l.carry_bytes);
        p.carry_rays = lane_alloc<PathRay>(l.carry_allocs, t_cap + b_cap, l.carry_bytes);
        for (int k = 0; k < N_CLASSES; k++)
            p.carry_queue[k] = (l.class_mask >> k) & 1u ? lane_alloc<uint32_t>(l.carry_allocs, q_cap, l.carry_bytes) : nullptr;
        l.carry_t_cap = t_cap; l.carry_b_cap = b_cap; l.carry_q_cap = q_cap;
        p.carry_t_cap = (uint32_t)t_cap; p.carry_b_cap = (uint32_t)b_cap; p.carry_slots = (uint32_t)(t_cap + b_cap); p.carry_queue_cap = (uint32_t)q_cap;
    }
    ensure_hits(l, std::max<uint64_t>(p.ray_slots, l.carry_t_cap));
}

// the descriptor the cold paths of k_trace_fast read through a pointer follows the host's
void publish_pool(DeviceCtx::Lane& l, uint32_t carry_bounce) {
    l.pool.carry_bounce = carry_bounce;
    if (std::memcmp(&l.pool, &l.host_copy, sizeof(PathPool)) == 0) return;
    CK(cudaStreamSynchronize(l.stream));
    CK(cudaMemcpyAsync(l.d_pool, &l.pool, sizeof(PathPool), cudaMemcpyHostToDevice, l.stream)); // ordered with the lane's kernels
    CK(cudaStreamSynchronize(l.stream));                                                          // (&l.pool may change after this call)
    l.host_copy = l.pool;
}

// hooks and small renders: one lane, no tail merging
void ensure_lane(DeviceCtx& c, size_t index, uint32_t capacity, int bounces) {
    ensure_lane_base(c, index, capacity, bounces, capacity, (1u << N_CLASSES) - 1u);
    publish_pool(c.lanes[index], 0xFFFFFFFFu);
}

// a failed device-side pool check (kernels.cuh: pool_guard, routed_append) surfaces here; the flag is cleared
void check_pool_error(DeviceCtx& c) {
    uint32_t e = 0;
    CK(cudaMemcpy(&e, c.d_pool_error, sizeof e, cudaMemcpyDeviceToHost));
    if (e == 0) return;
    CK(cudaMemset(c.d_pool_error, 0, sizeof e));
    CK(cudaDeviceSynchronize()); // the retry's kernels run on non-blocking streams and must see the cleared flag
    c.survival.valid = false;
    throw CudaError{"path pool overflow (code " + std::to_string(e) + "): more paths survived than the pilot sample predicted; "
                    "set RTGPU_TAIL_GROUP=1 to render without tail merging"};
}

CameraDev make_camera(const RtgpuCameraDesc& cam, int sqrt_spt) {
    CameraDev c{};
    c.width = cam.width; c.height = cam.height;
    for (int i = 0; i < 3; i++) {
        c.position[i] = cam.position[i]; c.first_pixel[i] = cam.first_pixel[i];
        c.du[i] = cam.pixel_delta_u[i]; c.dv[i] = cam.pixel_delta_v[i];
        c.bu[i] = cam.basis_u[i]; c.bv[i] = cam.basis_v[i];
        c.background[i] = (float)cam.background[i];
    }
    c.aperture_radius = cam.aperture_radius;
    c.sqrt_spt = sqrt_spt; c.inv_sqrt_spt = 1.0 / (double)sqrt_spt;
    return c;
}

// tuning knobs (tools/tune_trace.py); the defaults are the measured best on cornell_dragon
int env_int(const char* name, int dflt) {
    const char* e = std::getenv(name);
    return e ? std::atoi(e) : dflt;
}

// resident shade CTAs per SM the register allocation must allow: 7 (72 registers) — at 8 (64 registers) the kernels
// spill ~110 bytes per thread into an L1 they already overflow (12 KB less L1 per CTA costs 6 %), at 6 (80 registers)
// too few warps are left to hide the gathers (profiles/r01_tune17_shade_registers.log)
#ifndef RT_SHADE_MIN_CTAS
#define RT_SHADE_MIN_CTAS 6
#endif
template <int CLASS> void launch_shade(DeviceCtx& c, DeviceCtx::Lane& l, uint32_t class_mask, int grid, const RenderConst& rc, const CameraDev& cam,
                                       uint32_t bounce, float4* accum, uint32_t bypass) {
    if (!(class_mask & (1u << CLASS))) return; // no material of this class in the scene: its queue stays empty
    // (only Glossy / NormalDebug can carry a normal map: the other classes have one instantiation)
    if constexpr (CLASS == CLASS_GLOSSY || CLASS == CLASS_EMISSIVE) {
        if (c.has_normal_maps) {
            k_shade<CLASS, RT_SHADE_MIN_CTAS, true><<<grid, 128, 0, l.stream>>>(c.ds, l.pool, rc, cam, bounce, accum, bypass);
            c.launches++;
            return;
        }
    }
    const bool ctex = (c.const_tex_classes >> CLASS) & 1u;
    if (ctex && !c.has_spheres) k_shade<CLASS, RT_SHADE_MIN_CTAS, false, true, false><<<grid, 128, 0, l.stream>>>(c.ds, l.pool, rc, cam, bounce, accum, bypass);
    else if (ctex) k_shade<CLASS, RT_SHADE_MIN_CTAS, false, true><<<grid, 128, 0, l.stream>>>(c.ds, l.pool, rc, cam, bounce, accum, bypass);
    else k_shade<CLASS, RT_SHADE_MIN_CTAS, false><<<grid, 128, 0, l.stream>>>(c.ds, l.pool, rc, cam, bounce, accum, bypass);
    c.launches++;
}

// RTGPU_TRACE_VARIANT: 0 = one ray per thread (v1 kernel), otherwise k_trace_fast<FETCH_THRESH, POSTPONE_PCT, MIN_CTAS, COLD>
void launch_trace_fast(DeviceCtx& c, DeviceCtx::Lane& l, bool cold, int variant, int grid, const RenderConst& rc, uint32_t bounce, const DeviceScene* scene_override = nullptr) {
    const DeviceScene& ds = scene_override ? *scene_override : c.ds;
#define RT_LAUNCH_TF(FT, PP, MB, CD, PF, ...) k_trace_fast<FT, PP, MB, CD, PF, ##__VA_ARGS__><<<grid, RT_TRACE_THREADS, 0, l.stream>>>(ds, l.pool, c.d_ds, l.d_pool, rc, bounce)
    if (cold) { // f64 call paths present: 128 registers
        switch (variant) {
        default: RT_LAUNCH_TF(20, 0, 4, true, 0); break;
        case 2: RT_LAUNCH_TF(16, 0, 4, true, 0); break;
        case 20: RT_LAUNCH_TF(20, 0, 4, true, 0, 0, false, 1); break;
        case 32: RT_LAUNCH_TF(22, 0, 4, true, 0, 0, false, 1); break;
        case 8: case 9: case 10: case 11: RT_LAUNCH_TF(20, 0, 4, true, 0, 24); break;
        }
        return;
    }
    switch (variant) {
    default:
    case 4: RT_LAUNCH_TF(22, 0, 8, false, 0, 0, false, 1); break; // one leaf primitive per step, refill below 22 busy lanes (profiles/r02_tune10, r02_tune12)
    case 3: RT_LAUNCH_TF(20, 0, 8, false, 0); break;              // the default until then: whole leaf groups per step
    case 1: RT_LAUNCH_TF(24, 0, 8, false, 0); break;
    case 2: RT_LAUNCH_TF(16, 0, 8, false, 0); break;
    case 5: RT_LAUNCH_TF(24, 25, 8, false, 0); break;
    case 6: RT_LAUNCH_TF(20, 0, 8, false, 1); break;
    case 7: RT_LAUNCH_TF(20, 0, 8, false, 2); break;
    case 14: RT_LAUNCH_TF(20, 0, 8, false, 3); break;     // leaf-record line prefetched when its node arrives
    case 8: RT_LAUNCH_TF(20, 0, 8, false, 0, 1); break;   // cooperative leaf tests, flushed every step
    case 9: RT_LAUNCH_TF(20, 0, 8, false, 0, 12); break;  // deferred: flush at 12 / 24 / 32 queued pairs
    case 10: RT_LAUNCH_TF(20, 0, 8, false, 0, 24); break;
    case 11: RT_LAUNCH_TF(20, 0, 8, false, 0, 32); break;
    case 18: RT_LAUNCH_TF(20, 0, 8, false, 5); break;     // next 32 ray records of the warp's range prefetched to L1
    case 19: RT_LAUNCH_TF(20, 0, 8, false, 6); break;     // ... and the rest of a newly reserved chunk to L2
    case 15: RT_LAUNCH_TF(20, 0, 8, false, 0, 0, true); break; // FFMA2 slab planes
    case 16: RT_LAUNCH_TF(20, 0, 6, false, 0, 0, true); break; // ... under an 80-register cap
    case 17: RT_LAUNCH_TF(20, 0, 6, false, 0, 0, false); break;
    case 20: RT_LAUNCH_TF(20, 0, 8, false, 0, 0, false, 1); break; // at most 1 / 2 leaf primitives per step
    case 21: RT_LAUNCH_TF(20, 0, 8, false, 0, 0, false, 2); break;
    case 22: RT_LAUNCH_TF(16, 0, 8, false, 0, 0, false, 1); break;
    case 23: RT_LAUNCH_TF(24, 0, 8, false, 0, 0, false, 1); break;
    case 28: RT_LAUNCH_TF(20, 0, 8, false, 0, 0, false, 1, 3); break; // ... and a primitive phase only when 3 / 5 / 8 lanes wait for one
    case 29: RT_LAUNCH_TF(20, 0, 8, false, 0, 0, false, 1, 5); break;
    case 30: RT_LAUNCH_TF(20, 0, 8, false, 0, 0, false, 1, 8); break;
    case 31: RT_LAUNCH_TF(18, 0, 8, false, 0, 0, false, 1); break;
    case 32: RT_LAUNCH_TF(22, 0, 8, false, 0, 0, false, 1); break;
    case 33: RT_LAUNCH_TF(22, 0, 8, false, 0, 0, false, 1, 0, 1); break; // + prefetch of a waiting group's next primitive
    case 34: RT_LAUNCH_TF(22, 0, 8, false, 0, 0, false, 1, 0, 2); break; // + top of the stack in registers
    case 35: RT_LAUNCH_TF(22, 0, 8, false, 0, 0, false, 1, 0, 3); break;
    case 36: RT_LAUNCH_TF(22, 0, 6, false, 0, 0, false, 1, 0, 2); break; // ... under an 80-register cap
    case 37: RT_LAUNCH_TF(22, 0, 6, false, 0, 0, false, 1, 0, 0); break;
    case 38: RT_LAUNCH_TF(22, 0, 8, false, 5, 0, false, 1); break; // leaf step + the warp's next ray records prefetched to L1 (18) / L2 (19)
    case 39: RT_LAUNCH_TF(22, 0, 8, false, 6, 0, false, 1); break;
    case 40: RT_LAUNCH_TF(24, 0, 8, false, 5, 0, false, 1); break;
    case 24: RT_LAUNCH_TF(20, 0, 8, false, 0, 0, false, -1); break; // the group found by the previous step tested first, 1 / 2 primitives per step
    case 25: RT_LAUNCH_TF(20, 0, 8, false, 0, 0, false, -2); break;
    case 26: RT_LAUNCH_TF(16, 0, 8, false, 0, 0, false, -1); break;
    case 27: RT_LAUNCH_TF(24, 0, 8, false, 0, 0, false, -1); break;
    case 12: RT_LAUNCH_TF(24, 0, 8, false, 0, 24); break;
    case 13: RT_LAUNCH_TF(16, 0, 8, false, 0, 24); break;
    }
#undef RT_LAUNCH_TF
}

uint64_t fnv1a(const void* p, size_t n, uint64_t h = 1469598103934665603ull) {
    const unsigned char* b = (const unsigned char*)p;
    for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 1099511628211ull; }
    return h;
}

// Everything one render needs to launch the kernels of a batch.
struct RenderCtx {
    RtgpuScene* scene; DeviceCtx* c; CameraDev camd; RenderConst rc; float4* accum;
    uint32_t s_begin, class_mask, bypass;
    int variant, grid_trace, grid_shade, max_depth;
    bool exact, profile;
};

// One batch on lane `l`: camera paths [first, first + n) of the render (sample-major: path g = sample g / npix of pixel
// g % npix), bounces [0, b_end); counter rows [0, b_reset) are cleared first.
void launch_batch(const RenderCtx& r, DeviceCtx::Lane& l, uint64_t first, uint32_t n, int b_reset, int b_end) {
    DeviceCtx& c = *r.c;
    size_t span = 0;
    CK(cudaMemsetAsync(l.pool.counters, 0, (size_t)b_reset * RT_COUNTERS_PER_BOUNCE * sizeof(uint32_t), l.stream));
    if (r.profile) span = c.span_begin(0, l.stream);
    const uint64_t npix = (uint64_t)r.camd.width * (uint64_t)r.camd.height;
    if (first % npix != 0) throw CudaError{"launch_batch: a batch must start at a whole image of samples"};
    // (a pinhole camera gets the kernel without the lens code; RTGPU_RAYGEN_GENERIC=1 = A/B)
    const bool generic = env_int("RTGPU_RAYGEN_GENERIC", 0) != 0;
    const int raygen_ctas = std::max(1, std::min(8, env_int("RTGPU_RAYGEN_CTAS_PER_SM", 4)));
#define RT_LAUNCH_RAYGEN(DOF, SPH) k_raygen<DOF, SPH><<<c.sm_count * raygen_ctas, 256, 0, l.stream>>>(c.ds, l.pool, r.camd, r.rc, (uint32_t)(first / npix), n, r.s_begin, r.bypass, r.accum)
    if (generic || (r.camd.aperture_radius >= 0.0 && c.has_spheres)) RT_LAUNCH_RAYGEN(true, true);
    else if (r.camd.aperture_radius >= 0.0) RT_LAUNCH_RAYGEN(true, false);
    else if (c.has_spheres) RT_LAUNCH_RAYGEN(false, true);
    else RT_LAUNCH_RAYGEN(false, false);
#undef RT_LAUNCH_RAYGEN
    if (r.profile) c.span_end(span, l.stream);
    c.launches++;
    for (int b = 0; b < b_end; b++) {
        if (r.profile) span = c.span_begin(1, l.stream);
        if (r.exact) k_trace<true><<<r.grid_trace, 128, 0, l.stream>>>(c.ds, l.pool, r.rc, (uint32_t)b);
        else if (r.variant == 0) k_trace<false><<<r.grid_trace, 128, 0, l.stream>>>(c.ds, l.pool, r.rc, (uint32_t)b);
        else launch_trace_fast(c, l, r.scene->hs.trace_cold, r.variant, r.grid_trace, r.rc, (uint32_t)b);
        if (r.profile) { c.span_end(span, l.stream); span = c.span_begin(2, l.stream); }
        c.launches++; c.trace_launches++;
        launch_shade<CLASS_LAMBERTIAN>(c, l, r.class_mask, r.grid_shade, r.rc, r.camd, (uint32_t)b, r.accum, r.bypass);
        launch_shade<CLASS_METAL>(c, l, r.class_mask, r.grid_shade, r.rc, r.camd, (uint32_t)b, r.accum, r.bypass);
        launch_shade<CLASS_DIELECTRIC>(c, l, r.class_mask, r.grid_shade, r.rc, r.camd, (uint32_t)b, r.accum, r.bypass);
        launch_shade<CLASS_GLOSSY>(c, l, r.class_mask, r.grid_shade, r.rc, r.camd, (uint32_t)b, r.accum, r.bypass);
        launch_shade<CLASS_ISOTROPIC>(c, l, r.class_mask, r.grid_shade, r.rc, r.camd, (uint32_t)b, r.accum, r.bypass);
        launch_shade<CLASS_EMISSIVE>(c, l, r.class_mask, r.grid_shade, r.rc, r.camd, (uint32_t)b, r.accum, r.bypass);
        if (r.profile) c.span_end(span, l.stream);
    }
    // rows of this batch alone, and (last batch of a group) the merged rows of the whole group
    k_stats<<<1, 64, 0, l.stream>>>(l.pool.counters, 0, b_end, c.d_rays_per_bounce);
    c.launches++;
}

// Enqueue the render of global samples [s_begin, s_end) of every pixel into `accum` (fp32 sums).
// Work is ordered after everything already in the context's stream (or `stream_override`) and
// that stream waits for it: fork to the lanes, join back. The call itself does not wait for the kernels, with one
// exception: the first render of a (camera, depth, light bias) that is large enough for tail merging renders its first
// sample of every pixel alone and reads the per-bounce ray counts back (the pilot: ~1/1000 of the headline job).
void enqueue_render(RtgpuScene* scene, DeviceCtx& c, const RtgpuCameraDesc& cam, const RtgpuRenderParams& prm, int s_begin, int s_end,
                    float4* accum, cudaStream_t stream_override) {
    CK(cudaSetDevice(c.device));
    cudaStream_t main_stream = stream_override ? stream_override : c.stream;
    const uint64_t npix = (uint64_t)cam.width * cam.height;
    if ((uint64_t)(s_end - s_begin) * npix == 0) return;
    if (npix > 0x7FFF0000ull) throw CudaError{"image too large for 32-bit path indices"};
    const bool exact = (prm.flags & RTGPU_FLAG_EXACT) != 0;
    const bool profile = (prm.flags & RTGPU_FLAG_PROFILE) != 0;
    const int variant = env_int("RTGPU_TRACE_VARIANT", 4);
    const int shade_ctas = std::max(1, std::min(16, env_int("RTGPU_SHADE_CTAS_PER_SM", RT_SHADE_MIN_CTAS)));
    const int trace_ctas = variant == 0 ? 8 : std::max(1, std::min(16, env_int("RTGPU_TRACE_CTAS_PER_SM", 4)));
    const uint32_t class_mask = scene->hs.class_mask;
    RenderCtx r{};
    r.scene = scene; r.c = &c; r.camd = make_camera(cam, prm.sqrt_spt); r.rc = RenderConst{prm.seed, (float)prm.light_bias, prm.max_depth, prm.flags};
    r.accum = accum; r.s_begin = (uint32_t)s_begin; r.class_mask = class_mask;
    // rays whose closest hit the broadcast pre-test already knows (BVH bounds out of reach) go straight to
    // the shade queue of the next bounce; constant media would still compete, and the exact mode re-traces in f64
    r.bypass = !exact && scene->hs.volumes.empty() && env_int("RTGPU_BYPASS", 1) != 0 ? 1u : 0u;
    // misses contribute camera.rs:331's background colour: nothing at all when it is black and the scene has no sky / sun
    if (!exact && variant != 0 && scene->hs.infinite.empty() && cam.background[0] == 0.0 && cam.background[1] == 0.0 && cam.background[2] == 0.0 &&
        env_int("RTGPU_DROP_MISSES", 1) != 0)
        r.rc.flags |= RT_FLAG_DROP_MISSES;
    if (!exact && variant != 0 && r.bypass && env_int("RTGPU_INLINE_EMISSION", 1) != 0) r.rc.flags |= RT_FLAG_INLINE_EMISSION;
    r.variant = variant; r.grid_trace = c.sm_count * trace_ctas; r.grid_shade = c.sm_count * shade_ctas; r.max_depth = prm.max_depth;
    r.exact = exact; r.profile = profile;
    // unused ends of reserved record chunks (kernels.cuh: record_chunk): at most 255 records per warp, region (trace / bypass)
    // and kernel writing into the array — the shade kernels of the classes present, or k_raygen alone
    const uint64_t waste_shade = (uint64_t)popcount_u32(class_mask) * 2u * (uint64_t)r.grid_shade * 4u * 256u;
    const uint64_t waste_raygen = 2ull * (uint64_t)c.sm_count * 8u * 8u * 256u; // (up to 8 CTAs of 8 warps per SM: RTGPU_RAYGEN_CTAS_PER_SM)
    const uint32_t waste_max = (uint32_t)std::min<uint64_t>(std::max(waste_shade, waste_raygen), 0x7FFFFFFFull);

    // ---- batches and lanes within the memory budget
    size_t free_b = 0, total_b = 0;
    CK(cudaMemGetInfo(&free_b, &total_b));
    size_t owned = 0;
    for (auto& l : c.lanes) owned += l.bytes();
    const double avail = (double)free_b + (double)owned - 768.0 * 1048576.0;
    const double budget = std::min((double)env_int("RTGPU_POOL_MEM_PCT", 60) * 0.01 * (double)total_b, avail);
    // batch size: whole images of samples, about RTGPU_BATCH_MPATHS Mi paths (measured: tools/tune_trace.py, profiles/)
    const uint64_t batch_paths = (uint64_t)std::max(1, env_int("RTGPU_BATCH_MPATHS", 40)) << 20;
    // profiling times every kernel alone (one lane); otherwise a few batches are in flight
    const size_t want_lanes = profile ? 1 : (size_t)std::max(1, env_int("RTGPU_LANES", 4));
    const BatchPlan plan = plan_batches((uint64_t)(s_end - s_begin), npix, batch_paths, want_lanes, budget, [&](uint64_t paths) {
        return (double)base_bytes_for((uint32_t)std::min<uint64_t>(paths, 0xFFFFFFFFull), waste_max, class_mask, prm.max_depth);
    }); // (batch_plan.h: the rules and where they were measured)
    if (!plan.ok) throw CudaError{"not enough free device memory for one batch of one sample per pixel"};
    const uint64_t samples_per_batch = plan.samples_per_batch;
    uint64_t n_batches = plan.n_batches;
    const size_t n_lanes = plan.n_lanes;
    const uint32_t capacity = (uint32_t)(samples_per_batch * npix);
    for (size_t i = 0; i < n_lanes; i++) ensure_lane_base(c, i, capacity, prm.max_depth, waste_max, class_mask);
    const double base_total = (double)n_lanes * (double)base_bytes_for(capacity, waste_max, class_mask, prm.max_depth);

    // ---- tail merging: a lane runs its batches in groups of `tail_group`; the first ones stop after bounce tail_bounce - 1
    // and leave their surviving rays in the lane's carry arrays, the last one runs the remaining bounces on the
    // survivors of the whole group. The bounces past ~6 hold a few per cent of a batch's rays in launches whose
    // duration is the latency of the longest ray; merged, the same launches carry tail_group times the rays.
    // Same paths, same Philox streams (addressed by pixel, sample, bounce): the image changes by fp32 summation order only.
    const uint64_t batches_per_lane = (n_batches + n_lanes - 1) / n_lanes;
    uint32_t tail_group = (uint32_t)std::max(1, env_int("RTGPU_TAIL_GROUP", 5));
    const int tail_bounce = env_int("RTGPU_TAIL_BOUNCE", 4);
    if (c.no_merge || exact || variant == 0 || tail_bounce < 1 || tail_bounce >= prm.max_depth || batches_per_lane < 2) tail_group = 1;
    tail_group = (uint32_t)std::min<uint64_t>(tail_group, batches_per_lane);
    uint64_t sample_cursor = 0; // samples of [s_begin, s_end) already enqueued
    uint64_t carry_t = 0, carry_b = 0, carry_q = 0;
    if (tail_group > 1) {
        // How many paths are still alive at the tail bounce is a property of scene and camera; it is MEASURED: the
        // first sample of every pixel is rendered on its own (it is part of the job, nothing is wasted) and the
        // per-bounce counters are read back. The arrays that receive merged rays are sized from it with a margin of
        // RT_SURVIVAL_MARGIN, and every append on the device is bounds-checked (kernels.cuh) — a closed room, a
        // mirror hall or --light-bias=0 simply end up with a smaller group or none.
        struct { RtgpuCameraDesc cam; int32_t depth; double bias; uint32_t bypass; } keyed{};
        keyed.cam = cam; keyed.depth = prm.max_depth; keyed.bias = prm.light_bias; keyed.bypass = r.bypass | ((prm.flags & RTGPU_FLAG_FIXED) ? 2u : 0u);
        const uint64_t key = fnv1a(&keyed, sizeof keyed);
        DeviceCtx::Survival& sv = c.survival;
        if (!sv.valid || sv.key != key || (int)sv.all.size() < prm.max_depth) {
            DeviceCtx::Lane& l = c.lanes[0];
            publish_pool(l, 0xFFFFFFFFu);
            CK(cudaEventRecord(c.fork, main_stream));
            CK(cudaStreamWaitEvent(l.stream, c.fork, 0));
            launch_batch(r, l, 0, (uint32_t)npix, prm.max_depth + 2, prm.max_depth);
            std::vector<uint32_t> cnt((size_t)prm.max_depth * RT_COUNTERS_PER_BOUNCE);
            CK(cudaMemcpyAsync(cnt.data(), l.pool.counters, cnt.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost, l.stream));
            CK(cudaStreamSynchronize(l.stream));
            sv.all.assign((size_t)prm.max_depth, 0.0); sv.traced.assign((size_t)prm.max_depth, 0.0);
            for (int b = 0; b < prm.max_depth; b++) {
                const uint32_t* row = cnt.data() + (size_t)b * RT_COUNTERS_PER_BOUNCE;
                uint64_t all = 0;
                for (int k = 0; k < N_CLASSES; k++) all += row[RT_CNT_CLASS + k];
                sv.all[(size_t)b] = (double)all / (double)npix;
                sv.traced[(size_t)b] = (double)(row[0] - row[RT_CNT_DEAD]) / (double)npix;
            }
            sv.key = key; sv.valid = true;
            sample_cursor = 1;
            { std::lock_guard<std::mutex> lock(g_mutex); scene->stats.paths += npix; }
        }
        // RTGPU_SURVIVAL_SCALE (tests): pretend the pilot measured this multiple of the survival, e.g. 0.2 to provoke the
        // overflow checks and the un-merged retry of rtgpu_render
        const char* scale_env = std::getenv("RTGPU_SURVIVAL_SCALE");
        const double margin = RT_SURVIVAL_MARGIN * (scale_env ? std::atof(scale_env) : 1.0);
        const double alive = sv.all[(size_t)tail_bounce] * margin, traced = std::min(sv.traced[(size_t)tail_bounce], sv.all[(size_t)tail_bounce]) * margin;
        // the merged rays of bounces > tail_bounce live in the per-batch arrays: group x alive paths must fit one batch
        while (tail_group > 1 && (double)tail_group * alive > 1.0) tail_group--;
        for (; tail_group > 1; tail_group--) {
            // unused chunk ends per batch and region: at most chunk - 1 records per warp that wrote at all — bounded by the
            // rays the producers (the shade kernels of bounce tail_bounce - 1) read, and by their grids
            const double g = (double)tail_group, cap = (double)capacity;
            const double w = std::min(cap * sv.all[(size_t)tail_bounce - 1] * margin + 32.0 * popcount_u32(class_mask) + 64.0, (double)(waste_shade / 2));
            carry_t = (uint64_t)(g * (cap * traced + w)) + 64; carry_b = (uint64_t)(g * (cap * (alive - traced) + w)) + 64; carry_q = (uint64_t)(g * cap * alive) + 64;
            const double carry_bytes = (double)(carry_t + carry_b) * sizeof(PathRay) + (double)popcount_u32(class_mask) * (double)carry_q * 4.0 +
                                       std::max(0.0, (double)carry_t - (double)ray_slots_for(capacity, waste_max)) * sizeof(PathHit);
            if (base_total + (double)n_lanes * carry_bytes <= budget && carry_t + carry_b <= 0xFFFFFF00ull) break;
        }
    }
    const bool merge = tail_group > 1;
    if (merge) for (size_t i = 0; i < n_lanes; i++) ensure_lane_carry(c, i, carry_t, carry_b, carry_q);
    for (size_t i = 0; i < n_lanes; i++) publish_pool(c.lanes[i], merge ? (uint32_t)tail_bounce : 0xFFFFFFFFu);
    { std::lock_guard<std::mutex> lock(g_mutex);
      scene->stats.lanes = (uint32_t)n_lanes; scene->stats.batch_paths = capacity; scene->stats.tail_group = tail_group; scene->stats.tail_bounce = (uint32_t)tail_bounce; }
    if (std::getenv("RTGPU_POOL_REPORT")) {
        size_t bytes = 0;
        for (size_t i = 0; i < n_lanes; i++) bytes += c.lanes[i].bytes();
        std::fprintf(stderr, "[rtgpu pools] %zu lanes x %u paths, tail group %u from bounce %d, %.2f GB (budget %.1f GB)\n", n_lanes, capacity,
                     tail_group, tail_bounce, (double)bytes / 1073741824.0, budget / 1073741824.0);
    }

    CK(cudaEventRecord(c.fork, main_stream));
    for (size_t i = 0; i < n_lanes; i++) CK(cudaStreamWaitEvent(c.lanes[i].stream, c.fork, 0));
    const uint64_t total_samples = (uint64_t)(s_end - s_begin);
    const uint64_t todo = total_samples - sample_cursor;
    n_batches = (todo + samples_per_batch - 1) / samples_per_batch;
    uint64_t batch = 0;
    for (uint64_t s0 = sample_cursor; s0 < total_samples; s0 += samples_per_batch, batch++) {
        DeviceCtx::Lane& l = c.lanes[batch % n_lanes];
        const uint32_t n = (uint32_t)(std::min<uint64_t>(samples_per_batch, total_samples - s0) * npix);
        // position of the batch in its lane's current group
        const uint64_t in_lane = batch / n_lanes, lane_batches = (n_batches - (batch % n_lanes) + n_lanes - 1) / n_lanes;
        const bool group_first = !merge || in_lane % tail_group == 0;
        const bool group_last = !merge || in_lane % tail_group == tail_group - 1 || in_lane + 1 == lane_batches;
        const int b_reset = group_first ? prm.max_depth + 2 : tail_bounce;   // counter rows >= tail_bounce accumulate over the group
        const int b_end = group_last ? prm.max_depth : tail_bounce;
        launch_batch(r, l, s0 * npix, n, b_reset, b_end);
    }
    for (size_t i = 0; i < n_lanes; i++) {
        CK(cudaEventRecord(c.lanes[i].done, c.lanes[i].stream));
        CK(cudaStreamWaitEvent(main_stream, c.lanes[i].done, 0));
    }
    CK(cudaGetLastError());
    { std::lock_guard<std::mutex> lock(g_mutex); scene->stats.paths += todo * npix; } // (one host thread per device)
}

// fn(d) for every device index, one host thread each (a single device runs inline); the first CudaError is re-thrown.
// Context-bound work of different devices (uploads, pool allocation, ~600 launches per render) does not serialise
// on one host thread this way.
template <typename F> void for_each_device(int n_dev, F fn) {
    if (n_dev <= 1) { if (n_dev == 1) fn(0); return; }
    std::vector<std::string> errs((size_t)n_dev);
    std::vector<std::thread> pool;
    for (int d = 0; d < n_dev; d++)
        pool.emplace_back([&, d] { try { fn(d); } catch (CudaError& e) { errs[(size_t)d] = e.msg.empty() ? "CUDA error" : e.msg; } });
    for (auto& t : pool) t.join();
    for (auto& e : errs) if (!e.empty()) throw CudaError{e};
}

int check_params(const RtgpuCameraDesc* cam, const RtgpuRenderParams* p, int& s_begin, int& s_end) {
    if (!cam || !p) { rthost::set_error("null camera/params"); return RTGPU_ERR_INVALID; }
    if (cam->width <= 0 || cam->height <= 0) { rthost::set_error("camera: bad image size"); return RTGPU_ERR_INVALID; }
    if (p->replicas < 1 || p->sqrt_spt < 1 || p->max_depth < 0 || p->max_depth > 65535) { rthost::set_error("params: replicas/sqrt_spt must be >= 1 and 0 <= max_depth <= 65535"); return RTGPU_ERR_INVALID; }
    const int spp = p->replicas * p->sqrt_spt * p->sqrt_spt;
    s_begin = p->sample_begin; s_end = p->sample_end;
    if (s_begin == 0 && s_end == 0) s_end = spp;
    if (s_begin < 0 || s_end > spp || s_begin > s_end) { rthost::set_error("params: sample range out of [0, spp]"); return RTGPU_ERR_INVALID; }
    return 0;
}

// Resolved f64 Buffer on the device + pinned host staging, cached per device (a pageable 46 MB copy
// and a fresh cudaMalloc per call cost more than the 1200x1200 resolve itself).
struct ResolveStage {
    double* d_out = nullptr;
    double* h_stage = nullptr;
    const float4** d_ptrs = nullptr;
    size_t capacity = 0;
    float* h_acc = nullptr;        // single-GPU path: pinned copy of the fp32 accumulator
    size_t acc_capacity = 0;
    uint8_t* d_rgb = nullptr;      // rtgpu_render_srgb8: RGB8 image on the device and its pinned copy
    uint8_t* h_rgb = nullptr;
    size_t rgb_capacity = 0;
    cudaEvent_t chunk_done[16] = {};
};
std::vector<ResolveStage> g_stage(64);
std::mutex g_stage_mutex;

ResolveStage& resolve_stage(int device, size_t n_out) {
    ResolveStage& st = g_stage[(size_t)device % g_stage.size()];
    if (st.capacity < n_out) {
        if (st.h_stage) CK(cudaFreeHost(st.h_stage));
        if (st.d_out) CK(cudaFree(st.d_out));
        st.h_stage = nullptr; st.d_out = nullptr;
        CK(cudaMalloc((void**)&st.d_out, n_out * sizeof(double)));
        CK(cudaHostAlloc((void**)&st.h_stage, n_out * sizeof(double), cudaHostAllocDefault));
        if (!st.d_ptrs) CK(cudaMalloc((void**)&st.d_ptrs, 64 * sizeof(float4*)));
        st.capacity = n_out;
    }
    return st;
}

// camera.rs:247-253: render ADDS into the caller's buffer
void add_into(double* out_rgba, const double* src, size_t n_out) {
    const unsigned n_thr = std::max(1u, std::min(8u, std::thread::hardware_concurrency()));
    if (n_out < (1u << 18) || n_thr == 1) {
        for (size_t i = 0; i < n_out; i++) out_rgba[i] += src[i];
        return;
    }
    std::vector<std::thread> pool;
    for (unsigned t = 0; t < n_thr; t++)
        pool.emplace_back([=] { for (size_t i = n_out * t / n_thr, e = n_out * (t + 1) / n_thr; i < e; i++) out_rgba[i] += src[i]; });
    for (auto& th : pool) th.join();
}

// One GPU: the fp32 sums cross PCIe as they are (half the bytes of the f64 Buffer) in RT_RESOLVE_CHUNKS
// pieces; host threads convert, scale by 1 / spp in f64 (the arithmetic of k_reduce_resolve) and add piece c while piece
// c + 1 is still in flight.
constexpr int RT_RESOLVE_CHUNKS = 8;
void resolve_one_gpu_pipelined(ResolveStage& st, const float4* accum, uint32_t npix, double inv_spp, cudaStream_t stream, double* out_rgba) {
    const size_t n = (size_t)npix * 4;
    if (st.acc_capacity < n) {
        if (st.h_acc) CK(cudaFreeHost(st.h_acc));
        st.h_acc = nullptr;
        CK(cudaHostAlloc((void**)&st.h_acc, n * sizeof(float), cudaHostAllocDefault));
        st.acc_capacity = n;
    }
    for (int c = 0; c < RT_RESOLVE_CHUNKS; c++) {
        if (!st.chunk_done[c]) CK(cudaEventCreateWithFlags(&st.chunk_done[c], cudaEventDisableTiming));
        const size_t b = (size_t)npix * c / RT_RESOLVE_CHUNKS * 4, e = (size_t)npix * (c + 1) / RT_RESOLVE_CHUNKS * 4;
        if (e > b) CK(cudaMemcpyAsync(st.h_acc + b, reinterpret_cast<const float*>(accum) + b, (e - b) * sizeof(float), cudaMemcpyDeviceToHost, stream));
        CK(cudaEventRecord(st.chunk_done[c], stream));
    }
    const unsigned n_thr = npix < (1u << 16) ? 1u : std::max(1u, std::min(8u, std::thread::hardware_concurrency()));
    std::vector<cudaError_t> status(n_thr, cudaSuccess);
    auto work = [&](unsigned t) {
        for (int c = 0; c < RT_RESOLVE_CHUNKS; c++) {
            const cudaError_t rc = cudaEventSynchronize(st.chunk_done[c]);
            if (rc != cudaSuccess) { status[t] = rc; return; }
            const size_t p0 = (size_t)npix * c / RT_RESOLVE_CHUNKS, p1 = (size_t)npix * (c + 1) / RT_RESOLVE_CHUNKS;
            for (size_t i = p0 + (p1 - p0) * t / n_thr, e = p0 + (p1 - p0) * (t + 1) / n_thr; i < e; i++) {
                const float* a = st.h_acc + i * 4;
                double* o = out_rgba + i * 4;
                o[0] += (double)a[0] * inv_spp; o[1] += (double)a[1] * inv_spp; o[2] += (double)a[2] * inv_spp; // (alpha: += 0)
            }
        }
    };
    if (n_thr == 1) work(0);
    else {
        std::vector<std::thread> pool;
        for (unsigned t = 0; t < n_thr; t++) pool.emplace_back(work, t);
        for (auto& th : pool) th.join();
    }
    for (cudaError_t rc : status) CK(rc);
}

// sum of the devices' fp32 accumulators / spp -> f64 on `device`, D2H through pinned memory, += into out_rgba
void resolve_add_to_host(int device, const std::vector<const float4*>& accums, uint32_t npix, double inv_spp, cudaStream_t stream, double* out_rgba) {
    std::lock_guard<std::mutex> lock(g_stage_mutex);
    if (accums.size() > 64) throw CudaError{"more than 64 devices"};
    const size_t n_out = (size_t)npix * 4;
    if (accums.size() == 1) {
        resolve_one_gpu_pipelined(g_stage[(size_t)device % g_stage.size()], accums[0], npix, inv_spp, stream, out_rgba);
        return;
    }
    ResolveStage& st = resolve_stage(device, n_out);
    CK(cudaMemcpyAsync(st.d_ptrs, accums.data(), accums.size() * sizeof(float4*), cudaMemcpyHostToDevice, stream));
    k_reduce_resolve<<<(npix + 255) / 256, 256, 0, stream>>>(st.d_ptrs, (int)accums.size(), st.d_out, npix, inv_spp);
    CK(cudaMemcpyAsync(st.h_stage, st.d_out, n_out * sizeof(double), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    add_into(out_rgba, st.h_stage, n_out);
}

} // namespace

extern "C" {

int rtgpu_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

// Device selection + CUDA bring-up (driver init ~1.2 s and one primary context per device on the GPU box) run on a
// helper thread so that the host half of the "Ready" phase — flag parsing, scene DSL, OBJ parse, flatten, BVH build,
// none of which touches CUDA — overlaps with it. wait_init() joins before the first upload.
namespace {
// (joins at process exit: a still-joinable std::thread must not reach its destructor when the caller stops, e.g. on a
// scene-load error, before anything has waited for the bring-up)
struct JoiningThread {
    std::thread t;
    ~JoiningThread() { if (t.joinable()) t.join(); }
};
JoiningThread g_init_holder;
std::thread& g_init_thread = g_init_holder.t;
bool g_init_pending = false;
int g_init_rc = 0;
std::string g_init_err;

void init_worker(std::vector<int> want, bool use_default) {
    int count = rtgpu_device_count();
    if (count <= 0) { g_init_rc = RTGPU_ERR_NO_DEVICE; g_init_err = "no CUDA device available (this backend has no CPU fallback)"; return; }
    std::vector<int> sel;
    if (use_default) sel.push_back(0);
    else for (int d : want) {
        if (d < 0 || d >= count) { g_init_rc = RTGPU_ERR_INVALID; g_init_err = "rtgpu_init: device index out of range"; return; }
        sel.push_back(d);
    }
    // primary contexts, one helper thread per device (context creation is serial per device, parallel across devices)
    std::vector<std::thread> warm;
    for (int d : sel) warm.emplace_back([d] { if (cudaSetDevice(d) == cudaSuccess) cudaFree(nullptr); cudaGetLastError(); });
    for (auto& t : warm) t.join();
    std::lock_guard<std::mutex> lock(g_mutex);
    g_devices = sel;
    g_init_rc = 0;
}

int wait_init() {
    if (g_init_pending) {
        if (g_init_thread.joinable()) g_init_thread.join();
        g_init_pending = false;
        if (g_init_rc) rthost::set_error(g_init_err); // (the error string is thread-local: re-raise it on the caller's thread)
    }
    return g_init_rc;
}
} // namespace

int rtgpu_init_async(const int* devices, int n_devices) {
    wait_init();
    std::vector<int> want;
    const bool use_default = !devices || n_devices <= 0;
    if (!use_default) want.assign(devices, devices + n_devices);
    g_init_rc = 0; g_init_err.clear();
    g_init_pending = true;
    g_init_thread = std::thread(init_worker, want, use_default);
    return 0;
}

int rtgpu_init(const int* devices, int n_devices) {
    rtgpu_init_async(devices, n_devices);
    return wait_init();
}

int rtgpu_shutdown(void) {
    wait_init();
    std::lock_guard<std::mutex> lock(g_mutex);
    g_devices.clear();
    return 0;
}

int rtgpu_scene_create(const RtgpuSceneDesc* desc, RtgpuScene** out_scene) {
    if (!desc || !out_scene) { rthost::set_error("rtgpu_scene_create: null argument"); return RTGPU_ERR_INVALID; }
    if (!g_init_pending && g_devices.empty()) rtgpu_init_async(nullptr, 0);
    auto* sc = new RtgpuScene();
    std::string err;
    int rc;
    // The acceleration structure is built ON THE DEVICE (build_gpu.cuh: Morton sort, radix tree, DP collapse: 3-6 ms for the
    // 1.0 M-triangle dragon against ~290 ms for the host's binned-SAH build, and the tree traverses as fast); the flattening
    // around it is host code and still overlaps with a pending rtgpu_init_async — the builder waits for CUDA only when
    // it is reached. RTGPU_BUILD=host selects the host builder (bvh.cpp).
    const char* build_env = std::getenv("RTGPU_BUILD");
    const bool gpu_build = !(build_env && std::string(build_env) == "host");
    if (gpu_build) {
        std::lock_guard<std::mutex> lock(g_build_mutex); // (the builder hook is process-wide)
        g_last_gpu_build = GpuBuildTimes();
        try {
            g_wait_init_fn = wait_init;
            g_wide_bvh_builder = build_wide_bvh_gpu_entry;
            rc = flatten_scene(*desc, sc->hs, err);
            g_wide_bvh_builder = build_wide_bvh;
        } catch (CudaError& e) {
            g_wide_bvh_builder = build_wide_bvh;
            rthost::set_error(e.msg);
            delete sc;
            return e.code;
        }
        sc->stats_gpu_build_ms = g_last_gpu_build.ms_device;
    } else rc = flatten_scene(*desc, sc->hs, err); // host only: overlaps with a pending rtgpu_init_async
    if (rc) { rthost::set_error(err); delete sc; return rc; }
    if (sc->hs.bvh.max_depth > RT_STACK_SIZE) { rthost::set_error("BVH deeper than the traversal stack"); delete sc; return RTGPU_ERR_UNSUPPORTED; }
    {
        auto t_w = std::chrono::steady_clock::now();
        rc = wait_init();
        if (std::getenv("RTGPU_BUILD_TIMING")) std::fprintf(stderr, "[rtgpu build] waited for CUDA bring-up     %8.1f ms\n",
                                                           std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_w).count());
        if (rc != 0) { delete sc; return rc; }
    }
    try {
        const bool timing = std::getenv("RTGPU_BUILD_TIMING") != nullptr;
        auto t_up = std::chrono::steady_clock::now();
        sc->devs.resize(g_devices.size());
        for (size_t i = 0; i < g_devices.size(); i++) sc->devs[i].device = g_devices[i];
        for_each_device((int)g_devices.size(), [&](int i) { upload_scene(sc->hs, sc->devs[(size_t)i]); });
        if (timing) std::fprintf(stderr, "[rtgpu build] upload to %zu device(s)      %8.1f ms\n", g_devices.size(),
                                 std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_up).count());
    } catch (CudaError& e) {
        rthost::set_error(e.msg);
        for (auto& d : sc->devs) d.release();
        delete sc;
        return RTGPU_ERR_CUDA;
    }
    sc->stats.bvh_nodes = sc->hs.bvh.nodes.size(); sc->stats.bvh_triangles = sc->hs.n_triangles; sc->stats.bvh_prims = sc->hs.prims.size();
    sc->stats.ms_build = sc->hs.ms_build;
    sc->stats.ms_build_device = sc->stats_gpu_build_ms;
    *out_scene = sc;
    return 0;
}

int rtgpu_scene_destroy(RtgpuScene* scene) {
    if (!scene) return 0;
    for (auto& d : scene->devs) d.release();
    delete scene;
    return 0;
}

int rtgpu_render_device(RtgpuScene* scene, int device_index, const RtgpuCameraDesc* cam, const RtgpuRenderParams* params,
                        float* d_accum_rgba, void* stream) {
    if (!scene || !d_accum_rgba) { rthost::set_error("rtgpu_render_device: null argument"); return RTGPU_ERR_INVALID; }
    if (device_index < 0 || device_index >= (int)scene->devs.size()) { rthost::set_error("rtgpu_render_device: bad device index"); return RTGPU_ERR_INVALID; }
    int s_begin, s_end;
    int rc = check_params(cam, params, s_begin, s_end);
    if (rc) return rc;
    try {
        enqueue_render(scene, scene->devs[device_index], *cam, *params, s_begin, s_end, (float4*)d_accum_rgba, (cudaStream_t)stream);
    } catch (CudaError& e) { rthost::set_error(e.msg); return RTGPU_ERR_CUDA; }
    return 0;
}

int rtgpu_resolve_device(int device_index, const float* d_accum_rgba, int width, int height, int total_spp, double* out_rgba, void* stream) {
    if (!d_accum_rgba || !out_rgba || width <= 0 || height <= 0 || total_spp <= 0) { rthost::set_error("rtgpu_resolve_device: bad argument"); return RTGPU_ERR_INVALID; }
    wait_init();
    if (g_devices.empty() || device_index < 0 || device_index >= (int)g_devices.size()) { rthost::set_error("rtgpu_resolve_device: bad device index"); return RTGPU_ERR_INVALID; }
    try {
        CK(cudaSetDevice(g_devices[device_index]));
        const uint32_t npix = (uint32_t)width * (uint32_t)height;
        resolve_add_to_host(g_devices[device_index], {(const float4*)d_accum_rgba}, npix, 1.0 / (double)total_spp, (cudaStream_t)stream, out_rgba);
    } catch (CudaError& e) { rthost::set_error(e.msg); return RTGPU_ERR_CUDA; }
    return 0;
}

namespace {
// Camera::render on n_dev devices: every device renders its contiguous block of the sample range into its own fp32
// accumulator; returns their addresses (those of devices > 0 readable from device 0).
std::vector<const float4*> render_accumulators(RtgpuScene* scene, const RtgpuCameraDesc* cam, const RtgpuRenderParams* params, int s_begin, int s_end, int n_dev) {
    const uint32_t npix = (uint32_t)cam->width * (uint32_t)cam->height;
    const int total = s_end - s_begin;
    for_each_device(n_dev, [&](int d) {
        DeviceCtx& c = scene->devs[d];
        CK(cudaSetDevice(c.device));
        const int reader = scene->devs[0].device;
        if (c.accum_pixels < npix || (n_dev > 1 && d > 0 && (!c.accum_shared || c.shared_reader != reader))) {
            if (c.accum_from_pool && c.accum) { CK(cudaStreamSynchronize(c.stream)); CK(cudaFree(c.accum)); }
            c.accum = nullptr; c.accum_from_pool = false; c.accum_shared = false;
            if (n_dev > 1 && d > 0) { // read by device 0's k_reduce_resolve / k_resolve_tonemap through NVLink
                int can = 0, pools = 0;
                CK(cudaDeviceCanAccessPeer(&can, reader, c.device));
                CK(cudaDeviceGetAttribute(&pools, cudaDevAttrMemoryPoolsSupported, c.device));
                if (!can || !pools) throw CudaError{"peer access between the selected devices is not available"};
                if (c.shared_pool && c.shared_reader != reader) { CK(cudaMemPoolDestroy(c.shared_pool)); c.shared_pool = nullptr; }
                if (!c.shared_pool) {
                    cudaMemPoolProps props{};
                    props.allocType = cudaMemAllocationTypePinned;
                    props.handleTypes = cudaMemHandleTypeNone;
                    props.location.type = cudaMemLocationTypeDevice;
                    props.location.id = c.device;
                    CK(cudaMemPoolCreate(&c.shared_pool, &props));
                    cudaMemAccessDesc acc{};
                    acc.location.type = cudaMemLocationTypeDevice;
                    acc.location.id = reader;
                    acc.flags = cudaMemAccessFlagsProtReadWrite;
                    CK(cudaMemPoolSetAccess(c.shared_pool, &acc, 1));
                    c.shared_reader = reader;
                }
                void* p = nullptr;
                CK(cudaMallocFromPoolAsync(&p, (size_t)npix * sizeof(float4), c.shared_pool, c.stream));
                CK(cudaStreamSynchronize(c.stream));
                c.accum = (float4*)p; c.accum_from_pool = true; c.accum_shared = true;
            } else c.accum = c.alloc<float4>(npix);
            c.accum_pixels = npix;
        }
        const int b = s_begin + (int)((int64_t)total * d / n_dev), e = s_begin + (int)((int64_t)total * (d + 1) / n_dev);
        for (int attempt = 0; attempt < 2; attempt++) {
            CK(cudaMemsetAsync(c.accum, 0, (size_t)npix * sizeof(float4), c.stream));
            if (e > b) enqueue_render(scene, c, *cam, *params, b, e, c.accum, nullptr);
            CK(cudaStreamSynchronize(c.stream));
            try { check_pool_error(c); break; }
            catch (CudaError&) {
                // more paths survived than the pilot sample predicted (every append was bounds-checked: nothing was
                // overwritten, but rays were dropped): render this device's samples again without tail merging
                if (attempt == 1) throw;
                c.no_merge = true;
            }
        }
        c.no_merge = false;
    });
    DeviceCtx& c0 = scene->devs[0];
    CK(cudaSetDevice(c0.device));
    std::vector<const float4*> ptrs;
    for (int d = 0; d < n_dev; d++) ptrs.push_back(scene->devs[d].accum); // (d > 0: readable from device 0 through its pool's access list)
    return ptrs;
}
} // namespace

int rtgpu_render(RtgpuScene* scene, const RtgpuCameraDesc* cam, const RtgpuRenderParams* params, double* out_rgba) {
    if (!scene || !out_rgba) { rthost::set_error("rtgpu_render: null argument"); return RTGPU_ERR_INVALID; }
    int s_begin, s_end;
    int rc = check_params(cam, params, s_begin, s_end);
    if (rc) return rc;
    const int n_dev = std::max(1, std::min<int>(params->n_gpus > 0 ? params->n_gpus : 1, (int)scene->devs.size()));
    const uint32_t npix = (uint32_t)cam->width * (uint32_t)cam->height;
    const int spp = params->replicas * params->sqrt_spt * params->sqrt_spt;
    try {
        auto t0 = std::chrono::steady_clock::now();
        const std::vector<const float4*> ptrs = render_accumulators(scene, cam, params, s_begin, s_end, n_dev);
        // reduce + resolve on device 0 (peer reads over NVLink when several devices took part)
        DeviceCtx& c0 = scene->devs[0];
        resolve_add_to_host(c0.device, ptrs, npix, 1.0 / (double)spp, c0.stream, out_rgba);
        c0.launches++;
        scene->stats.ms_total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    } catch (CudaError& e) { rthost::set_error(e.msg); return RTGPU_ERR_CUDA; }
    return 0;
}

int rtgpu_render_srgb8(RtgpuScene* scene, const RtgpuCameraDesc* cam, const RtgpuRenderParams* params, int tonemap, uint8_t* out_rgb) {
    if (!scene || !out_rgb) { rthost::set_error("rtgpu_render_srgb8: null argument"); return RTGPU_ERR_INVALID; }
    int s_begin, s_end;
    int rc = check_params(cam, params, s_begin, s_end);
    if (rc) return rc;
    const int n_dev = std::max(1, std::min<int>(params->n_gpus > 0 ? params->n_gpus : 1, (int)scene->devs.size()));
    const uint32_t npix = (uint32_t)cam->width * (uint32_t)cam->height;
    const int spp = params->replicas * params->sqrt_spt * params->sqrt_spt;
    try {
        auto t0 = std::chrono::steady_clock::now();
        const std::vector<const float4*> ptrs = render_accumulators(scene, cam, params, s_begin, s_end, n_dev);
        if (ptrs.size() > 64) throw CudaError{"more than 64 devices"};
        DeviceCtx& c0 = scene->devs[0];
        std::lock_guard<std::mutex> lock(g_stage_mutex);
        ResolveStage& st = g_stage[(size_t)c0.device % g_stage.size()];
        if (!st.d_ptrs) CK(cudaMalloc((void**)&st.d_ptrs, 64 * sizeof(float4*)));
        if (st.rgb_capacity < (size_t)npix * 3) {
            if (st.d_rgb) CK(cudaFree(st.d_rgb));
            if (st.h_rgb) CK(cudaFreeHost(st.h_rgb));
            st.d_rgb = nullptr; st.h_rgb = nullptr;
            CK(cudaMalloc((void**)&st.d_rgb, (size_t)npix * 3));
            CK(cudaHostAlloc((void**)&st.h_rgb, (size_t)npix * 3, cudaHostAllocDefault));
            st.rgb_capacity = (size_t)npix * 3;
        }
        CK(cudaMemcpyAsync(st.d_ptrs, ptrs.data(), ptrs.size() * sizeof(float4*), cudaMemcpyHostToDevice, c0.stream));
        k_resolve_tonemap<<<(npix + 255) / 256, 256, 0, c0.stream>>>(st.d_ptrs, (int)ptrs.size(), st.d_rgb, npix, 1.0 / (double)spp, tonemap);
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(st.h_rgb, st.d_rgb, (size_t)npix * 3, cudaMemcpyDeviceToHost, c0.stream));
        CK(cudaStreamSynchronize(c0.stream));
        std::memcpy(out_rgb, st.h_rgb, (size_t)npix * 3);
        c0.launches++;
        scene->stats.ms_total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    } catch (CudaError& e) { rthost::set_error(e.msg); return RTGPU_ERR_CUDA; }
    return 0;
}

int rtgpu_tonemap_srgb8(const double* rgba, int width, int height, int tonemap, uint8_t* out_rgb) {
    if (!rgba || !out_rgb || width <= 0 || height <= 0) { rthost::set_error("rtgpu_tonemap_srgb8: bad argument"); return RTGPU_ERR_INVALID; }
    { int rc = wait_init(); if (rc) return rc; }
    if (g_devices.empty()) { int rc = rtgpu_init(nullptr, 0); if (rc) return rc; }
    try {
        CK(cudaSetDevice(g_devices[0]));
        const uint32_t npix = (uint32_t)width * (uint32_t)height;
        double* d_in = nullptr; uint8_t* d_out = nullptr;
        CK(cudaMalloc(&d_in, (size_t)npix * 4 * sizeof(double)));
        CK(cudaMalloc(&d_out, (size_t)npix * 3));
        CK(cudaMemcpy(d_in, rgba, (size_t)npix * 4 * sizeof(double), cudaMemcpyHostToDevice));
        k_tonemap<<<(npix + 255) / 256, 256>>>(d_in, d_out, npix, tonemap);
        CK(cudaGetLastError());
        CK(cudaMemcpy(out_rgb, d_out, (size_t)npix * 3, cudaMemcpyDeviceToHost));
        CK(cudaFree(d_in)); CK(cudaFree(d_out));
    } catch (CudaError& e) { rthost::set_error(e.msg); return RTGPU_ERR_CUDA; }
    return 0;
}

// ---- rtgpu_tonemap_srgb8_noalloc ------------------------------------------------------------------------------------
// Same maths and kernel as rtgpu_tonemap_srgb8, for callers that may not allocate (CUDA graph capture around them, tight
// serving loops, harnesses that forbid cudaMalloc). No counterpart in the reference.
// The device staging rtgpu_tonemap_srgb8 cudaMallocs per call (the f64 RGBA input and the RGB8 output) lives in
// ONE file-scope device array instead: RTGPU_NOALLOC_MAX_PIXELS x (32 + 3) bytes = 8.75 MiB, held by every GPU that loads
// the library until the process exits. Every call on a device shares it and it keeps the last call's contents; nothing
// in it has to start at zero (the kernel reads only the input region, which the call has just overwritten, and writes
// every output byte it copies back). Calls that overlap on the same GPU race on it: serialise them.
} // extern "C"
namespace {
constexpr size_t RT_NOALLOC_IN_BYTES = (size_t)RTGPU_NOALLOC_MAX_PIXELS * 4 * sizeof(double);
constexpr size_t RT_NOALLOC_OUT_BYTES = (size_t)RTGPU_NOALLOC_MAX_PIXELS * 3;
static __device__ __align__(256) unsigned char g_noalloc_scratch[RT_NOALLOC_IN_BYTES + RT_NOALLOC_OUT_BYTES];
// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address a kernel may use comes
// from cudaGetSymbolAddress, once per device, with that device current.
constexpr int RT_NOALLOC_MAX_DEVICES = 64;
unsigned char* g_noalloc_ptr[RT_NOALLOC_MAX_DEVICES] = {};
std::mutex g_noalloc_mutex;
} // namespace
extern "C" {

int rtgpu_tonemap_srgb8_noalloc(const double* rgba, int width, int height, int tonemap, uint8_t* out_rgb) {
    if (!rgba || !out_rgb || width <= 0 || height <= 0) { rthost::set_error("rtgpu_tonemap_srgb8_noalloc: bad argument"); return RTGPU_ERR_INVALID; }
    if ((uint64_t)width * (uint64_t)height > (uint64_t)RTGPU_NOALLOC_MAX_PIXELS) {
        rthost::set_error("rtgpu_tonemap_srgb8_noalloc: image larger than RTGPU_NOALLOC_MAX_PIXELS (262144 pixels): the static scratch "
                          "array cannot hold it; use rtgpu_tonemap_srgb8");
        return RTGPU_ERR_INVALID;
    }
    { int rc = wait_init(); if (rc) return rc; }
    if (g_devices.empty()) { int rc = rtgpu_init(nullptr, 0); if (rc) return rc; }
    const int dev = g_devices[0];
    if (dev < 0 || dev >= RT_NOALLOC_MAX_DEVICES) { rthost::set_error("rtgpu_tonemap_srgb8_noalloc: device ordinal out of range"); return RTGPU_ERR_INVALID; }
    std::lock_guard<std::mutex> lock(g_noalloc_mutex); // (also serialises callers of this process on the shared array)
    cudaError_t e = cudaSetDevice(dev);
    if (e == cudaSuccess && !g_noalloc_ptr[dev]) {
        unsigned char* ptr = nullptr;
        e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
        if (e == cudaSuccess) g_noalloc_ptr[dev] = ptr;
    }
    const uint32_t npix = (uint32_t)width * (uint32_t)height;
    unsigned char* base = g_noalloc_ptr[dev];
    if (e == cudaSuccess) e = cudaMemcpy(base, rgba, (size_t)npix * 4 * sizeof(double), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        k_tonemap<<<(npix + 255) / 256, 256>>>(reinterpret_cast<const double*>(base), base + RT_NOALLOC_IN_BYTES, npix, tonemap);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(out_rgb, base + RT_NOALLOC_IN_BYTES, (size_t)npix * 3, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) { rthost::set_error(std::string("rtgpu_tonemap_srgb8_noalloc: ") + cudaGetErrorString(e)); return RTGPU_ERR_CUDA; }
    return 0;
}

int rtgpu_trace_primary(RtgpuScene* scene, const double* rays, int64_t n, RtgpuHit* out_hits) {
    if (!scene || !rays || !out_hits || n < 0) { rthost::set_error("rtgpu_trace_primary: bad argument"); return RTGPU_ERR_INVALID; }
    if (n == 0) return 0;
    try {
        DeviceCtx& c = scene->devs[0];
        CK(cudaSetDevice(c.device));
        double* d_rays = nullptr; RtgpuHit* d_hits = nullptr;
        CK(cudaMalloc(&d_rays, (size_t)n * 6 * sizeof(double)));
        CK(cudaMalloc(&d_hits, (size_t)n * sizeof(RtgpuHit)));
        // on the kernel's stream: a synchronous cudaMemcpy from pageable memory may return before its DMA has landed, and
        // c.stream is a non-blocking stream the legacy default stream does not order with (seen as a few stale rays at the
        // end of the array when the allocation re-used dirty memory)
        CK(cudaMemcpyAsync(d_rays, rays, (size_t)n * 6 * sizeof(double), cudaMemcpyHostToDevice, c.stream));
        k_trace_exact<<<(unsigned)((n + 127) / 128), 128, 0, c.stream>>>(c.ds, d_rays, (long long)n, d_hits);
        c.launches++;
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(c.stream));
        CK(cudaMemcpy(out_hits, d_hits, (size_t)n * sizeof(RtgpuHit), cudaMemcpyDeviceToHost));
        CK(cudaFree(d_rays)); CK(cudaFree(d_hits));
    } catch (CudaError& e) { rthost::set_error(e.msg); return RTGPU_ERR_CUDA; }
    return 0;
}

int rtgpu_trace_fast(RtgpuScene* scene, const double* rays, int64_t n, RtgpuFastHit* out_hits) {
    if (!scene || !rays || !out_hits || n < 0 || n > (int64_t)1 << 30) { rthost::set_error("rtgpu_trace_fast: bad argument"); return RTGPU_ERR_INVALID; }
    if (n == 0) return 0;
    try {
        DeviceCtx& c = scene->devs[0];
        CK(cudaSetDevice(c.device));
        ensure_lane(c, 0, (uint32_t)n, 1);
        DeviceCtx::Lane& l = c.lanes[0];
        std::vector<PathRay> host((size_t)n);
        for (int64_t i = 0; i < n; i++) {
            PathRay r{};
            r.ox = rays[i * 6]; r.oy = rays[i * 6 + 1]; r.oz = rays[i * 6 + 2];
            r.dx = (float)rays[i * 6 + 3]; r.dy = (float)rays[i * 6 + 4]; r.dz = (float)rays[i * 6 + 5];
            r.pixel = (uint32_t)i; r.sample = 0; r.tr = r.tg = r.tb = 1.f;
            host[(size_t)i] = r;
        }
        CK(cudaMemcpyAsync(l.pool.rays[0], host.data(), host.size() * sizeof(PathRay), cudaMemcpyHostToDevice, l.stream)); // record i = queue entry i
        CK(cudaMemsetAsync(l.pool.counters, 0, (size_t)3 * RT_COUNTERS_PER_BOUNCE * sizeof(uint32_t), l.stream));
        const uint32_t n32 = (uint32_t)n;
        CK(cudaMemcpyAsync(l.pool.counters, &n32, sizeof n32, cudaMemcpyHostToDevice, l.stream));
        RenderConst rc{1, 0.f, 1, 0u};
        k_pretest_pool<<<(n32 + 127) / 128, 128, 0, l.stream>>>(c.ds, l.pool, n32); // what k_raygen / k_shade do when they write a ray
        const int time_hook = env_int("RTGPU_TIME_HOOK", 0); // tools/sort_experiment.py: device time of the trace launch alone
        cudaEvent_t ev0 = nullptr, ev1 = nullptr;
        if (time_hook) { CK(cudaEventCreate(&ev0)); CK(cudaEventCreate(&ev1)); CK(cudaEventRecord(ev0, l.stream)); }
        DeviceScene surfaces_only = c.ds; // constant media are stochastic: this hook reports the surface hit (like rtgpu_trace_primary)
        surfaces_only.n_volumes = 0;
        launch_trace_fast(c, l, scene->hs.trace_cold, env_int("RTGPU_TRACE_VARIANT", 4), c.sm_count * 8, rc, 0, &surfaces_only);
        if (time_hook) {
            CK(cudaEventRecord(ev1, l.stream)); CK(cudaEventSynchronize(ev1));
            float ms = 0.f; CK(cudaEventElapsedTime(&ms, ev0, ev1));
            c.ms_family[1] = ms; // read back through rtgpu_get_stats (ms_trace)
            cudaEventDestroy(ev0); cudaEventDestroy(ev1);
        }
        k_refine_hits<<<(n32 + 127) / 128, 128, 0, l.stream>>>(c.ds, l.pool, n32); // what surface_for_slot does before shading
        c.launches += 3;
        CK(cudaGetLastError());
        std::vector<PathHit> hits((size_t)n);
        CK(cudaMemcpyAsync(hits.data(), l.pool.hits, hits.size() * sizeof(PathHit), cudaMemcpyDeviceToHost, l.stream));
        CK(cudaStreamSynchronize(l.stream));
        for (int64_t i = 0; i < n; i++) {
            const PathHit& h = hits[(size_t)i];
            RtgpuFastHit o{};
            o.instance = -1; o.primitive = 0; o.t = h.t; o.u = h.u; o.v = h.v;
            if (h.prim < scene->hs.prims.size()) {
                const Prim& p = scene->hs.prims[h.prim];
                uint32_t inst, tri, flags;
                std::memcpy(&inst, &p.a.w, 4); std::memcpy(&tri, &p.b.w, 4); std::memcpy(&flags, &p.c.w, 4);
                o.instance = (int32_t)inst;
                o.primitive = (flags & PRIM_ANALYTIC) ? 0 : (int32_t)tri;
            }
            out_hits[i] = o;
        }
    } catch (CudaError& e) { rthost::set_error(e.msg); return RTGPU_ERR_CUDA; }
    return 0;
}

int rtgpu_bounce(RtgpuScene* scene, double light_bias, const double* rays, const double* xi, int n_xi, int64_t n, RtgpuBounce* out) {
    if (!scene || !rays || !xi || !out || n < 0 || n_xi < 1) { rthost::set_error("rtgpu_bounce: bad argument"); return RTGPU_ERR_INVALID; }
    if (n == 0) return 0;
    static_assert(sizeof(RtgpuBounce) == sizeof(BounceRecord), "RtgpuBounce layout");
    try {
        DeviceCtx& c = scene->devs[0];
        CK(cudaSetDevice(c.device));
        double *d_rays = nullptr, *d_xi = nullptr; BounceRecord* d_out = nullptr;
        CK(cudaMalloc(&d_rays, (size_t)n * 6 * sizeof(double)));
        CK(cudaMalloc(&d_xi, (size_t)n * n_xi * sizeof(double)));
        CK(cudaMalloc(&d_out, (size_t)n * sizeof(BounceRecord)));
        CK(cudaMemcpyAsync(d_rays, rays, (size_t)n * 6 * sizeof(double), cudaMemcpyHostToDevice, c.stream)); // (stream-ordered with the kernel, see rtgpu_trace_primary)
        CK(cudaMemcpyAsync(d_xi, xi, (size_t)n * n_xi * sizeof(double), cudaMemcpyHostToDevice, c.stream));
        k_bounce_hook<<<(unsigned)((n + 127) / 128), 128, 0, c.stream>>>(c.ds, (float)light_bias, d_rays, d_xi, n_xi, (long long)n, d_out);
        c.launches++;
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(c.stream));
        CK(cudaMemcpy(out, d_out, (size_t)n * sizeof(BounceRecord), cudaMemcpyDeviceToHost));
        CK(cudaFree(d_rays)); CK(cudaFree(d_xi)); CK(cudaFree(d_out));
    } catch (CudaError& e) { rthost::set_error(e.msg); return RTGPU_ERR_CUDA; }
    return 0;
}

int rtgpu_generate_rays(const RtgpuCameraDesc* cam, int sqrt_spt, const int32_t* px, const int32_t* sxy, const double* xi, int64_t n,
                        double* out_rays) {
    if (!cam || !px || !sxy || !xi || !out_rays || sqrt_spt < 1 || n < 0) { rthost::set_error("rtgpu_generate_rays: bad argument"); return RTGPU_ERR_INVALID; }
    if (n == 0) return 0;
    { int rc = wait_init(); if (rc) return rc; }
    if (g_devices.empty()) { int rc = rtgpu_init(nullptr, 0); if (rc) return rc; }
    try {
        CK(cudaSetDevice(g_devices[0]));
        int32_t *d_px = nullptr, *d_sxy = nullptr; double *d_xi = nullptr, *d_out = nullptr;
        CK(cudaMalloc(&d_px, (size_t)n * 2 * sizeof(int32_t))); CK(cudaMalloc(&d_sxy, (size_t)n * 2 * sizeof(int32_t)));
        CK(cudaMalloc(&d_xi, (size_t)n * 4 * sizeof(double))); CK(cudaMalloc(&d_out, (size_t)n * 6 * sizeof(double)));
        CK(cudaMemcpy(d_px, px, (size_t)n * 2 * sizeof(int32_t), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_sxy, sxy, (size_t)n * 2 * sizeof(int32_t), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_xi, xi, (size_t)n * 4 * sizeof(double), cudaMemcpyHostToDevice));
        k_generate_rays<<<(unsigned)((n + 255) / 256), 256>>>(make_camera(*cam, sqrt_spt), d_px, d_sxy, d_xi, (long long)n, d_out);
        CK(cudaGetLastError());
        CK(cudaMemcpy(out_rays, d_out, (size_t)n * 6 * sizeof(double), cudaMemcpyDeviceToHost));
        CK(cudaFree(d_px)); CK(cudaFree(d_sxy)); CK(cudaFree(d_xi)); CK(cudaFree(d_out));
    } catch (CudaError& e) { rthost::set_error(e.msg); return RTGPU_ERR_CUDA; }
    return 0;
}

int rtgpu_get_stats(RtgpuScene* scene, RtgpuStats* out) {
    if (!scene || !out) { rthost::set_error("rtgpu_get_stats: null argument"); return RTGPU_ERR_INVALID; }
    try {
        uint64_t total[128] = {0};
        uint64_t launches = 0;
        for (auto& c : scene->devs) {
            CK(cudaSetDevice(c.device));
            unsigned long long h[128];
            CK(cudaDeviceSynchronize());
            check_pool_error(c); // (rtgpu_render_device is asynchronous: a failed pool check of its kernels surfaces here)
            c.collect_spans();
            CK(cudaMemcpy(h, c.d_rays_per_bounce, sizeof h, cudaMemcpyDeviceToHost));
            for (int i = 0; i < 128; i++) total[i] += h[i];
            launches += c.launches;
        }
        scene->stats.rays = scene->stats.rays_traversed = 0;
        for (int i = 0; i < 64; i++) {
            scene->stats.rays_per_bounce[i] = total[i]; scene->stats.rays += total[i];
            scene->stats.rays_traversed_per_bounce[i] = total[64 + i]; scene->stats.rays_traversed += total[64 + i];
        }
        scene->stats.kernel_launches = launches;
        scene->stats.trace_launches = 0; scene->stats.pool_bytes = 0;
        for (auto& c : scene->devs) {
            scene->stats.trace_launches += c.trace_launches;
            for (auto& l : c.lanes) scene->stats.pool_bytes += l.bytes();
        }
        scene->stats.ms_raygen = scene->stats.ms_trace = scene->stats.ms_shade = 0;
        for (auto& c : scene->devs) { scene->stats.ms_raygen += c.ms_family[0]; scene->stats.ms_trace += c.ms_family[1]; scene->stats.ms_shade += c.ms_family[2]; }
        *out = scene->stats;
    } catch (CudaError& e) { rthost::set_error(e.msg); return RTGPU_ERR_CUDA; }
    return 0;
}

} // extern "C"
