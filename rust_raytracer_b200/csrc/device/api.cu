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
#include "kernels.cuh"

namespace rthost { void set_error(const std::string& msg); }

using namespace rt;

namespace {

std::vector<int> g_devices; // selected by rtgpu_init
std::mutex g_mutex;

struct CudaError { std::string msg; };
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
        PathPool* d_pool = nullptr;
        uint32_t capacity = 0;
        int bounces = 0;
        uint32_t carry_group = 0; // batches per tail-merging group the carry arrays were sized for (0 = none)
    };
    std::vector<Lane> lanes;
    char* geom_arena = nullptr; // nodes | prims
    size_t geom_nodes_bytes = 0, geom_bytes = 0, l2_persist_max = 0, l2_window_max = 0;
    cudaEvent_t fork = nullptr;
    cudaStream_t stream = nullptr;
    float4* accum = nullptr; // owned accumulation buffer (rtgpu_render)
    size_t accum_pixels = 0;
    unsigned long long* d_rays_per_bounce = nullptr;
    uint64_t launches = 0;
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

    template <typename T> T* upload(const std::vector<T>& v) {
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
        for (Lane& l : lanes) { if (l.stream) cudaStreamDestroy(l.stream); if (l.done) cudaEventDestroy(l.done); }
        lanes.clear();
        if (fork) cudaEventDestroy(fork);
        fork = nullptr;
        if (stream) cudaStreamDestroy(stream);
        stream = nullptr;
    }
};

} // namespace

struct RtgpuScene {
    HostScene hs;
    std::vector<DeviceCtx> devs;
    RtgpuStats stats{};
};

namespace {

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

// carry_bounce / carry_group: tail merging (enqueue_render); carry_group 0 = off
void ensure_lane(DeviceCtx& c, size_t index, uint32_t capacity, int bounces, uint32_t carry_bounce = 0xFFFFFFFFu, uint32_t carry_group = 0) {
    while (c.lanes.size() <= index) {
        DeviceCtx::Lane l;
        CK(cudaStreamCreateWithFlags(&l.stream, cudaStreamNonBlocking));
        set_l2_window(c, l.stream);
        CK(cudaEventCreateWithFlags(&l.done, cudaEventDisableTiming));
        l.d_pool = c.alloc<PathPool>(1);
        l.pool.carry_bounce = 0xFFFFFFFFu;
        c.lanes.push_back(l);
    }
    DeviceCtx::Lane& l = c.lanes[index];
    if (carry_group == 0) carry_bounce = 0xFFFFFFFFu;
    const bool fits = l.capacity >= capacity && l.bounces >= bounces && l.carry_group >= carry_group;
    if (fits && l.pool.carry_bounce == carry_bounce) return;
    CK(cudaStreamSynchronize(l.stream)); // no kernel of this lane may still read the old descriptor
    PathPool& p = l.pool;
    if (!fits) {
        // (re)allocate; old buffers stay in `allocs` until the scene is destroyed — pools only grow once per size
        capacity = std::max(capacity, l.capacity); bounces = std::max(bounces, l.bounces); carry_group = std::max(carry_group, l.carry_group);
        p.capacity = capacity;
        p.ray_slots = 2u * capacity + 65536u; // unused chunk ends: at most half of a kernel's output in each of the two regions (record_chunk)
        p.rays[0] = c.alloc<PathRay>(p.ray_slots); p.rays[1] = c.alloc<PathRay>(p.ray_slots);
        // worst case of a group: every path of every batch is still alive at the carry bounce
        const uint64_t carry_slots = carry_group ? (uint64_t)carry_group * (2ull * capacity) + 65536ull : 0ull;
        if (carry_slots > 0xFFFFFF00ull) throw CudaError{"tail-merging group too large for 32-bit record indices"};
        p.carry_slots = (uint32_t)carry_slots;
        p.carry_rays = carry_group ? c.alloc<PathRay>(p.carry_slots) : nullptr;
        p.hits = c.alloc<PathHit>(std::max(p.ray_slots, p.carry_slots)); // indexed like the trace region of the bounce's ray array
        for (int h = 0; h < 2; h++)
            for (int k = 0; k < N_CLASSES; k++) p.class_queue[h][k] = c.alloc<uint32_t>(capacity);
        for (int k = 0; k < N_CLASSES; k++) p.carry_queue[k] = carry_group ? c.alloc<uint32_t>((size_t)carry_group * capacity) : nullptr;
        p.counters = c.alloc<uint32_t>((size_t)(bounces + 2) * RT_COUNTERS_PER_BOUNCE);
        l.capacity = capacity; l.bounces = bounces; l.carry_group = carry_group;
    }
    p.carry_bounce = carry_bounce;
    CK(cudaMemcpy(l.d_pool, &p, sizeof p, cudaMemcpyHostToDevice));
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
    k_shade<CLASS, RT_SHADE_MIN_CTAS><<<grid, 128, 0, l.stream>>>(c.ds, l.pool, rc, cam, bounce, accum, bypass);
    c.launches++;
}

// RTGPU_TRACE_VARIANT: 0 = one ray per thread (v1 kernel), otherwise k_trace_fast<FETCH_THRESH, POSTPONE_PCT, MIN_CTAS, COLD>
void launch_trace_fast(DeviceCtx& c, DeviceCtx::Lane& l, bool cold, int variant, int grid, const RenderConst& rc, uint32_t bounce) {
#define RT_LAUNCH_TF(FT, PP, MB, CD, PF) k_trace_fast<FT, PP, MB, CD, PF><<<grid, RT_TRACE_THREADS, 0, l.stream>>>(c.ds, l.pool, c.d_ds, l.d_pool, rc, bounce)
    if (cold) { // f64 call paths present: 128 registers
        switch (variant) {
        default: RT_LAUNCH_TF(20, 0, 4, true, 0); break;
        case 2: RT_LAUNCH_TF(16, 0, 4, true, 0); break;
        }
        return;
    }
    switch (variant) {
    default:
    case 4: RT_LAUNCH_TF(20, 0, 8, false, 0); break;
    case 1: RT_LAUNCH_TF(24, 0, 8, false, 0); break;
    case 2: RT_LAUNCH_TF(16, 0, 8, false, 0); break;
    case 5: RT_LAUNCH_TF(24, 25, 8, false, 0); break;
    case 6: RT_LAUNCH_TF(20, 0, 8, false, 1); break;
    case 7: RT_LAUNCH_TF(20, 0, 8, false, 2); break;
    }
#undef RT_LAUNCH_TF
}

// Enqueue the render of global samples [s_begin, s_end) of every pixel into `accum` (fp32 sums).
// Work is ordered after everything already in the context's stream (or `stream_override`) and
// that stream waits for it: fork to the lanes, join back.
void enqueue_render(RtgpuScene* scene, DeviceCtx& c, const RtgpuCameraDesc& cam, const RtgpuRenderParams& prm, int s_begin, int s_end,
                    float4* accum, cudaStream_t stream_override) {
    CK(cudaSetDevice(c.device));
    cudaStream_t main_stream = stream_override ? stream_override : c.stream;
    const uint64_t npix = (uint64_t)cam.width * cam.height;
    const uint64_t total = npix * (uint64_t)(s_end - s_begin);
    if (total == 0) return;
    // batch size: whole images of samples, about RTGPU_BATCH_MPATHS Mi paths (measured: tools/tune_trace.py, profiles/)
    const uint64_t batch_paths = (uint64_t)std::max(1, env_int("RTGPU_BATCH_MPATHS", 24)) << 20;
    uint64_t samples_per_batch = std::max<uint64_t>(1, batch_paths / npix);
    uint32_t capacity = (uint32_t)std::min<uint64_t>(total, samples_per_batch * npix);
    const uint64_t n_batches = (total + capacity - 1) / capacity;
    const bool exact = (prm.flags & RTGPU_FLAG_EXACT) != 0;
    const bool profile = (prm.flags & RTGPU_FLAG_PROFILE) != 0;
    // profiling times every kernel alone (one lane); otherwise a few batches are in flight
    const size_t n_lanes = profile ? 1 : (size_t)std::min<uint64_t>(n_batches, (uint64_t)std::max(1, env_int("RTGPU_LANES", 4)));
    // Tail merging: a lane runs its batches in groups of `tail_group`; the first ones stop after bounce tail_bounce - 1
    // and leave their surviving rays in the lane's carry arrays, the last one runs the remaining bounces on the
    // survivors of the whole group. The bounces past ~6 hold a few per cent of a batch's rays in launches whose
    // duration is the latency of the longest ray; merged, the same launches carry tail_group times the rays.
    // Same paths, same Philox streams (addressed by pixel, sample, bounce): the image changes by fp32 summation order only.
    const uint64_t batches_per_lane = (n_batches + n_lanes - 1) / n_lanes;
    uint32_t tail_group = (uint32_t)std::max(1, env_int("RTGPU_TAIL_GROUP", 3));
    const int tail_bounce = env_int("RTGPU_TAIL_BOUNCE", 4);
    if (exact || tail_bounce < 1 || tail_bounce >= prm.max_depth || batches_per_lane < 2) tail_group = 1;
    tail_group = (uint32_t)std::min<uint64_t>(tail_group, batches_per_lane);
    {
        // the carry arrays are sized for the worst case (every path of every batch of a group still alive, plus the unused
        // chunk ends): 304 + 216 x group bytes per path of capacity and lane (368 without). Shrink the group until the pools of all
        // lanes stay below 60 % of the device's memory (B200, 24 Mi-path batches, 4 lanes: group 3 = 94 GB).
        size_t free_b = 0, total_b = 0;
        CK(cudaMemGetInfo(&free_b, &total_b));
        auto need = [&](uint32_t g) { return (double)n_lanes * (double)capacity * (g > 1 ? 304.0 + 216.0 * g : 368.0); };
        while (tail_group > 1 && need(tail_group) > (double)env_int("RTGPU_POOL_MEM_PCT", 60) * 0.01 * (double)total_b) tail_group--;
    }
    const bool merge = tail_group > 1;
    for (size_t i = 0; i < n_lanes; i++) ensure_lane(c, i, capacity, prm.max_depth, (uint32_t)tail_bounce, merge ? tail_group : 0u);
    const CameraDev camd = make_camera(cam, prm.sqrt_spt);
    RenderConst rc{prm.seed, (float)prm.light_bias, prm.max_depth};
    const int variant = env_int("RTGPU_TRACE_VARIANT", 4);
    const int grid_trace = c.sm_count * (variant == 0 ? 8 : env_int("RTGPU_TRACE_CTAS_PER_SM", 4)), grid_shade = c.sm_count * env_int("RTGPU_SHADE_CTAS_PER_SM", RT_SHADE_MIN_CTAS);
    const uint32_t class_mask = scene->hs.class_mask;
    // rays whose closest hit the broadcast pre-test already knows (BVH bounds out of reach) go straight to
    // the shade queue of the next bounce; constant media would still compete, and the exact mode re-traces in f64
    const uint32_t bypass = !exact && scene->hs.volumes.empty() && env_int("RTGPU_BYPASS", 1) != 0 ? 1u : 0u;
    CK(cudaEventRecord(c.fork, main_stream));
    for (size_t i = 0; i < n_lanes; i++) CK(cudaStreamWaitEvent(c.lanes[i].stream, c.fork, 0));
    size_t span = 0;
    uint64_t batch = 0;
    for (uint64_t first = 0; first < total; first += capacity, batch++) {
        DeviceCtx::Lane& l = c.lanes[batch % n_lanes];
        const uint32_t n = (uint32_t)std::min<uint64_t>(capacity, total - first);
        // position of the batch in its lane's current group
        const uint64_t in_lane = batch / n_lanes, lane_batches = (n_batches - (batch % n_lanes) + n_lanes - 1) / n_lanes;
        const bool group_first = !merge || in_lane % tail_group == 0;
        const bool group_last = !merge || in_lane % tail_group == tail_group - 1 || in_lane + 1 == lane_batches;
        const int b_reset = group_first ? prm.max_depth + 2 : tail_bounce;   // counter rows >= tail_bounce accumulate over the group
        const int b_end = group_last ? prm.max_depth : tail_bounce;
        CK(cudaMemsetAsync(l.pool.counters, 0, (size_t)b_reset * RT_COUNTERS_PER_BOUNCE * sizeof(uint32_t), l.stream));
        if (profile) span = c.span_begin(0, l.stream);
        k_raygen<<<c.sm_count * 4, 256, 0, l.stream>>>(c.ds, l.pool, camd, rc, first, n, (uint32_t)s_begin, bypass);
        if (profile) c.span_end(span, l.stream);
        c.launches++;
        for (int b = 0; b < b_end; b++) {
            if (profile) span = c.span_begin(1, l.stream);
            if (exact) k_trace<true><<<grid_trace, 128, 0, l.stream>>>(c.ds, l.pool, rc, (uint32_t)b);
            else if (variant == 0) k_trace<false><<<grid_trace, 128, 0, l.stream>>>(c.ds, l.pool, rc, (uint32_t)b);
            else launch_trace_fast(c, l, scene->hs.trace_cold, variant, grid_trace, rc, (uint32_t)b);
            if (profile) { c.span_end(span, l.stream); span = c.span_begin(2, l.stream); }
            c.launches++;
            launch_shade<CLASS_LAMBERTIAN>(c, l, class_mask, grid_shade, rc, camd, (uint32_t)b, accum, bypass);
            launch_shade<CLASS_METAL>(c, l, class_mask, grid_shade, rc, camd, (uint32_t)b, accum, bypass);
            launch_shade<CLASS_DIELECTRIC>(c, l, class_mask, grid_shade, rc, camd, (uint32_t)b, accum, bypass);
            launch_shade<CLASS_GLOSSY>(c, l, class_mask, grid_shade, rc, camd, (uint32_t)b, accum, bypass);
            launch_shade<CLASS_ISOTROPIC>(c, l, class_mask, grid_shade, rc, camd, (uint32_t)b, accum, bypass);
            launch_shade<CLASS_EMISSIVE>(c, l, class_mask, grid_shade, rc, camd, (uint32_t)b, accum, bypass);
            if (profile) c.span_end(span, l.stream);
        }
        // rows of this batch alone, and (last batch of a group) the merged rows of the whole group
        k_stats<<<1, 64, 0, l.stream>>>(l.pool.counters, 0, std::min(b_end, 64), c.d_rays_per_bounce);
        c.launches++;
    }
    for (size_t i = 0; i < n_lanes; i++) {
        CK(cudaEventRecord(c.lanes[i].done, c.lanes[i].stream));
        CK(cudaStreamWaitEvent(main_stream, c.lanes[i].done, 0));
    }
    CK(cudaGetLastError());
    { std::lock_guard<std::mutex> lock(g_mutex); scene->stats.paths += total; } // (one host thread per device)
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
    if (p->replicas < 1 || p->sqrt_spt < 1 || p->max_depth < 0 || p->max_depth > 62) { rthost::set_error("params: replicas/sqrt_spt must be >= 1 and 0 <= max_depth <= 62"); return RTGPU_ERR_INVALID; }
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
    int rc = flatten_scene(*desc, sc->hs, err); // host only: overlaps with a pending rtgpu_init_async
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
        // shard the sample range over the devices: contiguous blocks
        const int total = s_end - s_begin;
        for_each_device(n_dev, [&](int d) {
            DeviceCtx& c = scene->devs[d];
            CK(cudaSetDevice(c.device));
            if (c.accum_pixels < npix) { c.accum = c.alloc<float4>(npix); c.accum_pixels = npix; }
            CK(cudaMemsetAsync(c.accum, 0, (size_t)npix * sizeof(float4), c.stream));
            const int b = s_begin + (int)((int64_t)total * d / n_dev), e = s_begin + (int)((int64_t)total * (d + 1) / n_dev);
            if (e > b) enqueue_render(scene, c, *cam, *params, b, e, c.accum, nullptr);
            CK(cudaStreamSynchronize(c.stream));
        });
        // reduce + resolve on device 0 (peer reads over NVLink when several devices took part)
        DeviceCtx& c0 = scene->devs[0];
        CK(cudaSetDevice(c0.device));
        std::vector<const float4*> ptrs;
        for (int d = 0; d < n_dev; d++) {
            if (d > 0) {
                int can = 0;
                CK(cudaDeviceCanAccessPeer(&can, c0.device, scene->devs[d].device));
                if (!can) throw CudaError{"peer access between the selected devices is not available"};
                cudaError_t pe = cudaDeviceEnablePeerAccess(scene->devs[d].device, 0);
                if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled) throw CudaError{std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(pe)};
                cudaGetLastError();
            }
            ptrs.push_back(scene->devs[d].accum);
        }
        resolve_add_to_host(c0.device, ptrs, npix, 1.0 / (double)spp, c0.stream, out_rgba);
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

int rtgpu_trace_primary(RtgpuScene* scene, const double* rays, int64_t n, RtgpuHit* out_hits) {
    if (!scene || !rays || !out_hits || n < 0) { rthost::set_error("rtgpu_trace_primary: bad argument"); return RTGPU_ERR_INVALID; }
    if (n == 0) return 0;
    try {
        DeviceCtx& c = scene->devs[0];
        CK(cudaSetDevice(c.device));
        double* d_rays = nullptr; RtgpuHit* d_hits = nullptr;
        CK(cudaMalloc(&d_rays, (size_t)n * 6 * sizeof(double)));
        CK(cudaMalloc(&d_hits, (size_t)n * sizeof(RtgpuHit)));
        CK(cudaMemcpy(d_rays, rays, (size_t)n * 6 * sizeof(double), cudaMemcpyHostToDevice));
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
        RenderConst rc{1, 0.f, 1};
        k_pretest_pool<<<(n32 + 127) / 128, 128, 0, l.stream>>>(c.ds, l.pool, n32); // what k_raygen / k_shade do when they write a ray
        const int time_hook = env_int("RTGPU_TIME_HOOK", 0); // tools/sort_experiment.py: device time of the trace launch alone
        cudaEvent_t ev0 = nullptr, ev1 = nullptr;
        if (time_hook) { CK(cudaEventCreate(&ev0)); CK(cudaEventCreate(&ev1)); CK(cudaEventRecord(ev0, l.stream)); }
        launch_trace_fast(c, l, scene->hs.trace_cold, env_int("RTGPU_TRACE_VARIANT", 4), c.sm_count * 8, rc, 0);
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
        CK(cudaMemcpy(d_rays, rays, (size_t)n * 6 * sizeof(double), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_xi, xi, (size_t)n * n_xi * sizeof(double), cudaMemcpyHostToDevice));
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
        scene->stats.ms_raygen = scene->stats.ms_trace = scene->stats.ms_shade = 0;
        for (auto& c : scene->devs) { scene->stats.ms_raygen += c.ms_family[0]; scene->stats.ms_trace += c.ms_family[1]; scene->stats.ms_shade += c.ms_family[2]; }
        *out = scene->stats;
    } catch (CudaError& e) { rthost::set_error(e.msg); return RTGPU_ERR_CUDA; }
    return 0;
}

} // extern "C"
