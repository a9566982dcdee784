// build_gpu_host.cuh — host driver of the GPU acceleration-structure build (build_gpu.cuh); included by api.cu inside its
// anonymous namespace (uses CK / CudaError). Same contract as build_wide_bvh (bvh.h): padded boxes in, wide nodes +
// leaf order out. Runs on the current device's default stream, synchronously; the temporaries are freed before it returns.
#pragma once

struct GpuBuildTimes { double ms_total = 0, ms_device = 0; int levels = 0; };
GpuBuildTimes g_last_gpu_build;

void build_wide_bvh_gpu(const std::vector<PrimBox>& boxes, WideBvh& out) {
    using namespace gpubuild;
    out = WideBvh();
    const uint32_t n = (uint32_t)boxes.size();
    if (n < 2) { build_wide_bvh(boxes, out); return; } // nothing to build a radix tree over
    const auto t_begin = std::chrono::steady_clock::now();
    // the cube around the box centres (host: one pass over 24 bytes per primitive)
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (const PrimBox& b : boxes)
        for (int a = 0; a < 3; a++) { const float c = 0.5f * (b.lo[a] + b.hi[a]); lo[a] = std::min(lo[a], c); hi[a] = std::max(hi[a], c); }
    const float extent = std::max(std::max(hi[0] - lo[0], hi[1] - lo[1]), std::max(hi[2] - lo[2], 1e-30f));
    // ONE allocation for every temporary (17 cudaMalloc / cudaFree pairs cost ~10 ms): sizes first, then carve
    // wide nodes <= binary internal nodes (the DP may open even a 2-primitive subtree into a wide node): n - 1, never more
    const uint32_t n_all = 2 * n - 1, max_nodes = n + 16;
    size_t sort_bytes = 0;
    CK(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (uint64_t*)nullptr, (uint64_t*)nullptr, (uint32_t*)nullptr, (uint32_t*)nullptr, (int)n, 0, 63));
    size_t arena_bytes = 0;
    auto reserve = [&](size_t bytes) { const size_t at = arena_bytes; arena_bytes += (bytes + 255) & ~(size_t)255; return at; };
    const size_t o_boxes = reserve((size_t)n * sizeof(PrimBox)), o_keys0 = reserve((size_t)n * 8), o_keys1 = reserve((size_t)n * 8),
                 o_idx0 = reserve((size_t)n * 4), o_idx1 = reserve((size_t)n * 4), o_box = reserve((size_t)n_all * sizeof(Box6)),
                 o_left = reserve((size_t)n * 4), o_right = reserve((size_t)n * 4), o_parent = reserve((size_t)n_all * 4),
                 o_first = reserve((size_t)n_all * 4), o_count = reserve((size_t)n_all * 4), o_visits = reserve((size_t)n * 4),
                 o_dp = reserve((size_t)n_all * sizeof(Dp)), o_nodes = reserve((size_t)max_nodes * sizeof(BvhNode)), o_order = reserve((size_t)n * 4),
                 o_counters = reserve(8 * 4), o_tasks0 = reserve((size_t)max_nodes * sizeof(CollapseTask)),
                 o_tasks1 = reserve((size_t)max_nodes * sizeof(CollapseTask)), o_sort = reserve(sort_bytes);
    char* arena = nullptr;
    CK(cudaMalloc((void**)&arena, arena_bytes));
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    struct Guard { char* p; cudaEvent_t& a; cudaEvent_t& b; ~Guard() { cudaFree(p); if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); } } guard{arena, e0, e1};
    PrimBox* d_boxes = (PrimBox*)(arena + o_boxes);
    uint64_t* keys[2] = {(uint64_t*)(arena + o_keys0), (uint64_t*)(arena + o_keys1)};
    uint32_t* idx[2] = {(uint32_t*)(arena + o_idx0), (uint32_t*)(arena + o_idx1)};
    Tree2 t{};
    t.n = n;
    t.box = (Box6*)(arena + o_box);
    t.left = (int32_t*)(arena + o_left); t.right = (int32_t*)(arena + o_right);
    t.parent = (int32_t*)(arena + o_parent);
    t.first = (uint32_t*)(arena + o_first); t.count = (uint32_t*)(arena + o_count);
    t.visits = (uint32_t*)(arena + o_visits);
    Dp* dp = (Dp*)(arena + o_dp);
    BvhNode* d_nodes = (BvhNode*)(arena + o_nodes);
    uint32_t* d_order = (uint32_t*)(arena + o_order);
    uint32_t* counters = (uint32_t*)(arena + o_counters); // [0] nodes, [1] prims, [2] overflow, [4], [5] task counts (ping-pong)
    CollapseTask* tasks[2] = {(CollapseTask*)(arena + o_tasks0), (CollapseTask*)(arena + o_tasks1)};
    void* sort_tmp = arena + o_sort;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    CK(cudaMemcpy(d_boxes, boxes.data(), (size_t)n * sizeof(PrimBox), cudaMemcpyHostToDevice));
    CK(cudaEventRecord(e0, 0));
    const unsigned blocks = (n + 255) / 256;
    k_morton<<<blocks, 256>>>(d_boxes, n, lo[0], lo[1], lo[2], 1.0f / extent, keys[0], idx[0]);
    CK(cub::DeviceRadixSort::SortPairs(sort_tmp, sort_bytes, keys[0], keys[1], idx[0], idx[1], (int)n, 0, 63));
    CK(cudaMemsetAsync(t.visits, 0, (size_t)n * 4, 0));
    k_radix_tree<<<blocks, 256>>>(keys[1], t);
    // c_prim / c_node of the collapse: the host builder's measured 1.5 (bvh.cpp), RTGPU_BVH_CPRIM overrides both builders
    float c_prim = 1.5f;
    if (const char* e = std::getenv("RTGPU_BVH_CPRIM")) c_prim = (float)std::atof(e);
    k_refit_dp<<<blocks, 256>>>(d_boxes, idx[1], t, dp, c_prim, 1.0f);
    CK(cudaGetLastError());
    // collapse, level by level from the root (binary node 0 -> wide node 0)
    const uint32_t init[8] = {1u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
    CK(cudaMemcpyAsync(counters, init, sizeof init, cudaMemcpyHostToDevice, 0));
    const CollapseTask root{0, 0u};
    CK(cudaMemcpyAsync(tasks[0], &root, sizeof root, cudaMemcpyHostToDevice, 0));
    CollapseCtx ctx{t, dp, idx[1], d_nodes, d_order, counters, max_nodes};
    uint32_t n_tasks = 1;
    int level = 0;
    while (n_tasks > 0) {
        const int cur = level & 1, nxt = cur ^ 1;
        CK(cudaMemsetAsync(counters + 4 + nxt, 0, 4, 0));
        k_collapse<<<(n_tasks + 127) / 128, 128>>>(ctx, tasks[cur], n_tasks, tasks[nxt], counters + 4 + nxt);
        CK(cudaMemcpy(&n_tasks, counters + 4 + nxt, 4, cudaMemcpyDeviceToHost)); // (the default stream orders this after the kernel)
        level++;
        if (level > 4 * RT_STACK_SIZE) throw CudaError{"GPU BVH build: collapse does not terminate"};
    }
    CK(cudaEventRecord(e1, 0));
    uint32_t h_counters[4];
    CK(cudaMemcpy(h_counters, counters, sizeof h_counters, cudaMemcpyDeviceToHost));
    if (h_counters[2] != 0 || h_counters[0] > max_nodes) throw CudaError{"GPU BVH build: node arena overflow"};
    if (h_counters[1] != n) throw CudaError{"GPU BVH build: " + std::to_string(h_counters[1]) + " of " + std::to_string(n) + " primitives placed"};
    out.nodes.resize(h_counters[0]);
    out.prim_order.resize(n);
    CK(cudaMemcpy(out.nodes.data(), d_nodes, (size_t)h_counters[0] * sizeof(BvhNode), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(out.prim_order.data(), d_order, (size_t)n * 4, cudaMemcpyDeviceToHost));
    out.max_depth = level;
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    g_last_gpu_build.ms_device = ms; g_last_gpu_build.levels = level;
    g_last_gpu_build.ms_total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count();
    if (std::getenv("RTGPU_BUILD_TIMING"))
        std::fprintf(stderr, "[rtgpu build]   bvh (GPU): %u prims -> %u wide nodes, %d levels, device %.2f ms, with copies %.1f ms\n", n, h_counters[0], level,
                     ms, g_last_gpu_build.ms_total);
}

// the hook flatten_scene calls (bvh.h: g_wide_bvh_builder): CUDA bring-up is joined here, after the host-side walk
int (*g_wait_init_fn)() = nullptr; // = wait_init (defined further down in api.cu), set by rtgpu_scene_create
// Below RTGPU_BUILD_GPU_MIN primitives (default 131 072) the host's binned-SAH builder is used: there it costs 20-40 ms, and
// its trees are better where primitive sizes and densities are mixed — the default scene (Suzanne + ~440 small spheres,
// 16 k primitives) traces 16 % faster on the SAH tree, the Suzanne-only scenes 1 %; the 1.0 M-triangle dragon traces the
// same on both and builds in 3-6 ms instead of ~290 ms (profiles/r02_tune5_gpu_vs_host_tree_all_scenes.log).
void build_wide_bvh_gpu_entry(const std::vector<PrimBox>& boxes, WideBvh& out) {
    size_t gpu_min = 131072;
    if (const char* e = std::getenv("RTGPU_BUILD_GPU_MIN")) gpu_min = (size_t)std::max(0, std::atoi(e));
    if (boxes.size() < 2 || boxes.size() < gpu_min) { build_wide_bvh(boxes, out); return; }
    const int rc = g_wait_init_fn ? g_wait_init_fn() : RTGPU_ERR_NO_DEVICE;
    if (rc != 0 || g_devices.empty()) throw CudaError{"no CUDA device available (this backend has no CPU fallback)", rc ? rc : RTGPU_ERR_NO_DEVICE};
    CK(cudaSetDevice(g_devices[0]));
    build_wide_bvh_gpu(boxes, out);
    // a radix tree follows the bits of the codes: a pathological distribution (extents spaced exponentially) can make it
    // deeper than the traversal stack — the SAH builder's trees are not, so such a scene is rebuilt there instead of refused
    if (out.max_depth > RT_STACK_SIZE) build_wide_bvh(boxes, out);
}
