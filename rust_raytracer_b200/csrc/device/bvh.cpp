// bvh.cpp — see bvh.h. Host code, runs once per scene upload.
#include "bvh.h"

#include <algorithm>
#include <atomic>
#include <thread>
#include <cfloat>
#include <cmath>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace rt {
namespace {

struct Box {
    float lo[3], hi[3];
    void reset() { for (int i = 0; i < 3; i++) { lo[i] = FLT_MAX; hi[i] = -FLT_MAX; } }
    void grow(const float* l, const float* h) { for (int i = 0; i < 3; i++) { lo[i] = std::min(lo[i], l[i]); hi[i] = std::max(hi[i], h[i]); } }
    void grow_pt(const float* p) { for (int i = 0; i < 3; i++) { lo[i] = std::min(lo[i], p[i]); hi[i] = std::max(hi[i], p[i]); } }
    float area() const {
        float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        if (dx < 0 || dy < 0 || dz < 0) return 0.f;
        return 2.f * (dx * dy + dy * dz + dz * dx);
    }
};

struct Node2 { // binary node
    Box box;
    int32_t left = -1, right = -1; // children (internal)
    uint32_t first = 0, count = 0; // leaf range into `order`
    bool leaf() const { return left < 0; }
};

// host cores this process may count on: under torchrun (one process per GPU) every rank runs the same set-up at the same
// time, so the cores are divided by LOCAL_WORLD_SIZE (8 ranks x 16 threads on 32 cores made the "Ready" phase slower at N = 8)
static inline unsigned rt_host_cores() {
    unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    if (const char* e = std::getenv("LOCAL_WORLD_SIZE")) { const int w = std::atoi(e); if (w > 1) hw = std::max(2u, hw / (unsigned)w); }
    return hw;
}
static const int BINS = 16;
static const int MAX_LEAF = 3; // the wide node encodes at most 3 primitives per leaf slot

// RTGPU_BUILD_THREADS / RTGPU_BUILD_PAR_MIN: test knobs (the tree must not depend on either)
static int env_or(const char* name, int dflt) { const char* e = std::getenv(name); return e ? std::atoi(e) : dflt; }
static int build_threads() { return std::max(1, std::min(64, env_or("RTGPU_BUILD_THREADS", (int)std::min(32u, rt_host_cores())))); }

struct Builder2 {
    const std::vector<PrimBox>& boxes;
    std::vector<uint32_t> order;
    std::vector<float> cent; // 3 per prim
    std::vector<Node2> nodes;

    explicit Builder2(const std::vector<PrimBox>& b) : boxes(b) {
        order.resize(b.size());
        cent.resize(b.size() * 3);
        for (size_t i = 0; i < b.size(); i++) {
            order[i] = (uint32_t)i;
            for (int a = 0; a < 3; a++) cent[i * 3 + a] = 0.5f * (b[i].lo[a] + b[i].hi[a]);
        }
        nodes.resize(b.size() * 2 + 1); // a subtree over c primitives owns the index range [base, base + 2c - 1)
        max_threads = build_threads();
        par_min = (uint32_t)std::max(2, env_or("RTGPU_BUILD_PAR_MIN", 32768));
    }

    // Subtrees are independent (disjoint ranges of `order` and of `nodes`), so large ones are built on
    // their own threads; the result does not depend on the thread count.
    std::atomic<int> live_threads{1};
    int max_threads = 1;
    uint32_t par_min = 32768;

    int build(uint32_t first, uint32_t count, int idx = 0) {
        Box box, cbox;
        box.reset(); cbox.reset();
        for (uint32_t i = first; i < first + count; i++) {
            uint32_t p = order[i];
            box.grow(boxes[p].lo, boxes[p].hi);
            cbox.grow_pt(&cent[p * 3]);
        }
        nodes[idx].box = box;
        if (count <= 1) { nodes[idx].first = first; nodes[idx].count = count; return idx; }

        // binned SAH over the three axes
        int best_axis = -1, best_split = -1;
        float best_cost = FLT_MAX;
        for (int a = 0; a < 3; a++) {
            float cmin = cbox.lo[a], cmax = cbox.hi[a];
            if (!(cmax > cmin)) continue;
            Box bb[BINS]; uint32_t bc[BINS];
            for (int b = 0; b < BINS; b++) { bb[b].reset(); bc[b] = 0; }
            float scale = BINS / (cmax - cmin);
            for (uint32_t i = first; i < first + count; i++) {
                uint32_t p = order[i];
                int b = std::min(BINS - 1, std::max(0, (int)((cent[p * 3 + a] - cmin) * scale)));
                bb[b].grow(boxes[p].lo, boxes[p].hi); bc[b]++;
            }
            float right_area[BINS]; uint32_t right_cnt[BINS];
            Box acc; acc.reset(); uint32_t n = 0;
            for (int b = BINS - 1; b > 0; b--) { if (bc[b]) acc.grow(bb[b].lo, bb[b].hi); n += bc[b]; right_area[b] = acc.area(); right_cnt[b] = n; }
            acc.reset(); n = 0;
            for (int b = 0; b < BINS - 1; b++) {
                if (bc[b]) acc.grow(bb[b].lo, bb[b].hi);
                n += bc[b];
                if (n == 0 || right_cnt[b + 1] == 0) continue;
                float cost = acc.area() * n + right_area[b + 1] * right_cnt[b + 1];
                if (cost < best_cost) { best_cost = cost; best_axis = a; best_split = b; }
            }
        }
        float leaf_cost = box.area() * count;
        if (count <= (uint32_t)MAX_LEAF && (best_axis < 0 || best_cost + box.area() * 0.5f >= leaf_cost)) {
            nodes[idx].first = first; nodes[idx].count = count;
            return idx;
        }
        uint32_t mid;
        if (best_axis >= 0) {
            float cmin = cbox.lo[best_axis], scale = BINS / (cbox.hi[best_axis] - cmin);
            auto it = std::partition(order.begin() + first, order.begin() + first + count, [&](uint32_t p) {
                int b = std::min(BINS - 1, std::max(0, (int)((cent[p * 3 + best_axis] - cmin) * scale)));
                return b <= best_split;
            });
            mid = (uint32_t)(it - order.begin());
            if (mid == first || mid == first + count) mid = first + count / 2;
        } else mid = first + count / 2; // all centroids coincide
        const uint32_t lc = mid - first, rc = first + count - mid;
        const int li = idx + 1, ri = idx + 1 + (int)(2 * lc - 1);
        if (count >= par_min && live_threads.load(std::memory_order_relaxed) < max_threads) {
            live_threads.fetch_add(1, std::memory_order_relaxed);
            std::thread th([this, first, lc, li] { build(first, lc, li); live_threads.fetch_sub(1, std::memory_order_relaxed); });
            build(mid, rc, ri);
            th.join();
        } else {
            build(first, lc, li);
            build(mid, rc, ri);
        }
        nodes[idx].left = li; nodes[idx].right = ri;
        nodes[idx].first = first; nodes[idx].count = count; // range of the whole subtree (WideBuilder job split)
        return idx;
    }
};

struct WideBuilder {
    const Builder2& b2;
    WideBvh& out;

    bool greedy = false;
    struct Task { int node2; uint32_t wide_index; int depth; };
    struct Arena { std::vector<BvhNode> nodes; std::vector<uint32_t> prims; int max_depth = 0; };

    WideBuilder(const Builder2& b, WideBvh& o) : b2(b), out(o) {}

    // ---- which binary nodes become wide nodes, which leaf slots ------------------------------------
    // SAH-optimal collapse (Ylitie, Karras, Laine 2017, section 3.1) by dynamic programming over the
    // binary tree: c[i-1] = C(n, i), the cheapest way to represent the subtree of n as at most i children
    // of some wide node:
    //   C(n, 1) = min(A_n P_n c_prim  [one leaf slot, P_n <= 3],  A_n c_node + C_dist(n, 8)  [a wide node])
    //   C(n, i) = min(C_dist(n, i), C(n, i - 1)),   C_dist(n, j) = min_k C(left, k) + C(right, j - k)
    // The greedy collapse this replaces (open the child with the largest area until there are 8) left
    // 41 % of the dragon's wide nodes with two occupied slots: every visit pays the full 8-wide test.
    struct Dp { float c[7]; uint8_t split[7]; uint8_t leaf; }; // split[j-2] = k of C_dist(n, j), j = 2..8
    std::vector<Dp> dp;
    // c_prim / c_node = 1.5: measured on B200 (profiles/r01_tune16_bvh_collapse.log). The paper's 0.3 assumes a cheap
    // triangle test; here the leaf loop runs at 2.6 of 32 lanes, so a primitive costs more than a share of a node test.
    float c_prim = 1.5f, c_node = 1.0f;

    void dp_node(int idx) {
        const Node2& n = b2.nodes[idx];
        Dp& d = dp[(size_t)idx];
        const float A = n.box.area();
        if (n.leaf()) {
            const float c = A * (float)n.count * c_prim;
            for (int i = 0; i < 7; i++) { d.c[i] = c; d.split[i] = 0; }
            d.leaf = 1;
            return;
        }
        const Dp& L = dp[(size_t)n.left];
        const Dp& R = dp[(size_t)n.right];
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
        const float c_leaf = n.count <= (uint32_t)MAX_LEAF ? A * (float)n.count * c_prim : FLT_MAX;
        const float c_int = dist[8] + A * c_node;
        d.leaf = c_leaf <= c_int ? 1 : 0;
        d.c[0] = std::min(c_leaf, c_int);
        for (int i = 2; i <= 7; i++) d.c[i - 1] = std::min(dist[i], d.c[i - 2]);
    }

    // bottom-up over the index range of a subtree: children have larger indices than their parent
    void dp_range(int root) {
        const Node2& r = b2.nodes[root];
        const int end = root + (r.leaf() ? 1 : 2 * (int)r.count - 1);
        for (int i = end - 1; i >= root; i--)
            if (b2.nodes[(size_t)i].count > 0) dp_node(i);
    }

    void run_dp(int root2, uint32_t job_prims) {
        dp.resize(b2.nodes.size());
        std::vector<int> jobs, top, st;
        st.push_back(root2);
        while (!st.empty()) {
            const int i = st.back(); st.pop_back();
            const Node2& n = b2.nodes[(size_t)i];
            if (n.leaf() || n.count <= job_prims) { jobs.push_back(i); continue; }
            top.push_back(i);
            st.push_back(n.left); st.push_back(n.right);
        }
        const unsigned n_thr = jobs.size() < 2 ? 1u : (unsigned)build_threads();
        if (n_thr == 1) for (int j : jobs) dp_range(j);
        else {
            std::atomic<size_t> next{0};
            std::vector<std::thread> pool;
            for (unsigned k = 0; k < n_thr; k++)
                pool.emplace_back([&] { for (size_t i = next.fetch_add(1); i < jobs.size(); i = next.fetch_add(1)) dp_range(jobs[i]); });
            for (auto& th : pool) th.join();
        }
        std::sort(top.begin(), top.end());
        for (size_t i = top.size(); i-- > 0;) dp_node(top[i]);
    }

    // the children of the wide node made from binary node `idx`: follow the DP's decisions
    void collect(int idx, int j, int* kids, bool* kid_leaf, int& nk) const {
        const Node2& n = b2.nodes[(size_t)idx];
        if (n.leaf()) { kids[nk] = idx; kid_leaf[nk++] = true; return; }
        const Dp& d = dp[(size_t)idx];
        if (j == 1) { kids[nk] = idx; kid_leaf[nk++] = d.leaf != 0; return; }
        const int k = d.split[j - 2];
        const float dist = dp[(size_t)n.left].c[k - 1] + dp[(size_t)n.right].c[j - k - 1];
        if (j == 8 || dist < d.c[j - 2]) { collect(n.left, k, kids, kid_leaf, nk); collect(n.right, j - k, kids, kid_leaf, nk); }
        else collect(idx, j - 1, kids, kid_leaf, nk);
    }

    // The top of the tree is collapsed sequentially into the global arena; every subtree with at most
    // `job_prims` primitives becomes a job collapsed into its own arena (in parallel), appended
    // afterwards with its indices rebased. The split depends on the tree only, not on the thread count.
    void run(int root2) {
        Arena G;
        G.nodes.emplace_back();
        const uint32_t total = b2.nodes[root2].count;
        const uint32_t job_prims = std::max<uint32_t>((uint32_t)std::max(1, env_or("RTGPU_BUILD_JOB_MIN", 4096)), total / 128u);
        if (const char* e = std::getenv("RTGPU_BVH_CPRIM")) c_prim = (float)std::atof(e);
        greedy = env_or("RTGPU_BVH_GREEDY", 0) != 0;
        if (!greedy) run_dp(root2, job_prims);
        struct Job { Task task; Arena arena; BvhNode root; };
        std::vector<Job> jobs;
        std::vector<Task> stack;
        stack.push_back({root2, 0, 1});
        while (!stack.empty()) {
            Task t = stack.back(); stack.pop_back();
            if (t.wide_index != 0 && b2.nodes[t.node2].count <= job_prims) { jobs.emplace_back(); jobs.back().task = t; continue; }
            G.max_depth = std::max(G.max_depth, t.depth);
            emit(t, G, nullptr, stack);
        }
        auto run_job = [this](Job& j) {
            std::vector<Task> st;
            j.arena.max_depth = std::max(j.arena.max_depth, j.task.depth);
            emit(j.task, j.arena, &j.root, st);
            while (!st.empty()) {
                Task t = st.back(); st.pop_back();
                j.arena.max_depth = std::max(j.arena.max_depth, t.depth);
                emit(t, j.arena, nullptr, st);
            }
        };
        const unsigned n_thr = jobs.size() < 2 ? 1u : (unsigned)build_threads();
        if (n_thr == 1) for (Job& j : jobs) run_job(j);
        else {
            std::atomic<size_t> next{0};
            std::vector<std::thread> pool;
            for (unsigned k = 0; k < n_thr; k++)
                pool.emplace_back([&] { for (size_t i = next.fetch_add(1); i < jobs.size(); i = next.fetch_add(1)) run_job(jobs[i]); });
            for (auto& th : pool) th.join();
        }
        size_t n_nodes = G.nodes.size(), n_prims = G.prims.size();
        for (const Job& j : jobs) { n_nodes += j.arena.nodes.size(); n_prims += j.arena.prims.size(); }
        G.nodes.reserve(n_nodes); G.prims.reserve(n_prims);
        for (Job& j : jobs) {
            const uint32_t node_off = (uint32_t)G.nodes.size(), prim_off = (uint32_t)G.prims.size();
            for (BvhNode& n : j.arena.nodes) { n.child_base += node_off; n.prim_base += prim_off; }
            j.root.child_base += node_off; j.root.prim_base += prim_off;
            G.nodes.insert(G.nodes.end(), j.arena.nodes.begin(), j.arena.nodes.end());
            G.prims.insert(G.prims.end(), j.arena.prims.begin(), j.arena.prims.end());
            G.nodes[j.task.wide_index] = j.root;
            G.max_depth = std::max(G.max_depth, j.arena.max_depth);
        }
        out.nodes = std::move(G.nodes);
        out.prim_order = std::move(G.prims);
        out.max_depth = G.max_depth;
    }

    // Emits the wide node of task `t` (into *external when given, else A.nodes[t.wide_index]); children and
    // leaf primitives are allocated in arena A.
    void emit(const Task& t, Arena& A, BvhNode* external, std::vector<Task>& stack) {
        const Node2& n = b2.nodes[t.node2];
        int kids[8]; bool kid_leaf[8]; int nk = 0;
        if (n.leaf()) { kids[nk] = t.node2; kid_leaf[nk++] = true; } // degenerate: the whole (sub)tree is one leaf
        else if (!greedy) collect(t.node2, 8, kids, kid_leaf, nk);
        else { // RTGPU_BVH_GREEDY=1 (the round-1 collapse, kept for A/B): open the internal child with the largest area
            kids[nk++] = n.left; kids[nk++] = n.right;
            while (nk < 8) {
                int best = -1; float best_area = -1.f;
                for (int i = 0; i < nk; i++) {
                    const Node2& c = b2.nodes[kids[i]];
                    if (!c.leaf() && c.box.area() > best_area) { best_area = c.box.area(); best = i; }
                }
                if (best < 0) break;
                const Node2& c = b2.nodes[kids[best]];
                kids[best] = c.left; kids[nk++] = c.right;
            }
            for (int i = 0; i < nk; i++) kid_leaf[i] = b2.nodes[kids[i]].leaf();
        }
        // slot assignment: slot bit set <=> child lies towards +axis, so that a ray of octant o
        // (bit set = negative direction) meets slot s at rank (s ^ o)
        float pc[3];
        for (int a = 0; a < 3; a++) pc[a] = 0.5f * (n.box.lo[a] + n.box.hi[a]);
        float cost[8][8];
        for (int c = 0; c < nk; c++) {
            const Box& cb = b2.nodes[kids[c]].box;
            float d[3];
            for (int a = 0; a < 3; a++) d[a] = 0.5f * (cb.lo[a] + cb.hi[a]) - pc[a];
            for (int s = 0; s < 8; s++)
                cost[c][s] = ((s & 4) ? d[0] : -d[0]) + ((s & 2) ? d[1] : -d[1]) + ((s & 1) ? d[2] : -d[2]);
        }
        int slot_of[8]; bool slot_used[8] = {false, false, false, false, false, false, false, false};
        bool child_done[8] = {false, false, false, false, false, false, false, false};
        for (int k = 0; k < nk; k++) {
            int bc = -1, bs = -1; float best = -FLT_MAX;
            for (int c = 0; c < nk; c++) {
                if (child_done[c]) continue;
                for (int s = 0; s < 8; s++) {
                    if (slot_used[s]) continue;
                    if (cost[c][s] > best) { best = cost[c][s]; bc = c; bs = s; }
                }
            }
            slot_of[bc] = bs; slot_used[bs] = true; child_done[bc] = true;
        }
        int child_in_slot[8]; bool leaf_in_slot[8];
        for (int s = 0; s < 8; s++) { child_in_slot[s] = -1; leaf_in_slot[s] = false; }
        for (int c = 0; c < nk; c++) { child_in_slot[slot_of[c]] = kids[c]; leaf_in_slot[slot_of[c]] = kid_leaf[c]; }

        BvhNode w;
        std::memset(&w, 0, sizeof w);
        // quantisation frame
        w.px = n.box.lo[0]; w.py = n.box.lo[1]; w.pz = n.box.lo[2];
        uint8_t e[3];
        for (int a = 0; a < 3; a++) {
            double ext = (double)n.box.hi[a] - (double)n.box.lo[a];
            int ex = 0;
            if (ext > 0) ex = (int)std::ceil(std::log2(ext / 255.0));
            ex = std::max(-126, std::min(126, ex));
            while (std::ldexp(255.0, ex) < ext && ex < 126) ex++;
            e[a] = (uint8_t)(ex + 127);
        }
        w.ex = e[0]; w.ey = e[1]; w.ez = e[2];
        const float origin[3] = {w.px, w.py, w.pz};

        int n_internal = 0;
        for (int s = 0; s < 8; s++) if (child_in_slot[s] >= 0 && !leaf_in_slot[s]) n_internal++;
        w.child_base = (uint32_t)A.nodes.size();
        w.prim_base = (uint32_t)A.prims.size();
        A.nodes.resize(A.nodes.size() + n_internal);
        uint32_t next_child = w.child_base, prim_off = 0;
        uint8_t* qlo[3] = {w.qlox, w.qloy, w.qloz};
        uint8_t* qhi[3] = {w.qhix, w.qhiy, w.qhiz};
        for (int s = 0; s < 8; s++) {
            int c = child_in_slot[s];
            if (c < 0) { // empty slot: inverted box never intersects
                for (int a = 0; a < 3; a++) { qlo[a][s] = 255; qhi[a][s] = 0; }
                w.meta[s] = 0;
                continue;
            }
            const Node2& cn = b2.nodes[c];
            for (int a = 0; a < 3; a++) {
                double scale = std::ldexp(1.0, (int)e[a] - 127);
                double lo = ((double)cn.box.lo[a] - (double)origin[a]) / scale;
                double hi = ((double)cn.box.hi[a] - (double)origin[a]) / scale;
                int ql = (int)std::floor(lo), qh = (int)std::ceil(hi);
                ql = std::max(0, std::min(255, ql)); qh = std::max(0, std::min(255, qh));
                // conservative in exact arithmetic: origin + q*scale is exact in f64
                while (ql > 0 && (double)origin[a] + ql * scale > (double)cn.box.lo[a]) ql--;
                while (qh < 255 && (double)origin[a] + qh * scale < (double)cn.box.hi[a]) qh++;
                qlo[a][s] = (uint8_t)ql; qhi[a][s] = (uint8_t)qh;
            }
            if (!leaf_in_slot[s]) {
                w.imask |= (uint8_t)(1u << s);
                w.meta[s] = (uint8_t)((1u << 5) | (24u + (uint32_t)s));
                stack.push_back({c, next_child++, t.depth + 1});
            } else {
                uint32_t cnt = cn.count; // 1..3 (a binary leaf, or a small binary subtree the DP turned into one slot)
                uint32_t unary = cnt == 1 ? 1u : cnt == 2 ? 3u : 7u;
                w.meta[s] = (uint8_t)((unary << 5) | prim_off);
                for (uint32_t i = 0; i < cnt; i++) A.prims.push_back(b2.order[cn.first + i]);
                prim_off += cnt;
            }
        }
        if (external) *external = w; else A.nodes[t.wide_index] = w;
    }
};

} // namespace

WideBvhBuilderFn g_wide_bvh_builder = build_wide_bvh;

void build_wide_bvh(const std::vector<PrimBox>& boxes, WideBvh& out) {
    out = WideBvh();
    if (boxes.empty()) return;
    const bool timing = std::getenv("RTGPU_BUILD_TIMING") != nullptr;
    auto t0 = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (timing) std::fprintf(stderr, "[rtgpu build]   bvh: %-21s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    };
    Builder2 b2(boxes);
    int root = b2.build(0, (uint32_t)boxes.size());
    lap("binary SAH");
    WideBuilder wb(b2, out);
    wb.run(root);
    lap("collapse + quantise");
}

} // namespace rt
