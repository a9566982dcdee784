// build_gpu_collapse.cuh — step 5 of build_gpu.cuh (binary radix tree + DP tables -> 8-wide quantised nodes) and the
// host driver of the whole GPU build. Same node encoding as bvh.cpp: WideBuilder::emit.
#pragma once

namespace rt {
namespace gpubuild {

struct CollapseTask { int32_t node2; uint32_t wide_index; };

struct CollapseCtx {
    Tree2 t;
    const Dp* dp;
    const uint32_t* sorted_idx;
    BvhNode* nodes;          // [max_nodes]
    uint32_t* prim_order;    // [n]
    uint32_t* counters;      // [0] wide nodes allocated, [1] leaf primitives allocated, [2] overflow flag
    uint32_t max_nodes;
};

// the children of the wide node made from binary node `root`: follow the DP's decisions (bvh.cpp: WideBuilder::collect,
// recursion unrolled onto a small stack: at most 8 children, so at most 8 pending (node, j) pairs)
__device__ __forceinline__ int collect_children(const CollapseCtx& c, int root, int* kids, bool* kid_leaf) {
    const int n_leaf0 = (int)c.t.n - 1;
    int nk = 0, sp = 0;
    int st_node[16], st_j[16];
    st_node[sp] = root; st_j[sp++] = 8;
    while (sp > 0) {
        const int idx = st_node[--sp];
        int j = st_j[sp];
        if (idx >= n_leaf0) { kids[nk] = idx; kid_leaf[nk++] = true; continue; } // a single primitive
        const Dp d = c.dp[idx];
        for (;;) {
            if (j == 1) { kids[nk] = idx; kid_leaf[nk++] = d.leaf != 0; break; }
            const int k = d.split[j - 2];
            const int l = c.t.left[idx], r = c.t.right[idx];
            const float dist = c.dp[l].c[k - 1] + c.dp[r].c[j - k - 1];
            if (j == 8 || dist < d.c[j - 2]) { // split the budget between the two children (right first: popped last)
                st_node[sp] = r; st_j[sp++] = j - k;
                st_node[sp] = l; st_j[sp++] = k;
                break;
            }
            j--; // C(n, j) = C(n, j - 1)
        }
    }
    return nk;
}

__global__ void k_collapse(CollapseCtx c, const CollapseTask* tasks, uint32_t n_tasks, CollapseTask* next, uint32_t* next_count) {
    const uint32_t ti = blockIdx.x * blockDim.x + threadIdx.x;
    if (ti >= n_tasks) return;
    const CollapseTask task = tasks[ti];
    const int n_leaf0 = (int)c.t.n - 1;
    int kids[8]; bool kid_leaf[8];
    int nk;
    if (task.node2 >= n_leaf0 || c.t.count[task.node2] <= (uint32_t)MAX_LEAF) { kids[0] = task.node2; kid_leaf[0] = true; nk = 1; } // the whole tree is one leaf slot
    else nk = collect_children(c, task.node2, kids, kid_leaf);
    const Box6 nb = c.t.box[task.node2];
    // slot assignment: slot bit set <=> child lies towards +axis, so that a ray of octant o (bit set = negative
    // direction) meets slot s at rank (s ^ o); greedy on the projected centre offsets like the host builder
    float pc[3];
    for (int a = 0; a < 3; a++) pc[a] = 0.5f * (nb.lo[a] + nb.hi[a]);
    float dcen[8][3];
    for (int k = 0; k < nk; k++) {
        const Box6 cb = c.t.box[kids[k]];
        for (int a = 0; a < 3; a++) dcen[k][a] = 0.5f * (cb.lo[a] + cb.hi[a]) - pc[a];
    }
    int slot_of[8];
    uint32_t slot_used = 0, child_done = 0;
    for (int it = 0; it < nk; it++) {
        int bc = -1, bs = -1; float best = -FLT_MAX;
        for (int k = 0; k < nk; k++) {
            if ((child_done >> k) & 1u) continue;
            for (int s = 0; s < 8; s++) {
                if ((slot_used >> s) & 1u) continue;
                const float cost = ((s & 4) ? dcen[k][0] : -dcen[k][0]) + ((s & 2) ? dcen[k][1] : -dcen[k][1]) + ((s & 1) ? dcen[k][2] : -dcen[k][2]);
                if (cost > best) { best = cost; bc = k; bs = s; }
            }
        }
        slot_of[bc] = bs; slot_used |= 1u << bs; child_done |= 1u << bc;
    }
    int child_in_slot[8]; bool leaf_in_slot[8];
    for (int s = 0; s < 8; s++) { child_in_slot[s] = -1; leaf_in_slot[s] = false; }
    for (int k = 0; k < nk; k++) { child_in_slot[slot_of[k]] = kids[k]; leaf_in_slot[slot_of[k]] = kid_leaf[k]; }

    BvhNode w;
    memset(&w, 0, sizeof w);
    w.px = nb.lo[0]; w.py = nb.lo[1]; w.pz = nb.lo[2];
    int e[3];
    for (int a = 0; a < 3; a++) { // smallest exponent with 255 * 2^e >= extent
        const double ext = (double)nb.hi[a] - (double)nb.lo[a];
        int ex = 0;
        if (ext > 0) ex = (int)ceil(log2(ext / 255.0));
        ex = max(-126, min(126, ex));
        while (ldexp(255.0, ex) < ext && ex < 126) ex++;
        e[a] = ex + 127;
    }
    w.ex = (uint8_t)e[0]; w.ey = (uint8_t)e[1]; w.ez = (uint8_t)e[2];
    const float origin[3] = {w.px, w.py, w.pz};
    uint32_t n_internal = 0, n_prims = 0;
    for (int s = 0; s < 8; s++)
        if (child_in_slot[s] >= 0) { if (leaf_in_slot[s]) n_prims += c.t.count[child_in_slot[s]]; else n_internal++; }
    const uint32_t child_base = n_internal ? atomicAdd(&c.counters[0], n_internal) : 0u;
    const uint32_t prim_base = n_prims ? atomicAdd(&c.counters[1], n_prims) : 0u;
    if (child_base + n_internal > c.max_nodes) { atomicExch(&c.counters[2], 1u); return; }
    const uint32_t task_base = n_internal ? atomicAdd(next_count, n_internal) : 0u;
    w.child_base = child_base; w.prim_base = prim_base;
    uint8_t* qlo[3] = {w.qlox, w.qloy, w.qloz};
    uint8_t* qhi[3] = {w.qhix, w.qhiy, w.qhiz};
    uint32_t next_child = 0, prim_off = 0;
    for (int s = 0; s < 8; s++) {
        const int ch = child_in_slot[s];
        if (ch < 0) { // empty slot: inverted box never intersects
            for (int a = 0; a < 3; a++) { qlo[a][s] = 255; qhi[a][s] = 0; }
            w.meta[s] = 0;
            continue;
        }
        const Box6 cb = c.t.box[ch];
        for (int a = 0; a < 3; a++) {
            const double scale = ldexp(1.0, e[a] - 127);
            const double lo = ((double)cb.lo[a] - (double)origin[a]) / scale, hi = ((double)cb.hi[a] - (double)origin[a]) / scale;
            int ql = (int)floor(lo), qh = (int)ceil(hi);
            ql = max(0, min(255, ql)); qh = max(0, min(255, qh));
            // conservative in exact arithmetic: origin + q * scale is exact in f64
            while (ql > 0 && (double)origin[a] + ql * scale > (double)cb.lo[a]) ql--;
            while (qh < 255 && (double)origin[a] + qh * scale < (double)cb.hi[a]) qh++;
            qlo[a][s] = (uint8_t)ql; qhi[a][s] = (uint8_t)qh;
        }
        if (!leaf_in_slot[s]) {
            w.imask |= (uint8_t)(1u << s);
            w.meta[s] = (uint8_t)((1u << 5) | (24u + (uint32_t)s));
            next[task_base + next_child] = CollapseTask{ch, child_base + next_child};
            next_child++;
        } else {
            const uint32_t cnt = c.t.count[ch], first = c.t.first[ch]; // 1..3, contiguous in the sorted order
            const uint32_t unary = cnt == 1 ? 1u : cnt == 2 ? 3u : 7u;
            w.meta[s] = (uint8_t)((unary << 5) | prim_off);
            for (uint32_t i = 0; i < cnt; i++) c.prim_order[prim_base + prim_off + i] = c.sorted_idx[first + i];
            prim_off += cnt;
        }
    }
    c.nodes[task.wide_index] = w;
}

} // namespace gpubuild
} // namespace rt
