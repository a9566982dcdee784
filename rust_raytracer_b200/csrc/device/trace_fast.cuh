// trace_fast.cuh — K2 of the wavefront pipeline, production variant (fp32 traversal).
//
// Replaces world.test(ray, Interval(0.001, inf)) of camera.rs:289 (ObjectList / BVH / Transform /
// octree DFS: list.rs:58-74, bvh.rs:84-101, transform.rs:122-139, mesh.rs:165-197) by a
// persistent, warp-synchronous walk of ONE 8-wide quantised BVH:
//
//   * persistent warps with DYNAMIC RAY FETCH: a warp pulls rays from the bounce queue through a
//     device-side cursor; when fewer than FETCH_THRESH lanes still hold a live ray the finished
//     lanes are retired (hit record + material-class queue append) and refilled. On the
//     cornell_dragon workload the cost per ray is bimodal (wall rays: 1-2 nodes, dragon rays:
//     30+), and a static ray-per-thread mapping measured 4.8-5.3 active lanes of 32
//     (profiles/r01_k_trace_ncu_full_v1.txt); refilling keeps the warp populated.
//   * one explicit reconvergence point (ballot) per traversal step;
//   * child-box dequantisation without I2F: PRMT drops the byte into the mantissa of 2^15 and the
//     bias is folded into the plane offset: one PRMT + one FFMA per plane; min/max through FMNMX3;
//   * the 8 meta bytes are turned into hit-mask bit positions 4 at a time (SIMD-in-register);
//   * short traversal stack in shared memory (RT_SM_STACK entries per thread), remainder in local;
//   * quads (Plane, plane.rs:66-101) without ancestor gates are stored as world-space fp32
//     parallelograms in the 48-byte leaf record and share the Möller–Trumbore code of the
//     triangles (u, v in [0,1] instead of u + v <= 1): on Cornell-type scenes the walls are the
//     most tested primitives; the shade kernel refines t with the f64 object-space formula;
//   * rays are handed out in chunks per warp (one same-address atomic per RT_FETCH_CHUNK rays) and
//     the material-class queues are appended with one match_any + one atomic per class per warp;
//   * ONE leaf primitive per lane and step (LEAF_STEP = 1, the production setting): the rest of a leaf group waits in L.tg for the next
//     step, in which the lane skips its node phase — the primitive loop no longer runs as long as the lane with the most primitives
//     needs (17.2 -> 19.6 active lanes, -12 % on the kernel; profiles/r02_tune10_leaf_step.log). Postponed / deferred / cooperative
//     leaf tests, prefetches and a register-resident stack top are kept as measured, rejected variants (api.cu: launch_trace_fast).
//
// Results (closest primitive, t, barycentrics) are the same as closest_hit<false> of
// traverse.cuh up to the visiting order of equal-t candidates; tests/test_gpu_parity.py checks
// this kernel against the exact f64 path on primary and secondary rays.
#pragma once
#include "leaf_fast.cuh"

namespace rt {

// Entries of the traversal stack kept in shared memory (the rest lives in local memory, i.e. in L1 as well).
// Measured on the whole cornell_dragon job, 4 lanes: 8 entries 647 ms, 6: 649, 4: 635, 3: 630, 2: 625, 1: 626, 0: 620 —
// the shared memory a stack takes from the unified L1 costs more than the local-memory traffic it saves (the shade
// kernels of the other lanes share the SM and live in L1), so the default keeps the whole stack in local memory.
#ifndef RT_SM_STACK
#define RT_SM_STACK 0
#endif
#define RT_TRACE_THREADS 128

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
    return r;
}
// 32768 + byte `I` (0..3) of `w`, exactly, without I2F: PRMT drops the byte into mantissa bits
// 8..15 of 2^15. The bias is folded into the plane offset (node_hitmask), so a child plane costs
// one PRMT and one FFMA.
// `magic` = 0x47000000 held in a register the optimiser cannot see through: SASS PRMT takes one
// immediate, and with both the selector and the constant known ptxas re-materialised the selector
// into a register before each of the 48 PRMTs of a node test (39 extra IMADs per node).
template <int I> __device__ __forceinline__ float biased_byte(uint32_t w, uint32_t magic) {
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(w), "r"(magic), "n"(0x7404 | (I << 4)));
    return __uint_as_float(r);
}

struct TraceLane { // per-lane traversal registers
    float fox, foy, foz;    // fp32 origin
    float dx, dy, dz;
    float idx, idy, idz;
    uint32_t octinv;
    U2 ng, tg;
    int sp;
    float best_t;
    uint32_t best_prim, best_flags;
    float bu, bv;
    uint32_t slot;
};

__device__ __forceinline__ RayD pool_ray(const PathPool& pool, uint32_t bounce, uint32_t slot) {
    return ray_of(ray_rec_load<false>(rays_of(pool, bounce) + slot));
}

// Out-of-line f64 paths. They take POINTERS to the device-resident copies of the scene / pool
// descriptors (not references to the kernel parameters, which would force a per-thread copy of the
// parameter block to local memory) so that the fp32 loop keeps its registers.
// Nothing is passed by reference: an address-taken local would pin the caller's state to memory.
// Sphere / gated quad met during traversal: the reference formula in object space; returns t or -1.
__device__ __noinline__ float analytic_hit_cold(const DeviceScene* sp, const PathPool* pp, uint32_t bounce, uint32_t slot, uint32_t aprim, float t_hi) {
    const DeviceScene& s = *sp;
    const RayD ray = pool_ray(*pp, bounce, slot);
    double t;
    if (world_aprim_hit(s, s.aprims[aprim], ray, 0.001, (double)t_hi, t) && t < (double)t_hi) return (float)t;
    return -1.f;
}

// constant media competing with the surface hit of a finished ray; writes the hit record, returns the class
__device__ __noinline__ int retire_cold(const DeviceScene* sp, const PathPool* pp, uint64_t seed, uint32_t bounce, uint32_t slot,
                                        uint32_t prim, float t_fast, float hu, float hv, int cls) {
    const DeviceScene& s = *sp;
    const PathPool& pool = *pp;
    const RayD ray = pool_ray(pool, bounce, slot);
    double t = (double)t_fast;
    if (s.n_volumes > 0) { // constant media compete with the surfaces (volume.rs:33-71)
        const uint32_t pixel = rays_of(pool, bounce)[slot].pixel, sample = rays_of(pool, bounce)[slot].sample;
        for (int vi = 0; vi < s.n_volumes; vi++) {
            double tv;
            if (volume_hit(s, vi, ray, 0.001, t, seed, pixel, sample, bounce, tv)) { t = tv; prim = HIT_VOLUME; hu = __uint_as_float((uint32_t)vi); }
        }
    }
    if (prim == HIT_VOLUME) cls = material_class(s, s.volumes[__float_as_uint(hu)].material);
    hit_rec_store(pool.hits + slot, prim, (float)t, hu, hv, t);
    return cls;
}

// sm_100 packed fp32: one FFMA2 evaluates the near and the far plane of one axis of one child (two issue slots -> one)
__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void fma2(float& o0, float& o1, float a0, float a1, uint64_t b, uint64_t c) {
    uint64_t r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(pack2(a0, a1)), "l"(b), "l"(c));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(o0), "=f"(o1) : "l"(r));
}

// slab test of the 8 quantised child boxes of one node -> hit mask (bits 24..31 internal children
// in octant-permuted order, bits 0..23 leaf primitives)
template <bool PACK2>
__device__ __forceinline__ uint32_t node_hitmask(const uint4& n0, const uint4& n1, const uint4& n2, const uint4& n3, const uint4& n4,
                                                 const TraceLane& L, float tmin, float tmax, uint32_t magic) {
    const float px = __uint_as_float(n0.x), py = __uint_as_float(n0.y), pz = __uint_as_float(n0.z);
    const uint32_t e = n0.w;
    const float ax = __uint_as_float((e & 0xFFu) << 23) * L.idx;
    const float ay = __uint_as_float(((e >> 8) & 0xFFu) << 23) * L.idy;
    const float az = __uint_as_float(((e >> 16) & 0xFFu) << 23) * L.idz;
    // plane t = (32768 + q) * a + (b - 32768 a); folding the bias costs |a| / 512 of rounding in the
    // offset (0.002 quantisation steps), covered by widening every child box by 0.004 steps per side
    const float bx0 = fmaf(-32768.f, ax, (px - L.fox) * L.idx), by0 = fmaf(-32768.f, ay, (py - L.foy) * L.idy),
                bz0 = fmaf(-32768.f, az, (pz - L.foz) * L.idz);
    const float wx = fabsf(ax) * 0.004f, wy = fabsf(ay) * 0.004f, wz = fabsf(az) * 0.004f;
    const float nbx = bx0 - wx, fbx = bx0 + wx, nby = by0 - wy, fby = by0 + wy, nbz = bz0 - wz, fbz = bz0 + wz;
    const bool negx = L.idx < 0.f, negy = L.idy < 0.f, negz = L.idz < 0.f;
    // byte planes: n2 = qlox[0..7] qloy[0..7]; n3 = qloz[0..7] qhix[0..7]; n4 = qhiy[0..7] qhiz[0..7]
    const uint32_t nx[2] = {negx ? n3.z : n2.x, negx ? n3.w : n2.y}, fx[2] = {negx ? n2.x : n3.z, negx ? n2.y : n3.w};
    const uint32_t ny[2] = {negy ? n4.x : n2.z, negy ? n4.y : n2.w}, fy[2] = {negy ? n2.z : n4.x, negy ? n2.w : n4.y};
    const uint32_t nz[2] = {negz ? n4.z : n3.x, negz ? n4.w : n3.y}, fz[2] = {negz ? n3.x : n4.z, negz ? n3.y : n4.w};
    const uint32_t octinv4 = L.octinv * 0x01010101u;
    const float slack = 1.0f + 4e-7f;
    uint32_t hitmask = 0;
    const uint64_t ax2 = PACK2 ? pack2(ax, ax) : 0ull, ay2 = PACK2 ? pack2(ay, ay) : 0ull, az2 = PACK2 ? pack2(az, az) : 0ull;
    const uint64_t bx2 = PACK2 ? pack2(nbx, fbx) : 0ull, by2 = PACK2 ? pack2(nby, fby) : 0ull, bz2 = PACK2 ? pack2(nbz, fbz) : 0ull;
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const uint32_t meta4 = h ? n1.w : n1.z;
        const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
        const uint32_t inner_mask4 = prmt(is_inner4 << 3, 0u, 0xBA98u);            // 0xFF where the slot is internal
        const uint32_t bit_index4 = (meta4 ^ (octinv4 & inner_mask4)) & 0x1F1F1F1Fu;
        const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
#define RT_CHILD(J)                                                                                          \
        {                                                                                                    \
            float t0x, t1x, t0y, t1y, t0z, t1z;                                                              \
            if (PACK2) {                                                                                     \
                fma2(t0x, t1x, biased_byte<J>(nx[h], magic), biased_byte<J>(fx[h], magic), ax2, bx2);         \
                fma2(t0y, t1y, biased_byte<J>(ny[h], magic), biased_byte<J>(fy[h], magic), ay2, by2);         \
                fma2(t0z, t1z, biased_byte<J>(nz[h], magic), biased_byte<J>(fz[h], magic), az2, bz2);         \
            } else {                                                                                         \
                t0x = fmaf(biased_byte<J>(nx[h], magic), ax, nbx); t1x = fmaf(biased_byte<J>(fx[h], magic), ax, fbx); \
                t0y = fmaf(biased_byte<J>(ny[h], magic), ay, nby); t1y = fmaf(biased_byte<J>(fy[h], magic), ay, fby); \
                t0z = fmaf(biased_byte<J>(nz[h], magic), az, nbz); t1z = fmaf(biased_byte<J>(fz[h], magic), az, fbz); \
            }                                                                                                \
            const float cmin = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, tmin));                                     \
            const float cmax = fminf(fminf(t1x, t1y), fminf(t1z, tmax));                                     \
            if (cmin <= cmax * slack) hitmask |= ((child_bits4 >> (8 * J)) & 0xFFu) << ((bit_index4 >> (8 * J)) & 0xFFu); \
        }
        RT_CHILD(0) RT_CHILD(1) RT_CHILD(2) RT_CHILD(3)
#undef RT_CHILD
    }
    return hitmask;
}

// FETCH_THRESH: refill the warp when fewer lanes than this are busy; POSTPONE_PCT: postpone the
// triangle phase while fewer than this percentage of the busy lanes have triangles waiting
// (0 = never postpone); MIN_CTAS: resident CTAs per SM the register allocation must allow;
// COLD: the scene has spheres under non-similarity transforms, gated quads or constant media. Their f64 code
// is reached through calls, and with calls in the loop ptxas keeps the lane state in local memory
// under a 64-register cap — so scenes without them (COLD = false) get a call-free kernel at
// 8 CTAs/SM and the others run the same loop at MIN_CTAS = 4 (128 registers, no spills).
// PREFETCH: 0 none, 1 next child node to L1, 2 also the other hit children to L2.
// COOP: 0 = every lane tests the primitives of its own leaf group as it meets them (a loop that runs at ~2.6 of 32
// lanes: 3-4 lanes of a warp reach a leaf in the same step, with 1-3 primitives each, and every iteration waits for its
// own 48-byte record); N > 0 = DEFERRED, WARP-COOPERATIVE leaf tests: a lane that reaches a leaf appends its
// (lane, primitive) pairs to a per-warp queue in shared memory and waits; when N pairs are queued, or too few lanes can
// still walk, the queue is tested 32 pairs at a time — every lane loads ONE record (all loads in flight together), takes
// its owner's ray through shuffles, and the closest hit per owner goes back through a match_any + redux.min. Same
// candidates, same culling points (a waiting lane does not walk on with a stale bound), same hits.
#define RT_COOP_QCAP 96
// LEAF_STEP: 0 = a lane tests every primitive of its leaf group in the step that found it (the loop runs as long as the lane with the
// most primitives needs); N > 0 = at most N primitives per step, the rest of the group waits in L.tg for the next step, in which
// the lane skips its node phase (same candidates, same order per ray; fewer nearly empty loop iterations per warp).
// TWEAK bit 0: the next primitive of a group that waits for the following step is prefetched to L1; bit 1: the top of the traversal
// stack lives in registers (a pop takes it and reloads the new top from local memory, which is not needed before the next pop).
// LEAF_MIN (with LEAF_STEP > 0): a step runs its primitive phase only when at least this many lanes have primitives waiting, or no
// lane can open a node; the lanes that wait skip node phases until then.
template <int FETCH_THRESH, int POSTPONE_PCT, int MIN_CTAS, bool COLD, int PREFETCH, int COOP = 0, bool PACK2 = false, int LEAF_STEP = 0, int LEAF_MIN = 0, int TWEAK = 0>
__global__ void __launch_bounds__(RT_TRACE_THREADS, MIN_CTAS) k_trace_fast(DeviceScene s, PathPool pool, const DeviceScene* d_scene, const PathPool* d_pool, RenderConst rc, uint32_t bounce) {
#if RT_SM_STACK > 0
    __shared__ U2 sm_stack[RT_SM_STACK][RT_TRACE_THREADS];
#endif
    // COOP: per-warp pair queue and per-lane result slots
    __shared__ uint32_t s_q[COOP ? RT_TRACE_THREADS / 32 : 1][COOP ? RT_COOP_QCAP : 1];
    __shared__ uint32_t s_qn[RT_TRACE_THREADS / 32];
    __shared__ uint4 s_res[COOP ? RT_TRACE_THREADS / 32 : 1][COOP ? 32 : 1];    // u, v, prim, flags of the closest queued hit
    __shared__ uint32_t s_key[COOP ? RT_TRACE_THREADS / 32 : 1][COOP ? 32 : 1]; // bits of its t
    const int warp_in_cta = threadIdx.x >> 5;
    if (COOP) { if ((threadIdx.x & 31) == 0) s_qn[warp_in_cta] = 0; s_key[warp_in_cta][threadIdx.x & 31] = 0x7F800000u; __syncwarp(); }
    U2 lstack[RT_STACK_SIZE];
    uint32_t* cnt = pool.counters + bounce * RT_COUNTERS_PER_BOUNCE;
    const uint32_t n = *pool.error != 0u ? 0u : cnt[0]; // (after a failed pool check the counters no longer describe the arrays)
    // 0x47000000 in a form neither nvcc nor ptxas can fold (n < 2^31 always: pool capacity)
    const uint32_t magic = 0x47000000u | (n & 0x80000000u);
    uint32_t* cursor = cnt + 1;
    uint32_t* cnt_class = cnt + RT_CNT_CLASS;
    const PathRay* rays = rays_of(pool, bounce); // record i is entry i of the bounce's trace queue
    const uint32_t queue_cap = queue_cap_of(pool, bounce);
    const int lane = threadIdx.x & 31;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const float t_lo_f = 0.001f * 0.999f;
    // rays per cursor atomic: large when there is plenty of work, 32 when the bounce is nearly empty
    const uint32_t total_warps = gridDim.x * (RT_TRACE_THREADS / 32);
    uint32_t chunk = (n / total_warps / 2u) & ~31u;
    chunk = chunk < 32u ? 32u : (chunk > 128u ? 128u : chunk);

    uint32_t n_dropped = 0;  // (warp-uniform) rays of this warp that missed into blackness (RT_CNT_DROPPED)
    TraceLane L;
    bool alive = false;      // lane holds a ray that still traverses
    bool exhausted = false;  // (warp-uniform) the bounce queue has been handed out completely
    uint32_t r_next = 0, r_end = 0; // (warp-uniform) reserved range of queue indices
    L.sp = 0; L.slot = 0; L.ng.x = L.ng.y = L.tg.x = L.tg.y = 0; L.best_prim = HIT_MISS; L.best_flags = 0; L.best_t = 0.f; L.bu = L.bv = 0.f;

#if RT_SM_STACK > 0
#define RT_PUSH(E) { if (L.sp < RT_SM_STACK) sm_stack[L.sp][threadIdx.x] = (E); else lstack[L.sp - RT_SM_STACK] = (E); L.sp++; }
#define RT_POP(E)  { L.sp--; if (L.sp < RT_SM_STACK) (E) = sm_stack[L.sp][threadIdx.x]; else (E) = lstack[L.sp - RT_SM_STACK]; }
#else
#define RT_PUSH(E) { if (TWEAK & 2) { if (tos.y != 0u) { lstack[L.sp] = tos; L.sp++; } tos = (E); } else { lstack[L.sp] = (E); L.sp++; } }
#define RT_POP(E)  { if (TWEAK & 2) { (E) = tos; if (L.sp > 0) { L.sp--; tos = lstack[L.sp]; } else tos.y = 0u; } else { L.sp--; (E) = lstack[L.sp]; } }
#endif
#define RT_STACK_EMPTY ((TWEAK & 2) ? tos.y == 0u : L.sp == 0)
    U2 tos; tos.x = tos.y = 0u; // (TWEAK & 2: top of the stack, y == 0: empty)

// the primitive phase of one step: at most MAXK (0 = all) primitives of the lane's leaf group
#define RT_LEAF_PHASE(MAXK) \
                for (int lk = 0; L.tg.y != 0 && (MAXK == 0 || lk < MAXK); lk++) { \
                    const uint32_t bit = 31u - (uint32_t)__clz(L.tg.y); \
                    L.tg.y &= ~(1u << bit); \
                    const uint32_t pi = L.tg.x + bit; \
                    const float4* pp = reinterpret_cast<const float4*>(s.prims + pi); \
                    const float4 qa = __ldg(pp), qb = __ldg(pp + 1), qc = __ldg(pp + 2); \
                    const uint32_t flags = __float_as_uint(qc.w); \
                    float t, u = 0.f, v = 0.f; \
                    bool hit; \
                    if (COLD && (flags & (PRIM_ANALYTIC | PRIM_QUAD | PRIM_SPHERE)) == PRIM_ANALYTIC) { \
                        t = analytic_hit_cold(d_scene, d_pool, bounce, L.slot, __float_as_uint(qb.w), L.best_t); \
                        hit = t >= 0.f; \
                    } else if (flags & PRIM_SPHERE) \
                        hit = sphere_hit_fast(qa, qb, L, 0.001f, L.best_t, t); \
                    else \
                        hit = leaf_hit_fast(qa, qb, qc, flags, L, 0.001f, L.best_t, t, u, v); \
                    if (hit) { L.best_t = t; L.best_prim = pi; L.best_flags = flags; L.bu = u; L.bv = v; } \
                } \
                if ((TWEAK & 1) && L.tg.y != 0) asm volatile("prefetch.global.L1 [%0];" ::"l"(reinterpret_cast<const char*>(s.prims + (L.tg.x + 31u - (uint32_t)__clz(L.tg.y)))));
    for (;;) {
        // ---- fetch: idle lanes pull the next rays of the bounce queue
        if (!exhausted) {
            const unsigned want = __ballot_sync(0xFFFFFFFFu, !alive);
            if (want) {
                const uint32_t need = (uint32_t)__popc(want), avail = r_end - r_next;
                uint32_t nb = 0;
                if (avail < need) {
                    if (lane == 0) nb = atomicAdd(cursor, chunk);
                    nb = __shfl_sync(0xFFFFFFFFu, nb, 0);
                }
                const uint32_t rank = (uint32_t)__popc(want & lt_mask);
                const uint32_t i = rank < avail ? r_next + rank : nb + (rank - avail);
                if (avail < need) { r_next = nb + (need - avail); r_end = nb + chunk; } else r_next += need;
                if (PREFETCH >= 5) { // the records the NEXT fetch phases of this warp will read: towards L1 now (5), the rest of a new chunk towards L2 (6)
                    const uint32_t pi = r_next + (uint32_t)lane;
                    if (pi < r_end && pi < n) asm volatile("prefetch.global.L1 [%0];" ::"l"(reinterpret_cast<const char*>(rays + pi)));
                    if (PREFETCH == 6 && avail < need) {
                        for (uint32_t k = r_next + 32u + 2u * (uint32_t)lane; k < r_end && k < n; k += 64u)
                            asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char*>(rays + k)));
                    }
                }
                if (!alive && i < n) {
                    const uint32_t slot = i;
                    const RayRec rr = ray_rec_load<false>(rays + slot);
                    const uint2 pre = ray_rec_pretest(rays + slot);
                  if (rr.pixel != RT_DEAD_PIXEL) { // (dead: the unused end of a chunk of record positions, routed_flush) // broadcast-list hit found by the kernel that made the ray
                    const float pre_t = __uint_as_float(pre.y);
                    L.slot = slot;
                    L.fox = (float)rr.ox; L.foy = (float)rr.oy; L.foz = (float)rr.oz;
                    L.dx = rr.dx; L.dy = rr.dy; L.dz = rr.dz;
                    L.idx = safe_rcp(rr.dx); L.idy = safe_rcp(rr.dy); L.idz = safe_rcp(rr.dz);
                    const uint32_t oct = (rr.dx < 0.f ? 4u : 0u) | (rr.dy < 0.f ? 2u : 0u) | (rr.dz < 0.f ? 1u : 0u);
                    L.octinv = 7u - oct;
                    L.sp = 0; tos.y = 0u;
                    L.ng.x = 0u; L.ng.y = pre_t < 0.f ? 0u : 0x80000000u; // negative: the BVH bounds are out of reach
                    L.tg.x = L.tg.y = 0u;
                    L.best_t = fabsf(pre_t); L.best_prim = pre.x; L.bu = L.bv = 0.f;
                    L.best_flags = 0; // (of a broadcast-list hit: read when the ray retires)
                    if (COOP) s_key[warp_in_cta][lane] = __float_as_uint(L.best_t); // the bound the queued hits of THIS ray must beat
                    alive = true;
                  }
                }
                if (r_next >= n) exhausted = true;
            }
        }
        unsigned live = __ballot_sync(0xFFFFFFFFu, alive);
        if (live == 0) {
            if (exhausted) break;
            continue; // every record fetched was a dead one (unused chunk end): fetch again
        }
        const int thresh = exhausted ? 1 : FETCH_THRESH;

        // ---- traverse: warp-synchronous steps until too few lanes are busy
        bool busy = alive;
        bool pend = false; // COOP: the lane has queued leaf primitives (or could not: queue full) and waits for the flush
        do {
            // LEAF_STEP < 0: the group found by the previous step's node phase is tested FIRST, -LEAF_STEP primitives per step; a lane
            // whose group is finished walks on in the same step, the others skip this step's node phase
            if (LEAF_STEP < 0 && busy) { RT_LEAF_PHASE(-LEAF_STEP) }
            if (busy && ((!COOP && LEAF_STEP == 0) || (!pend && L.tg.y == 0)) && L.ng.y > 0x00FFFFFFu) {
                // node phase: take the nearest pending child of the current node group
                const uint32_t hits = L.ng.y, imask = hits & 0xFFu;
                const uint32_t child_bit = 31u - (uint32_t)__clz(hits);
                const uint32_t base = L.ng.x;
                L.ng.y &= ~(1u << child_bit);
                if (L.ng.y > 0x00FFFFFFu) RT_PUSH(L.ng)
                const uint32_t slot_i = (child_bit - 24u) ^ L.octinv;
                const uint32_t rel = (uint32_t)__popc(imask & ~(0xFFFFFFFFu << slot_i));
                const uint4* q = reinterpret_cast<const uint4*>(s.nodes + (base + rel));
                const uint4 n0 = __ldg(q), n1 = __ldg(q + 1), n2 = __ldg(q + 2), n3 = __ldg(q + 3), n4 = __ldg(q + 4);
                if (PREFETCH == 3 && (n0.w >> 24) != 0xFFu) { // the node has leaf slots: its primitives are contiguous from n1.y —
                    // start their first line towards L1 now, under the slab test (21 % of the stall samples sit on the first use of a leaf record)
                    asm volatile("prefetch.global.L1 [%0];" ::"l"(reinterpret_cast<const char*>(s.prims + n1.y)));
                }
                const uint32_t hm = node_hitmask<PACK2>(n0, n1, n2, n3, n4, L, t_lo_f, L.best_t * 1.000001f, magic);
                L.ng.x = n1.x; L.ng.y = (hm & 0xFF000000u) | (n0.w >> 24);
                L.tg.x = n1.y; L.tg.y = hm & 0x00FFFFFFu;
                if (PREFETCH > 0 && PREFETCH < 3 && L.ng.y > 0x00FFFFFFu) {
                    // the child the next step will open is known now: start its 80 bytes towards L1 while
                    // this step's leaf tests run (PREFETCH = 2: the other hit children towards L2 as well)
                    uint32_t pend = L.ng.y >> 24;
                    const uint32_t im = L.ng.y & 0xFFu;
                    bool first = true;
                    while (pend) {
                        const uint32_t cb = 31u - (uint32_t)__clz(pend);
                        pend &= ~(1u << cb);
                        const uint32_t sl = cb ^ L.octinv;
                        const char* a = reinterpret_cast<const char*>(s.nodes + (L.ng.x + (uint32_t)__popc(im & ~(0xFFFFFFFFu << sl))));
                        if (first) {
                            asm volatile("prefetch.global.L1 [%0];" ::"l"(a));
                            asm volatile("prefetch.global.L1 [%0];" ::"l"(a + 64));
                            first = false;
                            if (PREFETCH < 2) break;
                        } else {
                            asm volatile("prefetch.global.L2 [%0];" ::"l"(a));
                            asm volatile("prefetch.global.L2 [%0];" ::"l"(a + 64));
                        }
                    }
                }
            }
            if (POSTPONE_PCT > 0) {
                const unsigned with_tris = __ballot_sync(0xFFFFFFFFu, busy && L.tg.y != 0);
                const unsigned all_busy = __ballot_sync(0xFFFFFFFFu, busy);
                // postpone only lanes that still have node work, and only while the entry lands in the
                // shared part of the stack (bounds the number of postponed entries)
                const bool few = __popc(with_tris) * 100 < __popc(all_busy) * POSTPONE_PCT;
                if (busy && few && L.tg.y != 0 && L.ng.y > 0x00FFFFFFu && L.sp < RT_SM_STACK) { RT_PUSH(L.tg) L.tg.y = 0; }
            }
            if (COOP) {
                // ---- push: a lane that met leaf primitives queues them and waits for the next flush
                const bool has = busy && !pend && L.tg.y != 0;
                if (__ballot_sync(0xFFFFFFFFu, has)) {
                    if (has) {
                        uint32_t m = L.tg.y;
                        const uint32_t k = (uint32_t)__popc(m);
                        uint32_t pos = atomicAdd(&s_qn[warp_in_cta], k);
                        if (pos + k <= RT_COOP_QCAP) {
                            while (m) {
                                const uint32_t bit = 31u - (uint32_t)__clz(m);
                                m &= ~(1u << bit);
                                s_q[warp_in_cta][pos++] = ((uint32_t)lane << 27) | (L.tg.x + bit);
                            }
                            L.tg.y = 0;
                        } else atomicSub(&s_qn[warp_in_cta], k); // queue full: keeps its group, queues it after the flush
                        pend = true;
                    }
                    __syncwarp();
                }
                // ---- pop (lanes that wait keep their state)
                if (busy && !pend && L.ng.y <= 0x00FFFFFFu) {
                    if (RT_STACK_EMPTY) busy = false;
                    else {
                        U2 e; RT_POP(e)
                        if (e.y > 0x00FFFFFFu) L.ng = e; else { L.tg = e; L.ng.y = 0; }
                    }
                }
                // ---- flush: enough pairs queued, or too few lanes can still walk (they would leave the loop next)
                const uint32_t qn = s_qn[warp_in_cta];
                const unsigned walking = __ballot_sync(0xFFFFFFFFu, busy && !pend);
                if (qn >= (uint32_t)COOP || (__ballot_sync(0xFFFFFFFFu, pend) != 0u && __popc(walking) < thresh)) {
                    for (uint32_t base = 0; base < qn; base += 32u) {
                        const uint32_t g = base + (uint32_t)lane;
                        const bool valid = g < qn;
                        const uint32_t e = valid ? s_q[warp_in_cta][g] : 0u;
                        const int owner = (int)(e >> 27);
                        const uint32_t pi = e & 0x07FFFFFFu;
                        RayF r; // the owner's ray
                        r.fox = __shfl_sync(0xFFFFFFFFu, L.fox, owner); r.foy = __shfl_sync(0xFFFFFFFFu, L.foy, owner); r.foz = __shfl_sync(0xFFFFFFFFu, L.foz, owner);
                        r.dx = __shfl_sync(0xFFFFFFFFu, L.dx, owner); r.dy = __shfl_sync(0xFFFFFFFFu, L.dy, owner); r.dz = __shfl_sync(0xFFFFFFFFu, L.dz, owner);
                        const float t_hi = __shfl_sync(0xFFFFFFFFu, L.best_t, owner);
                        const uint32_t oslot = COLD ? __shfl_sync(0xFFFFFFFFu, L.slot, owner) : 0u;
                        float t = 0.f, u = 0.f, v = 0.f;
                        uint32_t flags = 0u;
                        bool hit = false;
                        if (valid) {
                            const float4* pp = reinterpret_cast<const float4*>(s.prims + pi);
                            const float4 qa = __ldg(pp), qb = __ldg(pp + 1), qc = __ldg(pp + 2);
                            flags = __float_as_uint(qc.w);
                            if (COLD && (flags & (PRIM_ANALYTIC | PRIM_QUAD | PRIM_SPHERE)) == PRIM_ANALYTIC) {
                                t = analytic_hit_cold(d_scene, d_pool, bounce, oslot, __float_as_uint(qb.w), t_hi);
                                hit = t >= 0.f;
                            } else if (flags & PRIM_SPHERE)
                                hit = sphere_hit_fast(qa, qb, r, 0.001f, t_hi, t);
                            else
                                hit = leaf_hit_fast(qa, qb, qc, flags, r, 0.001f, t_hi, t, u, v);
                        }
                        // closest hit per owner: a shared-memory atomicMin on the bits of t (t > 0: the patterns order like the
                        // values); the slot holds the owner's current bound from the moment it took its ray
                        const uint32_t tb = __float_as_uint(t);
                        if (hit) atomicMin(&s_key[warp_in_cta][owner], tb);
                        __syncwarp();
                        if (hit && s_key[warp_in_cta][owner] == tb) s_res[warp_in_cta][owner] = make_uint4(__float_as_uint(u), __float_as_uint(v), pi, flags);
                        __syncwarp();
                        {
                            const uint32_t kb = s_key[warp_in_cta][lane];
                            if (kb < __float_as_uint(L.best_t)) {
                                const uint4 rr = s_res[warp_in_cta][lane];
                                L.best_t = __uint_as_float(kb); L.bu = __uint_as_float(rr.x); L.bv = __uint_as_float(rr.y); L.best_prim = rr.z; L.best_flags = rr.w;
                            }
                        }
                        __syncwarp();
                    }
                    if (lane == 0) s_qn[warp_in_cta] = 0u;
                    pend = false;
                    __syncwarp();
                }
            } else {
            bool leaf_now = true;
            if (LEAF_MIN > 0) {
                const unsigned with_tris = __ballot_sync(0xFFFFFFFFu, busy && L.tg.y != 0);
                const unsigned can_walk = __ballot_sync(0xFFFFFFFFu, busy && L.tg.y == 0 && (L.ng.y > 0x00FFFFFFu || !RT_STACK_EMPTY));
                leaf_now = __popc(with_tris) >= LEAF_MIN || can_walk == 0u;
            }
            if (busy) {
                // primitive phase: every primitive of the current leaf group (LEAF_STEP > 0: at most that many)
                if (LEAF_STEP >= 0 && leaf_now) { RT_LEAF_PHASE(LEAF_STEP) }
                // pop
                if (L.ng.y <= 0x00FFFFFFu && (LEAF_STEP == 0 || L.tg.y == 0)) { // (nothing left to open here: back to the stack)
                    if (RT_STACK_EMPTY) busy = false;
                    else {
                        U2 e; RT_POP(e)
                        if (e.y > 0x00FFFFFFu) L.ng = e; else { L.tg = e; L.ng.y = 0; }
                    }
                }
            }
            }
            live = __ballot_sync(0xFFFFFFFFu, busy);
        } while (__popc(live) >= thresh);

        // ---- retire finished rays: f64 refinement, media, miss shader, hit record, class queues
        const bool done = alive && !busy;
        if (__ballot_sync(0xFFFFFFFFu, done) == 0) continue;
        int cls = -1;
        bool dropped = false;
        if (done) {
            const uint32_t prim = L.best_prim;
            cls = CLASS_EMISSIVE;
            // analytic hits carry an fp32-selected t here; the shade kernel recomputes it with the
            // reference's f64 formula when it rebuilds the HitRecord (surface_for_slot)
            if (prim != HIT_MISS) {
                const uint32_t flags = prim >= (uint32_t)s.global_first ? __float_as_uint(__ldg(&s.prims[prim].c.w)) : L.best_flags;
                cls = (int)((flags >> PRIM_CLASS_SHIFT) & 7u);
            }
            if (COLD && s.n_volumes > 0) // (a miss is resolved against sky / sun by the emissive shade kernel)
                cls = retire_cold(d_scene, d_pool, rc.seed, bounce, L.slot, prim, L.best_t, L.bu, L.bv, cls);
            else if (prim == HIT_MISS && (rc.flags & RT_FLAG_DROP_MISSES)) { cls = -1; dropped = true; } // nothing to shade (RT_CNT_DROPPED)
            else {
                hit_rec_store(pool.hits + L.slot, prim, L.best_t, L.bu, L.bv, (double)L.best_t);
            }
            alive = false;
        }
        n_dropped += (uint32_t)__popc(__ballot_sync(0xFFFFFFFFu, dropped)); // (warp-uniform; one atomic when the warp leaves)
        // one atomic per (warp, class): lanes of the same class elect a leader
        const unsigned peers = __match_any_sync(0xFFFFFFFFu, cls);
        const int leader = __ffs(peers) - 1;
        uint32_t qbase = 0;
        const uint32_t n_peers = (uint32_t)__popc(peers);
        if (cls >= 0 && lane == leader) qbase = atomicAdd(cnt_class + cls, n_peers);
        qbase = __shfl_sync(0xFFFFFFFFu, qbase, leader);
        if (cls >= 0) {
            if (qbase > queue_cap || n_peers > queue_cap - qbase) pool_fail(pool, RT_POOL_ERR_QUEUE); // (never an out-of-bounds write)
            else class_queue_of(pool, bounce, cls)[qbase + (uint32_t)__popc(peers & lt_mask)] = L.slot;
        }
    }
    if (n_dropped != 0u && lane == 0) atomicAdd(cnt + RT_CNT_DROPPED, n_dropped);
#undef RT_PUSH
#undef RT_POP
#undef RT_LEAF_PHASE
#undef RT_STACK_EMPTY
}

} // namespace rt
