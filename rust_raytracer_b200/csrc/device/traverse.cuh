// traverse.cuh — closest-hit search over the 8-wide quantised BVH (the GPU replacement of
// world.test(ray, Interval(0.001, inf)) = ObjectList / BVH / Transform / octree traversal,
// reference src/object/list.rs:58-74, bvh.rs:84-101, mesh.rs:165-197).
//
// One thread walks one ray. Traversal state is a (node group, triangle group) pair plus a
// short stack of node groups (Ylitie et al. 2017): children are visited front to back using
// the octant-ordered slot assignment made by the builder.
//
//   EXACT = false : fp32 slab tests + fp32 world-space triangles; analytic primitives in f64
//   EXACT = true  : f64 slab tests on the (conservative) quantised boxes + the reference's f64
//                   object-space primitive tests; ties resolved by reference visit order
#pragma once
#include <type_traits>

#include "bvh.h"
#include "intersect.cuh"

namespace rt {

struct U2 { uint32_t x, y; };

#ifndef RT_STACK_SIZE
#define RT_STACK_SIZE 32
#endif

struct Hit {
    double t;          // exact: f64 reference t; fast: widened fp32 t (triangles) or f64 t (analytic)
    uint32_t prim;     // index into prims (leaf order); 0xFFFFFFFF = miss
    float u, v;        // barycentrics (triangles)
    double ud, vd;     // exact barycentrics (EXACT only)
    uint64_t key;      // reference visit-order key for exact ties
};

RT_HD void load_node(const BvhNode* nodes, uint32_t idx, U4& n0, U4& n1, U4& n2, U4& n3, U4& n4) {
    const U4* p = reinterpret_cast<const U4*>(nodes + idx);
#if defined(__CUDA_ARCH__)
    const uint4* q = reinterpret_cast<const uint4*>(p);
    uint4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2), d = __ldg(q + 3), e = __ldg(q + 4);
    n0 = {a.x, a.y, a.z, a.w}; n1 = {b.x, b.y, b.z, b.w}; n2 = {c.x, c.y, c.z, c.w}; n3 = {d.x, d.y, d.z, d.w}; n4 = {e.x, e.y, e.z, e.w};
#else
    n0 = p[0]; n1 = p[1]; n2 = p[2]; n3 = p[3]; n4 = p[4];
#endif
}

RT_HD uint32_t byte_of(uint32_t lo, uint32_t hi, int i) { return ((i < 4 ? lo : hi) >> ((i & 3) * 8)) & 0xFFu; }

// Slab test of the 8 quantised child boxes; returns the hit mask (bits 24..31: internal
// children in octant-permuted order, bits 0..23: leaf primitives).
template <typename T>
RT_HD uint32_t intersect_children(const U4& n0, const U4& n1, const U4& n2, const U4& n3, const U4& n4,
                                  T ox, T oy, T oz, T idx, T idy, T idz, uint32_t octinv, T tmin, T tmax, T slack) {
    const float px = u2f(n0.x), py = u2f(n0.y), pz = u2f(n0.z);
    const uint32_t e = n0.w;
    const T sx = (T)bvh_exp2((uint8_t)(e & 0xFF)), sy = (T)bvh_exp2((uint8_t)((e >> 8) & 0xFF)), sz = (T)bvh_exp2((uint8_t)((e >> 16) & 0xFF));
    const T ax = sx * idx, ay = sy * idy, az = sz * idz;                 // 2^e / d
    const T bx = ((T)px - ox) * idx, by = ((T)py - oy) * idy, bz = ((T)pz - oz) * idz; // (p - o) / d
    const uint32_t meta_lo = n1.z, meta_hi = n1.w;
    // byte planes: n2 = qlox[0..3] qlox[4..7] qloy[0..3] qloy[4..7]; n3 = qloz.. qhix..; n4 = qhiy.. qhiz..
    const bool negx = idx < (T)0, negy = idy < (T)0, negz = idz < (T)0;
    const uint32_t nx_lo = negx ? n3.z : n2.x, nx_hi = negx ? n3.w : n2.y; // near planes
    const uint32_t fx_lo = negx ? n2.x : n3.z, fx_hi = negx ? n2.y : n3.w; // far planes
    const uint32_t ny_lo = negy ? n4.x : n2.z, ny_hi = negy ? n4.y : n2.w;
    const uint32_t fy_lo = negy ? n2.z : n4.x, fy_hi = negy ? n2.w : n4.y;
    const uint32_t nz_lo = negz ? n4.z : n3.x, nz_hi = negz ? n4.w : n3.y;
    const uint32_t fz_lo = negz ? n3.x : n4.z, fz_hi = negz ? n3.y : n4.w;
    uint32_t hitmask = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t meta = byte_of(meta_lo, meta_hi, i);
        T t0x = (T)byte_of(nx_lo, nx_hi, i) * ax + bx, t1x = (T)byte_of(fx_lo, fx_hi, i) * ax + bx;
        T t0y = (T)byte_of(ny_lo, ny_hi, i) * ay + by, t1y = (T)byte_of(fy_lo, fy_hi, i) * ay + by;
        T t0z = (T)byte_of(nz_lo, nz_hi, i) * az + bz, t1z = (T)byte_of(fz_lo, fz_hi, i) * az + bz;
        // an empty slot has qlo = 255 > qhi = 0 on every axis => near > far => no hit
        T cmin = t0x > t0y ? t0x : t0y; cmin = cmin > t0z ? cmin : t0z; cmin = cmin > tmin ? cmin : tmin;
        T cmax = t1x < t1y ? t1x : t1y; cmax = cmax < t1z ? cmax : t1z; cmax = cmax < tmax ? cmax : tmax;
        const bool empty = meta == 0;
        if (!empty && cmin <= cmax * slack) {
            const bool inner = (meta & 0x18u) == 0x18u;
            const uint32_t bit_index = inner ? (24u + ((meta & 7u) ^ octinv)) : (meta & 31u);
            hitmask |= (meta >> 5) << bit_index;
        }
    }
    return hitmask;
}

// Closest hit of `ray` with t in (t_lo, t_hi). Returns false on a miss.
// order_limit (EXACT only): consider only primitives of instances the reference's world walk visits BEFORE
// the instance with that visit order — the closest_t an object of the world list sees when its own test runs
// (list.rs:62-70 shrinks the interval object by object); used by the parity hook for Volume::test (hooks.cuh).
template <bool EXACT>
RT_HD bool closest_hit(const DeviceScene& s, const RayD& ray, double t_lo, double t_hi, Hit& best, U2* stack, uint32_t order_limit = 0xFFFFFFFFu) {
    const bool filtered = EXACT && order_limit != 0xFFFFFFFFu;
    best.prim = 0xFFFFFFFFu; best.t = t_hi; best.u = best.v = 0.f; best.ud = best.vd = 0.0; best.key = ~0ull;
    // the broadcast list: the few large analytic primitives kept out of the BVH (flatten.cpp)
    for (int g = 0; g < s.n_global; g++) {
        const uint32_t pi = (uint32_t)(s.global_first + g);
        const APrim& a = s.aprims[f2u(s.prims[pi].b.w)];
        if (filtered && (uint32_t)s.instances[a.instance].order >= order_limit) continue;
        double t;
        if (world_aprim_hit(s, a, ray, t_lo, EXACT ? t_hi : best.t, t)) {
            const uint64_t key = ((uint64_t)(uint32_t)s.instances[a.instance].order << 32);
            if (t < best.t || (EXACT && t == best.t && key < best.key)) { best.t = t; best.prim = pi; best.u = best.v = 0.f; best.key = key; }
        }
    }
    if (!s.has_bvh) return best.prim != 0xFFFFFFFFu;

    // octant: bit set = negative direction; slot s is met at rank (s ^ oct)
    const uint32_t oct = (ray.d.x < 0.0 ? 4u : 0u) | (ray.d.y < 0.0 ? 2u : 0u) | (ray.d.z < 0.0 ? 1u : 0u);
    const uint32_t octinv = 7u - oct;

    // fp32 copies for the fast path
    const f3 of((float)ray.o.x, (float)ray.o.y, (float)ray.o.z), df((float)ray.d.x, (float)ray.d.y, (float)ray.d.z);
    typedef typename std::conditional<EXACT, double, float>::type T;
    const T ox = (T)ray.o.x, oy = (T)ray.o.y, oz = (T)ray.o.z;
    const T tiny = EXACT ? (T)1e-300 : (T)1e-30f;
    const T dx = (T)ray.d.x, dy = (T)ray.d.y, dz = (T)ray.d.z;
    const T idx = (T)1 / (fabs((double)dx) > (double)tiny ? dx : (dx < 0 ? -tiny : tiny));
    const T idy = (T)1 / (fabs((double)dy) > (double)tiny ? dy : (dy < 0 ? -tiny : tiny));
    const T idz = (T)1 / (fabs((double)dz) > (double)tiny ? dz : (dz < 0 ? -tiny : tiny));
    const T slack = EXACT ? (T)(1.0 + 1e-12) : (T)(1.0f + 4e-7f);
    const T tmin_t = (T)t_lo * (EXACT ? (T)1 : (T)0.999f);

    // exact path: object-space ray of the instance being tested (cached)
    int cur_inst = -1;
    RayD obj_ray = ray;

    int sp = 0;
    U2 ngroup = {0u, 0x80000000u}, tgroup = {0u, 0u};
    for (;;) {
        if (ngroup.y > 0x00FFFFFFu) {
            const uint32_t hits = ngroup.y, imask = ngroup.y & 0xFFu;
            const uint32_t child_bit = (uint32_t)rt_bfind(hits);
            const uint32_t base = ngroup.x;
            ngroup.y &= ~(1u << child_bit);
            if (ngroup.y > 0x00FFFFFFu) {
                if (sp < RT_STACK_SIZE) stack[sp] = ngroup;
                sp++;
            }
            const uint32_t slot = (child_bit - 24u) ^ octinv;
            const uint32_t rel = (uint32_t)rt_popc(imask & ~(0xFFFFFFFFu << slot));
            U4 n0, n1, n2, n3, n4;
            load_node(s.nodes, base + rel, n0, n1, n2, n3, n4);
            T tmax_t = EXACT ? (T)best.t : (T)((float)best.t * 1.000001f);
            const uint32_t hm = intersect_children<T>(n0, n1, n2, n3, n4, ox, oy, oz, idx, idy, idz, octinv, tmin_t, tmax_t, slack);
            ngroup.x = n1.x;
            ngroup.y = (hm & 0xFF000000u) | (n0.w >> 24);
            tgroup.x = n1.y;
            tgroup.y = hm & 0x00FFFFFFu;
        }

        while (tgroup.y != 0) {
            const uint32_t bit = (uint32_t)rt_bfind(tgroup.y);
            tgroup.y &= ~(1u << bit);
            const uint32_t pi = tgroup.x + bit;
            const Prim* pp = s.prims + pi;
#if defined(__CUDA_ARCH__)
            const float4 qa = __ldg(reinterpret_cast<const float4*>(pp)), qb = __ldg(reinterpret_cast<const float4*>(pp) + 1),
                         qc = __ldg(reinterpret_cast<const float4*>(pp) + 2);
            const F4 pa = {qa.x, qa.y, qa.z, qa.w}, pb = {qb.x, qb.y, qb.z, qb.w}, pc = {qc.x, qc.y, qc.z, qc.w};
#else
            const F4 pa = pp->a, pb = pp->b, pc = pp->c;
#endif
            const uint32_t flags = f2u(pc.w);
            if (filtered && (uint32_t)s.instances[f2u(pa.w)].order >= order_limit) continue;
            if (flags & PRIM_ANALYTIC) {
                const APrim& a = s.aprims[f2u(pb.w)];
                double t;
                if (world_aprim_hit(s, a, ray, t_lo, EXACT ? t_hi : best.t, t)) {
                    const uint64_t key = ((uint64_t)(uint32_t)s.instances[a.instance].order << 32);
                    if (t < best.t || (EXACT && t == best.t && key < best.key)) { best.t = t; best.prim = pi; best.u = best.v = 0.f; best.key = key; }
                }
            } else if (EXACT) {
                const int inst = (int)f2u(pa.w);
                const Instance& in = s.instances[inst];
                if (inst != cur_inst) { obj_ray = ray_to_object(s, in.chain_first, in.chain_len, ray); cur_inst = inst; }
                const int tri = (int)f2u(pb.w);
                double t, u, v;
                if (tri_hit_exact(s.meshes[in.mesh], tri, (flags & PRIM_TWO_SIDED) != 0, obj_ray, t_lo, t_hi, t, u, v)) {
                    const uint64_t key = ((uint64_t)(uint32_t)in.order << 32) | (uint32_t)tri;
                    if (t < best.t || (t == best.t && key < best.key)) { best.t = t; best.prim = pi; best.ud = u; best.vd = v; best.u = (float)u; best.v = (float)v; best.key = key; }
                }
            } else {
                float t, u, v;
                if (tri_hit_fast(pa, pb, pc, of, df, (float)t_lo, (float)best.t, t, u, v)) { best.t = (double)t; best.prim = pi; best.u = u; best.v = v; }
            }
        }

        if (ngroup.y <= 0x00FFFFFFu) {
            if (sp == 0) break;
            sp--;
            if (sp < RT_STACK_SIZE) ngroup = stack[sp];
            else { ngroup.x = 0; ngroup.y = 0; } // overflowed entries are dropped (never happens for depth <= RT_STACK_SIZE)
        }
    }
    return best.prim != 0xFFFFFFFFu;
}

} // namespace rt
