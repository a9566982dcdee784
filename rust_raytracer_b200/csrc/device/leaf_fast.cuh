// leaf_fast.cuh — fp32 world-space primitive tests on the 48-byte leaf record, and the broadcast
// pre-test every ray gets where it is generated.
//
// The handful of large analytic primitives of a scene (room walls, ground sphere, area lights:
// flatten.cpp keeps up to 16 fp32 quads / spheres out of the BVH) would be met by nearly every ray
// in the first node of the BVH, one divergent leaf test at a time (7 of 32 lanes active in the
// primitive loop of k_trace_fast, profiles/r01_k_trace_fast_v4_ncu_summary.txt). Instead the kernel that
// CREATES a ray (k_raygen, k_shade) tests the whole list right there — all lanes together,
// the records broadcast from L1 — and stores (closest list primitive, t) in the spare 8 bytes of
// the 64-byte ray record. It also slab-tests the bounds of what IS in the BVH: a ray that cannot reach
// them before its list hit is flagged (sign of t) and skips the traversal altogether.
// Same semantics as ObjectList::test over the whole world (list.rs:58-74): closest hit in (0.001, inf).
#pragma once
#include "traverse.cuh"

namespace rt {

struct RayF { float fox, foy, foz, dx, dy, dz; }; // fp32 ray: the fields the leaf tests read of a TraceLane

// MUFU.RCP (1 ulp): an IEEE division would bring an out-of-line slow path (a CALL) into the loop;
// the slab test carries 4e-7 of slack and the leaf test needs t only to fp32 accuracy
__device__ __forceinline__ float fast_rcp(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float safe_rcp(float d) {
    const float tiny = 1e-30f;
    return fast_rcp(fabsf(d) > tiny ? d : (d < 0.f ? -tiny : tiny));
}

// fp32 world-space Möller–Trumbore on the leaf record (mesh.rs:62-100); a quad (plane.rs:66-101)
// is the same computation with v in [0,1] instead of u + v <= 1 and a hair of edge tolerance so
// that the seams between walls stay closed in fp32.
template <typename R>
__device__ __forceinline__ bool leaf_hit_fast(const float4& qa, const float4& qb, const float4& qc, uint32_t flags, const R& L,
                                              float t_lo, float t_hi, float& t_out, float& u_out, float& v_out) {
    const f3 e1(qb.x, qb.y, qb.z), e2(qc.x, qc.y, qc.z), d(L.dx, L.dy, L.dz);
    const f3 pvec = cross(d, e2);
    const float det = dot(e1, pvec);
    const float dd = (flags & PRIM_TWO_SIDED) ? fabsf(det) : det;
    if (!(dd > 0.f)) return false; // back faces culled (mesh.rs:76-80, plane.rs:70-74)
    const float inv_det = fast_rcp(det);
    const f3 b(L.fox - qa.x, L.foy - qa.y, L.foz - qa.z);
    const float u = dot(b, pvec) * inv_det;
    const bool quad = (flags & PRIM_QUAD) != 0;
    const float eps = quad ? 1e-6f : 0.f;
    if (u < -eps || u > 1.f + eps) return false;
    const f3 qvec = cross(b, e1);
    const float v = dot(d, qvec) * inv_det;
    if (v < -eps || (quad ? v : u + v) > 1.f + eps) return false;
    const float t = dot(e2, qvec) * inv_det;
    if (t <= t_lo || t_hi <= t) return false;
    t_out = t; u_out = u; v_out = v;
    return true;
}

// fp32 world-space sphere (sphere.rs:40-62: near root, then far root, both against the open interval).
// Centre-to-ray form (l = oc - (d.oc / d.d) d, roots -k +- sqrt((r^2 - l.l) / d.d)): no cancellation
// between |oc|^2 and r^2 when the origin is far from a small sphere.
template <typename R>
__device__ __forceinline__ bool sphere_hit_fast(const float4& qa, const float4& qb, const R& L, float t_lo, float t_hi, float& t_out) {
    const f3 d(L.dx, L.dy, L.dz), oc(L.fox - qa.x, L.foy - qa.y, L.foz - qa.z);
    const float inv_a = fast_rcp(dot(d, d));
    const float k = dot(d, oc) * inv_a;
    const f3 l = oc - d * k;
    const float dp = qb.x * qb.x - dot(l, l);
    if (dp < 0.f) return false;
    const float sq = sqrtf(dp * inv_a);
    // A root is the difference of -k and sq: its ABSOLUTE error is a few ulps of the larger of the two. For a ray that
    // leaves a large sphere (radius-1000 ground) at a grazing angle the root at the origin is 0 +- 1e-3 in fp32 and could
    // pass t > 0.001 as a false self-intersection the reference (f64: 0 +- 1e-13) never sees: such a root must clear t_lo
    // by its own error bound.
    const float lo = t_lo + 4.8e-7f * (fabsf(k) + sq);
    float root = -k - sq;
    if (root <= lo || t_hi <= root) {
        root = -k + sq;
        if (root <= lo || t_hi <= root) return false;
    }
    t_out = root;
    return true;
}

// Broadcast pre-test of a new ray: closest primitive of the broadcast list (HIT_MISS / +inf when
// none) and whether the BVH can still matter. Returned t is NEGATED when the traversal can be skipped.
// SPH = false: the caller guarantees a scene without analytic spheres — the sphere test is compiled out of the loop.
template <bool SPH = true>
__device__ __forceinline__ void pretest_ray(const DeviceScene& s, const RayF& L, uint32_t& prim_out, float& t_signed) {
    float best_t = INFINITY;
    uint32_t best = 0xFFFFFFFFu; // HIT_MISS
    const f3 d(L.dx, L.dy, L.dz);
#ifndef RT_PRETEST_UNROLL
#define RT_PRETEST_UNROLL 1
#endif
    constexpr int unroll_n = RT_PRETEST_UNROLL;
#pragma unroll(unroll_n)
    for (int g = 0; g < s.n_global; g++) {
        const float4* pp = reinterpret_cast<const float4*>(s.bcast + g); // warp-uniform address: one L1 broadcast per load
        const float4 q = __ldg(pp), n = __ldg(pp + 1), w1 = __ldg(pp + 2), w2 = __ldg(pp + 3); // (all four before the branch: one wait)
        const uint32_t flags = __float_as_uint(q.w);
        float t;
        bool hit;
        if (SPH && (flags & PRIM_SPHERE)) hit = sphere_hit_fast(q, n, L, 0.001f, best_t, t);
        else { // same acceptance rules as leaf_hit_fast (culling, 1e-6 edge tolerance, open t interval)
            const f3 b(L.fox - q.x, L.foy - q.y, L.foz - q.z), nn(n.x, n.y, n.z);
            const float den = dot(d, nn);
            const float dd = (flags & PRIM_TWO_SIDED) ? fabsf(den) : -den;
            t = -dot(b, nn) * fast_rcp(den);
            const f3 p = b + d * t;
            const float u = dot(p, f3(w1.x, w1.y, w1.z)), v = dot(p, f3(w2.x, w2.y, w2.z));
            hit = dd > 0.f && u >= -1e-6f && u <= 1.f + 1e-6f && v >= -1e-6f && v <= 1.f + 1e-6f && t > 0.001f && t < best_t;
        }
        if (hit) { best_t = t; best = (uint32_t)(s.global_first + g); }
    }
    bool skip = !s.has_bvh;
    if (!skip) { // slab test of the BVH bounds, with the slack of the node test
        const float ix = safe_rcp(L.dx), iy = safe_rcp(L.dy), iz = safe_rcp(L.dz);
        const float ax = (s.bvh_lo[0] - L.fox) * ix, bx = (s.bvh_hi[0] - L.fox) * ix;
        const float ay = (s.bvh_lo[1] - L.foy) * iy, by = (s.bvh_hi[1] - L.foy) * iy;
        const float az = (s.bvh_lo[2] - L.foz) * iz, bz = (s.bvh_hi[2] - L.foz) * iz;
        const float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.001f * 0.999f));
        const float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), best_t * 1.000001f));
        skip = !(tn <= tf * (1.0f + 4e-6f)); // (tn, tf are never NaN: safe_rcp is finite, fmaxf / fminf drop a NaN operand)
    }
    prim_out = best;
    t_signed = skip ? -best_t : best_t;
}

} // namespace rt
