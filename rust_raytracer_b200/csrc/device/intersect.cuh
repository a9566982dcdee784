// intersect.cuh — primitive intersection routines of the trace kernels.
//
//   fast path : world-space fp32 Möller–Trumbore on the 48-byte leaf primitive
//   exact path: the reference's f64 arithmetic, operation for operation, in OBJECT space
//               (TriangleMesh::test_tri mesh.rs:62-163, Sphere::test_impl sphere.rs:40-94,
//               Plane::test plane.rs:66-101, Transform::test transform.rs:122-139,
//               test_bounding_box aabb.rs:50-87), using never-contracted IEEE ops so that
//               t, u, v are bit-identical to the CPU reference for the same ray.
#pragma once
#include "dmath.cuh"

namespace rt {

#define RT_F64_EPSILON 2.220446049250313e-16

struct RayD { // canonical ray: f64 origin and direction (direction NOT normalised)
    X3 o, d;
};

// Transform::test (transform.rs:124-127): ray_obj = (inv * origin, inv * dir), chain outermost first
RT_HD RayD ray_to_object(const DeviceScene& s, int chain_first, int chain_len, RayD r) {
    for (int i = 0; i < chain_len; i++) {
        const double* inv = s.xforms[s.chain[chain_first + i]].inv;
        r.o = xmat_apply(inv, r.o, 1.0);
        r.d = xmat_apply(inv, r.d, 0.0);
    }
    return r;
}

// aabb.rs:50-87 with the reference's comparison order (NaN behaves identically)
RT_HD bool gate_test(const GateBox& g, const RayD& r, double t_lo, double t_hi) {
    double ix = xdiv(1.0, r.d.x), iy = xdiv(1.0, r.d.y), iz = xdiv(1.0, r.d.z);
    int sx = ix < 0.0, sy = iy < 0.0, sz = iz < 0.0;
    double t_min = xmul(xsub(sx ? g.hi[0] : g.lo[0], r.o.x), ix);
    double t_max = xmul(xsub(sx ? g.lo[0] : g.hi[0], r.o.x), ix);
    double ty_min = xmul(xsub(sy ? g.hi[1] : g.lo[1], r.o.y), iy);
    double ty_max = xmul(xsub(sy ? g.lo[1] : g.hi[1], r.o.y), iy);
    if ((t_min > ty_max) || (ty_min > t_max)) return false;
    if (ty_min > t_min) t_min = ty_min;
    if (ty_max < t_max) t_max = ty_max;
    double tz_min = xmul(xsub(sz ? g.hi[2] : g.lo[2], r.o.z), iz);
    double tz_max = xmul(xsub(sz ? g.lo[2] : g.hi[2], r.o.z), iz);
    if ((t_min > tz_max) || (tz_min > t_max)) return false;
    if (tz_min > t_min) t_min = tz_min;
    if (tz_max < t_max) t_max = tz_max;
    return t_min < t_hi && t_max > t_lo;
}

// Sphere::test_impl roots (sphere.rs:41-62) / Plane::test (plane.rs:67-90); object-space ray.
// Returns true and t when the primitive is hit with t_lo < t < t_hi.
// SPH = false: the caller guarantees a scene without analytic spheres (api.cu: has_spheres) — the sphere code is compiled out.
template <bool SPH = true>
RT_HD bool aprim_hit(const APrim& a, const RayD& r, double t_lo, double t_hi, double& t_out) {
    if (SPH && a.kind == APRIM_SPHERE) {
        X3 c = x3(a.c[0], a.c[1], a.c[2]);
        double radius = a.c[3];
        X3 cd = xsub3(r.o, c);
        double aa = xdot3(r.d, r.d);
        double half_b = xdot3(r.d, cd);
        double cc = xsub(xdot3(cd, cd), xmul(radius, radius));
        double disc = xsub(xmul(half_b, half_b), xmul(aa, cc));
        if (disc < 0.0) return false;
        double d_sqrt = xsqrt(disc);
        double root = xdiv(xsub(-half_b, d_sqrt), aa);
        if (root <= t_lo || t_hi <= root) {
            root = xdiv(xadd(-half_b, d_sqrt), aa);
            if (root <= t_lo || t_hi <= root) return false;
        }
        t_out = root;
        return true;
    }
    X3 n = x3(a.n[0], a.n[1], a.n[2]);
    X3 corner = x3(a.c[0], a.c[1], a.c[2]);
    double dot_ray_normal = xdot3(n, r.d);
    double dd = (a.flags & APRIM_BACKFACE) ? fabs(dot_ray_normal) : -dot_ray_normal;
    if (dd < RT_F64_EPSILON) return false;
    double hit_t = xdiv(xdot3(n, xsub3(corner, r.o)), dot_ray_normal);
    if (hit_t <= t_lo || t_hi <= hit_t) return false;
    X3 hit_pos = xadd3(r.o, xscale3(r.d, hit_t));
    X3 local = xsub3(hit_pos, corner);
    double u = xdot3(local, x3(a.iu[0], a.iu[1], a.iu[2]));
    double v = xdot3(local, x3(a.iv[0], a.iv[1], a.iv[2]));
    if (u < 0.0 || u > 1.0 || v < 0.0 || v > 1.0) return false;
    t_out = hit_t;
    return true;
}

// analytic primitive of the world BVH: transform chain + reference gate boxes + primitive
template <bool SPH = true>
RT_HD bool world_aprim_hit(const DeviceScene& s, const APrim& a, const RayD& world_ray, double t_lo, double t_hi, double& t_out) {
    const Instance& in = s.instances[a.instance];
    RayD r = ray_to_object(s, in.chain_first, in.chain_len, world_ray);
    for (int g = 0; g < a.n_gates; g++)
        if (!gate_test(s.gates[a.gate_first + g], r, t_lo, t_hi)) return false;
    return aprim_hit<SPH>(a, r, t_lo, t_hi, t_out);
}

// TriangleMesh::test_tri (mesh.rs:62-100), f64 object space, exact op order
RT_HD bool tri_hit_exact(const ExactMesh& m, int tri, bool two_sided, const RayD& r, double t_lo, double t_hi,
                         double& t_out, double& u_out, double& v_out) {
    const int32_t* vi = m.tri_pos + (size_t)tri * 3;
    const double* p0 = m.positions + (size_t)vi[0] * 3;
    const double* p1 = m.positions + (size_t)vi[1] * 3;
    const double* p2 = m.positions + (size_t)vi[2] * 3;
    X3 v0 = x3(p0[0], p0[1], p0[2]), v1 = x3(p1[0], p1[1], p1[2]), v2 = x3(p2[0], p2[1], p2[2]);
    X3 edge1 = xsub3(v1, v0), edge2 = xsub3(v2, v0);
    X3 ray_x_edge2 = xcross3(r.d, edge2);
    double det = xdot3(edge1, ray_x_edge2);
    double dd = two_sided ? fabs(det) : det;
    if (dd < RT_F64_EPSILON) return false;
    double inv_det = xdiv(1.0, det);
    X3 b = xsub3(r.o, v0);
    double u = xmul(xdot3(b, ray_x_edge2), inv_det);
    if (u < 0.0 || u > 1.0) return false;
    X3 b_x_edge1 = xcross3(b, edge1);
    double v = xmul(xdot3(r.d, b_x_edge1), inv_det);
    if (v < 0.0 || xadd(u, v) > 1.0) return false;
    double t = xmul(xdot3(edge2, b_x_edge1), inv_det);
    if (t <= t_lo || t_hi <= t) return false;
    t_out = t; u_out = u; v_out = v;
    return true;
}

// fp32 world-space Möller–Trumbore on the leaf primitive (same structure as mesh.rs:62-100)
RT_HD bool tri_hit_fast(const F4& pa, const F4& pb, const F4& pc, f3 o, f3 d, float t_lo, float t_hi,
                        float& t_out, float& u_out, float& v_out) {
    f3 v0(pa.x, pa.y, pa.z), e1(pb.x, pb.y, pb.z), e2(pc.x, pc.y, pc.z);
    f3 pvec = cross(d, e2);
    float det = dot(e1, pvec);
    bool two_sided = (f2u(pc.w) & PRIM_TWO_SIDED) != 0;
    float dd = two_sided ? fabsf(det) : det;
    if (!(dd > 0.f)) return false; // back faces culled (mesh.rs:76-80)
    float inv_det = 1.0f / det;
    f3 b = o - v0;
    float u = dot(b, pvec) * inv_det;
    if (u < 0.f || u > 1.f) return false;
    f3 qvec = cross(b, e1);
    float v = dot(d, qvec) * inv_det;
    if (v < 0.f || u + v > 1.f) return false;
    float t = dot(e2, qvec) * inv_det;
    if (t <= t_lo || t_hi <= t) return false;
    t_out = t; u_out = u; v_out = v;
    return true;
}

} // namespace rt
