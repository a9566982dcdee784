// surface.cuh — HitRecord reconstruction (reference src/object.rs:32-72) from a closest-hit
// result: position, shading normal (flipped against the ray), front_face, uv and the
// object-space tangent frame, following each primitive's own rules:
//   TriangleMesh::test_tri  mesh.rs:103-161     Sphere::test_impl  sphere.rs:64-93
//   Plane::test             plane.rs:84-100     Transform::test    transform.rs:130-135
//   Sky::test               sky.rs:28-52        Sun::test          sun.rs:33-61
//   Volume::test            volume.rs:58-70
// R = float for the render kernels, double for the parity hook.
#pragma once
#include "rtgpu.h"
#include "traverse.cuh"

namespace rt {

enum : uint32_t { HIT_MISS = 0xFFFFFFFFu, HIT_VOLUME = 0xFFFFFFFEu, HIT_INFINITE = 0xFFFFFFFDu };

template <typename R> struct Surface {
    d3 pos;            // world position (always f64)
    V3<R> normal;      // world-space, flipped so that it opposes the ray
    V3<R> tangent, bitangent; // object space
    R u, v;
    double t;
    int instance;
    int material;      // -1: sky / sun (private Emissive over `texture`)
    int texture;
    int primitive;
    bool front_face;
};

// Transform::test: hit.normal = (transform * hit.normal).to_unit(), innermost level first
template <typename R>
RT_HD V3<R> normal_to_world(const DeviceScene& s, const Instance& in, V3<R> n) {
    for (int i = in.chain_len - 1; i >= 0; i--) {
        const double* m = s.xforms[s.chain[in.chain_first + i]].m;
        V3<R> r((R)m[0] * n.x + (R)m[1] * n.y + (R)m[2] * n.z, (R)m[4] * n.x + (R)m[5] * n.y + (R)m[6] * n.z,
                (R)m[8] * n.x + (R)m[9] * n.y + (R)m[10] * n.z);
        R len = (R)sqrt((double)dot(r, r));
        n = r / len;
    }
    return n;
}

RT_HD d3 point_to_world(const DeviceScene& s, const Instance& in, d3 p) {
    for (int i = in.chain_len - 1; i >= 0; i--) {
        const double* m = s.xforms[s.chain[in.chain_first + i]].m;
        p = d3(m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3], m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7], m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]);
    }
    return p;
}

// Attributes of a finite-primitive hit. `bu`, `bv` are the barycentrics for triangles.
template <typename R, bool SPH = true>
RT_HD void surface_from_prim(const DeviceScene& s, const RayD& ray, uint32_t prim, double t, double bu, double bv,
                             bool need_tangents, Surface<R>& out) {
    const Prim& p = s.prims[prim];
    const uint32_t flags = f2u(p.c.w);
    const int inst = (int)f2u(p.a.w);
    const Instance& in = s.instances[inst];
    out.instance = inst; out.material = in.material; out.texture = -1; out.t = t;
    out.tangent = V3<R>((R)1, (R)0, (R)0); out.bitangent = V3<R>((R)1, (R)0, (R)0);
    if (sizeof(R) == 4 && !(flags & PRIM_ANALYTIC)) {
        // Render kernels, triangle hit: nothing of the HitRecord needs the object-space POSITION (normals, uvs and tangents
        // come from the barycentrics), so the f64 round trip world -> object -> world of Transform::test
        // (transform.rs:124-132: two 4x4 f64 products per level) reduces to the object-space DIRECTION for the front-face
        // test, in fp32, and pos = ray.at(t) in world space (equal to M (M^-1 o + t M^-1 d) up to rounding).
        const ShadeTri& st = s.shade_tris[prim];
        out.primitive = (int)f2u(p.b.w);
        V3<R> od((R)ray.d.x, (R)ray.d.y, (R)ray.d.z);
        for (int i = 0; i < in.chain_len; i++) {
            const double* m = s.xforms[s.chain[in.chain_first + i]].inv;
            od = V3<R>((R)m[0] * od.x + (R)m[1] * od.y + (R)m[2] * od.z, (R)m[4] * od.x + (R)m[5] * od.y + (R)m[6] * od.z,
                       (R)m[8] * od.x + (R)m[9] * od.y + (R)m[10] * od.z);
        }
        const R u = (R)bu, v = (R)bv, w = (R)1 - u - v;
        V3<R> n_obj((R)st.n0[0] * w + (R)st.n1[0] * u + (R)st.n2[0] * v, (R)st.n0[1] * w + (R)st.n1[1] * u + (R)st.n2[1] * v,
                    (R)st.n0[2] * w + (R)st.n1[2] * u + (R)st.n2[2] * v); // unnormalised blend; flat: three equal normals
        if (st.has_uv) {
            out.u = (R)st.uv0[0] * w + (R)st.uv1[0] * u + (R)st.uv2[0] * v;
            out.v = (R)st.uv0[1] * w + (R)st.uv1[1] * u + (R)st.uv2[1] * v;
            if (need_tangents) { // mesh.rs:131-145
                const V3<R> e1((R)st.e1o[0], (R)st.e1o[1], (R)st.e1o[2]), e2((R)st.e2o[0], (R)st.e2o[1], (R)st.e2o[2]);
                const R du1 = (R)st.uv1[0] - (R)st.uv0[0], dv1 = (R)st.uv1[1] - (R)st.uv0[1];
                const R du2 = (R)st.uv2[0] - (R)st.uv0[0], dv2 = (R)st.uv2[1] - (R)st.uv0[1];
                const V3<R> e1p = cross(n_obj, e1), e2p = cross(e2, n_obj);
                V3<R> tg = e2p * du1 + e1p * du2, bt = e2p * dv1 + e1p * dv2;
                const R lt = dot(tg, tg), lb = dot(bt, bt);
                const R inv_max = (R)rt_rsqrtf((float)(lt > lb ? lt : lb));
                out.tangent = tg * (-inv_max);
                out.bitangent = bt * inv_max;
            }
        } else { out.u = (R)0; out.v = (R)0; }
        out.front_face = dot(od, n_obj) < (R)0;
        if (!out.front_face) n_obj = -n_obj;
        out.normal = normal_to_world(s, in, n_obj);
        out.pos = d3(ray.o.x + ray.d.x * t, ray.o.y + ray.d.y * t, ray.o.z + ray.d.z * t);
        return;
    }
    const RayD ro = ray_to_object(s, in.chain_first, in.chain_len, ray);
    const d3 oo(ro.o.x, ro.o.y, ro.o.z), od(ro.d.x, ro.d.y, ro.d.z);
    const d3 pos_obj = oo + od * t; // ray.at(t) in object space
    V3<R> n_obj;
    if (flags & PRIM_ANALYTIC) {
        const APrim& a = s.aprims[f2u(p.b.w)];
        out.primitive = 0;
        if (SPH && a.kind == APRIM_SPHERE) {
            const d3 c(a.c[0], a.c[1], a.c[2]);
            const d3 nd = (pos_obj - c) / a.c[3];
            n_obj = V3<R>(nd);
            const double theta = acos(nd.y), phi = atan2(-nd.z, nd.x) + RT_PI_D;
            out.u = (R)(phi / (2.0 * RT_PI_D)); out.v = (R)(theta / RT_PI_D);
            if (need_tangents) {
                out.tangent = V3<R>((R)-nd.z, (R)0, (R)-nd.x);
                out.bitangent = cross(n_obj, out.tangent);
            }
        } else {
            const d3 corner(a.c[0], a.c[1], a.c[2]);
            const d3 local = pos_obj - corner;
            out.u = (R)dot(local, d3(a.iu[0], a.iu[1], a.iu[2]));
            out.v = (R)dot(local, d3(a.iv[0], a.iv[1], a.iv[2]));
            n_obj = V3<R>((R)a.n[0], (R)a.n[1], (R)a.n[2]);
            if (need_tangents) {
                out.tangent = V3<R>(normalize(d3(a.u[0], a.u[1], a.u[2])));
                out.bitangent = V3<R>(normalize(d3(a.v[0], a.v[1], a.v[2])));
            }
        }
    } else {
        const ShadeTri& st = s.shade_tris[prim];
        out.primitive = (int)f2u(p.b.w);
        const R u = (R)bu, v = (R)bv, w = (R)1 - u - v;
        const bool flat = (in.flags & RTGPU_OBJF_FLAT_SHADING) != 0;
        if (sizeof(R) == 8 && !flat) { // parity hook: f64 vertex normals / uvs of the reference layout (mesh.rs:110-116,147)
            const ExactMesh& em = s.meshes[in.mesh];
            const int32_t* ni = em.tri_nrm + (size_t)out.primitive * 3;
            const double *a0 = em.normals + (size_t)ni[0] * 3, *a1 = em.normals + (size_t)ni[1] * 3, *a2 = em.normals + (size_t)ni[2] * 3;
            n_obj = V3<R>((R)(a0[0] * w + a1[0] * u + a2[0] * v), (R)(a0[1] * w + a1[1] * u + a2[1] * v), (R)(a0[2] * w + a1[2] * u + a2[2] * v));
        } else
        n_obj = V3<R>((R)st.n0[0] * w + (R)st.n1[0] * u + (R)st.n2[0] * v, (R)st.n0[1] * w + (R)st.n1[1] * u + (R)st.n2[1] * v,
                      (R)st.n0[2] * w + (R)st.n1[2] * u + (R)st.n2[2] * v); // unnormalised blend
        if (st.has_uv) {
            if (sizeof(R) == 8) {
                const ExactMesh& em = s.meshes[in.mesh];
                const int32_t* ti = em.tri_uv + (size_t)out.primitive * 3;
                const double *t0 = em.uvs + (size_t)ti[0] * 2, *t1 = em.uvs + (size_t)ti[1] * 2, *t2 = em.uvs + (size_t)ti[2] * 2;
                out.u = (R)(t0[0] * w + t1[0] * u + t2[0] * v);
                out.v = (R)(t0[1] * w + t1[1] * u + t2[1] * v);
            } else {
            out.u = (R)st.uv0[0] * w + (R)st.uv1[0] * u + (R)st.uv2[0] * v;
            out.v = (R)st.uv0[1] * w + (R)st.uv1[1] * u + (R)st.uv2[1] * v;
            }
            if (need_tangents) { // mesh.rs:131-145
                const V3<R> e1((R)st.e1o[0], (R)st.e1o[1], (R)st.e1o[2]), e2((R)st.e2o[0], (R)st.e2o[1], (R)st.e2o[2]);
                const R du1 = (R)st.uv1[0] - (R)st.uv0[0], dv1 = (R)st.uv1[1] - (R)st.uv0[1];
                const R du2 = (R)st.uv2[0] - (R)st.uv0[0], dv2 = (R)st.uv2[1] - (R)st.uv0[1];
                const V3<R> e1p = cross(n_obj, e1), e2p = cross(e2, n_obj);
                V3<R> tg = e2p * du1 + e1p * du2, bt = e2p * dv1 + e1p * dv2;
                const R lt = dot(tg, tg), lb = dot(bt, bt);
                const R inv_max = (R)1 / (R)sqrt((double)(lt > lb ? lt : lb));
                out.tangent = tg * (-inv_max);
                out.bitangent = bt * inv_max;
            }
        } else { out.u = (R)0; out.v = (R)0; }
    }
    // HitRecord::new (object.rs:55-60) in object space, then Transform::test outward
    out.front_face = dot(V3<R>(od), n_obj) < (R)0;
    if (!out.front_face) n_obj = -n_obj;
    out.normal = normal_to_world(s, in, n_obj);
    out.pos = point_to_world(s, in, pos_obj);
}

// Sky / Sun "hits" at infinity: normal = -unit_dir, front face always (sky.rs:36-51, sun.rs:47-60)
template <typename R>
RT_HD void surface_from_infinite(const DeviceScene& s, const RayD& ray, int inst, Surface<R>& out) {
    const Instance& in = s.instances[inst];
    const d3 d(ray.d.x, ray.d.y, ray.d.z);
    const d3 unit = normalize(d);
    out.instance = inst; out.material = -1; out.texture = in.texture; out.primitive = 0;
    out.normal = V3<R>(-unit);
    out.tangent = V3<R>((R)1, (R)0, (R)0); out.bitangent = V3<R>((R)1, (R)0, (R)0);
    out.front_face = true;
    if (in.kind == INST_SKY) {
        out.t = INFINITY;
        out.u = (R)(atan2(unit.x, unit.z) / (2.0 * RT_PI_D) + 0.5);
        out.v = (R)(unit.y / 2.0 + 0.5);
    } else {
        out.t = 1.7976931348623157e308;
        out.u = (R)0; out.v = (R)0;
    }
    out.pos = d3(ray.o.x + ray.d.x * out.t, ray.o.y + ray.d.y * out.t, ray.o.z + ray.d.z * out.t);
}

// Volume::test hit record (volume.rs:64-70): position on the ray, an ARBITRARY frame — normal (1,0,0) flipped against
// the ray by HitRecord::new, uv (0,0); the only material a medium carries in practice (Isotropic) reads none of it.
template <typename R>
RT_HD void surface_from_volume(const DeviceScene& s, const RayD& ray, int vi, double t, Surface<R>& out) {
    const Volume& v = s.volumes[vi];
    out.instance = v.instance; out.material = v.material; out.texture = -1; out.primitive = 0; out.t = t;
    out.pos = d3(ray.o.x + ray.d.x * t, ray.o.y + ray.d.y * t, ray.o.z + ray.d.z * t);
    out.front_face = ray.d.x < 0.0;
    out.normal = out.front_face ? V3<R>((R)1, (R)0, (R)0) : V3<R>((R)-1, (R)0, (R)0);
    out.tangent = V3<R>((R)1, (R)0, (R)0); out.bitangent = V3<R>((R)1, (R)0, (R)0); out.u = (R)0; out.v = (R)0;
}

// Which sky / sun (if any) the world list reports for a ray with no finite hit, following the
// interval logic of ObjectList::test with Sky::test (`hit_t > t.max => None`) and Sun::test
// (`hit_t >= t.max => None`) in list order.
RT_HD int resolve_infinite(const DeviceScene& s, const RayD& ray) {
    int winner = -1;
    bool closest_is_inf = true; // closest_t == +inf
    d3 unit(0, 0, 0);
    bool have_unit = false;
    for (int k = 0; k < s.n_infinite; k++) {
        const int inst = s.infinite[k];
        const Instance& in = s.instances[inst];
        if (in.kind == INST_SKY) {
            if (closest_is_inf) winner = inst; // inf > inf is false => hit, closest stays inf
        } else {
            if (!have_unit) { unit = normalize(d3(ray.d.x, ray.d.y, ray.d.z)); have_unit = true; }
            const double c = in.dir[0] * unit.x + in.dir[1] * unit.y + in.dir[2] * unit.z;
            if (fabs(c - 1.0) > 0.001) continue;
            if (closest_is_inf) { winner = inst; closest_is_inf = false; } // f64::MAX >= closest fails only for closest = inf
        }
    }
    return winner;
}

} // namespace rt
