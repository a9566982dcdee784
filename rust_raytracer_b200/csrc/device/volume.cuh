// volume.cuh — constant media: Volume::test (reference src/object/volume.rs:33-71) against a boundary
// made of analytic primitives (spheres / quads, possibly under transforms; flatten.cpp: walk_boundary).
//
//   enter = boundary.test(ray, (-inf, inf)); exit = boundary.test(ray, (enter.t + 1e-4, inf))      volume.rs:39-44
//   [t_min, t_max] = [max(enter, t.lo), min(exit, t.hi)], empty => no hit, t_min = max(t_min, 0)   volume.rs:46-55
//   hit_dist = -1/density * ln(xi), ONE uniform drawn only when the segment is not empty            volume.rs:57-62
//   hit at t_min + hit_dist / |dir| when hit_dist <= (t_max - t_min) * |dir|                        volume.rs:64-70
//
// The draw sits INSIDE the intersection routine, before any of the material's draws: the per-event parity hook
// (hooks.cuh) consumes the injected stream in exactly that position; the render kernels give every (path vertex,
// volume) its own Philox domain, which is the same distribution.
#pragma once
#include "intersect.cuh"
#include "rng.cuh"

namespace rt {

RT_HD bool boundary_closest(const DeviceScene& s, const Volume& v, const RayD& ray, double t_lo, double t_hi, double& t_out) {
    bool any = false;
    double best = t_hi;
    for (int k = 0; k < v.bprim_count; k++) {
        const int bi = v.bprim_first + k;
        const APrim& a = s.vol_aprims[bi];
        const RayD r = ray_to_object(s, s.vol_chain_first[bi * 2], s.vol_chain_first[bi * 2 + 1], ray);
        double t;
        if (aprim_hit(a, r, t_lo, best, t)) { best = t; any = true; }
    }
    t_out = best;
    return any;
}

// The part of Volume::test before the draw: false = the reference returns None WITHOUT drawing.
// seg[0] = t_min, seg[1] = distance inside the boundary, seg[2] = |dir| in the volume's object space.
RT_HD bool volume_segment(const DeviceScene& s, int vi, const RayD& ray, double t_lo, double t_hi, double seg[3]) {
    const Volume& v = s.volumes[vi];
    double t_enter, t_exit;
    if (!boundary_closest(s, v, ray, -INFINITY, INFINITY, t_enter)) return false;
    if (!boundary_closest(s, v, ray, t_enter + 0.0001, INFINITY, t_exit)) return false;
    double t_min = fmax(t_enter, t_lo);
    const double t_max = fmin(t_exit, t_hi);
    if (t_min >= t_max) return false;
    t_min = fmax(t_min, 0.0);
    // ray.dir.length() of the ray Volume::test receives: under a Transform that is the object-space direction
    const Instance& in = s.instances[v.instance];
    const RayD ro = in.chain_len > 0 ? ray_to_object(s, in.chain_first, in.chain_len, ray) : ray;
    const double ray_len = sqrt(ro.d.x * ro.d.x + ro.d.y * ro.d.y + ro.d.z * ro.d.z);
    seg[0] = t_min; seg[1] = (t_max - t_min) * ray_len; seg[2] = ray_len;
    return true;
}

// the part after the draw
RT_HD bool volume_free_flight(const DeviceScene& s, int vi, const double seg[3], double xi, double& t_out) {
    const double hit_dist = s.volumes[vi].neg_inv_density * log(xi);
    if (hit_dist > seg[1]) return false;
    t_out = seg[0] + hit_dist / seg[2];
    return true;
}

// render kernels: the uniform of (path vertex, volume vi) comes from Philox domain 2 + vi
RT_HD bool volume_hit(const DeviceScene& s, int vi, const RayD& ray, double t_lo, double t_hi, uint64_t seed,
                      uint32_t pixel, uint32_t sample, uint32_t bounce, double& t_out) {
    double seg[3];
    if (!volume_segment(s, vi, ray, t_lo, t_hi, seg)) return false;
    PhiloxStream rng(seed, pixel, sample, bounce, 2u + (uint32_t)vi);
    return volume_free_flight(s, vi, seg, rng.uniform_open_d(), t_out);
}

} // namespace rt
