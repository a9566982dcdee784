// hooks.cuh — the per-event parity hook: ONE bounce of ray_color (camera.rs:294-328) for one ray with injected
// random numbers, consumed in the reference's draw order:
//
//   world.test(ray, (0.001, inf))    draws ONE uniform per Volume whose boundary segment is not empty, in world-list
//                                    visit order, inside Volume::test (volume.rs:57) — before any material draw
//   material.scatter / MixPDF        shade_bounce (shade.cuh)
//
// RT_HD: the CUDA kernel k_bounce_hook (kernels.cuh) and the host-simulation unit tests (tests/hostsim) run the same
// function; the product never runs it on the CPU.
#pragma once
#include "shade.cuh"
#include "volume.cuh"

namespace rt {

struct BounceRecord { // mirrors OracleBounce (oracle/oracle.h)
    int32_t kind, reserved;
    double emission[3], weight[3], next_origin[3], next_dir[3], pdf, scattering_pdf;
};

RT_HD void bounce_hook_one(const DeviceScene& s, float light_bias, const double* r6, const double* xi, int n_xi, BounceRecord& o) {
    RayD ray;
    ray.o = x3(r6[0], r6[1], r6[2]);
    ray.d = x3(r6[3], r6[4], r6[5]);
    memset(&o, 0, sizeof o);
    U2 stack[RT_STACK_SIZE];
    Hit h;
    Surface<float> sf;
    InjectedStream rng(xi, n_xi);
    const bool surface_hit = closest_hit<true>(s, ray, 0.001, INFINITY, h, stack);
    double t_closest = surface_hit ? h.t : INFINITY;
    // Constant media. ObjectList::test hands every object the interval (0.001, closest so far) (list.rs:62-70), so a
    // Volume sees the closest hit among the objects visited BEFORE it — that decides whether its segment is empty,
    // i.e. whether it draws. Volumes are stored in visit order (flatten.cpp walks the tree in DFS pre-order).
    int vol = -1;
    double t_vol = INFINITY;
    for (int vi = 0; vi < s.n_volumes; vi++) {
        const uint32_t order = (uint32_t)s.instances[s.volumes[vi].instance].order;
        double t_before = INFINITY; // closest surface among the instances visited before this volume
        if (surface_hit) {
            const int inst = (int)f2u(s.prims[h.prim].a.w);
            if ((uint32_t)s.instances[inst].order < order) t_before = h.t; // the overall closest surface is one of them
            else {
                Hit hb;
                if (closest_hit<true>(s, ray, 0.001, INFINITY, hb, stack, order)) t_before = hb.t;
            }
        }
        double seg[3], tv;
        if (!volume_segment(s, vi, ray, 0.001, fmin(t_before, t_vol), seg)) continue;
        if (volume_free_flight(s, vi, seg, rng.uniform_open_d(), tv)) { t_vol = tv; vol = vi; }
    }
    bool have = false;
    if (vol >= 0 && t_vol < t_closest) { // (a surface visited after the volume wins only when it is strictly closer)
        surface_from_volume<float>(s, ray, vol, t_vol, sf);
        have = true;
    } else if (surface_hit) {
        const int inst = (int)f2u(s.prims[h.prim].a.w);
        const int mat = s.instances[inst].material;
        surface_from_prim<float>(s, ray, h.prim, h.t, h.ud, h.vd, mat >= 0 && s.materials[mat].normal_map >= 0, sf);
        have = true;
    } else {
        const int inf = resolve_infinite(s, ray);
        if (inf >= 0) { surface_from_infinite<float>(s, ray, inf, sf); have = true; }
    }
    if (!have) return;
    Bounce b;
    shade_bounce<-1, true>(s, ray, sf, light_bias, rng, b);
    o.kind = b.kind == BOUNCE_EMISSIVE ? 4 : b.kind;
    o.emission[0] = b.emission.x; o.emission[1] = b.emission.y; o.emission[2] = b.emission.z;
    if (b.kind == BOUNCE_PDF || b.kind == BOUNCE_RAY) {
        o.weight[0] = b.weight.x; o.weight[1] = b.weight.y; o.weight[2] = b.weight.z;
        o.next_origin[0] = b.origin.x; o.next_origin[1] = b.origin.y; o.next_origin[2] = b.origin.z;
        o.next_dir[0] = b.dir.x; o.next_dir[1] = b.dir.y; o.next_dir[2] = b.dir.z;
    }
    o.pdf = b.pdf; o.scattering_pdf = b.spdf;
}

} // namespace rt
