// kernels.cuh — the wavefront pipeline (sm_100a). One launch sequence per batch of camera
// paths, no host synchronisation inside a batch: every kernel is a persistent grid-stride
// kernel that reads its element count from device memory.
//
//   K1 k_raygen        Camera::get_ray (camera.rs:260-280,334-349) for (pixel, sample) slots
//   K2 k_trace         world.test closest hit + constant media + sky/sun miss resolution,
//                      then warp-aggregated append of the path to its material-class queue
//   K3 k_shade<class>  Material::emit/scatter + MixPDF for one material class; accumulates
//                      throughput * emission, writes the next ray, appends to the next trace queue
//   K5 k_reduce_resolve  N GPUs: sums of the peers' accumulators / spp -> f64 Buffer (camera.rs:229,243-255);
//                      one GPU: the fp32 sums go to the host as they are, scaled and added there (api.cu)
//      k_tonemap       Writer::save per-pixel maths (output.rs:23-50, tonemapping/aces.rs)
//   parity hooks       k_trace_exact (RtgpuHit records), k_bounce_hook (injected draws)
#pragma once
#include <cuda_runtime.h>

#include "rtgpu.h"
#include "shade.cuh"
#include "leaf_fast.cuh"
#include "volume.cuh"

namespace rt {

// counters per bounce: [0] = rays to trace, [1] = fetch cursor of the persistent trace kernel,
// [RT_CNT_CLASS + class] = paths queued per material class (by k_trace_fast, and by k_raygen / k_shade of
// the previous bounce for rays the broadcast pre-test resolved completely: they bypass the trace kernel)
// [2] = rays of the bounce that bypass the trace kernel (their records fill the ray array from its far end)
#define RT_CNT_BYPASS 2
// [3] = of [0], record positions a warp had reserved and did not use (filled with RT_DEAD_PIXEL records)
#define RT_CNT_DEAD 3
// [4] = rays of the bounce that MISSED everything in a scene without sky / sun and with a black background: they contribute
// nothing (camera.rs:331 returns the background colour) and are dropped where the miss is found — by the kernel that made
// the ray when the broadcast pre-test already knows it, else by the trace kernel when the ray retires — instead of being
// written, queued and read again by the emissive shade kernel just to die there. Counted here so that the ray statistics
// (world.test-equivalent queries) stay exact.
#define RT_CNT_DROPPED 4
#define RT_CNT_CLASS 5
#define RT_FLAG_DROP_MISSES 0x80000000u   // internal bit of RenderConst.flags, set by enqueue_render
#define RT_FLAG_INLINE_EMISSION 0x40000000u // internal: a new ray the pre-test resolves onto an Emissive surface ends where it is made
#define RT_DEAD_PIXEL 0xFFFFFFFFu
#define RT_COUNTERS_PER_BOUNCE (RT_CNT_CLASS + N_CLASSES)

struct CameraDev { // RtgpuCameraDesc + sampling parameters
    int width, height;
    double position[3], first_pixel[3], du[3], dv[3], bu[3], bv[3];
    double aperture_radius;
    float background[3];
    int sqrt_spt;
    double inv_sqrt_spt;
};

// Path state is addressed through index queues in an order that is random after the first bounce,
// so it is kept as two sector-aligned records per slot (one 64-byte ray record, one 32-byte hit
// record): a gather touches 2 + 1 (trace) or 3 (shade) full 32-byte sectors instead of one
// mostly wasted sector per 4..16-byte field of a structure of arrays.
struct __align__(16) PathRay {   // 64 B
    double ox, oy;               // [0,16)   ray origin (f64: robust self-intersection at Cornell scale)
    double oz; float dx, dy;     // [16,32)  direction, not normalised
    float dz; uint32_t pixel, sample; float tr; // [32,48) stream address, throughput r
    float tg, tb; uint32_t pre_prim; float pre_t; // [48,64) throughput g b, broadcast pre-test (RayPre)
};
static_assert(sizeof(PathRay) == 64, "PathRay must be 64 bytes");
struct __align__(16) PathHit {   // 32 B
    uint32_t prim; float t32, u, v; // HIT_* or leaf index, fp32 t, barycentrics / volume or infinite id in u
    double t;                       // f64 t for analytic / volume hits
    uint32_t pad0, pad1;
};
static_assert(sizeof(PathHit) == 32, "PathHit must be 32 bytes");

// Ray records are STREAMED, not updated in place: the rays of bounce b live in rays[b & 1], written by
// the kernel that made them (k_raygen / k_shade of bounce b - 1) at the position a warp-aggregated atomic
// hands out — rays for the trace kernel from index 0 upwards (record i IS entry i of the trace queue: no
// index array, and k_trace_fast reads its chunks as contiguous memory), rays that bypass it from index
// capacity - 1 downwards. A path has no fixed slot; what identifies it (pixel, sample) travels in the record.
// Writes are contiguous per warp, and the shade kernels' gathers follow the retirement order of the trace
// kernel, which is its (sequential) fetch order shuffled only locally.
struct PathPool {
    PathRay* rays[2];         // ping-pong per bounce
    PathHit* hits;            // hits[i] belongs to rays[b & 1][i] (traced rays only)
    uint32_t* class_queue[2][N_CLASSES]; // record indices per material class; ping-pong per bounce: k_shade of bounce b appends bypassing paths to [(b + 1) & 1]
    uint32_t* counters;       // [max_bounces + 1][RT_COUNTERS_PER_BOUNCE]
    uint32_t capacity;        // paths per batch
    uint32_t ray_slots;       // records per rays[] array: capacity + room for the unused ends of reserved chunks
    // Tail merging (api.cu: enqueue_render): the rays of bounce `carry_bounce` do not go to rays[b & 1] but to a
    // separate, larger array that is NOT reset between the batches of a group; only the last batch of the
    // group runs bounces >= carry_bounce, on the survivors of all its batches together (0xFFFFFFFF = off).
    uint32_t carry_bounce;
    uint32_t carry_slots;     // = carry_t_cap + carry_b_cap
    PathRay* carry_rays;
    uint32_t* carry_queue[N_CLASSES];
    // Bounds (api.cu sizes the arrays from the survival a pilot sample measured, not for the worst case):
    //   rays[]      a bounce with N rays needs at most N + min(N, waste_max) records (unused chunk ends, record_chunk):
    //               checked ONCE per producer kernel against the complete counters of its input bounce (pool_guard);
    //   carry_rays  filled by several batches: the trace region [0, carry_t_cap) and the bypass region (the last
    //               carry_b_cap records) are separate, every chunk reservation is checked against its region;
    //   queues      every append is checked against queue_cap / carry_queue_cap.
    // A failed check drops the rays, raises *error and the render call reports it — never an out-of-bounds write.
    uint32_t carry_t_cap, carry_b_cap;
    uint32_t queue_cap, carry_queue_cap;
    uint32_t waste_max;
    uint32_t* error;          // RT_POOL_ERR_* bits
};
enum : uint32_t { RT_POOL_ERR_RAYS = 1u, RT_POOL_ERR_QUEUE = 2u, RT_POOL_ERR_CARRY_RAYS = 4u, RT_POOL_ERR_CARRY_QUEUE = 8u };
__device__ __forceinline__ PathRay* rays_of(const PathPool& p, uint32_t bounce) { return bounce == p.carry_bounce ? p.carry_rays : p.rays[bounce & 1]; }
__device__ __forceinline__ uint32_t slots_of(const PathPool& p, uint32_t bounce) { return bounce == p.carry_bounce ? p.carry_slots : p.ray_slots; }
__device__ __forceinline__ uint32_t* class_queue_of(const PathPool& p, uint32_t bounce, int cls) {
    return bounce == p.carry_bounce ? p.carry_queue[cls] : p.class_queue[bounce & 1][cls];
}
__device__ __forceinline__ uint32_t queue_cap_of(const PathPool& p, uint32_t bounce) { return bounce == p.carry_bounce ? p.carry_queue_cap : p.queue_cap; }
__device__ __forceinline__ void pool_fail(const PathPool& p, uint32_t code) { if ((*p.error & code) == 0u) atomicOr(p.error, code); }
// Producer-side check for a bounce whose rays go to rays[] (not the carry arrays): `n_in` = the paths the producers of
// that bounce read in total (an upper bound of the rays they can emit). Warp-uniform, same answer in every thread:
// the counters it is computed from are complete before the kernel starts.
__device__ __forceinline__ bool pool_guard(const PathPool& p, uint32_t out_bounce, uint32_t n_in) {
    if (out_bounce == p.carry_bounce) return true; // checked per reservation instead (routed_append)
    const uint32_t need = n_in + (n_in < p.waste_max ? n_in : p.waste_max);
    if (need < n_in || need > p.ray_slots) { pool_fail(p, RT_POOL_ERR_RAYS); return false; }
    if (n_in > p.queue_cap) { pool_fail(p, RT_POOL_ERR_QUEUE); return false; }
    return true;
}

struct RenderConst {
    uint64_t seed;
    float light_bias;
    int max_depth;
    uint32_t flags;   // RTGPU_FLAG_* of the render
};

// Path records are written once and read once or twice, far apart: with RT_STREAM_CS they go through the
// evict-first (streaming) cache policy so that they do not push the BVH out of L2 while a trace kernel of
// another lane runs (__ldcs / __stcs = ld/st.global.cs).
#ifndef RT_STREAM_CS
#define RT_STREAM_CS 1
#endif
__device__ __forceinline__ uint4 rec_ld(const uint4* p) {
#if RT_STREAM_CS
    return __ldcs(p);
#else
    return *p;
#endif
}
__device__ __forceinline__ void rec_st(uint4* p, uint4 v) {
#if RT_STREAM_CS
    __stcs(p, v);
#else
    *p = v;
#endif
}
struct RayRec { double ox, oy, oz; float dx, dy, dz; uint32_t pixel, sample; float tr, tg, tb; };
// the ray record's last 8 bytes: closest broadcast-list primitive and its t, negated when the BVH
// cannot be reached (pretest_ray, leaf_fast.cuh). {HIT_MISS, +inf} = nothing known.
struct RayPre { uint32_t prim; float t_signed; };

// geometry + stream address (first 48 bytes); the throughput g/b lanes are only read when FULL
template <bool FULL> __device__ __forceinline__ RayRec ray_rec_load(const PathRay* p) {
    const uint4* q = reinterpret_cast<const uint4*>(p);
    const uint4 a = rec_ld(q), b = rec_ld(q + 1), c = rec_ld(q + 2);
    RayRec r;
    r.ox = __hiloint2double((int)a.y, (int)a.x); r.oy = __hiloint2double((int)a.w, (int)a.z);
    r.oz = __hiloint2double((int)b.y, (int)b.x); r.dx = __uint_as_float(b.z); r.dy = __uint_as_float(b.w);
    r.dz = __uint_as_float(c.x); r.pixel = c.y; r.sample = c.z; r.tr = __uint_as_float(c.w);
    r.tg = r.tb = 0.f;
    if (FULL) { const uint4 d = rec_ld(q + 3); r.tg = __uint_as_float(d.x); r.tb = __uint_as_float(d.y); }
    return r;
}
// all 64 bytes: the ray, the full throughput and the broadcast pre-test of the kernel that made the ray
__device__ __forceinline__ RayRec ray_rec_load_pre(const PathRay* p, RayPre& pre) {
    const uint4* q = reinterpret_cast<const uint4*>(p);
    const uint4 a = rec_ld(q), b = rec_ld(q + 1), c = rec_ld(q + 2), d = rec_ld(q + 3);
    RayRec r;
    r.ox = __hiloint2double((int)a.y, (int)a.x); r.oy = __hiloint2double((int)a.w, (int)a.z);
    r.oz = __hiloint2double((int)b.y, (int)b.x); r.dx = __uint_as_float(b.z); r.dy = __uint_as_float(b.w);
    r.dz = __uint_as_float(c.x); r.pixel = c.y; r.sample = c.z; r.tr = __uint_as_float(c.w);
    r.tg = __uint_as_float(d.x); r.tb = __uint_as_float(d.y);
    pre.prim = d.z; pre.t_signed = __uint_as_float(d.w);
    return r;
}
__device__ __forceinline__ uint2 ray_rec_pretest(const PathRay* p) {
    return *reinterpret_cast<const uint2*>(reinterpret_cast<const char*>(p) + 56);
}
template <bool SPH = true>
__device__ __forceinline__ RayPre pretest_of(const DeviceScene& s, const RayRec& r) {
    RayF f;
    f.fox = (float)r.ox; f.foy = (float)r.oy; f.foz = (float)r.oz; f.dx = r.dx; f.dy = r.dy; f.dz = r.dz;
    RayPre p;
    pretest_ray<SPH>(s, f, p.prim, p.t_signed);
    return p;
}
__device__ __forceinline__ void ray_rec_store(PathRay* p, const RayRec& r, const RayPre& pre) {
    uint4* q = reinterpret_cast<uint4*>(p);
    rec_st(q, make_uint4((uint32_t)__double2loint(r.ox), (uint32_t)__double2hiint(r.ox), (uint32_t)__double2loint(r.oy), (uint32_t)__double2hiint(r.oy)));
    rec_st(q + 1, make_uint4((uint32_t)__double2loint(r.oz), (uint32_t)__double2hiint(r.oz), __float_as_uint(r.dx), __float_as_uint(r.dy)));
    rec_st(q + 2, make_uint4(__float_as_uint(r.dz), r.pixel, r.sample, __float_as_uint(r.tr)));
    rec_st(q + 3, make_uint4(__float_as_uint(r.tg), __float_as_uint(r.tb), pre.prim, __float_as_uint(pre.t_signed)));
}
__device__ __forceinline__ void hit_rec_store(PathHit* p, uint32_t prim, float t32, float u, float v, double t) {
    uint4* q = reinterpret_cast<uint4*>(p);
    rec_st(q, make_uint4(prim, __float_as_uint(t32), __float_as_uint(u), __float_as_uint(v)));
    rec_st(q + 1, make_uint4((uint32_t)__double2loint(t), (uint32_t)__double2hiint(t), 0u, 0u));
}

__device__ __forceinline__ RayD ray_of(const RayRec& r) {
    RayD ray;
    ray.o = x3(r.ox, r.oy, r.oz);
    ray.d = x3((double)r.dx, (double)r.dy, (double)r.dz);
    return ray;
}

// warp-aggregated append: one atomicAdd per (warp, destination counter)
__device__ __forceinline__ void queue_append(uint32_t* queue, uint32_t* counter, uint32_t value, bool active) {
    const unsigned mask = __ballot_sync(0xFFFFFFFFu, active);
    if (mask == 0) return;
    const int lane = threadIdx.x & 31;
    const int leader = __ffs(mask) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(counter, (uint32_t)__popc(mask));
    base = __shfl_sync(0xFFFFFFFFu, base, leader);
    if (active) queue[base + (uint32_t)__popc(mask & ((1u << lane) - 1u))] = value;
}

// A new ray goes to the trace queue of `bounce` (dest 0) or, when the broadcast pre-test already has its
// closest hit and the BVH is out of reach, straight to the shade queue of the hit's material class
// (dest 1 + class; its hit IS the pre-test in the ray record, no hit record is written). One match_any + one
// atomic per (warp, destination).
__device__ __forceinline__ int route_new_ray(const DeviceScene& s, const RayPre& pre, bool bypass) {
    if (!bypass || !(pre.t_signed < 0.f)) return 0;
    int cls = CLASS_EMISSIVE; // a miss: background / sky / sun are handled by the emissive kernel
    if (pre.prim != HIT_MISS) cls = (int)((__float_as_uint(__ldg(&s.prims[pre.prim].c.w)) >> PRIM_CLASS_SHIFT) & 7u);
    // no hit record: the shade kernel finds (primitive, t) in the ray record it reads anyway (surface_for_slot)
    return 1 + cls;
}
// Record positions are reserved per warp in chunks (one same-address atomic per chunk, not per warp
// iteration): the position of a new ray is known from warp-local cursors, so its 64-byte store does not
// wait for an atomic round trip. Cursor state is warp-uniform.
struct RecordCursor {
    uint32_t t_next = 0, t_end = 0; // reserved range in the trace region (from index 0 upwards)
    uint32_t b_next = 0, b_end = 0; // reserved range in the bypass region (logical; stored from the far end downwards)
    uint32_t chunk = 32;
    uint32_t dropped = 0;           // rays dropped as known misses into blackness (count_dropped)
};
__device__ __forceinline__ uint32_t record_chunk(uint32_t n, uint32_t total_warps) {
    // at most half of a kernel's output can be lost to unused chunk ends (ray_slots = 1.5 x capacity covers it)
    const uint32_t c = (n / total_warps / 2u) & ~31u;
    return c < 32u ? 32u : (c > 256u ? 256u : c);
}
// `limit`: records the region holds (0xFFFFFFFF = not checked here). A chunk that would end past it is not used:
// the lanes it would have served get no position (0xFFFFFFFF) and `overflow` is set.
__device__ __forceinline__ uint32_t cursor_take(uint32_t* counter, unsigned mask, bool mine, uint32_t chunk, uint32_t& next, uint32_t& end,
                                                uint32_t limit, bool& overflow) {
    const int lane = threadIdx.x & 31;
    const uint32_t need = (uint32_t)__popc(mask), avail = end - next;
    uint32_t nb = 0;
    if (avail < need) {
        if (lane == 0) nb = atomicAdd(counter, chunk);
        nb = __shfl_sync(0xFFFFFFFFu, nb, 0);
    }
    const uint32_t rank = (uint32_t)__popc(mask & ((1u << lane) - 1u));
    uint32_t at = rank < avail ? next + rank : nb + (rank - avail);
    if (avail < need) {
        if (nb > limit || chunk > limit - nb) { // (warp-uniform) the region is full
            overflow = true;
            if (rank >= avail) at = 0xFFFFFFFFu;
            next = end = 0;
        } else { next = nb + (need - avail); end = nb + chunk; }
    } else next += need;
    return mine ? at : 0xFFFFFFFFu;
}
// Position of a new ray of `bounce` in rays[bounce & 1] (0xFFFFFFFF when dest < 0, or when the pool is full: the
// caller must not store then); rays that bypass the trace kernel (dest 1 + class) also get their shade-queue entry
// (one atomic per class present in the warp). All lanes of the warp must call.
__device__ __forceinline__ uint32_t routed_append(const PathPool& pool, uint32_t bounce, int dest, RecordCursor& rc) {
    const int lane = threadIdx.x & 31;
    uint32_t* cnt = pool.counters + bounce * RT_COUNTERS_PER_BOUNCE;
    const unsigned to_trace = __ballot_sync(0xFFFFFFFFu, dest == 0), to_shade = __ballot_sync(0xFFFFFFFFu, dest > 0);
    const bool carry = bounce == pool.carry_bounce; // rays[]: bounded by pool_guard at kernel start
    uint32_t idx = 0xFFFFFFFFu;
    bool overflow = false;
    if (to_trace) {
        const uint32_t at = cursor_take(cnt, to_trace, dest == 0, rc.chunk, rc.t_next, rc.t_end, carry ? pool.carry_t_cap : 0xFFFFFFFFu, overflow);
        if (dest == 0) idx = at;
    }
    if (to_shade) {
        const uint32_t at = cursor_take(cnt + RT_CNT_BYPASS, to_shade, dest > 0, rc.chunk, rc.b_next, rc.b_end, carry ? pool.carry_b_cap : 0xFFFFFFFFu, overflow);
        if (dest > 0 && at != 0xFFFFFFFFu) idx = slots_of(pool, bounce) - 1u - at;
        const bool queued = dest > 0 && idx != 0xFFFFFFFFu;
        const unsigned peers = __match_any_sync(0xFFFFFFFFu, queued ? dest : 0);
        const int qleader = __ffs(peers) - 1;
        const uint32_t n_peers = (uint32_t)__popc(peers);
        uint32_t qbase = 0;
        if (queued && lane == qleader) qbase = atomicAdd(cnt + RT_CNT_CLASS + (dest - 1), n_peers);
        qbase = __shfl_sync(0xFFFFFFFFu, qbase, qleader);
        if (queued) {
            const uint32_t qcap = queue_cap_of(pool, bounce);
            if (qbase > qcap || n_peers > qcap - qbase) { pool_fail(pool, carry ? RT_POOL_ERR_CARRY_QUEUE : RT_POOL_ERR_QUEUE); idx = 0xFFFFFFFFu; }
            else class_queue_of(pool, bounce, dest - 1)[qbase + (uint32_t)__popc(peers & ((1u << lane) - 1u))] = idx;
        }
    }
    if (overflow && lane == 0) pool_fail(pool, RT_POOL_ERR_CARRY_RAYS);
    return idx;
}
__device__ __forceinline__ void accumulate(float4* accum, uint32_t pixel, f3 v) {
    if (v.x != 0.f) atomicAdd(&accum[pixel].x, v.x);
    if (v.y != 0.f) atomicAdd(&accum[pixel].y, v.y);
    if (v.z != 0.f) atomicAdd(&accum[pixel].z, v.z);
}

// A new ray whose closest hit the broadcast pre-test knows completely (BVH out of reach) and that lands on a quad / sphere
// with an Emissive material and a CONSTANT emission map: Emissive::scatter ends the path (emissive.rs:24-34), so all that is
// left of it is `throughput * emission` if the hit is on the front face (HitRecord::new, object.rs:55-60). The kernel that
// made the ray adds that itself and the ray is never written (counted with RT_CNT_DROPPED: queried, not queued).
__device__ __forceinline__ bool inline_emission(const DeviceScene& s, const RayRec& r, const RayPre& pre, f3& emission) {
    // (the record says whether the material is Emissive over a constant map, and carries the value: flatten.cpp)
    const float4* bp = reinterpret_cast<const float4*>(s.bcast + (pre.prim - (uint32_t)s.global_first));
    const float4 q = __ldg(bp), n = __ldg(bp + 1);
    const uint32_t flags = __float_as_uint(q.w);
    if (!(flags & BCAST_EMIT_CONST)) return false;
    bool front;
    if (flags & PRIM_SPHERE) { // outward normal (p - c) / radius: a negative radius turns it inward (sphere.rs:64-66)
        const float t = -pre.t_signed;
        const f3 pc((float)r.ox + r.dx * t - q.x, (float)r.oy + r.dy * t - q.y, (float)r.oz + r.dz * t - q.z);
        front = (dot(f3(r.dx, r.dy, r.dz), pc) < 0.f) != ((flags & BCAST_INWARD) != 0u);
    } else front = dot(f3(r.dx, r.dy, r.dz), f3(n.x, n.y, n.z)) < 0.f; // n = e1 x e2 = 4 u x v: the Plane's normal direction
    emission = front ? f3(n.w, __ldg(&bp[2].w), __ldg(&bp[3].w)) : f3(0.f, 0.f, 0.f);
    return true;
}

// rays dropped as known misses (RT_CNT_DROPPED): counted in a warp-uniform register, ONE atomic per warp when the kernel
// ends (routed_flush) — a same-address atomic per loop iteration of every warp cost 4 ms per step; all lanes must call
__device__ __forceinline__ void count_dropped(RecordCursor& rc, bool dropped) {
    rc.dropped += (uint32_t)__popc(__ballot_sync(0xFFFFFFFFu, dropped));
}
// end of the kernel: the unused end of the warp's trace-region chunk becomes dead records the trace kernel skips
__device__ __forceinline__ void routed_flush(const PathPool& pool, uint32_t bounce, const RecordCursor& rc) {
    const int lane = threadIdx.x & 31;
    if (rc.dropped != 0u && lane == 0) atomicAdd(pool.counters + bounce * RT_COUNTERS_PER_BOUNCE + RT_CNT_DROPPED, rc.dropped);
    if (rc.t_end == rc.t_next) return;
    PathRay* rays = rays_of(pool, bounce);
    for (uint32_t p = rc.t_next + (uint32_t)lane; p < rc.t_end; p += 32u)
        reinterpret_cast<uint4*>(rays + p)[2] = make_uint4(0u, RT_DEAD_PIXEL, 0u, 0u);
    if (lane == 0) atomicAdd(pool.counters + bounce * RT_COUNTERS_PER_BOUNCE + RT_CNT_DEAD, rc.t_end - rc.t_next);
}

// Camera::get_ray (camera.rs:260-280) for pixel (px, py), stratum cell (sx, sy), jitter (rx, ry)
// and the two N(0,1) lens draws (nx, ny)
// DOF = false: the caller guarantees a pinhole camera (aperture_radius < 0): the lens code is compiled out
template <bool DOF = true>
__device__ __forceinline__ void camera_ray(const CameraDev& cam, uint32_t px, uint32_t py, uint32_t sx, uint32_t sy, double rx, double ry,
                                           double nx, double ny, double o[3], double d[3]) {
    // pixel_sample_square (camera.rs:334-341)
    const double jx = ((double)sx + rx) * cam.inv_sqrt_spt - 0.5, jy = ((double)sy + ry) * cam.inv_sqrt_spt - 0.5;
    double ps[3];
    for (int a = 0; a < 3; a++)
        ps[a] = cam.first_pixel[a] + cam.du[a] * (double)px + cam.dv[a] * (double)py + (cam.du[a] * jx + cam.dv[a] * jy);
    if (DOF && cam.aperture_radius >= 0.0) { // defocus_disk_sample (camera.rs:345-349): ON the unit circle
        const double len = sqrt(nx * nx + ny * ny);
        const double vx = nx / len, vy = ny / len;
        for (int a = 0; a < 3; a++) o[a] = cam.position[a] + (cam.bu[a] * vx + cam.bv[a] * vy) * cam.aperture_radius;
    } else for (int a = 0; a < 3; a++) o[a] = cam.position[a];
    for (int a = 0; a < 3; a++) d[a] = ps[a] - o[a];
}

// ---- K1 ---------------------------------------------------------------------------------------
// paths [0, n) of the launch are sample (first_sample + i / npix) of pixel (i % npix): a batch is whole images of samples,
// which keeps the index arithmetic in 32 bits (a 64-bit division per path was ~15 % of this kernel's instructions)
template <bool DOF, bool SPH = true>
__global__ void __launch_bounds__(256) k_raygen(DeviceScene scene, PathPool pool, CameraDev cam, RenderConst rc, uint32_t first_sample, uint32_t n,
                                                uint32_t sample_begin, uint32_t bypass, float4* accum) {
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t npix = (uint32_t)cam.width * (uint32_t)cam.height;
    const uint32_t n_round = (n + 31u) & ~31u; // whole warps stay together for the queue append
    RecordCursor cur;
    cur.chunk = record_chunk(n, stride / 32u);
    if (*pool.error != 0u || !pool_guard(pool, 0u, n)) return;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
        int dest = -1;
        RayRec r;
        RayPre pre;
        if (i < n) {
            const uint32_t si = i / npix;
            const uint32_t s = sample_begin + first_sample + si; // global sample index: tid*k*k + sy*k + sx
            const uint32_t pix = i - si * npix;
            const uint32_t px = pix % (uint32_t)cam.width, py = pix / (uint32_t)cam.width;
            const uint32_t k = (uint32_t)cam.sqrt_spt;
            const uint32_t cell = s % (k * k), sy = cell / k, sx = cell % k;
            PhiloxStreamT<!DOF> rng(rc.seed, pix, s, 0u, 0u); // (pinhole: two draws, plenty of registers)
            const double rx = (double)rng.uniform(), ry = (double)rng.uniform();
            double nx = 0.0, ny = 0.0;
            if (DOF && cam.aperture_radius >= 0.0) { nx = (double)rng.normal(); ny = (double)rng.normal(); }
            double o[3], d[3];
            camera_ray<DOF>(cam, px, py, sx, sy, rx, ry, nx, ny, o, d);
            r.ox = o[0]; r.oy = o[1]; r.oz = o[2]; r.dx = (float)d[0]; r.dy = (float)d[1]; r.dz = (float)d[2];
            r.pixel = pix; r.sample = s; r.tr = r.tg = r.tb = 1.f;
            pre = pretest_of<SPH>(scene, r);
            dest = route_new_ray(scene, pre, bypass != 0);
            if ((rc.flags & RT_FLAG_DROP_MISSES) && dest == 1 + CLASS_EMISSIVE && pre.prim == HIT_MISS) dest = -2; // a known miss into blackness
            else if ((rc.flags & RT_FLAG_INLINE_EMISSION) && rc.max_depth >= 1 && dest == 1 + CLASS_EMISSIVE && pre.prim != HIT_MISS) {
                // (max_depth 0: ray_color(depth == 0) returns black before looking at the hit, camera.rs:290-292)
                f3 em;
                if (inline_emission(scene, r, pre, em)) { accumulate(accum, r.pixel, em); dest = -2; } // a light seen directly
            }
        }
        count_dropped(cur, dest == -2);
        const uint32_t at = routed_append(pool, 0u, dest, cur);
        if (at != 0xFFFFFFFFu) ray_rec_store(pool.rays[0] + at, r, pre);
    }
    routed_flush(pool, 0u, cur);
}

// parity hook: Camera::get_ray with injected random numbers
__global__ void k_generate_rays(CameraDev cam, const int32_t* px, const int32_t* sxy, const double* xi, long long n, double* out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double o[3], d[3];
    camera_ray(cam, (uint32_t)px[i * 2], (uint32_t)px[i * 2 + 1], (uint32_t)sxy[i * 2], (uint32_t)sxy[i * 2 + 1], xi[i * 4], xi[i * 4 + 1],
               xi[i * 4 + 2], xi[i * 4 + 3], o, d);
    for (int a = 0; a < 3; a++) { out[i * 6 + a] = o[a]; out[i * 6 + 3 + a] = d[a]; }
}

__device__ __forceinline__ int material_class(const DeviceScene& s, int material) {
    if (material < 0) return CLASS_EMISSIVE;
    return material_class_of_kind(s.materials[material].kind);
}

// ---- K2 ---------------------------------------------------------------------------------------
template <bool EXACT>
__global__ void __launch_bounds__(128) k_trace(DeviceScene s, PathPool pool, RenderConst rc, uint32_t bounce) {
    const uint32_t* cnt_in = pool.counters + bounce * RT_COUNTERS_PER_BOUNCE;
    uint32_t* cnt_class = pool.counters + bounce * RT_COUNTERS_PER_BOUNCE + RT_CNT_CLASS;
    const uint32_t n = *pool.error != 0u ? 0u : *cnt_in;
    const PathRay* rays = rays_of(pool, bounce);
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t n_round = (n + 31u) & ~31u; // whole warps stay together for the ballots
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
        const bool active = i < n;
        int cls = -1;
        uint32_t slot = 0;
        if (active) {
            slot = i;
            const RayRec rr = ray_rec_load<false>(rays + slot);
          if (rr.pixel != RT_DEAD_PIXEL) { // (unused end of a reserved chunk)
            const RayD ray = ray_of(rr);
            U2 stack[RT_STACK_SIZE];
            Hit h;
            bool hit = closest_hit<EXACT>(s, ray, 0.001, INFINITY, h, stack);
            uint32_t prim = hit ? h.prim : HIT_MISS;
            double t = h.t;
            float hu = h.u, hv = h.v;
            int material = -1;
            for (int vi = 0; vi < s.n_volumes; vi++) { // constant media compete with the surfaces
                double tv;
                if (volume_hit(s, vi, ray, 0.001, t, rc.seed, rr.pixel, rr.sample, bounce, tv)) {
                    t = tv; prim = HIT_VOLUME; hu = __uint_as_float((uint32_t)vi); hit = true;
                }
            }
            if (prim == HIT_VOLUME) material = s.volumes[__float_as_uint(hu)].material;
            else if (hit) material = s.instances[__float_as_uint(s.prims[prim].a.w)].material;
            else {
                const int inf = resolve_infinite(s, ray);
                if (inf >= 0) { prim = HIT_INFINITE; hu = __uint_as_float((uint32_t)inf); }
            }
            hit_rec_store(pool.hits + slot, prim, (float)t, hu, hv, t);
            cls = material_class(s, material);
          }
        }
#pragma unroll
        for (int c = 0; c < N_CLASSES; c++) queue_append(class_queue_of(pool, bounce, c), cnt_class + c, slot, active && cls == c);
    }
}

} // namespace rt
#include "trace_fast.cuh"
namespace rt {

// The fp32 traversal only SELECTS quads and spheres (world-space fp32 copies in the leaf record); the t the
// integrator uses comes from the reference's f64 object-space formula (plane.rs:66-90, sphere.rs:40-62),
// accepted when it confirms the fp32 value (same root, same side of the interval).
template <bool SPH = true>
__device__ __forceinline__ double refine_analytic_t(const DeviceScene& s, const RayD& ray, uint32_t prim, double t) {
    if (!(__float_as_uint(s.prims[prim].c.w) & (PRIM_QUAD | PRIM_SPHERE))) return t;
    const APrim& a = s.aprims[__float_as_uint(s.prims[prim].b.w)];
    double t64;
    if (world_aprim_hit<SPH>(s, a, ray, 0.001, INFINITY, t64) && fabs(t64 - t) <= 1e-3 * fabs(t) + 1e-3) return t64;
    return t;
}

// parity hook: the broadcast pre-test for host-supplied rays (rtgpu_trace_fast)
__global__ void k_pretest_pool(DeviceScene s, PathPool pool, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const RayRec r = ray_rec_load<true>(pool.rays[0] + i);
    ray_rec_store(pool.rays[0] + i, r, pretest_of(s, r));
}

// parity hook: the refinement above applied to the hit records of a traced queue (rtgpu_trace_fast)
__global__ void k_refine_hits(DeviceScene s, PathPool pool, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    PathHit& h = pool.hits[i];
    if (h.prim >= (uint32_t)s.n_prims) return;
    h.t = refine_analytic_t(s, ray_of(ray_rec_load<false>(pool.rays[0] + i)), h.prim, h.t);
}

// rebuild the HitRecord of a traced path
// `bypassed`: the path never entered the trace kernel (route_new_ray); its hit is the broadcast pre-test
// carried by the ray record, and the 32-byte hit record is neither written nor read (one gather and
// one scattered store less for the ~73 % of cornell_dragon's rays that take this route).
// NMAP: the hit's material may carry a normal map (only Glossy and NormalDebug can, and only when the scene has one at all):
// the tangent-frame code of surface_from_prim is compiled out of every other shade kernel.
// INF: the queue this kernel reads can hold misses and sky / sun hits (they are routed to CLASS_EMISSIVE only: route_new_ray, k_trace_fast);
// the other classes' kernels are compiled without the sky / sun resolution.
template <bool NMAP, bool INF = true, bool SPH = true>
__device__ __forceinline__ bool surface_for_slot(const DeviceScene& s, const PathPool& pool, uint32_t slot, const RayD& ray, const RayPre& pre,
                                                 bool bypassed, Surface<float>& sf) {
    uint4 h0, h1;
    if (bypassed) {
        const float tp = -pre.t_signed;
        const double td = (double)tp;
        h0 = make_uint4(pre.prim, __float_as_uint(tp), 0u, 0u);
        h1 = make_uint4((uint32_t)__double2loint(td), (uint32_t)__double2hiint(td), 0u, 0u);
    } else {
        const uint4* hq = reinterpret_cast<const uint4*>(pool.hits + slot);
        h0 = rec_ld(hq); h1 = rec_ld(hq + 1);
    }
    const float4 h = make_float4(__uint_as_float(h0.x), __uint_as_float(h0.y), __uint_as_float(h0.z), __uint_as_float(h0.w));
    const uint32_t prim = h0.x;
    if (!INF && prim >= HIT_INFINITE && prim != HIT_VOLUME) return false; // (never queued for this class)
    if (prim == HIT_MISS) { // k_trace_fast leaves the sky / sun decision (list-order rules, sky.rs / sun.rs) to this kernel
        if (s.n_infinite == 0) return false;
        const int inf = resolve_infinite(s, ray);
        if (inf < 0) return false;
        surface_from_infinite<float>(s, ray, inf, sf);
        return true;
    }
    if (prim == HIT_INFINITE) { surface_from_infinite<float>(s, ray, (int)__float_as_uint(h.z), sf); return true; }
    const double t = __hiloint2double((int)h1.y, (int)h1.x);
    if (prim == HIT_VOLUME) { surface_from_volume<float>(s, ray, (int)__float_as_uint(h.z), t, sf); return true; }
    const int inst = (int)__float_as_uint(s.prims[prim].a.w);
    const int mat = s.instances[inst].material;
    const bool need_tangents = NMAP && mat >= 0 && s.materials[mat].normal_map >= 0;
    const double t_hit = refine_analytic_t<SPH>(s, ray, prim, t);
    surface_from_prim<float, SPH>(s, ray, prim, t_hit, (double)h.z, (double)h.w, need_tangents, sf);
    return true;
}


// ---- K3 ---------------------------------------------------------------------------------------
// SPH = false: a scene without analytic spheres (primitives or lights): their f64 code (roots, acos / atan2 of the uv) is compiled out.
template <int CLASS, int MIN_CTAS, bool NMAP = true, bool CTEX = false, bool SPH = true>
__global__ void __launch_bounds__(128, MIN_CTAS) k_shade(DeviceScene s, PathPool pool, RenderConst rc, CameraDev cam, uint32_t bounce, float4* accum,
                                                         uint32_t bypass) {
    const uint32_t n = pool.counters[bounce * RT_COUNTERS_PER_BOUNCE + RT_CNT_CLASS + CLASS];
    const uint32_t* q_in = class_queue_of(pool, bounce, CLASS);
    const PathRay* rays_in = rays_of(pool, bounce);
    PathRay* rays_out = rays_of(pool, bounce + 1u);
    const bool last = (int)bounce + 1 >= rc.max_depth; // ray_color(depth == 0) contributes nothing
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t n_round = (n + 31u) & ~31u;
    RecordCursor cur;
    cur.chunk = record_chunk(n, stride / 32u);
    if (n == 0 || *pool.error != 0u) return; // (after a failed pool check the counters no longer describe the arrays)
    if (!last) { // room for what ALL the shade kernels of this bounce can emit (the counters of the input bounce are complete)
        const uint32_t* cin = pool.counters + bounce * RT_COUNTERS_PER_BOUNCE + RT_CNT_CLASS;
        uint32_t n_in = 0;
#pragma unroll
        for (int c = 0; c < N_CLASSES; c++) n_in += cin[c];
        if (!pool_guard(pool, bounce + 1u, n_in)) return;
    }
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
        const bool active = i < n;
        int dest = -1; // of the continuing path: trace queue or (bypass) a shade queue of the next bounce
        uint32_t slot = 0;
        RayRec rr;
        RayPre pre;
        if (active) {
            slot = q_in[i];
            RayPre made;
            rr = ray_rec_load_pre(rays_in + slot, made);
            const RayD ray = ray_of(rr);
            const f3 thr(rr.tr, rr.tg, rr.tb);
            const uint32_t pixel = rr.pixel;
            Surface<float> sf;
            if (!surface_for_slot<NMAP && (CLASS == CLASS_GLOSSY || CLASS == CLASS_EMISSIVE), CLASS == CLASS_EMISSIVE, SPH>(s, pool, slot, ray, made, bypass != 0 && made.t_signed < 0.f, sf)) { // camera.rs:331: background colour
                accumulate(accum, pixel, thr * f3(cam.background[0], cam.background[1], cam.background[2]));
            } else {
                PhiloxStreamT<CTEX> rng(rc.seed, pixel, rr.sample, bounce, 1u);
                Bounce b;
                shade_bounce<CLASS, NMAP, CTEX, SPH>(s, ray, sf, rc.light_bias, rng, b, (rc.flags & RTGPU_FLAG_FIXED) != 0);
                if (b.kind != BOUNCE_ABSORBED) accumulate(accum, pixel, thr * b.emission);
                if ((b.kind == BOUNCE_PDF || b.kind == BOUNCE_RAY) && !last) {
                    const f3 nt = thr * b.weight;
                    if (nt.x > 0.f || nt.y > 0.f || nt.z > 0.f) {
                        rr.ox = b.origin.x; rr.oy = b.origin.y; rr.oz = b.origin.z;
                        rr.dx = b.dir.x; rr.dy = b.dir.y; rr.dz = b.dir.z;
                        rr.tr = nt.x; rr.tg = nt.y; rr.tb = nt.z;
                        pre = pretest_of<SPH>(s, rr);
                        dest = route_new_ray(s, pre, bypass != 0);
                        if ((rc.flags & RT_FLAG_DROP_MISSES) && dest == 1 + CLASS_EMISSIVE && pre.prim == HIT_MISS) dest = -2;
                        else if ((rc.flags & RT_FLAG_INLINE_EMISSION) && dest == 1 + CLASS_EMISSIVE && pre.prim != HIT_MISS) {
                            f3 em;
                            if (inline_emission(s, rr, pre, em)) { accumulate(accum, pixel, nt * em); dest = -2; } // the path ends on a light
                        }
                    }
                }
            }
        }
        count_dropped(cur, dest == -2);
        const uint32_t at = routed_append(pool, bounce + 1u, dest, cur);
        if (at != 0xFFFFFFFFu) ray_rec_store(rays_out + at, rr, pre);
    }
    routed_flush(pool, bounce + 1u, cur);
}

// adds the per-bounce ray counts of a finished batch into the running statistics
// (the statistics have 64 rows: bounces past 62 are summed into row 63 — --max-depth itself is not limited, config.rs:76)
__global__ void k_stats(const uint32_t* counters, int b_begin, int n_bounces, unsigned long long* rays_per_bounce) {
    for (int b = b_begin + (int)threadIdx.x; b < n_bounces; b += (int)blockDim.x) {
        const uint32_t* cnt = counters + b * RT_COUNTERS_PER_BOUNCE;
        unsigned long long all = 0; // every ray of the bounce ends in exactly one class queue
        for (int c = 0; c < N_CLASSES; c++) all += cnt[RT_CNT_CLASS + c];
        all += cnt[RT_CNT_DROPPED]; // misses into blackness: queried, never queued
        const int row = b < 63 ? b : 63;
        atomicAdd(&rays_per_bounce[row], all); // (atomics: lanes run concurrently)
        atomicAdd(&rays_per_bounce[64 + row], (unsigned long long)(cnt[0] - cnt[RT_CNT_DEAD])); // of those, the rays that walked the BVH
    }
}

// ---- K5 ---------------------------------------------------------------------------------------
// fused multi-GPU reduction + resolve on device 0: sums the peers' fp32 accumulators through
// NVLink peer pointers (replaces the sequential `buf += thread_buf` of camera.rs:243-255)
__global__ void k_reduce_resolve(const float4* const* accums, int n_dev, double* out_rgba, uint32_t npix, double inv_spp) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix) return;
    double r = 0, g = 0, b = 0;
    for (int d = 0; d < n_dev; d++) { const float4 a = accums[d][i]; r += a.x; g += a.y; b += a.z; }
    out_rgba[(size_t)i * 4 + 0] = r * inv_spp; out_rgba[(size_t)i * 4 + 1] = g * inv_spp; out_rgba[(size_t)i * 4 + 2] = b * inv_spp;
    out_rgba[(size_t)i * 4 + 3] = 0.0;
}

// output.rs:23-50 + tonemapping/aces.rs:5-33 / clamp.rs in f64 like the reference
__device__ __forceinline__ double srgb_oetf(double x) { return x < 0.0031308 ? x * 12.92 : pow(x, 1.0 / 2.4) * 1.055 - 0.055; }
__device__ __forceinline__ double clamp01(double x) { return x < 0.0 ? 0.0 : (x > 1.0 ? 1.0 : x); }
__device__ __forceinline__ uint8_t as_u8(double x) { return !(x > 0.0) ? 0 : (x >= 255.0 ? 255 : (uint8_t)x); }
__device__ __forceinline__ void tonemap_pixel(double c0, double c1, double c2, int tonemap, uint8_t* out) {
    double v[3];
    if (tonemap == 1) {
        const double a[3] = {__dadd_rn(__dadd_rn(__dmul_rn(0.59719, c0), __dmul_rn(0.35458, c1)), __dmul_rn(0.04823, c2)),
                             __dadd_rn(__dadd_rn(__dmul_rn(0.07600, c0), __dmul_rn(0.90834, c1)), __dmul_rn(0.01566, c2)),
                             __dadd_rn(__dadd_rn(__dmul_rn(0.02840, c0), __dmul_rn(0.13383, c1)), __dmul_rn(0.83777, c2))};
        double b[3];
        for (int r = 0; r < 3; r++) {
            const double x = a[r];
            const double num = __dsub_rn(__dmul_rn(x, __dadd_rn(x, 0.0245786)), 0.000090537);
            const double den = __dadd_rn(__dmul_rn(x, __dadd_rn(__dmul_rn(x, 0.983729), 0.4329510)), 0.238081);
            b[r] = __ddiv_rn(num, den);
        }
        v[0] = clamp01(__dadd_rn(__dadd_rn(__dmul_rn(1.60475, b[0]), __dmul_rn(-0.53108, b[1])), __dmul_rn(-0.07367, b[2])));
        v[1] = clamp01(__dadd_rn(__dadd_rn(__dmul_rn(-0.10208, b[0]), __dmul_rn(1.10813, b[1])), __dmul_rn(-0.00605, b[2])));
        v[2] = clamp01(__dadd_rn(__dadd_rn(__dmul_rn(-0.00327, b[0]), __dmul_rn(-0.07276, b[1])), __dmul_rn(1.07602, b[2])));
    } else { v[0] = clamp01(c0); v[1] = clamp01(c1); v[2] = clamp01(c2); }
    for (int r = 0; r < 3; r++) out[r] = as_u8(__dmul_rn(srgb_oetf(v[r]), 255.999));
}
__global__ void k_tonemap(const double* rgba, uint8_t* out_rgb, uint32_t npix, int tonemap) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix) return;
    tonemap_pixel(rgba[(size_t)i * 4], rgba[(size_t)i * 4 + 1], rgba[(size_t)i * 4 + 2], tonemap, out_rgb + (size_t)i * 3);
}
// K5 without the f64 Buffer in between: peers' sums (NVLink peer reads) -> / spp -> tonemap -> RGB8 (rtgpu_render_srgb8)
__global__ void k_resolve_tonemap(const float4* const* accums, int n_dev, uint8_t* out_rgb, uint32_t npix, double inv_spp, int tonemap) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix) return;
    double r = 0, g = 0, b = 0;
    for (int d = 0; d < n_dev; d++) { const float4 a = accums[d][i]; r += a.x; g += a.y; b += a.z; }
    tonemap_pixel(r * inv_spp, g * inv_spp, b * inv_spp, tonemap, out_rgb + (size_t)i * 3);
}

// ---- parity hooks -----------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_trace_exact(DeviceScene s, const double* rays, long long n, RtgpuHit* out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    RayD ray;
    ray.o = x3(rays[i * 6], rays[i * 6 + 1], rays[i * 6 + 2]);
    ray.d = x3(rays[i * 6 + 3], rays[i * 6 + 4], rays[i * 6 + 5]);
    U2 stack[RT_STACK_SIZE];
    Hit h;
    RtgpuHit o;
    memset(&o, 0, sizeof o);
    Surface<double> sf;
    bool have = false;
    if (closest_hit<true>(s, ray, 0.001, INFINITY, h, stack)) {
        const bool analytic = (__float_as_uint(s.prims[h.prim].c.w) & PRIM_ANALYTIC) != 0;
        surface_from_prim<double>(s, ray, h.prim, h.t, analytic ? 0.0 : h.ud, analytic ? 0.0 : h.vd, true, sf);
        have = true;
    } else {
        const int inf = resolve_infinite(s, ray);
        if (inf >= 0) { surface_from_infinite<double>(s, ray, inf, sf); have = true; }
    }
    if (have) {
        o.object = s.instances[sf.instance].object; o.instance = sf.instance; o.primitive = sf.primitive;
        o.front_face = sf.front_face ? 1 : 0; o.t = sf.t;
        o.pos[0] = sf.pos.x; o.pos[1] = sf.pos.y; o.pos[2] = sf.pos.z;
        o.normal[0] = sf.normal.x; o.normal[1] = sf.normal.y; o.normal[2] = sf.normal.z;
        o.uv[0] = sf.u; o.uv[1] = sf.v;
    } else { o.object = -1; o.instance = -1; }
    out[i] = o;
}

} // namespace rt
#include "hooks.cuh"
namespace rt {

__global__ void __launch_bounds__(128) k_bounce_hook(DeviceScene s, float light_bias, const double* rays, const double* xi, int n_xi,
                                                     long long n, BounceRecord* out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    BounceRecord o;
    bounce_hook_one(s, light_bias, rays + i * 6, xi + i * n_xi, n_xi, o);
    out[i] = o;
}

} // namespace rt
