// flatten.cpp — walks the reference-shaped object tree once (DFS pre-order, the same
// order the oracle numbers leaf instances in) and emits the flat tables of scene_types.h.
// Compile with -ffp-contract=off: the derived plane/sphere constants must be bit-equal to
// the reference's f64 arithmetic (src/object/plane.rs:29-62).
#include "flatten.h"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <limits>
#include <cstdlib>
#include <thread>

// host cores this process may count on: under torchrun (one process per GPU) every rank runs the same set-up at the same
// time, so the cores are divided by LOCAL_WORLD_SIZE (8 ranks x 16 threads on 32 cores made the "Ready" phase slower at N = 8)
static inline unsigned rt_host_cores() {
    unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    if (const char* e = std::getenv("LOCAL_WORLD_SIZE")) { const int w = std::atoi(e); if (w > 1) hw = std::max(2u, hw / (unsigned)w); }
    return hw;
}


namespace rt {
namespace {

struct V3 { double x, y, z; };
static V3 sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static V3 add(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static V3 mul(V3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
static V3 divs(V3 a, double s) { return {a.x / s, a.y / s, a.z / s}; }
static double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
static double length(V3 a) { return std::sqrt(dot(a, a)); }

// Mat4 * point (w = 1), row-major (mat4.rs:345-352)
static V3 xform_point(const double* m, V3 p) {
    return {m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3] * 1.0, m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7] * 1.0,
            m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11] * 1.0};
}

static float round_down(double v) { float f = (float)v; if ((double)f > v) f = std::nextafterf(f, -INFINITY); return f; }
static float round_up(double v) { float f = (float)v; if ((double)f < v) f = std::nextafterf(f, INFINITY); return f; }

struct Walker {
    const RtgpuSceneDesc& d;
    HostScene& out;
    std::string& err;
    std::vector<int32_t> chain_stack;             // xform indices, outermost first
    std::vector<GateBox> gate_stack;              // reference list/bvh boxes since the last Transform
    std::vector<PrimBox> boxes;                   // per prim (input order)
    PodVector<Prim> prims_in;                     // input order
    PodVector<ShadeTri> shade_in;
    std::vector<int> mesh_of_desc;                // desc mesh index -> out.meshes index
    int depth = 0;
    int order = 0;
    double scene_pad = 0;

    Walker(const RtgpuSceneDesc& dd, HostScene& o, std::string& e) : d(dd), out(o), err(e) { mesh_of_desc.assign(dd.n_meshes, -1); }

    bool fail(const std::string& m) { err = m; return false; }

    V3 to_world(V3 p) const { // innermost transform first
        for (int i = (int)chain_stack.size() - 1; i >= 0; i--) p = xform_point(out.xforms[chain_stack[i]].m, p);
        return p;
    }
    int push_chain() {
        int first = (int)out.chain.size();
        for (int32_t x : chain_stack) out.chain.push_back(x);
        return first;
    }
    int new_instance(int kind, int obj_index, const RtgpuObject& o) {
        Instance in;
        std::memset(&in, 0, sizeof in);
        in.kind = kind; in.object = obj_index; in.material = o.material; in.texture = o.texture;
        in.chain_first = push_chain(); in.chain_len = (int)chain_stack.size();
        in.mesh = -1; in.flags = o.flags; in.aprim = -1; in.volume = -1; in.order = order++;
        out.instances.push_back(in);
        return (int)out.instances.size() - 1;
    }
    bool check_material(int m) { return m >= 0 && m < d.n_materials; }

    // the reference's own bounding boxes (for gates); object space of the current level
    bool ref_bounds(int idx, double lo[3], double hi[3], int guard = 0) const {
        const double INF = std::numeric_limits<double>::infinity();
        if (idx < 0 || idx >= d.n_objects || guard > 64) return false;
        const RtgpuObject& o = d.objects[idx];
        const int32_t* kids = d.children + o.first_child;
        for (int i = 0; i < 3; i++) { lo[i] = INF; hi[i] = -INF; }
        switch (o.kind) {
        case RTGPU_OBJ_SPHERE: for (int i = 0; i < 3; i++) { lo[i] = o.p0[i] - o.p1[0]; hi[i] = o.p0[i] + o.p1[0]; } return true;
        case RTGPU_OBJ_PLANE:
            for (int su = -1; su <= 1; su += 2) for (int sv = -1; sv <= 1; sv += 2) for (int i = 0; i < 3; i++) {
                double c = o.p0[i] + su * o.p1[i] + sv * o.p2[i];
                lo[i] = std::min(lo[i], c); hi[i] = std::max(hi[i], c);
            }
            for (int i = 0; i < 3; i++) { lo[i] -= 0.001; hi[i] += 0.001; }
            return true;
        case RTGPU_OBJ_MESH: {
            if (o.mesh < 0 || o.mesh >= d.n_meshes) return false;
            const RtgpuMesh& m = d.meshes[o.mesh];
            for (int v = 0; v < m.n_positions; v++) for (int i = 0; i < 3; i++) {
                lo[i] = std::min(lo[i], m.positions[v * 3 + i]); hi[i] = std::max(hi[i], m.positions[v * 3 + i]);
            }
            for (int i = 0; i < 3; i++) { lo[i] -= 0.001; hi[i] += 0.001; }
            return true;
        }
        case RTGPU_OBJ_LIST:
            for (int k = 0; k < o.n_children; k++) {
                double cl[3], ch[3];
                if (!ref_bounds(kids[k], cl, ch, guard + 1)) return false;
                for (int i = 0; i < 3; i++) { lo[i] = std::min(lo[i], cl[i]) - 0.001; hi[i] = std::max(hi[i], ch[i]) + 0.001; }
            }
            return true;
        case RTGPU_OBJ_BVH: {
            if (o.n_children != 2) return false;
            double al[3], ah[3], bl[3], bh[3];
            if (!ref_bounds(kids[0], al, ah, guard + 1)) return false;
            if (d.objects[kids[1]].kind == RTGPU_OBJ_NULL) { for (int i = 0; i < 3; i++) { lo[i] = al[i]; hi[i] = ah[i]; } return true; }
            if (!ref_bounds(kids[1], bl, bh, guard + 1)) return false;
            for (int i = 0; i < 3; i++) { lo[i] = std::min(al[i], bl[i]) - 0.001; hi[i] = std::max(ah[i], bh[i]) + 0.001; }
            return true;
        }
        case RTGPU_OBJ_TRANSFORM: {
            if (o.n_children != 1) return false;
            double cl[3], ch[3];
            if (!ref_bounds(kids[0], cl, ch, guard + 1)) return false;
            for (int c = 0; c < 8; c++) {
                V3 p = {(c & 4) ? ch[0] : cl[0], (c & 2) ? ch[1] : cl[1], (c & 1) ? ch[2] : cl[2]};
                V3 q = xform_point(o.transform, p);
                lo[0] = std::min(lo[0], q.x); lo[1] = std::min(lo[1], q.y); lo[2] = std::min(lo[2], q.z);
                hi[0] = std::max(hi[0], q.x); hi[1] = std::max(hi[1], q.y); hi[2] = std::max(hi[2], q.z);
            }
            for (int i = 0; i < 3; i++) { lo[i] -= 0.001; hi[i] += 0.001; }
            return true;
        }
        case RTGPU_OBJ_VOLUME: return o.n_children == 1 && ref_bounds(kids[0], lo, hi, guard + 1);
        case RTGPU_OBJ_SKY:
        case RTGPU_OBJ_SUN: for (int i = 0; i < 3; i++) { lo[i] = -std::numeric_limits<double>::max(); hi[i] = std::numeric_limits<double>::max(); } return true;
        default: return true; // NullObject: [inf, -inf]
        }
    }

    // Plane::new (plane.rs:29-62) / Sphere::new constants in object space
    APrim make_aprim(const RtgpuObject& o, int instance) {
        APrim a;
        std::memset(&a, 0, sizeof a);
        a.instance = instance; a.gate_first = 0; a.n_gates = 0;
        if (o.kind == RTGPU_OBJ_SPHERE) {
            a.kind = APRIM_SPHERE;
            a.c[0] = o.p0[0]; a.c[1] = o.p0[1]; a.c[2] = o.p0[2]; a.c[3] = o.p1[0];
        } else {
            a.kind = APRIM_PLANE;
            if (o.flags & RTGPU_OBJF_BACKFACE) a.flags |= APRIM_BACKFACE;
            V3 center = {o.p0[0], o.p0[1], o.p0[2]}, u = {o.p1[0], o.p1[1], o.p1[2]}, v = {o.p2[0], o.p2[1], o.p2[2]};
            V3 u_unit = divs(u, length(u)), v_unit = divs(v, length(v));
            V3 n = cross(u, v);
            a.area = length(n) * 4.0;
            V3 normal = divs(n, length(n));
            V3 corner = sub(sub(center, u), v);
            V3 iu = divs(mul(u_unit, 0.5), length(u)), iv = divs(mul(v_unit, 0.5), length(v));
            a.c[0] = corner.x; a.c[1] = corner.y; a.c[2] = corner.z;
            a.n[0] = normal.x; a.n[1] = normal.y; a.n[2] = normal.z;
            a.iu[0] = iu.x; a.iu[1] = iu.y; a.iu[2] = iu.z;
            a.iv[0] = iv.x; a.iv[1] = iv.y; a.iv[2] = iv.z;
            a.u[0] = u.x; a.u[1] = u.y; a.u[2] = u.z;
            a.v[0] = v.x; a.v[1] = v.y; a.v[2] = v.z;
        }
        return a;
    }

    // true extent of an analytic primitive in its object space
    static void aprim_extent(const RtgpuObject& o, double lo[3], double hi[3]) {
        if (o.kind == RTGPU_OBJ_SPHERE) {
            double r = std::fabs(o.p1[0]);
            for (int i = 0; i < 3; i++) { lo[i] = o.p0[i] - r; hi[i] = o.p0[i] + r; }
        } else {
            for (int i = 0; i < 3; i++) { lo[i] = 1e300; hi[i] = -1e300; }
            for (int su = -1; su <= 1; su += 2) for (int sv = -1; sv <= 1; sv += 2) for (int i = 0; i < 3; i++) {
                double c = o.p0[i] + su * o.p1[i] + sv * o.p2[i];
                lo[i] = std::min(lo[i], c); hi[i] = std::max(hi[i], c);
            }
        }
    }

    void add_prim_box(const double lo[3], const double hi[3]) { // object-space box -> padded world fp32 box
        double wl[3] = {1e300, 1e300, 1e300}, wh[3] = {-1e300, -1e300, -1e300};
        for (int c = 0; c < 8; c++) {
            V3 q = to_world({(c & 4) ? hi[0] : lo[0], (c & 2) ? hi[1] : lo[1], (c & 1) ? hi[2] : lo[2]});
            wl[0] = std::min(wl[0], q.x); wl[1] = std::min(wl[1], q.y); wl[2] = std::min(wl[2], q.z);
            wh[0] = std::max(wh[0], q.x); wh[1] = std::max(wh[1], q.y); wh[2] = std::max(wh[2], q.z);
        }
        PrimBox b;
        for (int i = 0; i < 3; i++) {
            double pad = 1e-6 * (std::fabs(wl[i]) + std::fabs(wh[i])) + 1e-9;
            b.lo[i] = round_down(wl[i] - pad); b.hi[i] = round_up(wh[i] + pad);
        }
        boxes.push_back(b);
    }

    bool walk(int idx); // defined below
    bool walk_boundary(int idx, Volume& vol);
    bool add_mesh(int idx, const RtgpuObject& o);
};

bool Walker::add_mesh(int idx, const RtgpuObject& o) {
    if (o.mesh < 0 || o.mesh >= d.n_meshes || !check_material(o.material)) return fail("mesh: bad mesh/material index");
    const RtgpuMesh& m = d.meshes[o.mesh];
    if (mesh_of_desc[o.mesh] < 0) {
        HostMesh hm;
        hm.positions.assign(m.positions, m.positions + (size_t)m.n_positions * 3);
        hm.normals.assign(m.normals, m.normals + (size_t)m.n_normals * 3);
        if (m.uvs && m.n_uvs > 0) hm.uvs.assign(m.uvs, m.uvs + (size_t)m.n_uvs * 2);
        hm.tri_pos.assign(m.tri_pos, m.tri_pos + (size_t)m.n_triangles * 3);
        hm.tri_nrm.assign(m.tri_nrm, m.tri_nrm + (size_t)m.n_triangles * 3);
        if (m.tri_uv) hm.tri_uv.assign(m.tri_uv, m.tri_uv + (size_t)m.n_triangles * 3);
        for (size_t i = 0; i < hm.tri_pos.size(); i++) {
            if (hm.tri_pos[i] < 0 || hm.tri_pos[i] >= m.n_positions || hm.tri_nrm[i] < 0 || hm.tri_nrm[i] >= m.n_normals ||
                (!hm.tri_uv.empty() && hm.tri_uv[i] >= m.n_uvs))
                return fail("mesh: vertex index out of range");
        }
        out.meshes.push_back(std::move(hm));
        mesh_of_desc[o.mesh] = (int)out.meshes.size() - 1;
    }
    int inst = new_instance(INST_MESH, idx, o);
    out.instances[inst].mesh = mesh_of_desc[o.mesh];
    const HostMesh& hm = out.meshes[mesh_of_desc[o.mesh]];
    uint32_t flags = (o.flags & RTGPU_OBJF_HIT_BACK_FACES) ? PRIM_TWO_SIDED : 0u;
    bool flat = (o.flags & RTGPU_OBJF_FLAT_SHADING) != 0;
    auto bits = [](uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; };
    // one record per triangle, written in place by a few threads (pure per-triangle work)
    const size_t base = prims_in.size();
    prims_in.resize(base + (size_t)m.n_triangles);
    shade_in.resize(base + (size_t)m.n_triangles);
    boxes.resize(base + (size_t)m.n_triangles);
    auto convert = [&, base, inst, flags, flat](int t_begin, int t_end) {
    for (int t = t_begin; t < t_end; t++) {
        V3 p[3], w[3];
        for (int k = 0; k < 3; k++) {
            const double* v = &hm.positions[(size_t)hm.tri_pos[t * 3 + k] * 3];
            p[k] = {v[0], v[1], v[2]};
            w[k] = to_world(p[k]);
        }
        V3 e1 = sub(w[1], w[0]), e2 = sub(w[2], w[0]);
        Prim pr;
        pr.a = {(float)w[0].x, (float)w[0].y, (float)w[0].z, bits((uint32_t)inst)};
        pr.b = {(float)e1.x, (float)e1.y, (float)e1.z, bits((uint32_t)t)};
        pr.c = {(float)e2.x, (float)e2.y, (float)e2.z, bits(flags)};
        prims_in[base + (size_t)t] = pr;
        ShadeTri st;
        std::memset(&st, 0, sizeof st);
        V3 e1o = sub(p[1], p[0]), e2o = sub(p[2], p[0]);
        st.e1o[0] = (float)e1o.x; st.e1o[1] = (float)e1o.y; st.e1o[2] = (float)e1o.z;
        st.e2o[0] = (float)e2o.x; st.e2o[1] = (float)e2o.y; st.e2o[2] = (float)e2o.z;
        float* ns[3] = {st.n0, st.n1, st.n2};
        if (flat) { // mesh.rs:107-108
            V3 fn = cross(e1o, e2o); fn = divs(fn, length(fn));
            for (int k = 0; k < 3; k++) { ns[k][0] = (float)fn.x; ns[k][1] = (float)fn.y; ns[k][2] = (float)fn.z; }
        } else for (int k = 0; k < 3; k++) {
            const double* nn = &hm.normals[(size_t)hm.tri_nrm[t * 3 + k] * 3];
            ns[k][0] = (float)nn[0]; ns[k][1] = (float)nn[1]; ns[k][2] = (float)nn[2];
        }
        bool has_uv = !hm.tri_uv.empty() && hm.tri_uv[t * 3] >= 0 && hm.tri_uv[t * 3 + 1] >= 0 && hm.tri_uv[t * 3 + 2] >= 0;
        st.has_uv = has_uv ? 1u : 0u;
        if (has_uv) {
            float* us[3] = {st.uv0, st.uv1, st.uv2};
            for (int k = 0; k < 3; k++) {
                const double* uv = &hm.uvs[(size_t)hm.tri_uv[t * 3 + k] * 2];
                us[k][0] = (float)uv[0]; us[k][1] = (float)uv[1];
            }
        }
        shade_in[base + (size_t)t] = st;
        // padded world box of the fp32 triangle actually intersected
        PrimBox b;
        for (int i = 0; i < 3; i++) {
            double c0 = i == 0 ? w[0].x : i == 1 ? w[0].y : w[0].z;
            double c1 = i == 0 ? w[1].x : i == 1 ? w[1].y : w[1].z;
            double c2 = i == 0 ? w[2].x : i == 1 ? w[2].y : w[2].z;
            double lo = std::min(c0, std::min(c1, c2)), hi = std::max(c0, std::max(c1, c2));
            double pad = 1e-6 * (std::fabs(lo) + std::fabs(hi)) + 1e-9;
            b.lo[i] = round_down(lo - pad); b.hi[i] = round_up(hi + pad);
        }
        boxes[base + (size_t)t] = b;
    }
    };
    const int n_thr = m.n_triangles < 65536 ? 1 : (int)std::max(1u, std::min(16u, rt_host_cores()));
    if (n_thr == 1) convert(0, m.n_triangles);
    else {
        std::vector<std::thread> pool;
        for (int k = 0; k < n_thr; k++)
            pool.emplace_back(convert, (int)((int64_t)m.n_triangles * k / n_thr), (int)((int64_t)m.n_triangles * (k + 1) / n_thr));
        for (auto& th : pool) th.join();
    }
    out.n_triangles += (uint64_t)m.n_triangles;
    return true;
}

bool Walker::walk_boundary(int idx, Volume& vol) {
    if (idx < 0 || idx >= d.n_objects) return fail("volume boundary: object index out of range");
    if (++depth > 64) return fail("object graph too deep (cycle?)");
    const RtgpuObject& o = d.objects[idx];
    const int32_t* kids = d.children + o.first_child;
    bool ok = true;
    switch (o.kind) {
    case RTGPU_OBJ_LIST:
    case RTGPU_OBJ_BVH:
        for (int k = 0; k < o.n_children && ok; k++) ok = walk_boundary(kids[k], vol);
        break;
    case RTGPU_OBJ_TRANSFORM: {
        if (o.n_children != 1) return fail("transform needs one child");
        Xform x; std::memcpy(x.m, o.transform, sizeof x.m); std::memcpy(x.inv, o.inv_transform, sizeof x.inv);
        out.xforms.push_back(x);
        chain_stack.push_back((int32_t)out.xforms.size() - 1);
        ok = walk_boundary(kids[0], vol);
        chain_stack.pop_back();
        break;
    }
    case RTGPU_OBJ_SPHERE:
    case RTGPU_OBJ_PLANE: {
        int inst = new_instance(o.kind == RTGPU_OBJ_SPHERE ? INST_SPHERE : INST_PLANE, idx, o);
        out.vol_aprims.push_back(make_aprim(o, inst));
        out.vol_chain.push_back(out.instances[inst].chain_first);
        out.vol_chain.push_back(out.instances[inst].chain_len);
        vol.bprim_count++;
        break;
    }
    case RTGPU_OBJ_NULL: new_instance(INST_NULL, idx, o); break;
    default: return fail("volume boundary may only contain lists, transforms, spheres and planes on the GPU path");
    }
    depth--;
    return ok;
}

bool Walker::walk(int idx) {
    if (idx < 0 || idx >= d.n_objects) return fail("object index out of range");
    if (++depth > 64) return fail("object graph too deep (cycle?)");
    const RtgpuObject& o = d.objects[idx];
    if (o.first_child < 0 || o.n_children < 0 || o.first_child + o.n_children > d.n_children) return fail("object children out of range");
    const int32_t* kids = d.children + o.first_child;
    bool ok = true;
    switch (o.kind) {
    case RTGPU_OBJ_LIST:
    case RTGPU_OBJ_BVH: {
        if (o.kind == RTGPU_OBJ_BVH && o.n_children != 2) return fail("bvh node needs two children");
        bool gated = o.kind == RTGPU_OBJ_BVH || !(o.flags & RTGPU_OBJF_DISABLE_BOUNDS_CHECK);
        if (gated) {
            GateBox g;
            if (!ref_bounds(idx, g.lo, g.hi)) return fail("cannot compute reference bounds");
            gate_stack.push_back(g);
        }
        for (int k = 0; k < o.n_children && ok; k++) ok = walk(kids[k]);
        if (gated) gate_stack.pop_back();
        break;
    }
    case RTGPU_OBJ_NULL: new_instance(INST_NULL, idx, o); break;
    case RTGPU_OBJ_TRANSFORM: {
        if (o.n_children != 1) return fail("transform needs one child");
        Xform x; std::memcpy(x.m, o.transform, sizeof x.m); std::memcpy(x.inv, o.inv_transform, sizeof x.inv);
        out.xforms.push_back(x);
        chain_stack.push_back((int32_t)out.xforms.size() - 1);
        std::vector<GateBox> saved; saved.swap(gate_stack); // gates above a Transform contain its bounds
        ok = walk(kids[0]);
        gate_stack.swap(saved);
        chain_stack.pop_back();
        break;
    }
    case RTGPU_OBJ_MESH: ok = add_mesh(idx, o); break;
    case RTGPU_OBJ_SPHERE:
    case RTGPU_OBJ_PLANE: {
        if (!check_material(o.material)) return fail("sphere/plane: bad material index");
        if (o.kind == RTGPU_OBJ_SPHERE && o.p1[0] == 0.0) return fail("sphere: zero radius");
        int inst = new_instance(o.kind == RTGPU_OBJ_SPHERE ? INST_SPHERE : INST_PLANE, idx, o);
        APrim a = make_aprim(o, inst);
        double lo[3], hi[3];
        aprim_extent(o, lo, hi);
        // gates: reference ancestor boxes that do not contain the primitive (only possible for
        // inverted boxes of negative-radius spheres)
        a.gate_first = (int32_t)out.gates.size();
        bool never = false;
        for (const GateBox& g : gate_stack) {
            bool contains = true;
            int inverted = 0;
            for (int i = 0; i < 3; i++) { if (g.lo[i] > lo[i] || g.hi[i] < hi[i]) contains = false; if (g.lo[i] > g.hi[i]) inverted++; }
            if (contains) continue;
            if (inverted >= 2) never = true; // the slab test can never pass
            out.gates.push_back(g); a.n_gates++;
        }
        if (never) { out.gates.resize(a.gate_first); break; } // unreachable primitive: keep the ordinal, drop the geometry
        out.instances[inst].aprim = (int32_t)out.aprims.size();
        auto bits = [](uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; };
        Prim pr;
        pr.a = {0, 0, 0, bits((uint32_t)inst)};
        pr.b = {0, 0, 0, bits((uint32_t)out.aprims.size())};
        pr.c = {0, 0, 0, bits(PRIM_ANALYTIC)};
        if (a.kind == APRIM_PLANE && a.n_gates == 0) {
            // world-space parallelogram (corner, 2u, 2v) for the fp32 traversal; the front side is the
            // side of the object-space normal u x v, so a mirroring transform swaps the edges
            V3 corner = {a.c[0], a.c[1], a.c[2]}, eu = {2.0 * a.u[0], 2.0 * a.u[1], 2.0 * a.u[2]}, ev = {2.0 * a.v[0], 2.0 * a.v[1], 2.0 * a.v[2]};
            V3 nrm = {a.n[0], a.n[1], a.n[2]};
            V3 cw = to_world(corner);
            V3 e1 = sub(to_world({corner.x + eu.x, corner.y + eu.y, corner.z + eu.z}), cw);
            V3 e2 = sub(to_world({corner.x + ev.x, corner.y + ev.y, corner.z + ev.z}), cw);
            V3 nw = sub(to_world({corner.x + nrm.x, corner.y + nrm.y, corner.z + nrm.z}), cw);
            V3 cr = cross(e1, e2);
            if (cr.x * nw.x + cr.y * nw.y + cr.z * nw.z < 0.0) std::swap(e1, e2);
            uint32_t qf = PRIM_ANALYTIC | PRIM_QUAD | ((a.flags & APRIM_BACKFACE) ? PRIM_TWO_SIDED : 0u);
            pr.a = {(float)cw.x, (float)cw.y, (float)cw.z, bits((uint32_t)inst)};
            pr.b = {(float)e1.x, (float)e1.y, (float)e1.z, bits((uint32_t)out.aprims.size())};
            pr.c = {(float)e2.x, (float)e2.y, (float)e2.z, bits(qf)};
        }
        if (a.kind == APRIM_SPHERE && a.n_gates == 0) {
            // a sphere stays a sphere under rotation + uniform scale + translation: world centre and radius
            // for the fp32 traversal (anything else keeps the f64 object-space call)
            V3 c0 = {a.c[0], a.c[1], a.c[2]};
            V3 cw = to_world(c0);
            V3 ex = sub(to_world({c0.x + 1.0, c0.y, c0.z}), cw), ey = sub(to_world({c0.x, c0.y + 1.0, c0.z}), cw),
               ez = sub(to_world({c0.x, c0.y, c0.z + 1.0}), cw);
            auto dot3 = [](V3 p, V3 q) { return p.x * q.x + p.y * q.y + p.z * q.z; };
            const double sx = length(ex), sy = length(ey), sz = length(ez);
            const double tol = 1e-9 * sx * sx;
            const bool similarity = sx > 0 && std::fabs(sx - sy) <= 1e-9 * sx && std::fabs(sx - sz) <= 1e-9 * sx &&
                                    std::fabs(dot3(ex, ey)) <= tol && std::fabs(dot3(ex, ez)) <= tol && std::fabs(dot3(ey, ez)) <= tol;
            if (similarity) {
                pr.a = {(float)cw.x, (float)cw.y, (float)cw.z, bits((uint32_t)inst)};
                pr.b = {(float)(std::fabs(a.c[3]) * sx), 0.f, 0.f, bits((uint32_t)out.aprims.size())};
                pr.c = {0.f, 0.f, 0.f, bits(PRIM_ANALYTIC | PRIM_SPHERE)};
            }
        }
        out.aprims.push_back(a);
        prims_in.push_back(pr);
        ShadeTri st; std::memset(&st, 0, sizeof st);
        shade_in.push_back(st);
        add_prim_box(lo, hi);
        break;
    }
    case RTGPU_OBJ_VOLUME: {
        if (o.n_children != 1 || !check_material(o.material)) return fail("volume: bad boundary/material");
        int inst = new_instance(INST_VOLUME, idx, o);
        Volume v; std::memset(&v, 0, sizeof v);
        v.instance = inst; v.material = o.material; v.neg_inv_density = -1.0 / o.p1[0];
        v.bprim_first = (int32_t)out.vol_aprims.size();
        out.instances[inst].volume = (int32_t)out.volumes.size();
        ok = walk_boundary(kids[0], v);
        out.volumes.push_back(v);
        break;
    }
    case RTGPU_OBJ_SKY:
    case RTGPU_OBJ_SUN: {
        if (o.texture < 0 || o.texture >= d.n_textures) return fail("sky/sun: bad texture index");
        if (!chain_stack.empty()) { err = "sky/sun under a Transform is not supported on the GPU path"; return false; }
        int inst = new_instance(o.kind == RTGPU_OBJ_SKY ? INST_SKY : INST_SUN, idx, o);
        for (int i = 0; i < 4; i++) out.instances[inst].dir[i] = o.p0[i];
        out.infinite.push_back(inst);
        break;
    }
    default: return fail("unknown object kind");
    }
    depth--;
    return ok;
}

} // namespace

namespace {

struct ShadingBuilder {
    const RtgpuSceneDesc& d;
    HostScene& out;
    std::string& err;

    bool emit_tex(int t, int depth, int& stack_now, int& stack_max) {
        if (t < 0 || t >= d.n_textures) { err = "texture index out of range"; return false; }
        if (depth > 16) { err = "texture expression too deep (cycle?)"; return false; }
        const RtgpuTexture& tx = d.textures[t];
        TexOp op;
        std::memset(&op, 0, sizeof op);
        switch (tx.kind) {
        case RTGPU_TEX_CONSTANT:
            op.op = TOP_CONST;
            for (int i = 0; i < 4; i++) op.value[i] = (float)tx.value[i];
            stack_now++;
            break;
        case RTGPU_TEX_CHECKER:
        case RTGPU_TEX_CHECKER_SOLID:
            if (!emit_tex(tx.a, depth + 1, stack_now, stack_max) || !emit_tex(tx.b, depth + 1, stack_now, stack_max)) return false;
            op.op = tx.kind == RTGPU_TEX_CHECKER ? TOP_CHECKER : TOP_CHECKER_SOLID;
            op.value[0] = (float)tx.scale;
            stack_now -= 1;
            break;
        case RTGPU_TEX_IMAGE:
            if (tx.image < 0 || tx.image >= d.n_images) { err = "texture: bad image index"; return false; }
            op.op = TOP_IMAGE; op.arg = tx.image; op.repeat = tx.repeat;
            stack_now++;
            break;
        case RTGPU_TEX_LERP:
            if (!emit_tex(tx.a, depth + 1, stack_now, stack_max) || !emit_tex(tx.b, depth + 1, stack_now, stack_max) ||
                !emit_tex(tx.c, depth + 1, stack_now, stack_max)) return false;
            op.op = TOP_LERP;
            stack_now -= 2;
            break;
        case RTGPU_TEX_NOISE_SOLID:
            if (tx.noise < 0 || tx.noise >= d.n_noises) { err = "texture: bad noise index"; return false; }
            op.op = TOP_NOISE_SOLID; op.arg = tx.noise; op.samples = tx.samples; op.value[0] = (float)tx.scale;
            stack_now++;
            break;
        case RTGPU_TEX_CHANNEL:
            if (!emit_tex(tx.a, depth + 1, stack_now, stack_max)) return false;
            if (tx.channel < 0 || tx.channel > 3) { err = "channel index out of range"; return false; }
            op.op = TOP_CHANNEL; op.arg = tx.channel;
            break;
        case RTGPU_TEX_UV_DEBUG: op.op = TOP_UV_DEBUG; stack_now++; break;
        default: err = "unknown texture kind"; return false;
        }
        stack_max = std::max(stack_max, stack_now);
        out.tex_ops.push_back(op);
        return true;
    }

    bool run() {
        for (int t = 0; t < d.n_textures; t++) {
            TexProgram p; p.first = (int32_t)out.tex_ops.size();
            int now = 0, mx = 0;
            if (!emit_tex(t, 0, now, mx)) return false;
            if (mx > 8) { err = "texture expression needs more than 8 stack slots"; return false; }
            p.count = (int32_t)out.tex_ops.size() - p.first;
            out.tex_programs.push_back(p);
        }
        for (int i = 0; i < d.n_images; i++) {
            const RtgpuImage& im = d.images[i];
            if (im.width <= 0 || im.height <= 0 || !im.data) { err = "image: empty"; return false; }
            HostImage h; h.width = im.width; h.height = im.height; h.format = im.format;
            size_t n = (size_t)im.width * im.height;
            size_t bpc = im.format == RTGPU_IMG_RGB8 ? 1 : im.format == RTGPU_IMG_RGB16 ? 2 : im.format == RTGPU_IMG_RGBF32 ? 4 : 0;
            if (!bpc) { err = "image: unknown format"; return false; }
            h.texels.assign(n * 4 * bpc, 0);
            const uint8_t* src = (const uint8_t*)im.data;
            for (size_t p = 0; p < n; p++) std::memcpy(&h.texels[p * 4 * bpc], src + p * 3 * bpc, 3 * bpc);
            out.images.push_back(std::move(h));
        }
        for (int i = 0; i < d.n_noises; i++) {
            DevNoise n;
            for (int k = 0; k < 256; k++) {
                n.grad[k] = {(float)d.noises[i].gradients[k][0], (float)d.noises[i].gradients[k][1], (float)d.noises[i].gradients[k][2], 0.f};
                n.perm_x[k] = (uint8_t)(d.noises[i].perm_x[k] & 255);
                n.perm_y[k] = (uint8_t)(d.noises[i].perm_y[k] & 255);
                n.perm_z[k] = (uint8_t)(d.noises[i].perm_z[k] & 255);
            }
            out.noises.push_back(n);
        }
        for (int i = 0; i < d.n_materials; i++) {
            const RtgpuMaterial& m = d.materials[i];
            Material o; std::memset(&o, 0, sizeof o);
            o.kind = m.kind; o.albedo = m.albedo; o.roughness = m.roughness; o.normal_map = m.normal_map;
            o.ior = (float)m.ior; o.inv_ior = (float)(1.0 / m.ior);
            auto ok = [&](int t) { return t >= 0 && t < d.n_textures; };
            bool good = true;
            switch (m.kind) {
            case RTGPU_MAT_LAMBERTIAN: case RTGPU_MAT_EMISSIVE: case RTGPU_MAT_ISOTROPIC: good = ok(m.albedo); break;
            case RTGPU_MAT_METAL: good = ok(m.albedo) && ok(m.roughness); break;
            case RTGPU_MAT_GLOSSY: good = ok(m.albedo) && ok(m.roughness) && (m.normal_map < 0 || ok(m.normal_map)); break;
            case RTGPU_MAT_DIELECTRIC: break;
            case RTGPU_MAT_NORMAL_DEBUG: good = m.normal_map < 0 || ok(m.normal_map); break;
            default: good = false;
            }
            if (!good) { err = "material " + std::to_string(i) + ": bad kind or texture index"; return false; }
            if (o.normal_map >= d.n_textures) o.normal_map = -1;
            out.materials.push_back(o);
        }
        return true;
    }
};

// HittablePDF over the `lights` root (pdf/hittable.rs, list.rs:80-100), flattened to weighted leaves
static bool flatten_lights(const RtgpuSceneDesc& d, HostScene& out, int idx, double weight, int depth, std::string& err, Walker& w) {
    if (idx < 0 || idx >= d.n_objects) { err = "lights: object index out of range"; return false; }
    if (depth > 32) { err = "lights: graph too deep"; return false; }
    const RtgpuObject& o = d.objects[idx];
    Light l; std::memset(&l, 0, sizeof l);
    l.weight = (float)weight; l.aprim = -1;
    switch (o.kind) {
    case RTGPU_OBJ_LIST:
        if (o.n_children == 0) { l.kind = LIGHT_NULL; out.lights.push_back(l); return true; }
        for (int k = 0; k < o.n_children; k++)
            if (!flatten_lights(d, out, d.children[o.first_child + k], weight / (double)o.n_children, depth + 1, err, w)) return false;
        return true;
    case RTGPU_OBJ_PLANE:
    case RTGPU_OBJ_SPHERE:
        l.kind = o.kind == RTGPU_OBJ_PLANE ? LIGHT_PLANE : LIGHT_SPHERE;
        l.aprim = (int32_t)out.light_aprims.size();
        out.light_aprims.push_back(w.make_aprim(o, -1));
        break;
    case RTGPU_OBJ_SKY: l.kind = LIGHT_SKY; break;
    case RTGPU_OBJ_SUN: l.kind = LIGHT_SUN; for (int i = 0; i < 4; i++) l.dir[i] = o.p0[i]; break;
    default: l.kind = LIGHT_NULL; break; // Transform, BVH, mesh, volume, null: pdf 0, direction (1,0,0)
    }
    out.lights.push_back(l);
    return true;
}

} // namespace

int flatten_scene(const RtgpuSceneDesc& d, HostScene& out, std::string& err) {
    auto t0 = std::chrono::steady_clock::now();
    if (d.abi_version != RTGPU_ABI_VERSION) { err = "abi version mismatch"; return RTGPU_ERR_INVALID; }
    if (!d.objects || d.n_objects <= 0) { err = "scene has no objects"; return RTGPU_ERR_INVALID; }
    out = HostScene();
    ShadingBuilder sb{d, out, err};
    if (!sb.run()) return RTGPU_ERR_INVALID;
    const bool timing = std::getenv("RTGPU_BUILD_TIMING") != nullptr;
    auto lap = [&](const char* what) {
        if (timing) std::fprintf(stderr, "[rtgpu build] %-28s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    };
    lap("shading tables");
    Walker w(d, out, err);
    if (!w.walk(d.world_root)) return err.find("not supported") != std::string::npos || err.find("may only") != std::string::npos ? RTGPU_ERR_UNSUPPORTED : RTGPU_ERR_INVALID;
    if (!flatten_lights(d, out, d.lights_root, 1.0, 0, err, w)) return RTGPU_ERR_INVALID;

    lap("walk (prims, shade records)");
    // Broadcast list: when a scene has only a handful of fp32 quads / spheres (room walls, a ground sphere,
    // area lights) every ray tests them once, convergently, where it is generated (k_raygen / k_shade) and
    // enters the BVH with that distance as its far bound; they are kept out of the BVH, whose root then
    // bounds only the meshes. RTGPU_BROADCAST_MAX=0 turns the list off.
    std::vector<uint32_t> global_ids, bvh_ids;
    {
        int max_global = 16;
        if (const char* e = std::getenv("RTGPU_BROADCAST_MAX")) max_global = std::max(0, std::min(64, std::atoi(e)));
        for (size_t i = 0; i < w.prims_in.size() && (int)global_ids.size() <= max_global; i++) {
            uint32_t flags; std::memcpy(&flags, &w.prims_in[i].c.w, 4);
            if (flags & (PRIM_QUAD | PRIM_SPHERE)) global_ids.push_back((uint32_t)i);
        }
        if ((int)global_ids.size() > max_global) global_ids.clear();
    }
    if (global_ids.empty()) g_wide_bvh_builder(w.boxes, out.bvh);
    else {
        std::vector<PrimBox> boxes;
        boxes.reserve(w.boxes.size() - global_ids.size());
        bvh_ids.reserve(w.boxes.size() - global_ids.size());
        size_t g = 0;
        for (size_t i = 0; i < w.boxes.size(); i++) {
            if (g < global_ids.size() && global_ids[g] == i) { g++; continue; }
            boxes.push_back(w.boxes[i]); bvh_ids.push_back((uint32_t)i);
        }
        g_wide_bvh_builder(boxes, out.bvh);
        for (uint32_t& o : out.bvh.prim_order) o = bvh_ids[o];
    }
    lap("wide BVH");
    out.prims.resize(w.prims_in.size());
    out.shade_tris.resize(w.prims_in.size());
    // material class of every leaf primitive (the trace kernel routes paths to shade queues with it)
    auto class_of = [&](int material) {
        return material < 0 || material >= (int)out.materials.size() ? (int)CLASS_EMISSIVE : material_class_of_kind(out.materials[material].kind);
    };
    auto finish_prim = [&](Prim& p, uint32_t& class_mask, bool& cold) { // class bits into the flags word
        uint32_t inst, flags;
        std::memcpy(&inst, &p.a.w, 4); std::memcpy(&flags, &p.c.w, 4);
        const int cls = class_of(out.instances[inst].material);
        flags |= (uint32_t)cls << PRIM_CLASS_SHIFT;
        std::memcpy(&p.c.w, &flags, 4);
        class_mask |= 1u << cls;
        if ((flags & (PRIM_ANALYTIC | PRIM_QUAD | PRIM_SPHERE)) == PRIM_ANALYTIC) cold = true;
    };
    out.class_mask = 1u << CLASS_EMISSIVE; // misses and sky / sun hits are shaded by the emissive kernel
    {
        // one threaded pass over the leaf order: gather the two records, add the class bits, reduce the BVH bounds
        const size_t n = out.bvh.prim_order.size();
        struct Part { uint32_t class_mask = 0; bool cold = false; float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f}; };
        const size_t n_thr = n < 65536 ? 1 : std::max(1u, std::min(16u, rt_host_cores()));
        std::vector<Part> parts(n_thr);
        auto gather = [&](size_t k, size_t b, size_t e) {
            Part& pt = parts[k];
            for (size_t i = b; i < e; i++) {
                const uint32_t src = out.bvh.prim_order[i];
                out.prims[i] = w.prims_in[src];
                out.shade_tris[i] = w.shade_in[src];
                finish_prim(out.prims[i], pt.class_mask, pt.cold);
                const PrimBox& bx = w.boxes[src];
                for (int a = 0; a < 3; a++) { pt.lo[a] = std::min(pt.lo[a], bx.lo[a]); pt.hi[a] = std::max(pt.hi[a], bx.hi[a]); }
            }
        };
        if (n_thr == 1) gather(0, 0, n);
        else {
            std::vector<std::thread> pool;
            for (size_t k = 0; k < n_thr; k++) pool.emplace_back(gather, k, n * k / n_thr, n * (k + 1) / n_thr);
            for (auto& th : pool) th.join();
        }
        for (int a = 0; a < 3; a++) { out.bvh_lo[a] = 3.0e38f; out.bvh_hi[a] = -3.0e38f; }
        for (const Part& pt : parts) {
            out.class_mask |= pt.class_mask;
            if (pt.cold) out.trace_cold = true;
            for (int a = 0; a < 3; a++) { out.bvh_lo[a] = std::min(out.bvh_lo[a], pt.lo[a]); out.bvh_hi[a] = std::max(out.bvh_hi[a], pt.hi[a]); }
        }
        out.global_first = (int32_t)n; out.n_global = (int32_t)global_ids.size();
        for (size_t k = 0; k < global_ids.size(); k++) {
            const Prim& p = w.prims_in[global_ids[k]];
            out.prims[n + k] = p;
            out.shade_tris[n + k] = w.shade_in[global_ids[k]];
            bool cold = false;
            finish_prim(out.prims[n + k], out.class_mask, cold);
            if (cold) out.trace_cold = true;
            // w-vector form of the fp32 parallelogram in the leaf record (f64 here, rounded once)
            BroadcastPrim bp{};
            uint32_t flags; std::memcpy(&flags, &p.c.w, 4);
            bp.q = p.a; // (corner | centre, instance): the .w lane is rewritten with the final flags below
            if (flags & PRIM_SPHERE) bp.n.x = p.b.x;
            else {
                const double e1[3] = {p.b.x, p.b.y, p.b.z}, e2[3] = {p.c.x, p.c.y, p.c.z};
                const double nn[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
                const double n2 = nn[0] * nn[0] + nn[1] * nn[1] + nn[2] * nn[2];
                const double a1[3] = {e2[1] * nn[2] - e2[2] * nn[1], e2[2] * nn[0] - e2[0] * nn[2], e2[0] * nn[1] - e2[1] * nn[0]};
                const double a2[3] = {nn[1] * e1[2] - nn[2] * e1[1], nn[2] * e1[0] - nn[0] * e1[2], nn[0] * e1[1] - nn[1] * e1[0]};
                bp.n.x = (float)nn[0]; bp.n.y = (float)nn[1]; bp.n.z = (float)nn[2];
                bp.w1.x = (float)(a1[0] / n2); bp.w1.y = (float)(a1[1] / n2); bp.w1.z = (float)(a1[2] / n2);
                bp.w2.x = (float)(a2[0] / n2); bp.w2.y = (float)(a2[1] / n2); bp.w2.z = (float)(a2[2] / n2);
            }
            out.bcast.push_back(bp);
        }
    }
    for (size_t k = 0; k < out.bcast.size(); k++) {
        const Prim& pr = out.prims[(size_t)out.global_first + k];
        uint32_t flags; std::memcpy(&flags, &pr.c.w, 4); // flags incl. class bits
        // a light on the list: its constant emission travels in the record, so the kernel that finds a new ray ending on it
        // (inline_emission) needs no instance -> material -> texture-program look-ups
        uint32_t inst_id, aprim_id; std::memcpy(&inst_id, &pr.a.w, 4); std::memcpy(&aprim_id, &pr.b.w, 4);
        const int mat = inst_id < out.instances.size() ? out.instances[inst_id].material : -1;
        if (((flags >> PRIM_CLASS_SHIFT) & 7u) == (uint32_t)CLASS_EMISSIVE && mat >= 0 && out.materials[(size_t)mat].kind == MAT_EMISSIVE &&
            out.materials[(size_t)mat].albedo >= 0) {
            const TexProgram prog = out.tex_programs[(size_t)out.materials[(size_t)mat].albedo];
            if (prog.count == 1 && out.tex_ops[(size_t)prog.first].op == TOP_CONST) {
                const TexOp& op = out.tex_ops[(size_t)prog.first];
                flags |= BCAST_EMIT_CONST;
                out.bcast[k].n.w = op.value[0]; out.bcast[k].w1.w = op.value[1]; out.bcast[k].w2.w = op.value[2];
                if ((flags & PRIM_SPHERE) && aprim_id < out.aprims.size() && out.aprims[aprim_id].c[3] < 0.0) flags |= BCAST_INWARD;
            }
        }
        std::memcpy(&out.bcast[k].q.w, &flags, 4);
    }
    if (!out.volumes.empty()) out.trace_cold = true; // sky / sun misses are resolved by the emissive shade kernel
    for (const Volume& v : out.volumes) out.class_mask |= 1u << class_of(v.material);
    lap("leaf-order tables");
    out.ms_build = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return 0;
}

} // namespace rt
