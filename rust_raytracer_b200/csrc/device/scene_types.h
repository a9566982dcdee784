// scene_types.h — the flattened, device-resident scene (shared by the host flattener /
// BVH builder and the CUDA kernels). Everything is plain data, 16-byte aligned where the
// kernels use 128-bit loads.
//
// Layout in HBM (per GPU), see DESIGN.md §3:
//   nodes      80 B  8-wide BVH node, quantised child boxes (5 x 16 B loads)
//   prims      48 B  BVH leaf primitive in leaf order: world-space fp32 triangle
//                    (v0, e1, e2) or a reference to an analytic primitive (3 x 16 B loads)
//   shade_tris 96 B  per-triangle shading record in the same order (object-space normals, uvs)
//   aprims           analytic primitives (sphere / quad), f64, object space
//   xforms           Transform matrices (f64 M and M^-1 as accumulated by the reference)
//   instances        one per leaf occurrence of the world tree (DFS pre-order)
//   exact mesh data  f64 positions + index triples, only touched in exact mode
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define RT_HD __host__ __device__ __forceinline__
#define RT_ALIGN(n) __align__(n)
#else
#define RT_HD inline
#define RT_ALIGN(n) __attribute__((aligned(n)))
#endif

namespace rt {

struct RT_ALIGN(16) F4 { float x, y, z, w; };
struct RT_ALIGN(16) U4 { uint32_t x, y, z, w; };

// ---- 8-wide compressed BVH node (Ylitie, Karras, Laine 2017 layout), 80 bytes ----------
struct RT_ALIGN(16) BvhNode {
    float px, py, pz;      // quantisation origin
    uint8_t ex, ey, ez;    // per-axis exponent (biased, as in an IEEE float)
    uint8_t imask;         // bit s set: child slot s is an internal node
    uint32_t child_base;   // index of the first internal child node
    uint32_t prim_base;    // index of the first leaf primitive of this node
    uint8_t meta[8];       // per slot: empty 0 | internal 0b001_11sss | leaf (unary count << 5) | first prim offset
    uint8_t qlox[8], qloy[8], qloz[8];
    uint8_t qhix[8], qhiy[8], qhiz[8];
};
static_assert(sizeof(BvhNode) == 80, "BvhNode must be 80 bytes");

// ---- BVH leaf primitive, 48 bytes -----------------------------------------------------------
// triangle: a = (v0.xyz, bits(instance)), b = (e1.xyz, bits(local triangle index)), c = (e2.xyz, bits(flags))
// analytic: a.w = bits(instance), b.w = bits(aprim index), c.w = bits(flags | PRIM_ANALYTIC);
//           quads additionally carry their world-space parallelogram for the fp32 traversal (PRIM_QUAD)
enum : uint32_t {
    PRIM_ANALYTIC = 1u << 0,
    PRIM_TWO_SIDED = 1u << 1, // TriangleMesh.hit_back_faces
    PRIM_QUAD = 1u << 2,      // analytic quad without gates: a/b/c.xyz = world-space fp32 corner, edge 2u, edge 2v
    PRIM_SPHERE = 1u << 3,    // analytic sphere without gates under a similarity transform: a.xyz = world centre, b.x = |radius|
    PRIM_CLASS_SHIFT = 4,     // bits 4..6: material class of the owning instance (shade queue id)
    // BroadcastPrim.q.w only (not in Prim.c.w): the primitive's material is Emissive with a CONSTANT emission map, whose value sits
    // in the .w lanes of n / w1 / w2 (inline_emission); a sphere with a negative radius (inward normal, sphere.rs:64-66)
    BCAST_EMIT_CONST = 1u << 8,
    BCAST_INWARD = 1u << 9,
};
struct RT_ALIGN(16) Prim { F4 a, b, c; };
static_assert(sizeof(Prim) == 48, "Prim must be 48 bytes");

// ---- broadcast-list primitive, 64 bytes (leaf_fast.cuh: tested by every ray where it is generated) --------
// quad:   q = (corner, bits(flags)), n = (e1 x e2, 0), w1 = (e2 x n) / n.n, w2 = (n x e1) / n.n:
//         with b = o - corner: t = -b.n / d.n, p = b + t d, u = p.w1, v = p.w2 (the w-vector form of plane.rs:76-90)
// sphere: q = (centre, bits(flags)), n.x = |radius|
// n.w, w1.w, w2.w = constant emission (r, g, b) when BCAST_EMIT_CONST is set in the flags
struct RT_ALIGN(16) BroadcastPrim { F4 q, n, w1, w2; };
static_assert(sizeof(BroadcastPrim) == 64, "BroadcastPrim must be 64 bytes");

// per-triangle shading record (same index as Prim): object-space vertex normals, uvs
struct RT_ALIGN(16) ShadeTri {
    float n0[3], n1[3], n2[3]; // object-space normals (already unit, obj.rs:48)
    float uv0[2], uv1[2], uv2[2];
    uint32_t has_uv;
    float e1o[3], e2o[3]; // object-space edges (tangent frame, mesh.rs:119-150)
    uint32_t pad[2];
};
static_assert(sizeof(ShadeTri) == 96, "ShadeTri must be 96 bytes");

enum : int32_t { APRIM_SPHERE = 0, APRIM_PLANE = 1 };
enum : uint32_t { APRIM_BACKFACE = 1u << 0 };

// analytic primitive in its object space (f64, reference formulas)
struct RT_ALIGN(16) APrim {
    int32_t kind;
    int32_t instance;
    uint32_t flags;
    int32_t n_gates;   // ancestor boxes that do not contain the primitive (inverted sphere boxes)
    int32_t gate_first;
    int32_t pad[3];
    // sphere: c[0..2] centre, c[3] radius
    // plane : c[0..2] corner, n[0..2] unit normal, iu/iv = inv_u / inv_v, u/v = edge vectors, area
    double c[4];
    double n[3];
    double iu[3], iv[3];
    double u[3], v[3];
    double area;
};

struct RT_ALIGN(16) Xform { double m[16]; double inv[16]; };

struct RT_ALIGN(16) GateBox { double lo[3], hi[3]; };

enum : int32_t { INST_MESH = 0, INST_SPHERE = 1, INST_PLANE = 2, INST_VOLUME = 3, INST_SKY = 4, INST_SUN = 5, INST_NULL = 6 };

// one per leaf occurrence in the world tree, DFS pre-order == RtgpuHit.instance
struct RT_ALIGN(16) Instance {
    int32_t kind;
    int32_t object;      // index into RtgpuSceneDesc.objects
    int32_t material;    // material index (mesh / sphere / plane / volume); sky, sun: -1
    int32_t texture;     // sky / sun emission texture
    int32_t chain_first; // first entry in `chain` (outermost transform first)
    int32_t chain_len;
    int32_t mesh;        // exact-mode mesh record (INST_MESH)
    uint32_t flags;      // RTGPU_OBJF_*
    int32_t aprim;       // analytic primitive index (sphere / plane)
    int32_t volume;      // volume record index
    int32_t order;       // visit order among the world list (ties, sky/sun precedence)
    int32_t pad;
    double dir[4];       // sun: unit direction (w kept)
};

// exact-mode mesh (f64, reference layout mesh.rs:16-35)
struct ExactMesh {
    const double* positions; // n x 3
    const double* normals;   // n x 3
    const double* uvs;       // n x 2
    const int32_t* tri_pos;  // n_tri x 3
    const int32_t* tri_nrm;
    const int32_t* tri_uv;   // n_tri x 3 or null
    int32_t n_triangles;
    int32_t pad;
};

struct RT_ALIGN(16) Volume {
    int32_t instance;
    int32_t material;
    int32_t bprim_first; // boundary primitives: indices into `vol_aprims`
    int32_t bprim_count;
    double neg_inv_density;
};

// ---- materials, textures, lights -------------------------------------------------------------
enum : int32_t { MAT_LAMBERTIAN = 0, MAT_METAL, MAT_DIELECTRIC, MAT_GLOSSY, MAT_EMISSIVE, MAT_ISOTROPIC, MAT_NORMAL_DEBUG };

// shade queues of the wavefront pipeline: one per Material::scatter implementation
enum : int { CLASS_LAMBERTIAN = 0, CLASS_METAL, CLASS_DIELECTRIC, CLASS_GLOSSY, CLASS_ISOTROPIC, CLASS_EMISSIVE, N_CLASSES };
RT_HD int material_class_of_kind(int32_t kind) {
    switch (kind) {
    case MAT_LAMBERTIAN: return CLASS_LAMBERTIAN;
    case MAT_METAL: return CLASS_METAL;
    case MAT_DIELECTRIC: return CLASS_DIELECTRIC;
    case MAT_GLOSSY: return CLASS_GLOSSY;
    case MAT_ISOTROPIC: return CLASS_ISOTROPIC;
    default: return CLASS_EMISSIVE; // emissive, normal_debug (emit only)
    }
}

struct RT_ALIGN(16) Material {
    int32_t kind;
    int32_t albedo, roughness, normal_map; // texture program ids, -1 = none
    float ior, inv_ior;
    int32_t pad[2];
};

// texture programs: postfix op lists evaluated on a small value stack
enum : int32_t { TOP_CONST = 0, TOP_CHECKER, TOP_CHECKER_SOLID, TOP_IMAGE, TOP_LERP, TOP_NOISE_SOLID, TOP_CHANNEL, TOP_UV_DEBUG };
struct RT_ALIGN(16) TexOp {
    int32_t op;
    int32_t arg;     // image / noise index, channel
    int32_t samples; // noise octaves
    int32_t repeat;
    float value[4];  // constant value; [0] = scale for checker / noise
};
struct TexProgram { int32_t first, count; }; // into tex_ops

struct DevImage {
    const void* data; // u8x4 / u16x4 / f32x4 texels (RGB + pad), row-major
    int32_t width, height, format, pad;
};

struct RT_ALIGN(16) DevNoise {
    F4 grad[256];
    uint8_t perm_x[256], perm_y[256], perm_z[256];
};

enum : int32_t { LIGHT_NULL = 0, LIGHT_PLANE, LIGHT_SPHERE, LIGHT_SKY, LIGHT_SUN };
struct RT_ALIGN(16) Light {
    int32_t kind;
    int32_t aprim;  // LIGHT_PLANE / LIGHT_SPHERE: index into light_aprims (object space == world: lights are tested untransformed)
    float weight;   // product of 1/len over the enclosing lists (list.rs:80-89)
    int32_t pad;
    double dir[4];  // sun direction
};

// ---- the whole scene as device pointers -------------------------------------------------------
struct DeviceScene {
    const BvhNode* nodes;
    const Prim* prims;
    const ShadeTri* shade_tris;
    const APrim* aprims;
    const Xform* xforms;
    const int32_t* chain;
    const GateBox* gates;
    const Instance* instances;
    const ExactMesh* meshes;
    const Volume* volumes;
    const APrim* vol_aprims;
    const int32_t* vol_chain_first; // per vol_aprim: chain first / len (pairs)
    const Material* materials;
    const TexProgram* tex_programs;
    const TexOp* tex_ops;
    const DevImage* images;
    const DevNoise* noises;
    const Light* lights;
    const APrim* light_aprims;
    const int32_t* infinite; // instance ids of sky / sun objects in world-list visit order
    const BroadcastPrim* bcast; // n_global entries, same order as prims[global_first ..]
    int32_t n_nodes, n_prims, n_aprims, n_instances, n_volumes, n_lights, n_infinite, n_materials;
    int32_t has_bvh; // 0: the BVH is empty (no finite primitive, or all of them on the broadcast list)
    // broadcast list: prims[global_first .. global_first + n_global) are the few fp32 quads / spheres that
    // every ray tests when it is generated (k_raygen / k_shade) instead of meeting them in the BVH
    int32_t n_global, global_first;
    int32_t pad;
    float bvh_lo[3], bvh_hi[3]; // world bounds of the primitives that are in the BVH
    int32_t pad2[2];
};

} // namespace rt
