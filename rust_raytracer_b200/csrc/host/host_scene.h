// host_scene.h — owning, growable scene tables behind RtgpuLoadedScene, plus the small
// f64 math the one-shot set-up code needs (Mat4 products for Transform, bounds for the
// object BVH builder). Host only; no CUDA.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <deque>
#include <string>
#include <vector>

#include "rtgpu_host.h"

namespace rthost {

// ---- RNG for scene construction -----------------------------------------------------
// The reference seeds a rand_pcg::Pcg64Mcg from OS entropy (src/main.rs:44,
// scene/golden_monkey.rs:83), so no stream is pinned ("parity unpinned").
// We restate the published PCG-XSL-RR 128/64 (MCG) generator of rand_pcg 0.3.1 and the
// rand 0.8.5 conversions as documented, seeded explicitly.
struct Pcg64Mcg {
    unsigned __int128 state;
    explicit Pcg64Mcg(uint64_t seed) {
        // rand_core::SeedableRng::seed_from_u64: a PCG32 stream fills the 16 seed bytes.
        uint64_t s = seed;
        uint32_t w[4];
        for (int i = 0; i < 4; i++) {
            s = s * 6364136223846793005ULL + 11634580027462260723ULL;
            uint32_t xorshifted = (uint32_t)(((s >> 18u) ^ s) >> 27u);
            uint32_t rot = (uint32_t)(s >> 59u);
            w[i] = (xorshifted >> rot) | (xorshifted << ((32 - rot) & 31));
        }
        unsigned __int128 lo = (uint64_t)w[0] | ((uint64_t)w[1] << 32);
        unsigned __int128 hi = (uint64_t)w[2] | ((uint64_t)w[3] << 32);
        state = (lo | (hi << 64)) | 1; // Mcg128Xsl64::from_seed forces the state odd
    }
    uint64_t next_u64() {
        const unsigned __int128 mult =
            ((unsigned __int128)0x2360ED051FC65DA4ULL << 64) | 0x4385DF649FCCF645ULL;
        state = state * mult;
        uint64_t hi = (uint64_t)(state >> 64), lo = (uint64_t)state;
        uint32_t rot = (uint32_t)(hi >> 58);
        uint64_t x = hi ^ lo;
        return (x >> rot) | (x << ((64 - rot) & 63));
    }
    // rand `Standard` for f64: 53 random bits scaled into [0,1).
    double uniform() { return (double)(next_u64() >> 11) * (1.0 / 9007199254740992.0); }
    // gen_range(lo..hi) for f64
    double range(double lo, double hi) { return lo + (hi - lo) * uniform(); }
    // gen_range(0..n) for integers (widening-multiply with rejection)
    uint64_t below(uint64_t n) {
        uint64_t zone = (n << __builtin_clzll(n)) - 1;
        for (;;) {
            unsigned __int128 m = (unsigned __int128)next_u64() * n;
            if ((uint64_t)m <= zone) return (uint64_t)(m >> 64);
        }
    }
    // N(0,1): rand_distr uses a ziggurat; only the distribution matters (unpinned), so a
    // Marsaglia polar pair is used.
    bool have_spare = false;
    double spare = 0;
    double normal() {
        if (have_spare) { have_spare = false; return spare; }
        for (;;) {
            double u = 2 * uniform() - 1, v = 2 * uniform() - 1, s = u * u + v * v;
            if (s > 0 && s < 1) {
                double m = std::sqrt(-2 * std::log(s) / s);
                spare = v * m; have_spare = true;
                return u * m;
            }
        }
    }
};

// ---- Mat4 helpers (row-major, src/mat4.rs) --------------------------------------------
struct M4 { double m[16]; };
inline M4 m4_identity() { M4 r{}; r.m[0] = r.m[5] = r.m[10] = r.m[15] = 1; return r; }
inline M4 m4_mul(const M4& a, const M4& b) { // mat4.rs:49-66
    M4 r;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++)
            r.m[i * 4 + j] = a.m[i * 4 + 0] * b.m[0 + j] + a.m[i * 4 + 1] * b.m[4 + j] +
                             a.m[i * 4 + 2] * b.m[8 + j] + a.m[i * 4 + 3] * b.m[12 + j];
    return r;
}
inline void m4_apply(const M4& a, const double v[4], double out[4]) { // mat4.rs:69-79
    for (int i = 0; i < 4; i++)
        out[i] = a.m[i * 4] * v[0] + a.m[i * 4 + 1] * v[1] + a.m[i * 4 + 2] * v[2] + a.m[i * 4 + 3] * v[3];
}
inline M4 m4_translation(double x, double y, double z) { M4 r = m4_identity(); r.m[3] = x; r.m[7] = y; r.m[11] = z; return r; }
inline M4 m4_scale(double x, double y, double z) { M4 r = m4_identity(); r.m[0] = x; r.m[5] = y; r.m[10] = z; return r; }
inline M4 m4_rot_x(double t) { M4 r = m4_identity(); double s = std::sin(t), c = std::cos(t); r.m[5] = c; r.m[6] = -s; r.m[9] = s; r.m[10] = c; return r; }
inline M4 m4_rot_y(double t) { M4 r = m4_identity(); double s = std::sin(t), c = std::cos(t); r.m[0] = c; r.m[2] = s; r.m[8] = -s; r.m[10] = c; return r; }
inline M4 m4_rot_z(double t) { M4 r = m4_identity(); double s = std::sin(t), c = std::cos(t); r.m[0] = c; r.m[1] = -s; r.m[4] = s; r.m[5] = c; return r; }

struct Box3 { double lo[3], hi[3]; };

// ---- owning tables ---------------------------------------------------------------------
struct MeshData {
    std::vector<double> positions, normals, uvs;
    std::vector<int32_t> tri_pos, tri_nrm, tri_uv;
};
struct ImageData {
    int width = 0, height = 0, format = 0;
    std::vector<uint8_t> bytes;
};

struct SceneTables {
    std::vector<RtgpuObject> objects;
    std::vector<std::vector<int32_t>> object_children; // per object, flattened in finish()
    std::deque<MeshData> meshes;
    std::vector<RtgpuMaterial> materials;
    std::vector<RtgpuTexture> textures;
    std::deque<ImageData> images;
    std::vector<RtgpuNoise> noises;
    int world_root = -1, lights_root = -1;
    std::string warnings;

    // flattened views (valid after finish())
    std::vector<int32_t> children;
    std::vector<RtgpuMesh> mesh_views;
    std::vector<RtgpuImage> image_views;
    RtgpuSceneDesc desc{};

    int add_object(const RtgpuObject& o, const std::vector<int32_t>& kids = {}) {
        objects.push_back(o);
        object_children.push_back(kids);
        return (int)objects.size() - 1;
    }
    static RtgpuObject blank(int kind) {
        RtgpuObject o;
        std::memset(&o, 0, sizeof o);
        o.kind = kind; o.material = o.texture = o.mesh = -1; o.first_child = 0;
        M4 id = m4_identity();
        std::memcpy(o.transform, id.m, sizeof id.m);
        std::memcpy(o.inv_transform, id.m, sizeof id.m);
        return o;
    }
    int add_texture_const_color(double r, double g, double b) {
        RtgpuTexture t{}; t.kind = RTGPU_TEX_CONSTANT; t.is_float = 0; t.a = t.b = t.c = t.image = t.noise = -1;
        t.value[0] = r; t.value[1] = g; t.value[2] = b; t.scale = 1; t.samples = 7; t.repeat = 1;
        textures.push_back(t); return (int)textures.size() - 1;
    }
    int add_texture_const_float(double k) {
        RtgpuTexture t{}; t.kind = RTGPU_TEX_CONSTANT; t.is_float = 1; t.a = t.b = t.c = t.image = t.noise = -1;
        t.value[0] = k; t.scale = 1; t.samples = 7; t.repeat = 1;
        textures.push_back(t); return (int)textures.size() - 1;
    }
    int add_texture(int kind, bool is_float, int a, int b, int c, double scale) {
        RtgpuTexture t{}; t.kind = kind; t.is_float = is_float; t.a = a; t.b = b; t.c = c; t.image = t.noise = -1;
        t.scale = scale; t.samples = 7; t.repeat = 1;
        textures.push_back(t); return (int)textures.size() - 1;
    }
    int add_material(int kind, int albedo, int roughness, int normal_map, double ior) {
        RtgpuMaterial m{}; m.kind = kind; m.albedo = albedo; m.roughness = roughness; m.normal_map = normal_map; m.ior = ior;
        materials.push_back(m); return (int)materials.size() - 1;
    }
    int add_sphere(const double c[3], double r, int mat) {
        RtgpuObject o = blank(RTGPU_OBJ_SPHERE);
        o.material = mat; o.p0[0] = c[0]; o.p0[1] = c[1]; o.p0[2] = c[2]; o.p0[3] = 1; o.p1[0] = r;
        return add_object(o);
    }
    int add_plane(const double c[3], const double u[3], const double v[3], int mat, bool backface) {
        RtgpuObject o = blank(RTGPU_OBJ_PLANE);
        o.material = mat;
        for (int i = 0; i < 3; i++) { o.p0[i] = c[i]; o.p1[i] = u[i]; o.p2[i] = v[i]; }
        o.p0[3] = 1;
        if (backface) o.flags |= RTGPU_OBJF_BACKFACE;
        return add_object(o);
    }
    int add_list(const std::vector<int32_t>& kids, bool disable_bounds = false) {
        RtgpuObject o = blank(RTGPU_OBJ_LIST);
        if (disable_bounds) o.flags |= RTGPU_OBJF_DISABLE_BOUNDS_CHECK;
        return add_object(o, kids);
    }
    // make_box (src/object/obj_box.rs:8-48): six planes in this order
    int add_box(const double c[3], const double size[3], int mat, bool disable_bounds = false) {
        double hx = size[0] / 2, hy = size[1] / 2, hz = size[2] / 2;
        auto P = [&](double ox, double oy, double oz, double ux, double uy, double uz, double vx, double vy, double vz) {
            double cc[3] = {c[0] + ox, c[1] + oy, c[2] + oz}, u[3] = {ux, uy, uz}, v[3] = {vx, vy, vz};
            return add_plane(cc, u, v, mat, false);
        };
        std::vector<int32_t> k;
        k.push_back(P(0, hy, 0, hx, 0, 0, 0, 0, -hz));
        k.push_back(P(0, -hy, 0, -hx, 0, 0, 0, 0, -hz));
        k.push_back(P(-hx, 0, 0, 0, 0, hz, 0, hy, 0));
        k.push_back(P(hx, 0, 0, 0, 0, -hz, 0, hy, 0));
        k.push_back(P(0, 0, -hz, -hx, 0, 0, 0, hy, 0));
        k.push_back(P(0, 0, hz, hx, 0, 0, 0, hy, 0));
        return add_list(k, disable_bounds);
    }
    int add_transform(int child, const M4& t, const M4& inv) {
        RtgpuObject o = blank(RTGPU_OBJ_TRANSFORM);
        std::memcpy(o.transform, t.m, sizeof t.m);
        std::memcpy(o.inv_transform, inv.m, sizeof inv.m);
        return add_object(o, {child});
    }

    Box3 bounds_of(int obj) const; // reference get_bounding_box semantics (for the BVH builder)
    int build_bvh(std::vector<int32_t> objs, const bool axes[3], Pcg64Mcg& rng); // bvh.rs:32-81
    void finish();
};

// Transform builder that mirrors Transform::{translate,rotate_*,scale} (transform.rs:44-96):
// transform = Op * transform; inv_transform = inv_transform * Op^-1.
struct TransformOps {
    M4 t = m4_identity(), inv = m4_identity();
    void translate(double x, double y, double z) { t = m4_mul(m4_translation(x, y, z), t); inv = m4_mul(inv, m4_translation(-x, -y, -z)); }
    void rotate_x(double a) { t = m4_mul(m4_rot_x(a), t); inv = m4_mul(inv, m4_rot_x(-a)); }
    void rotate_y(double a) { t = m4_mul(m4_rot_y(a), t); inv = m4_mul(inv, m4_rot_y(-a)); }
    void rotate_z(double a) { t = m4_mul(m4_rot_z(a), t); inv = m4_mul(inv, m4_rot_z(-a)); }
    void scale(double x, double y, double z) { t = m4_mul(m4_scale(x, y, z), t); inv = m4_mul(inv, m4_scale(1.0 / x, 1.0 / y, 1.0 / z)); }
};

inline double deg_to_rad(double d) { return d / 180.0 * M_PI; } // utils.rs:5-7

// loaders
int load_obj(const std::string& path, MeshData& out, std::string& err);   // obj_loader.cpp
int decode_image(const std::string& path, ImageData& out, std::string& err); // png.cpp (+ callback)
int load_dsl_scene(const std::string& path, RtgpuConfig& cfg, SceneTables& st, Pcg64Mcg& rng, std::string& err);
int load_glb_scene(const std::string& path, RtgpuConfig& cfg, SceneTables& st, std::string& err); // gltf_loader.cpp (loaders/assimp.rs)
int build_builtin_scene(const std::string& name, RtgpuConfig& cfg, const std::string& asset_root,
                        SceneTables& st, Pcg64Mcg& rng, std::string& err);
void merge_scene_config(const RtgpuSceneConfig& base, const RtgpuSceneConfig& over, RtgpuSceneConfig& out);
RtgpuSceneConfig default_scene_config();
void set_error(const std::string& msg);

} // namespace rthost
