// host_scene.cpp — scene tables, object-BVH builder, config merge, Camera::init and the
// C entry points of rtgpu_host.h. One-shot set-up code; semantics follow the reference.
#include "host_scene.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <limits>

namespace rthost {

static thread_local std::string g_host_error;
void set_error(const std::string& msg) { g_host_error = msg; }
const std::string& host_error() { return g_host_error; }

static const double EPS_PAD = 0.001; // aabb.rs:9

// get_bounding_box of every Hit implementor, as the reference computes it.
Box3 SceneTables::bounds_of(int idx) const {
    const RtgpuObject& o = objects[idx];
    const double INF = std::numeric_limits<double>::infinity();
    Box3 b;
    switch (o.kind) {
    case RTGPU_OBJ_SPHERE: { // sphere.rs:27-30 (negative radius => inverted box)
        double r = o.p1[0];
        for (int i = 0; i < 3; i++) { b.lo[i] = o.p0[i] - r; b.hi[i] = o.p0[i] + r; }
        return b;
    }
    case RTGPU_OBJ_PLANE: { // plane.rs:41-48 via aabb::get_bounding_box (padded)
        for (int i = 0; i < 3; i++) { b.lo[i] = INF; b.hi[i] = -INF; }
        for (int su = -1; su <= 1; su += 2)
            for (int sv = -1; sv <= 1; sv += 2)
                for (int i = 0; i < 3; i++) {
                    double c = o.p0[i] + su * o.p1[i] + sv * o.p2[i];
                    b.lo[i] = std::min(b.lo[i], c); b.hi[i] = std::max(b.hi[i], c);
                }
        for (int i = 0; i < 3; i++) { b.lo[i] -= EPS_PAD; b.hi[i] += EPS_PAD; }
        return b;
    }
    case RTGPU_OBJ_MESH: { // mesh.rs:45
        const MeshData& m = meshes[o.mesh];
        for (int i = 0; i < 3; i++) { b.lo[i] = INF; b.hi[i] = -INF; }
        for (size_t v = 0; v + 2 < m.positions.size(); v += 3)
            for (int i = 0; i < 3; i++) {
                b.lo[i] = std::min(b.lo[i], m.positions[v + i]); b.hi[i] = std::max(b.hi[i], m.positions[v + i]);
            }
        for (int i = 0; i < 3; i++) { b.lo[i] -= EPS_PAD; b.hi[i] += EPS_PAD; }
        return b;
    }
    case RTGPU_OBJ_LIST: { // list.rs:51-55: bounds = combine([bounds, child]) per add => cumulative padding
        for (int i = 0; i < 3; i++) { b.lo[i] = INF; b.hi[i] = -INF; }
        for (int c : object_children[idx]) {
            Box3 cb = bounds_of(c);
            for (int i = 0; i < 3; i++) {
                b.lo[i] = std::min(b.lo[i], cb.lo[i]) - EPS_PAD;
                b.hi[i] = std::max(b.hi[i], cb.hi[i]) + EPS_PAD;
            }
        }
        return b;
    }
    case RTGPU_OBJ_BVH: { // bvh.rs:49-76
        const auto& k = object_children[idx];
        if (objects[k[1]].kind == RTGPU_OBJ_NULL) return bounds_of(k[0]);
        Box3 a = bounds_of(k[0]), c = bounds_of(k[1]);
        for (int i = 0; i < 3; i++) { b.lo[i] = std::min(a.lo[i], c.lo[i]) - EPS_PAD; b.hi[i] = std::max(a.hi[i], c.hi[i]) + EPS_PAD; }
        return b;
    }
    case RTGPU_OBJ_TRANSFORM: { // transform.rs:98-118
        Box3 cb = bounds_of(object_children[idx][0]);
        M4 t; std::memcpy(t.m, o.transform, sizeof t.m);
        for (int i = 0; i < 3; i++) { b.lo[i] = INF; b.hi[i] = -INF; }
        for (int c = 0; c < 8; c++) {
            double p[4] = {(c & 4) ? cb.hi[0] : cb.lo[0], (c & 2) ? cb.hi[1] : cb.lo[1], (c & 1) ? cb.hi[2] : cb.lo[2], 1.0}, q[4];
            m4_apply(t, p, q);
            for (int i = 0; i < 3; i++) { b.lo[i] = std::min(b.lo[i], q[i]); b.hi[i] = std::max(b.hi[i], q[i]); }
        }
        for (int i = 0; i < 3; i++) { b.lo[i] -= EPS_PAD; b.hi[i] += EPS_PAD; }
        return b;
    }
    case RTGPU_OBJ_VOLUME: return bounds_of(object_children[idx][0]); // volume.rs:73-75
    case RTGPU_OBJ_SKY:
    case RTGPU_OBJ_SUN: // sky.rs:54-59, sun.rs:63-68
        for (int i = 0; i < 3; i++) { b.lo[i] = std::numeric_limits<double>::lowest(); b.hi[i] = std::numeric_limits<double>::max(); }
        return b;
    default: // NullObject: [INFINITY, -INFINITY]
        for (int i = 0; i < 3; i++) { b.lo[i] = INF; b.hi[i] = -INF; }
        return b;
    }
}

// BoundingVolumeHierarchyNode::from (bvh.rs:32-81): random allowed axis (drawn before the
// size check), leaf cases pop from the back, otherwise sort by AABB min and split at n/2.
int SceneTables::build_bvh(std::vector<int32_t> objs, const bool axes[3], Pcg64Mcg& rng) {
    int axis = (int)rng.below(3);
    while (!axes[axis]) axis = (int)rng.below(3);
    RtgpuObject node = blank(RTGPU_OBJ_BVH);
    size_t n = objs.size();
    if (n == 1) {
        int nul = add_object(blank(RTGPU_OBJ_NULL));
        return add_object(node, {objs[0], nul});
    }
    if (n == 2) return add_object(node, {objs[1], objs[0]});
    std::vector<std::pair<double, int32_t>> keyed;
    for (int32_t o : objs) keyed.push_back({bounds_of(o).lo[axis], o});
    std::stable_sort(keyed.begin(), keyed.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
    size_t mid = n / 2;
    std::vector<int32_t> first, second;
    for (size_t i = 0; i < n; i++) (i < mid ? first : second).push_back(keyed[i].second);
    int c0 = build_bvh(first, axes, rng);
    int c1 = build_bvh(second, axes, rng);
    return add_object(node, {c0, c1});
}

void SceneTables::finish() {
    children.clear();
    for (size_t i = 0; i < objects.size(); i++) {
        objects[i].first_child = (int32_t)children.size();
        objects[i].n_children = (int32_t)object_children[i].size();
        for (int32_t c : object_children[i]) children.push_back(c);
    }
    mesh_views.clear();
    for (const MeshData& m : meshes) {
        RtgpuMesh v{};
        v.positions = m.positions.data(); v.normals = m.normals.data();
        v.uvs = m.uvs.empty() ? nullptr : m.uvs.data();
        v.tri_pos = m.tri_pos.data(); v.tri_nrm = m.tri_nrm.data();
        v.tri_uv = m.tri_uv.empty() ? nullptr : m.tri_uv.data();
        v.n_positions = (int32_t)(m.positions.size() / 3); v.n_normals = (int32_t)(m.normals.size() / 3);
        v.n_uvs = (int32_t)(m.uvs.size() / 2); v.n_triangles = (int32_t)(m.tri_pos.size() / 3);
        mesh_views.push_back(v);
    }
    image_views.clear();
    for (const ImageData& im : images) {
        RtgpuImage v{}; v.width = im.width; v.height = im.height; v.format = im.format; v.data = im.bytes.data();
        image_views.push_back(v);
    }
    desc.abi_version = RTGPU_ABI_VERSION;
    desc.world_root = world_root; desc.lights_root = lights_root;
    desc.n_objects = (int32_t)objects.size(); desc.objects = objects.data();
    desc.n_children = (int32_t)children.size(); desc.children = children.data();
    desc.n_meshes = (int32_t)mesh_views.size(); desc.meshes = mesh_views.data();
    desc.n_materials = (int32_t)materials.size(); desc.materials = materials.data();
    desc.n_textures = (int32_t)textures.size(); desc.textures = textures.data();
    desc.n_images = (int32_t)image_views.size(); desc.images = image_views.data();
    desc.n_noises = (int32_t)noises.size(); desc.noises = noises.data();
}

// ---- SceneConfig (config.rs:20-43) ------------------------------------------------------
RtgpuSceneConfig default_scene_config() {
    RtgpuSceneConfig c{};
    c.has_output_width = 1; c.output_width = 600;
    c.has_aspect_ratio = 1; c.aspect_ratio = 1.5;
    c.has_focal_length = 1; c.focal_length = 50.0;
    c.has_f_number = 0; c.has_focus_distance = 0;
    c.has_camera_pos = 1; c.camera_pos[0] = 0; c.camera_pos[1] = 0; c.camera_pos[2] = 1;
    c.has_camera_target = 1; c.camera_target[0] = c.camera_target[1] = c.camera_target[2] = 0;
    c.has_background = 1; c.background[0] = c.background[1] = c.background[2] = 0;
    return c;
}

void merge_scene_config(const RtgpuSceneConfig& base, const RtgpuSceneConfig& over, RtgpuSceneConfig& out) {
    RtgpuSceneConfig r = base;
#define MERGE_SCALAR(f) if (over.has_##f) { r.has_##f = 1; r.f = over.f; }
#define MERGE_VEC(f) if (over.has_##f) { r.has_##f = 1; for (int i = 0; i < 3; i++) r.f[i] = over.f[i]; }
    MERGE_SCALAR(output_width) MERGE_SCALAR(aspect_ratio) MERGE_SCALAR(focal_length)
    MERGE_SCALAR(f_number) MERGE_SCALAR(focus_distance)
    MERGE_VEC(camera_pos) MERGE_VEC(camera_target) MERGE_VEC(background)
#undef MERGE_SCALAR
#undef MERGE_VEC
    out = r;
}

} // namespace rthost

using namespace rthost;

struct RtgpuLoadedScene { SceneTables st; };

static RtgpuImageDecodeFn g_decode_fn = nullptr;
static void* g_decode_user = nullptr;
namespace rthost {
bool image_decoder_registered() { return g_decode_fn != nullptr; }
int decode_image_builtin(const std::string& path, ImageData& out, std::string& err); // png.cpp
int decode_via_callback(const std::string& path, ImageData& out, std::string& err) {
    if (!g_decode_fn) { err = "no decoder registered for " + path; return RTGPU_ERR_UNSUPPORTED; }
    RtgpuImage im{};
    int rc = g_decode_fn(path.c_str(), &im, g_decode_user);
    if (rc != 0 || !im.data) { err = "image decoder callback failed for " + path; return RTGPU_ERR_IO; }
    size_t bpc = im.format == RTGPU_IMG_RGB8 ? 1 : im.format == RTGPU_IMG_RGB16 ? 2 : 4;
    size_t n = (size_t)im.width * im.height * 3 * bpc;
    out.width = im.width; out.height = im.height; out.format = im.format;
    out.bytes.assign((const uint8_t*)im.data, (const uint8_t*)im.data + n);
    free((void*)im.data);
    return 0;
}
}

extern "C" {

int rtgpu_set_image_decoder(RtgpuImageDecodeFn fn, void* user) { g_decode_fn = fn; g_decode_user = user; return 0; }

// Buffer::from_file (buffer.rs:30-48) up to the integer texels: what `image <file>` textures hold
int rtgpu_decode_image(const char* path, int builtin_only, RtgpuImage* out) {
    if (!path || !out) { rthost::set_error("rtgpu_decode_image: null argument"); return RTGPU_ERR_INVALID; }
    ImageData im;
    std::string err;
    const int rc = builtin_only ? decode_image_builtin(path, im, err) : decode_image(path, im, err);
    if (rc != 0) { rthost::set_error(err); return rc; }
    void* mem = std::malloc(im.bytes.size() ? im.bytes.size() : 1);
    if (!mem) { rthost::set_error("rtgpu_decode_image: out of memory"); return RTGPU_ERR_IO; }
    std::memcpy(mem, im.bytes.data(), im.bytes.size());
    out->width = im.width; out->height = im.height; out->format = im.format; out->reserved = 0; out->data = mem;
    return 0;
}

// one error slot per thread for the whole library (front-end and CUDA side)
const char* rtgpu_last_error(void) { return host_error().c_str(); }

int rtgpu_scene_load(RtgpuConfig* cfg, const char* asset_root, RtgpuLoadedScene** out) {
    if (!cfg || !out) { set_error("rtgpu_scene_load: null argument"); return RTGPU_ERR_INVALID; }
    auto* ls = new RtgpuLoadedScene();
    Pcg64Mcg rng(cfg->scene_seed);
    std::string err, name = cfg->scene_name, root = asset_root ? asset_root : "";
    int rc;
    static const char* builtin[] = {"", "golden_monkey", "earth", "perlin", "light_test", "cornell", "cornell_smoke", "tonemap_test"};
    bool is_builtin = false;
    for (const char* b : builtin) if (name == b) is_builtin = true;
    if (is_builtin) rc = build_builtin_scene(name, *cfg, root, ls->st, rng, err);
    else if (name.rfind("model:", 0) == 0) rc = load_glb_scene(name.substr(6), *cfg, ls->st, err); // main.rs:38-41 (binary glTF only: no Assimp)
    else rc = load_dsl_scene(name, *cfg, ls->st, rng, err);
    if (rc != 0) { set_error(err); delete ls; return rc; }
    ls->st.finish();
    *out = ls;
    return 0;
}

int rtgpu_scene_free(RtgpuLoadedScene* s) { delete s; return 0; }
const RtgpuSceneDesc* rtgpu_loaded_scene_desc(const RtgpuLoadedScene* s) { return s ? &s->st.desc : nullptr; }
const char* rtgpu_loaded_scene_warnings(const RtgpuLoadedScene* s) { return s ? s->st.warnings.c_str() : ""; }

// Camera::new + init (camera.rs:47-130)
int rtgpu_camera_init(const RtgpuConfig* cfg, RtgpuCameraDesc* cam, RtgpuRenderParams* prm) {
    if (!cfg || !cam) { set_error("rtgpu_camera_init: null argument"); return RTGPU_ERR_INVALID; }
    const RtgpuSceneConfig& sc = cfg->scene;
    if (!sc.has_output_width || !sc.has_aspect_ratio || !sc.has_focal_length || !sc.has_camera_pos || !sc.has_camera_target) {
        set_error("rtgpu_camera_init: scene config not merged (missing required field)");
        return RTGPU_ERR_INVALID;
    }
    int W = sc.output_width;
    int H = (int)((double)W / sc.aspect_ratio);      // camera.rs:87 (`as usize` truncates)
    if (H < 1) H = 1;
    double dir[3], w[3], u[3], v[3];
    for (int i = 0; i < 3; i++) dir[i] = sc.camera_pos[i] - sc.camera_target[i];
    double len = std::sqrt(dir[0] * dir[0] + dir[1] * dir[1] + dir[2] * dir[2]);
    double focus_dist = sc.has_focus_distance ? sc.focus_distance : len;   // :90
    double h = 24.0 / sc.focal_length;                                     // :93
    double real_aspect = (double)W / (double)H;                            // :96
    double vh = focus_dist * h, vw = vh * real_aspect;                     // :97-98
    for (int i = 0; i < 3; i++) w[i] = dir[i] / len;                       // :101
    const double up[3] = {0, 1, 0};
    u[0] = up[1] * w[2] - up[2] * w[1]; u[1] = up[2] * w[0] - up[0] * w[2]; u[2] = up[0] * w[1] - up[1] * w[0]; // :102 (not normalised)
    v[0] = w[1] * u[2] - w[2] * u[1]; v[1] = w[2] * u[0] - w[0] * u[2]; v[2] = w[0] * u[1] - w[1] * u[0];       // :103
    double vu[3], vv[3];
    for (int i = 0; i < 3; i++) { vu[i] = u[i] * vw; vv[i] = -v[i] * vh; }  // :107-108
    for (int i = 0; i < 3; i++) { cam->pixel_delta_u[i] = vu[i] / (double)W; cam->pixel_delta_v[i] = vv[i] / (double)H; } // :110-113
    for (int i = 0; i < 3; i++) {
        double ul = sc.camera_pos[i] - w[i] * focus_dist - vu[i] / 2.0 - vv[i] / 2.0;          // :117-118
        cam->first_pixel[i] = ul + (cam->pixel_delta_u[i] + cam->pixel_delta_v[i]) * 0.5;      // :121
        cam->position[i] = sc.camera_pos[i];
        cam->basis_u[i] = u[i]; cam->basis_v[i] = v[i];
        cam->background[i] = sc.has_background ? sc.background[i] : 0.0;
    }
    cam->width = W; cam->height = H;
    cam->aperture_radius = sc.has_f_number ? (sc.focal_length / 1000.0) / sc.f_number : -1.0;   // :125-129
    if (prm) {
        std::memset(prm, 0, sizeof *prm);
        prm->replicas = cfg->thread_count; prm->sqrt_spt = cfg->sqrt_samples_per_thread;
        prm->max_depth = cfg->max_depth; prm->light_bias = cfg->light_bias;
        prm->seed = cfg->seed; prm->n_gpus = cfg->n_gpus > 0 ? cfg->n_gpus : 1;
        if (cfg->fixed) prm->flags |= RTGPU_FLAG_FIXED;
    }
    return 0;
}

} // extern "C"
