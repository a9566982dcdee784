// builtin_scenes.cpp — the reference's hard-coded scenes (src/scene/*.rs) as flat tables.
// golden_monkey is BASELINE config C2 (the default scene); its random spheres and BVH
// axis picks come from the explicitly seeded scene RNG (the reference uses OS entropy,
// golden_monkey.rs:83, so no stream is pinned).
#include "host_scene.h"

namespace rthost {

static void apply_defaults(RtgpuConfig& cfg, const RtgpuSceneConfig& scene_defaults) {
    // SceneConfig::merge(&SceneConfig::merge(&DEFAULT, &scene_defaults), &config.scene)
    RtgpuSceneConfig a, b;
    merge_scene_config(default_scene_config(), scene_defaults, a);
    merge_scene_config(a, cfg.scene, b);
    cfg.scene = b;
}

static RtgpuSceneConfig defaults(int w, double aspect, double focal, double f_number,
                                 double px, double py, double pz, double tx, double ty, double tz) {
    RtgpuSceneConfig c{};
    c.has_output_width = 1; c.output_width = w;
    c.has_aspect_ratio = 1; c.aspect_ratio = aspect;
    c.has_focal_length = 1; c.focal_length = focal;
    if (f_number > 0) { c.has_f_number = 1; c.f_number = f_number; }
    c.has_camera_pos = 1; c.camera_pos[0] = px; c.camera_pos[1] = py; c.camera_pos[2] = pz;
    c.has_camera_target = 1; c.camera_target[0] = tx; c.camera_target[1] = ty; c.camera_target[2] = tz;
    return c; // background: None
}

static int add_mesh(SceneTables& st, const std::string& path, int mat, std::string& err) {
    st.meshes.emplace_back();
    int rc = load_obj(path, st.meshes.back(), err);
    if (rc) { st.meshes.pop_back(); return -1; }
    RtgpuObject o = SceneTables::blank(RTGPU_OBJ_MESH);
    o.mesh = (int)st.meshes.size() - 1; o.material = mat;
    return st.add_object(o);
}

static int add_sun(SceneTables& st, double x, double y, double z, int tex) { // Vec4::vec => w stays 0
    RtgpuObject o = SceneTables::blank(RTGPU_OBJ_SUN);
    double len = std::sqrt(x * x + y * y + z * z);
    o.p0[0] = x / len; o.p0[1] = y / len; o.p0[2] = z / len; o.p0[3] = 0.0;
    o.texture = tex;
    return st.add_object(o);
}
static int add_sky(SceneTables& st, int tex) {
    RtgpuObject o = SceneTables::blank(RTGPU_OBJ_SKY);
    o.texture = tex;
    return st.add_object(o);
}
static int plane(SceneTables& st, double cx, double cy, double cz, double ux, double uy, double uz,
                 double vx, double vy, double vz, int mat, bool backface = false) {
    double c[3] = {cx, cy, cz}, u[3] = {ux, uy, uz}, v[3] = {vx, vy, vz};
    return st.add_plane(c, u, v, mat, backface);
}

// scene/golden_monkey.rs:23-138
static int golden_monkey(RtgpuConfig& cfg, const std::string& root, SceneTables& st, Pcg64Mcg& rng, std::string& err) {
    apply_defaults(cfg, defaults(600, 1.5, 50.0, 2.8, 5, 2, 9, 0, 0.5, 0));
    int tex_ground = st.add_texture(RTGPU_TEX_CHECKER, false, st.add_texture_const_color(0.2, 0.3, 0.1),
                                    st.add_texture_const_color(0.9, 0.9, 0.9), -1, 0.02);
    int mat_ground = st.add_material(RTGPU_MAT_LAMBERTIAN, tex_ground, -1, -1, 1.0);
    int mat_metal = st.add_material(RTGPU_MAT_METAL, st.add_texture_const_color(0.8, 0.6, 0.2), st.add_texture_const_float(0.05), -1, 1.0);
    int mat_glass = st.add_material(RTGPU_MAT_DIELECTRIC, -1, -1, -1, 1.5);
    int sky = add_sky(st, st.add_texture_const_color(0.2, 0.6, 2.0));
    int sun = add_sun(st, -1.0, 1.0, 0.0, st.add_texture_const_color(20.0, 20.0, 20.0));
    int floor = plane(st, 0, 0, 0, 20, 0, 0, 0, 0, -20, mat_ground);
    int mesh = add_mesh(st, root + "scenes/resource/monkey.obj", mat_metal, err);
    if (mesh < 0) return RTGPU_ERR_IO;
    TransformOps ops; ops.translate(0.0, 1.0, 0.0);
    int mesh_t = st.add_transform(mesh, ops.t, ops.inv);

    std::vector<int32_t> spheres;
    for (int i = -10; i < 11; i++)
        for (int j = -10; j < 11; j++) {
            double c[3] = {(double)i + rng.range(0.0, 0.9), 0.2, (double)j + rng.range(0.0, 0.9)};
            double dx = c[0], dy = c[1] - 0.2, dz = c[2];
            if (dx * dx + dy * dy + dz * dz < 1.0) continue;
            double mat_type = rng.range(0.0, 1.0);
            if (mat_type < 0.95) {
                double a[3], b[3];
                for (int k = 0; k < 3; k++) a[k] = rng.uniform(); // Vec4::random_vec x2, multiplied
                for (int k = 0; k < 3; k++) b[k] = rng.uniform();
                int mat = st.add_material(RTGPU_MAT_GLOSSY, st.add_texture_const_color(a[0] * b[0], a[1] * b[1], a[2] * b[2]),
                                          st.add_texture_const_float(0.1), -1, 1.5);
                spheres.push_back(st.add_sphere(c, 0.2, mat));
            } else {
                spheres.push_back(st.add_sphere(c, 0.2, mat_glass));
                spheres.push_back(st.add_sphere(c, -0.18, mat_glass));
            }
        }
    const bool axes_xz[3] = {true, false, true};
    int bvh = st.build_bvh(spheres, axes_xz, rng);
    st.world_root = st.add_list({mesh_t, floor, bvh, sky, sun});
    st.lights_root = st.add_list({sky, sun});
    return 0;
}

// scene/cornell_box.rs (same geometry as scenes/cornell)
static int cornell(RtgpuConfig& cfg, SceneTables& st) {
    apply_defaults(cfg, defaults(600, 1.0, 33.0, -1, 277.5, 277.5, -800, 277.5, 277.5, 0));
    int white = st.add_material(RTGPU_MAT_LAMBERTIAN, st.add_texture_const_color(0.73, 0.73, 0.73), -1, -1, 1);
    int green = st.add_material(RTGPU_MAT_LAMBERTIAN, st.add_texture_const_color(0.12, 0.45, 0.15), -1, -1, 1);
    int red = st.add_material(RTGPU_MAT_LAMBERTIAN, st.add_texture_const_color(0.65, 0.05, 0.05), -1, -1, 1);
    int light_m = st.add_material(RTGPU_MAT_EMISSIVE, st.add_texture_const_color(15, 15, 15), -1, -1, 1);
    int glass = st.add_material(RTGPU_MAT_DIELECTRIC, -1, -1, -1, 1.5);
    int checker = st.add_texture(RTGPU_TEX_CHECKER, true, st.add_texture_const_float(0.0), st.add_texture_const_float(1.0), -1, 0.25);
    int gloss = st.add_material(RTGPU_MAT_GLOSSY, st.add_texture_const_color(0.95, 0.95, 0.95), checker, -1, 1.5);
    int floor = plane(st, 277.5, 0, 277.5, 277.5, 0, 0, 0, 0, -277.5, gloss);
    int ceiling = plane(st, 277.5, 555, 277.5, 277.5, 0, 0, 0, 0, 277.5, white);
    int back = plane(st, 277.5, 277.5, 555, 0, 277.5, 0, 277.5, 0, 0, white);
    int left = plane(st, 555, 277.5, 277.5, 0, 277.5, 0, 0, 0, -277.5, green);
    int right = plane(st, 0, 277.5, 277.5, 0, 277.5, 0, 0, 0, 277.5, red);
    int light = plane(st, 277.5, 554.9, 277.5, -65, 0, 0, 0, 0, -52.5, light_m, true);
    double c0[3] = {0, 0, 0}, size[3] = {165, 330, 165};
    int box = st.add_box(c0, size, white);
    TransformOps ops; ops.translate(82.5, 165.0, 82.5); ops.rotate_y(deg_to_rad(18.0)); ops.translate(265.0, 0.0, 295.0);
    int box_t = st.add_transform(box, ops.t, ops.inv);
    double bc[3] = {212.5, 82.51, 147.5};
    int ball = st.add_sphere(bc, 82.5, glass);
    st.world_root = st.add_list({floor, ceiling, back, left, right, light, box_t, ball});
    st.lights_root = st.add_list({light, ball});
    return 0;
}

// scene/cornell_smoke.rs:14-138 (volumes bounded by one-sided boxes)
static int cornell_smoke(RtgpuConfig& cfg, SceneTables& st) {
    apply_defaults(cfg, defaults(600, 1.0, 35.0, -1, 0, 0, 110, 0, 0, 0));
    int white = st.add_material(RTGPU_MAT_LAMBERTIAN, st.add_texture_const_color(0.73, 0.73, 0.73), -1, -1, 1);
    int green = st.add_material(RTGPU_MAT_LAMBERTIAN, st.add_texture_const_color(0.12, 0.45, 0.15), -1, -1, 1);
    int red = st.add_material(RTGPU_MAT_LAMBERTIAN, st.add_texture_const_color(0.65, 0.05, 0.05), -1, -1, 1);
    int light_m = st.add_material(RTGPU_MAT_EMISSIVE, st.add_texture_const_color(15, 15, 15), -1, -1, 1);
    int smoke = st.add_material(RTGPU_MAT_ISOTROPIC, st.add_texture_const_color(0, 0, 0), -1, -1, 1);
    int fog = st.add_material(RTGPU_MAT_ISOTROPIC, st.add_texture_const_color(1, 1, 1), -1, -1, 1);
    int floor = plane(st, 0, -27.5, 0, -27.5, 0, 0, 0, 0, 27.5, white);
    int ceiling = plane(st, 0, 27.5, 0, 27.5, 0, 0, 0, 0, -27.5, white);
    int back = plane(st, 0, 0, -27.5, 0, 27.5, 0, -27.5, 0, 0, white);
    int left = plane(st, -27.5, 0, 0, 0, 27.5, 0, 0, 0, -27.5, green);
    int right = plane(st, 27.5, 0, 0, 0, 27.5, 0, 0, 0, 27.5, red);
    int light = plane(st, 0, 27.49, 0, 13, 0, 0, 0, 0, 10.5, light_m);
    double c0[3] = {0, 0, 0}, s1[3] = {16.5, 16.5, 16.5}, s2[3] = {16.5, 33.0, 16.5};
    int box1 = st.add_box(c0, s1, white, true);
    TransformOps o1; o1.rotate_y(deg_to_rad(-15.0)); o1.translate(27.5 - 21.25, 8.25 - 27.5, 27.5 - 14.75);
    RtgpuObject v1 = SceneTables::blank(RTGPU_OBJ_VOLUME); v1.material = smoke; v1.p1[0] = 0.15;
    int vol1 = st.add_object(v1, {st.add_transform(box1, o1.t, o1.inv)});
    int box2 = st.add_box(c0, s2, white, true);
    TransformOps o2; o2.rotate_y(deg_to_rad(18.0)); o2.translate(27.5 - 34.75, 16.5 - 27.5, 27.5 - 37.75);
    RtgpuObject v2 = SceneTables::blank(RTGPU_OBJ_VOLUME); v2.material = fog; v2.p1[0] = 0.15;
    int vol2 = st.add_object(v2, {st.add_transform(box2, o2.t, o2.inv)});
    st.world_root = st.add_list({floor, ceiling, back, left, right, light, vol1, vol2});
    st.lights_root = st.add_list({light});
    return 0;
}

// scene/light_test.rs
static int light_test(RtgpuConfig& cfg, const std::string& root, SceneTables& st, std::string& err) {
    apply_defaults(cfg, defaults(600, 1.5, 70.0, 4.0, 10, 1, 6, 0, 0, 0));
    int tex_ground = st.add_texture(RTGPU_TEX_CHECKER, false, st.add_texture_const_color(0.2, 0.3, 0.1),
                                    st.add_texture_const_color(0.9, 0.9, 0.9), -1, 0.02);
    int mat_ground = st.add_material(RTGPU_MAT_LAMBERTIAN, tex_ground, -1, -1, 1.0);
    int mat_metal = st.add_material(RTGPU_MAT_METAL, st.add_texture_const_color(0.8, 0.6, 0.2), st.add_texture_const_float(0.05), -1, 1.0);
    int l1 = st.add_material(RTGPU_MAT_EMISSIVE, st.add_texture_const_color(7, 1, 7), -1, -1, 1);
    int l2 = st.add_material(RTGPU_MAT_EMISSIVE, st.add_texture_const_color(1, 6, 8), -1, -1, 1);
    double c1[3] = {-1, 0, 1}, c2[3] = {2, 0.5, -1.2};
    int s1 = st.add_sphere(c1, 0.5, l1), s2 = st.add_sphere(c2, 0.4, l2);
    st.objects[s1].p0[3] = 0.0; st.objects[s2].p0[3] = 0.0; // built with Vec4::vec (light_test.rs)
    int floor = plane(st, 0, -1, 0, -10, 0, 0, 0, 0, 10, mat_ground);
    int mesh = add_mesh(st, root + "scenes/resource/monkey.obj", mat_metal, err);
    if (mesh < 0) return RTGPU_ERR_IO;
    TransformOps ops; ops.translate(0.0, 0.0, -1.5);
    int mesh_t = st.add_transform(mesh, ops.t, ops.inv);
    st.world_root = st.add_list({s1, s2, floor, mesh_t});
    st.lights_root = st.add_list({s1, s2});
    return 0;
}

// scene/earth.rs (lambertian earth, unlike scenes/earth which is glossy)
static int earth(RtgpuConfig& cfg, const std::string& root, SceneTables& st, std::string& err) {
    apply_defaults(cfg, defaults(600, 1.5, 70.0, -1, 13, 2, 3, 0, 0, 0));
    st.images.emplace_back();
    if (decode_image(root + "scenes/resource/earthmap.jpg", st.images.back(), err) != 0) { st.images.pop_back(); return RTGPU_ERR_IO; }
    int tex = st.add_texture(RTGPU_TEX_IMAGE, false, -1, -1, -1, 1.0);
    st.textures[tex].image = (int)st.images.size() - 1;
    int mat_earth = st.add_material(RTGPU_MAT_LAMBERTIAN, tex, -1, -1, 1);
    int mat_floor = st.add_material(RTGPU_MAT_LAMBERTIAN, st.add_texture_const_color(0.5, 0.5, 0.5), -1, -1, 1);
    int sun = add_sun(st, 0, 1, 2, st.add_texture_const_color(10, 10, 10));
    double c[3] = {0, 0, 0};
    int sphere = st.add_sphere(c, 1.5, mat_earth);
    int floor = plane(st, 0, -1.5, 0, -10, 0, 0, 0, 0, 10, mat_floor);
    st.world_root = st.add_list({sphere, floor, sun});
    st.lights_root = st.add_list({sun});
    return 0;
}

int build_builtin_scene(const std::string& name, RtgpuConfig& cfg, const std::string& asset_root,
                        SceneTables& st, Pcg64Mcg& rng, std::string& err) {
    std::string root = asset_root;
    if (!root.empty() && root.back() != '/') root += '/';
    if (name.empty() || name == "golden_monkey") return golden_monkey(cfg, root, st, rng, err);
    if (name == "cornell") return cornell(cfg, st);
    if (name == "cornell_smoke") return cornell_smoke(cfg, st);
    if (name == "light_test") return light_test(cfg, root, st, err);
    if (name == "earth") return earth(cfg, root, st, err);
    err = "built-in scene '" + name + "' is not restated; use the scenes/" + name + " DSL file";
    return RTGPU_ERR_UNSUPPORTED;
}

} // namespace rthost
