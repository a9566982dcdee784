// scene_loader.cpp — SceneLoader (reference src/loaders/scene.rs:65-869), restated.
// Line-based DSL: `label: type params`, `$ref`, `(inline)`, `@config key = value`.
// Errors on a line are warnings and the line is skipped (scene.rs:127-135); the scene
// fails only when `world` or `lights` is missing (scene.rs:138-155).
#include <fstream>
#include <map>
#include <sstream>

#include "host_scene.h"

namespace rthost {

bool parse_f64(const std::string& s, double& out);
int parse_vec3(const std::string& s, double out[3]);
int decode_via_callback(const std::string& path, ImageData& out, std::string& err);

namespace {

static std::string trim(const std::string& s) {
    size_t a = 0, b = s.size();
    while (a < b && std::isspace((unsigned char)s[a])) a++;
    while (b > a && std::isspace((unsigned char)s[b - 1])) b--;
    return s.substr(a, b - a);
}

enum class Ent { Object, Material, TexColor, TexFloat, Noise };
struct Entity { Ent type; int index; };
struct Fail { std::string msg; };

struct Loader {
    SceneTables& st;
    Pcg64Mcg& rng;
    std::string asset_path;
    RtgpuSceneConfig scene_config = default_scene_config();
    std::map<std::string, int> objects, materials, color_tex, float_tex, noise;

    Loader(SceneTables& s, Pcg64Mcg& r, const std::string& ap) : st(s), rng(r), asset_path(ap) {}

    // scene.rs:214-245
    static std::vector<std::string> parse_params(const std::string& decl) {
        std::vector<std::string> params;
        std::string cur;
        int nest = 0;
        for (char ch : decl) {
            if (ch == '(') { cur.push_back('('); nest++; }
            else if (ch == ')') { cur.push_back(')'); nest--; }
            else if (ch == ' ') {
                if (nest > 0) cur.push_back(' ');
                else { params.push_back(cur); cur.clear(); }
            } else cur.push_back(ch);
        }
        params.push_back(cur);
        return params;
    }

    struct Params {
        std::vector<std::string> v; size_t i = 0;
        bool next(std::string& out) { if (i < v.size()) { out = v[i++]; return true; } return false; }
    };

    // get_* (scene.rs:293-452): `$label` looks the typed map up, `(…)` parses inline.
    int resolve(const std::string& expr, Ent want, const char* what) {
        bool is_ref = !expr.empty() && expr[0] == '$';
        bool is_inline = expr.size() >= 2 && expr.front() == '(' && expr.back() == ')';
        if (is_ref) {
            std::string label = expr.substr(1);
            std::map<std::string, int>& m = want == Ent::Object ? objects : want == Ent::Material ? materials
                                          : want == Ent::TexColor ? color_tex : want == Ent::TexFloat ? float_tex : noise;
            auto it = m.find(label);
            if (it == m.end()) throw Fail{std::string("Invalid ") + what + " reference " + label};
            return it->second;
        }
        if (is_inline) {
            Entity e = parse_declaration(expr.substr(1, expr.size() - 2));
            if (e.type != want) throw Fail{std::string("Expression evaluates to a different entity type, expected ") + what};
            return e.index;
        }
        throw Fail{"Expected a reference or inline declaration"};
    }
    int get_color_texture(const std::string& e) { return resolve(e, Ent::TexColor, "color texture"); }
    int get_float_texture(const std::string& e) { return resolve(e, Ent::TexFloat, "float texture"); }
    int get_material(const std::string& e) { return resolve(e, Ent::Material, "material"); }
    int get_object(const std::string& e) { return resolve(e, Ent::Object, "object"); }
    int get_noise(const std::string& e) { return resolve(e, Ent::Noise, "noise gen"); }
    // get_texture (scene.rs:352-362): colour first, then float
    Entity get_texture(const std::string& e) {
        try { return {Ent::TexColor, get_color_texture(e)}; }
        catch (Fail&) { return {Ent::TexFloat, get_float_texture(e)}; }
    }

    static double opt_f64(Params& p, double dflt) {
        std::string s;
        if (!p.next(s)) return dflt;
        double v;
        if (!parse_f64(s, v)) throw Fail{"expected a number, got '" + s + "'"}; // unwrap() panic in the reference
        return v;
    }
    static void need_vec(const std::string& s, double out[3]) {
        int rc = parse_vec3(s, out);
        if (rc == 1) return;
        throw Fail{rc == 0 ? "Vector must have three components" : "Vector component must be a number"};
    }

    Entity parse_declaration(const std::string& decl) { // scene.rs:247-290
        Params p; p.v = parse_params(decl);
        std::string type;
        if (!p.next(type)) throw Fail{"Constant texture missing parameters"};
        std::string a, b, c, d;
        // ---- textures
        if (type == "constant") { // scene.rs:458-479
            if (!p.next(a)) throw Fail{"Constant texture missing parameters"};
            double v[3];
            int rc = parse_vec3(a, v);
            if (rc == 1) return {Ent::TexColor, st.add_texture_const_color(v[0], v[1], v[2])};
            if (rc < 0) throw Fail{"Vector component must be a number"};
            double k;
            if (!parse_f64(a, k)) throw Fail{"invalid float literal"};
            return {Ent::TexFloat, st.add_texture_const_float(k)};
        }
        if (type == "checker" || type == "checker_solid") { // scene.rs:481-515
            if (!p.next(a) || !p.next(b)) throw Fail{"Checkerboard texture missing parameters"};
            double scale = opt_f64(p, 1.0);
            int kind = type == "checker" ? RTGPU_TEX_CHECKER : RTGPU_TEX_CHECKER_SOLID;
            Entity t1 = get_texture(a);
            if (t1.type == Ent::TexColor) return {Ent::TexColor, st.add_texture(kind, false, t1.index, get_color_texture(b), -1, scale)};
            return {Ent::TexFloat, st.add_texture(kind, true, t1.index, get_float_texture(b), -1, scale)};
        }
        if (type == "lerp") { // scene.rs:517-540 (t is resolved first)
            if (!p.next(a) || !p.next(b) || !p.next(c)) throw Fail{"Interpolate texture missing parameters"};
            int t = get_float_texture(c);
            Entity t1 = get_texture(a);
            if (t1.type == Ent::TexColor) return {Ent::TexColor, st.add_texture(RTGPU_TEX_LERP, false, t1.index, get_color_texture(b), t, 1.0)};
            return {Ent::TexFloat, st.add_texture(RTGPU_TEX_LERP, true, t1.index, get_float_texture(b), t, 1.0)};
        }
        if (type == "noise") throw Fail{"Not implemented"}; // scene.rs:257
        if (type == "noise_solid") { // scene.rs:555-570
            if (!p.next(a)) throw Fail{"Noise texture missing parameters"};
            int n = get_noise(a);
            double scale = opt_f64(p, 1.0);
            int samples = (int)opt_f64(p, 7.0);
            int t = st.add_texture(RTGPU_TEX_NOISE_SOLID, true, -1, -1, -1, scale);
            st.textures[t].noise = n; st.textures[t].samples = samples;
            return {Ent::TexFloat, t};
        }
        if (type == "image") { // scene.rs:542-553
            if (!p.next(a)) throw Fail{"Image texture missing parameters"};
            std::string path = asset_path + a, err;
            st.images.emplace_back();
            int rc = decode_image(path, st.images.back(), err);
            if (rc != 0) { st.images.pop_back(); throw Fail{err}; }
            int t = st.add_texture(RTGPU_TEX_IMAGE, false, -1, -1, -1, 1.0);
            st.textures[t].image = (int)st.images.size() - 1;
            return {Ent::TexColor, t};
        }
        if (type == "channel") { // scene.rs:572-584
            if (!p.next(a) || !p.next(b)) throw Fail{"Channel texture missing parameters"};
            int col = get_color_texture(a);
            double ch;
            if (!parse_f64(b, ch) || ch < 0 || ch != (int)ch) throw Fail{"invalid digit found in string"};
            if (ch > 3) throw Fail{"channel index out of range"}; // would panic at render time
            int t = st.add_texture(RTGPU_TEX_CHANNEL, true, col, -1, -1, 1.0);
            st.textures[t].channel = (int)ch;
            return {Ent::TexFloat, t};
        }
        if (type == "uv_debug") return {Ent::TexColor, st.add_texture(RTGPU_TEX_UV_DEBUG, false, -1, -1, -1, 1.0)};
        // ---- materials (scene.rs:590-674)
        if (type == "lambertian") {
            if (!p.next(a)) throw Fail{"LambertianDiffuse material missing parameters"};
            return {Ent::Material, st.add_material(RTGPU_MAT_LAMBERTIAN, get_color_texture(a), -1, -1, 1.0)};
        }
        if (type == "metal") {
            if (!p.next(a) || !p.next(b)) throw Fail{"Metal material missing parameters"};
            int al = get_color_texture(a), ro = get_float_texture(b);
            return {Ent::Material, st.add_material(RTGPU_MAT_METAL, al, ro, -1, 1.0)};
        }
        if (type == "glass") return {Ent::Material, st.add_material(RTGPU_MAT_DIELECTRIC, -1, -1, -1, opt_f64(p, 1.5))};
        if (type == "glossy") {
            if (!p.next(a) || !p.next(b)) throw Fail{"Glossy material missing parameters"};
            int al = get_color_texture(a), ro = get_float_texture(b);
            double ior = opt_f64(p, 1.5);
            int nm = -1;
            if (p.next(c)) nm = get_color_texture(c);
            return {Ent::Material, st.add_material(RTGPU_MAT_GLOSSY, al, ro, nm, ior)};
        }
        if (type == "emissive") {
            if (!p.next(a)) throw Fail{"Emissive material missing parameters"};
            return {Ent::Material, st.add_material(RTGPU_MAT_EMISSIVE, get_color_texture(a), -1, -1, 1.0)};
        }
        if (type == "isotropic") {
            if (!p.next(a)) throw Fail{"Isotropic material missing parameters"};
            return {Ent::Material, st.add_material(RTGPU_MAT_ISOTROPIC, get_color_texture(a), -1, -1, 1.0)};
        }
        if (type == "normal_debug") {
            int nm = -1;
            if (p.next(a)) nm = get_color_texture(a);
            return {Ent::Material, st.add_material(RTGPU_MAT_NORMAL_DEBUG, -1, -1, nm, 1.0)};
        }
        // ---- objects (scene.rs:680-868)
        if (type == "sphere") {
            if (!p.next(a) || !p.next(b) || !p.next(c)) throw Fail{"Sphere missing parameters"};
            double o[3], r;
            need_vec(a, o);
            if (!parse_f64(b, r)) throw Fail{"invalid float literal"};
            return {Ent::Object, st.add_sphere(o, r, get_material(c))};
        }
        if (type == "plane") {
            if (!p.next(a) || !p.next(b) || !p.next(c) || !p.next(d)) throw Fail{"Plane missing parameters"};
            double o[3], u[3], v[3];
            need_vec(a, o); need_vec(b, u); need_vec(c, v);
            int mat = get_material(d);
            if (u[0] * v[0] + u[1] * v[1] + u[2] * v[2] != 0.0) throw Fail{"The UV vectors must be orthogonal!"}; // plane.rs:30-32 panic
            std::string extra;
            bool backface = p.next(extra) && extra == "backface";
            return {Ent::Object, st.add_plane(o, u, v, mat, backface)};
        }
        if (type == "box") {
            if (!p.next(a) || !p.next(b) || !p.next(c)) throw Fail{"Box missing parameters"};
            double o[3], s[3];
            need_vec(a, o); need_vec(b, s);
            return {Ent::Object, st.add_box(o, s, get_material(c))};
        }
        if (type == "mesh") {
            if (!p.next(a) || !p.next(b)) throw Fail{"Mesh missing parameters"};
            std::string path = asset_path + a, err;
            { std::ifstream probe(path); if (!probe.good()) throw Fail{"cannot open mesh file " + path}; } // File::open precedes get_material
            int mat = get_material(b);
            st.meshes.emplace_back();
            if (load_obj(path, st.meshes.back(), err) != 0) { st.meshes.pop_back(); throw Fail{err}; }
            RtgpuObject o = SceneTables::blank(RTGPU_OBJ_MESH);
            o.mesh = (int)st.meshes.size() - 1; o.material = mat;
            return {Ent::Object, st.add_object(o)};
        }
        if (type == "transform") { // scene.rs:755-798: `key=value` ops applied left to right
            if (!p.next(a)) throw Fail{"Transform missing parameters"};
            int child = get_object(a);
            TransformOps ops;
            std::string prm;
            while (p.next(prm)) {
                size_t eq = prm.find('=');
                if (eq == std::string::npos || eq == 0 || eq + 1 >= prm.size() || prm.find('=', eq + 1) != std::string::npos) continue;
                std::string key = prm.substr(0, eq), value = prm.substr(eq + 1);
                double v[3], s;
                if (key == "t") { need_vec(value, v); ops.translate(v[0], v[1], v[2]); }
                else if (key == "s") {
                    int rc = parse_vec3(value, v);
                    if (rc == 1) ops.scale(v[0], v[1], v[2]);
                    else if (rc == 0) { if (!parse_f64(value, s)) throw Fail{"invalid float literal"}; ops.scale(s, s, s); }
                    else throw Fail{"Vector component must be a number"};
                }
                else if (key == "rx") { if (!parse_f64(value, s)) throw Fail{"invalid float literal"}; ops.rotate_x(deg_to_rad(s)); }
                else if (key == "ry") { if (!parse_f64(value, s)) throw Fail{"invalid float literal"}; ops.rotate_y(deg_to_rad(s)); }
                else if (key == "rz") { if (!parse_f64(value, s)) throw Fail{"invalid float literal"}; ops.rotate_z(deg_to_rad(s)); }
            }
            return {Ent::Object, st.add_transform(child, ops.t, ops.inv)};
        }
        if (type == "list") {
            std::vector<int32_t> kids;
            while (p.next(a)) kids.push_back(get_object(a));
            return {Ent::Object, st.add_list(kids)};
        }
        if (type == "bvh") { // scene.rs:811-827
            if (!p.next(a)) throw Fail{"BVH missing parameters"};
            bool axes[3] = {a.find('x') != std::string::npos, a.find('y') != std::string::npos, a.find('z') != std::string::npos};
            std::vector<int32_t> kids;
            while (p.next(b)) kids.push_back(get_object(b));
            if (kids.empty()) throw Fail{"BVH needs at least one object"};          // the reference recurses forever
            if (!axes[0] && !axes[1] && !axes[2]) throw Fail{"BVH needs an axis"};   // the reference loops forever
            return {Ent::Object, st.build_bvh(kids, axes, rng)};
        }
        if (type == "sky") {
            if (!p.next(a)) throw Fail{"Sky missing parameters"};
            RtgpuObject o = SceneTables::blank(RTGPU_OBJ_SKY);
            o.texture = get_color_texture(a);
            return {Ent::Object, st.add_object(o)};
        }
        if (type == "sun") { // scene.rs:840-852: Vec4::point(x,y,z) then Sun::new -> to_unit() on all 4 components
            if (!p.next(a) || !p.next(b)) throw Fail{"Sun missing parameters"};
            double v[3];
            need_vec(a, v);
            RtgpuObject o = SceneTables::blank(RTGPU_OBJ_SUN);
            o.texture = get_color_texture(b);
            double len = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
            o.p0[0] = v[0] / len; o.p0[1] = v[1] / len; o.p0[2] = v[2] / len; o.p0[3] = 1.0 / len;
            return {Ent::Object, st.add_object(o)};
        }
        if (type == "volume") {
            if (!p.next(a) || !p.next(b) || !p.next(c)) throw Fail{"Volume missing parameters"};
            int boundary = get_object(a), mat = get_material(b);
            double density;
            if (!parse_f64(c, density)) throw Fail{"invalid float literal"};
            RtgpuObject o = SceneTables::blank(RTGPU_OBJ_VOLUME);
            o.material = mat; o.p1[0] = density;
            return {Ent::Object, st.add_object(o, {boundary})};
        }
        if (type == "perlin") { // PerlinNoise3D::new (noise/perlin.rs:18-52)
            RtgpuNoise n;
            for (int i = 0; i < 256; i++) { // Vec4::random_unit: normalised 3-D Gaussian (vec4.rs:42-48)
                double x = rng.normal(), y = rng.normal(), z = rng.normal();
                double len = std::sqrt(x * x + y * y + z * z);
                n.gradients[i][0] = x / len; n.gradients[i][1] = y / len; n.gradients[i][2] = z / len;
            }
            int32_t* perms[3] = {n.perm_x, n.perm_y, n.perm_z};
            for (int k = 0; k < 3; k++) {
                for (int i = 0; i < 256; i++) perms[k][i] = i;
                for (int i = 255; i >= 1; i--) { // perlin.rs:47-52: gen_range(0..=i)
                    int target = (int)rng.below((uint64_t)i + 1);
                    std::swap(perms[k][i], perms[k][target]);
                }
            }
            st.noises.push_back(n);
            return {Ent::Noise, (int)st.noises.size() - 1};
        }
        throw Fail{"Unknown object type"};
    }

    void parse_config_directive(const std::string& content) { // scene.rs:158-212
        size_t eq = content.find('=');
        if (eq == std::string::npos) throw Fail{"@config " + content};
        std::string key = trim(content.substr(0, eq)), value = trim(content.substr(eq + 1));
        double v[3], d;
        auto num = [&](const std::string& s) { if (!parse_f64(s, d)) throw Fail{"invalid float literal"}; return d; };
        if (key == "output_width") {
            double w = num(value);
            if (w < 0 || w != (long long)w) throw Fail{"invalid digit found in string"};
            scene_config.has_output_width = 1; scene_config.output_width = (int)w;
        } else if (key == "aspect_ratio") {
            double r;
            size_t slash = value.find('/');
            if (slash != std::string::npos) { double a = num(trim(value.substr(0, slash))); double b = num(trim(value.substr(slash + 1))); r = a / b; }
            else r = num(value);
            scene_config.has_aspect_ratio = 1; scene_config.aspect_ratio = r;
        } else if (key == "focal_length") { scene_config.has_focal_length = 1; scene_config.focal_length = num(value); }
        else if (key == "f_number") { scene_config.has_f_number = 1; scene_config.f_number = num(value); }
        else if (key == "focus_distance") { scene_config.has_focus_distance = 1; scene_config.focus_distance = num(value); }
        else if (key == "camera_pos") { need_vec(value, v); scene_config.has_camera_pos = 1; for (int i = 0; i < 3; i++) scene_config.camera_pos[i] = v[i]; }
        else if (key == "camera_target") { need_vec(value, v); scene_config.has_camera_target = 1; for (int i = 0; i < 3; i++) scene_config.camera_target[i] = v[i]; }
    }
};

} // namespace

int load_dsl_scene(const std::string& path, RtgpuConfig& cfg, SceneTables& st, Pcg64Mcg& rng, std::string& err) {
    std::ifstream in(path);
    if (!in.good()) { err = "cannot open scene file " + path; return RTGPU_ERR_IO; }
    // main.rs:46-56: assets are relative to the scene file's directory
    std::string asset_path;
    size_t slash = path.rfind('/');
    if (slash != std::string::npos) asset_path = path.substr(0, slash + 1);
    Loader L(st, rng, asset_path);
    std::string line;
    size_t line_number = 0;
    for (; std::getline(in, line); line_number++) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        if (line.empty() || line[0] == '#') continue;
        if (line[0] == '@') {
            size_t sp = line.find(' ');
            if (sp != std::string::npos && line.substr(1, sp - 1) == "config") {
                try { L.parse_config_directive(line.substr(sp + 1)); }
                catch (Fail& f) { st.warnings += "Warning: invalid @config directive\n\t" + f.msg + "\n"; }
            }
            continue;
        }
        // scene.rs:104-105: split on every ':' and use the first two parts
        size_t c1 = line.find(':');
        if (c1 == std::string::npos) { st.warnings += "Warning: parse failed on line " + std::to_string(line_number) + ", skipped\n\t" + line + "\n"; continue; }
        size_t c2 = line.find(':', c1 + 1);
        std::string label = trim(line.substr(0, c1));
        std::string decl = trim(line.substr(c1 + 1, c2 == std::string::npos ? std::string::npos : c2 - c1 - 1));
        try {
            Entity e = L.parse_declaration(decl);
            switch (e.type) {
            case Ent::Object: L.objects[label] = e.index; break;
            case Ent::Material: L.materials[label] = e.index; break;
            case Ent::TexColor: L.color_tex[label] = e.index; break;
            case Ent::TexFloat: L.float_tex[label] = e.index; break;
            case Ent::Noise: L.noise[label] = e.index; break;
            }
        } catch (Fail& f) {
            st.warnings += "Warning: error on line " + std::to_string(line_number) + ", skipped\n\t" + f.msg + "\n";
        }
    }
    auto w = L.objects.find("world"), l = L.objects.find("lights");
    if (w == L.objects.end() || l == L.objects.end()) { err = "ParseError: No world/lights object"; return RTGPU_ERR_PARSE; }
    st.world_root = w->second; st.lights_root = l->second;
    RtgpuSceneConfig merged;
    merge_scene_config(L.scene_config, cfg.scene, merged); // scene.rs:144: CLI overrides the file
    cfg.scene = merged;
    return 0;
}

} // namespace rthost
