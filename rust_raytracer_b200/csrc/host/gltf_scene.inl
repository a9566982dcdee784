// gltf_scene.inl — second half of gltf_loader.cpp: the glTF document -> SceneTables, following
// AssimpLoader::load / load_node / load_mesh / load_material (src/loaders/assimp.rs:37-189).
namespace rthost {
namespace {

// Mat4::inverse (mat4.rs:148-275): adjugate / determinant; None when the determinant is 0
bool m4_inverse(const M4& a, M4& out) {
    auto minor3 = [&](int r, int c) { // determinant of the 3x3 matrix left after deleting row r, column c
        int rr[3], cc[3];
        for (int i = 0, k = 0; i < 4; i++) if (i != r) rr[k++] = i;
        for (int i = 0, k = 0; i < 4; i++) if (i != c) cc[k++] = i;
        auto e = [&](int i, int j) { return a.m[rr[i] * 4 + cc[j]]; };
        return e(0, 0) * (e(1, 1) * e(2, 2) - e(1, 2) * e(2, 1)) - e(0, 1) * (e(1, 0) * e(2, 2) - e(1, 2) * e(2, 0)) +
               e(0, 2) * (e(1, 0) * e(2, 1) - e(1, 1) * e(2, 0));
    };
    double cof[16];
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++) cof[r * 4 + c] = ((r + c) & 1 ? -1.0 : 1.0) * minor3(r, c);
    const double det = a.m[0] * cof[0] + a.m[1] * cof[1] + a.m[2] * cof[2] + a.m[3] * cof[3];
    if (det == 0.0) return false;
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++) out.m[r * 4 + c] = cof[c * 4 + r] / det; // adjugate = transposed cofactors
    return true;
}

struct GltfBuilder {
    const Glb& g;
    SceneTables& st;
    std::string err;
    std::vector<int32_t> lights;
    std::map<int, std::pair<int, bool>> material_cache; // glTF material index -> (material id, is_emissive)

    bool fail(const std::string& m) { if (err.empty()) err = "model: " + m; return false; }
    const Json* table(const char* name) const { const Json* t = g.json.get(name); return t && t->kind == Json::Arr ? t : nullptr; }
    static double f32(double x) { return (double)(float)x; } // Assimp keeps every number as a float (ai_real)

    // node.transformation as Assimp builds it (aiMatrix4x4 of floats): `matrix` (column-major in the file), or T * R * S
    M4 node_matrix(const Json& n) const {
        M4 out = m4_identity();
        const Json* mj = n.get("matrix");
        if (mj && mj->arr.size() == 16) {
            for (int r = 0; r < 4; r++)
                for (int c = 0; c < 4; c++) out.m[r * 4 + c] = f32(mj->arr[(size_t)(c * 4 + r)].num);
            return out;
        }
        float t[3] = {0, 0, 0}, q[4] = {0, 0, 0, 1}, s[3] = {1, 1, 1};
        if (const Json* j = n.get("translation")) for (size_t i = 0; i < 3 && i < j->arr.size(); i++) t[i] = (float)j->arr[i].num;
        if (const Json* j = n.get("rotation")) for (size_t i = 0; i < 4 && i < j->arr.size(); i++) q[i] = (float)j->arr[i].num;
        if (const Json* j = n.get("scale")) for (size_t i = 0; i < 3 && i < j->arr.size(); i++) s[i] = (float)j->arr[i].num;
        const float x = q[0], y = q[1], z = q[2], w = q[3];
        const float R[9] = {1.f - 2.f * (y * y + z * z), 2.f * (x * y - z * w), 2.f * (x * z + y * w),
                            2.f * (x * y + z * w), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - x * w),
                            2.f * (x * z - y * w), 2.f * (y * z + x * w), 1.f - 2.f * (x * x + y * y)};
        for (int r = 0; r < 3; r++) {
            for (int c = 0; c < 3; c++) out.m[r * 4 + c] = (double)(R[r * 3 + c] * s[c]);
            out.m[r * 4 + 3] = (double)t[r];
        }
        return out;
    }

    // load_material (assimp.rs:143-189)
    bool load_material(int index, int& mat_out, bool& emissive_out) {
        auto it = material_cache.find(index);
        if (it != material_cache.end()) { mat_out = it->second.first; emissive_out = it->second.second; return true; }
        static const Json empty;
        const Json* mats = table("materials");
        const Json& m = (mats && index >= 0 && index < (int)mats->arr.size()) ? mats->arr[(size_t)index] : empty; // no material: the importer's default
        double em[3] = {0, 0, 0};
        if (const Json* e = m.get("emissiveFactor")) for (size_t i = 0; i < 3 && i < e->arr.size(); i++) em[i] = f32(e->arr[i].num);
        if (em[0] * em[0] + em[1] * em[1] + em[2] * em[2] > 0.0) { // assimp.rs:145-151
            double intensity = 1.0;
            if (const Json* ext = m.get("extensions"))
                if (const Json* k = ext->get("KHR_materials_emissive_strength")) intensity = f32(k->number("emissiveStrength", 1.0));
            const int tex = st.add_texture_const_color(em[0] * intensity, em[1] * intensity, em[2] * intensity);
            mat_out = st.add_material(RTGPU_MAT_EMISSIVE, tex, -1, -1, 1.0);
            emissive_out = true;
        } else { // assimp.rs:153-188: Glossy(base colour, roughness, ior 1.5)
            double base[3] = {1, 1, 1}, rough = 1.0; // the glTF defaults the importer always writes
            if (const Json* pbr = m.get("pbrMetallicRoughness")) {
                if (pbr->get("baseColorTexture") || pbr->get("metallicRoughnessTexture"))
                    return fail("textured materials of a .glb are embedded images; the reference reads `<dir>/<name>.<ext>` files "
                                "(assimp.rs:160-182) and panics on them: not supported");
                if (const Json* c = pbr->get("baseColorFactor")) for (size_t i = 0; i < 3 && i < c->arr.size(); i++) base[i] = f32(c->arr[i].num);
                rough = f32(pbr->number("roughnessFactor", 1.0));
            }
            const int albedo = st.add_texture_const_color(base[0], base[1], base[2]);
            const int roughness = st.add_texture_const_float(rough);
            mat_out = st.add_material(RTGPU_MAT_GLOSSY, albedo, roughness, -1, 1.5);
            emissive_out = false;
        }
        material_cache[index] = {mat_out, emissive_out};
        return true;
    }

    // load_mesh (assimp.rs:101-141) for one primitive (= one aiMesh)
    bool load_primitive(const Json& prim, const double pos[3], int& obj_out) {
        if (prim.integer("mode", 4) != 4) return fail("only triangle primitives (mode 4) are supported");
        const Json* attr = prim.get("attributes");
        if (!attr) return fail("primitive without attributes");
        const int a_pos = attr->integer("POSITION", -1), a_nrm = attr->integer("NORMAL", -1), a_uv = attr->integer("TEXCOORD_0", -1);
        if (a_pos < 0) return fail("primitive without POSITION");
        if (a_nrm < 0) return fail("primitive without NORMAL (the reference asks Assimp for no generated normals and would index an empty array)");
        MeshData md;
        size_t nv = 0, nn = 0, nt = 0;
        if (!read_accessor(g, a_pos, 3, md.positions, nv, err) || !read_accessor(g, a_nrm, 3, md.normals, nn, err)) return false;
        if (nn != nv) return fail("NORMAL and POSITION counts differ");
        if (a_uv >= 0) {
            if (!read_accessor(g, a_uv, 2, md.uvs, nt, err)) return false;
            if (nt != nv) return fail("TEXCOORD_0 and POSITION counts differ");
            for (size_t i = 0; i < nt; i++) md.uvs[i * 2 + 1] = (double)(1.0f - (float)md.uvs[i * 2 + 1]); // the importer flips v
        }
        std::vector<double> idx;
        size_t ni = 0;
        if (prim.get("indices")) {
            if (!read_accessor(g, prim.integer("indices", -1), 1, idx, ni, err)) return false;
        } else { ni = nv; idx.resize(nv); for (size_t i = 0; i < nv; i++) idx[i] = (double)i; }
        if (ni % 3) return fail("index count is not a multiple of 3");
        md.tri_pos.resize(ni); md.tri_nrm.resize(ni); md.tri_uv.resize(ni);
        for (size_t i = 0; i < ni; i++) {
            if (idx[i] < 0 || idx[i] >= (double)nv) return fail("vertex index out of range");
            md.tri_pos[i] = md.tri_nrm[i] = (int32_t)idx[i];             // assimp.rs:118-128: one index triple for all three
            md.tri_uv[i] = a_uv >= 0 ? (int32_t)idx[i] : -1;
        }
        int mat; bool emissive;
        if (!load_material(prim.integer("material", -1), mat, emissive)) return false;
        if (emissive) { // assimp.rs:123-129: invisible sphere for light sampling only
            double radius = INFINITY;
            for (size_t i = 0; i < nv; i++) {
                const double* v = &md.positions[i * 3];
                radius = std::fmin(std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]), radius);
            }
            lights.push_back(st.add_sphere(pos, radius, mat));
        }
        st.meshes.push_back(std::move(md));
        RtgpuObject o = SceneTables::blank(RTGPU_OBJ_MESH);
        o.mesh = (int32_t)st.meshes.size() - 1; o.material = mat;
        obj_out = st.add_object(o);
        return true;
    }

    // load_node (assimp.rs:77-99)
    bool load_node(int index, const double pos_in[3], int depth, int& obj_out) {
        const Json* nodes = table("nodes");
        if (!nodes || index < 0 || index >= (int)nodes->arr.size()) return fail("node index out of range");
        if (depth > 48) return fail("node hierarchy too deep (cycle?)");
        const Json& n = nodes->arr[(size_t)index];
        const M4 t = node_matrix(n);
        const double pos[3] = {pos_in[0] + t.m[3], pos_in[1] + t.m[7], pos_in[2] + t.m[11]}; // pos + column(3), w zeroed
        std::vector<int32_t> kids;
        const int mesh = n.integer("mesh", -1);
        if (mesh >= 0) {
            const Json* meshes = table("meshes");
            if (!meshes || mesh >= (int)meshes->arr.size()) return fail("mesh index out of range");
            const Json* prims = meshes->arr[(size_t)mesh].get("primitives");
            if (prims) for (const Json& p : prims->arr) { int o; if (!load_primitive(p, pos, o)) return false; kids.push_back(o); }
        }
        if (const Json* ch = n.get("children"))
            for (const Json& c : ch->arr) { int o; if (!load_node((int)c.num, pos, depth + 1, o)) return false; kids.push_back(o); }
        M4 inv;
        if (!m4_inverse(t, inv)) return fail("singular node transform (Transform::from_matrix(...).unwrap() panics in the reference)");
        obj_out = st.add_transform(st.add_list(kids), t, inv);
        return true;
    }
};

} // namespace

// main.rs:38-41: `model:<path>` -> AssimpLoader::new(path)?.load(config)
int load_glb_scene(const std::string& path, RtgpuConfig& cfg, SceneTables& st, std::string& err) {
    Glb g;
    int rc = read_glb(path, g, err);
    if (rc) return rc;
    GltfBuilder b{g, st, {}, {}, {}};
    // assimp.rs:39-52: DEFAULT_SCENE_CONFIG <- first camera <- command line
    RtgpuSceneConfig sc = default_scene_config();
    if (const Json* cams = b.table("cameras"))
        if (!cams->arr.empty())
            if (const Json* p = cams->arr[0].get("perspective")) {
                const float yfov = (float)p->number("yfov", 0.0), aspect = (float)p->number("aspectRatio", 0.0);
                const float hfov = 2.0f * std::atan(std::tan(yfov * 0.5f) * (aspect == 0.f ? 1.f : aspect)); // the importer's mHorizontalFOV
                sc.has_camera_pos = 1; sc.camera_pos[0] = sc.camera_pos[1] = sc.camera_pos[2] = 0.0;           // aiCamera.mPosition: node-local
                sc.has_camera_target = 1; sc.camera_target[0] = sc.camera_target[1] = 0.0; sc.camera_target[2] = -1.0; // aiCamera.mLookAt
                sc.has_aspect_ratio = 1; sc.aspect_ratio = (double)aspect;
                sc.has_focal_length = 1; sc.focal_length = 18.0 / std::tan((double)hfov / 2.0);
            }
    RtgpuSceneConfig merged;
    merge_scene_config(sc, cfg.scene, merged);
    cfg.scene = merged;
    // the importer's root: the scene's only root node, or an identity "ROOT" node over several
    const Json* scenes = b.table("scenes");
    const int scene = g.json.integer("scene", 0);
    if (!scenes || scene < 0 || scene >= (int)scenes->arr.size()) { err = "model: Assimp load fail (no scene)"; return RTGPU_ERR_PARSE; }
    const Json* roots = scenes->arr[(size_t)scene].get("nodes");
    if (!roots || roots->arr.empty()) { err = "model: Assimp load fail (no root node)"; return RTGPU_ERR_PARSE; } // assimp.rs:72-74
    const double origin[3] = {0, 0, 0};
    int world = -1;
    bool ok = true;
    if (roots->arr.size() == 1) ok = b.load_node((int)roots->arr[0].num, origin, 0, world);
    else {
        std::vector<int32_t> kids;
        for (const Json& r : roots->arr) { int o; if (!(ok = b.load_node((int)r.num, origin, 1, o))) break; kids.push_back(o); }
        if (ok) world = st.add_transform(st.add_list(kids), m4_identity(), m4_identity());
    }
    if (!ok) { err = b.err; return RTGPU_ERR_UNSUPPORTED; }
    st.world_root = world;
    st.lights_root = st.add_list(b.lights); // ObjectList::from(lights), assimp.rs:68
    return 0;
}

} // namespace rthost
