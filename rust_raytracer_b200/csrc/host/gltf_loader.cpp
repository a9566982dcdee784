// gltf_loader.cpp — `model:<path>.glb` scenes: the reference's AssimpLoader
// (src/loaders/assimp.rs:29-203) restated for binary glTF 2.0 only, so that no Assimp is needed.
//
// What AssimpLoader does with what Assimp 5.x's glTF2 importer hands it (russimp 3.2.0 / russimp-sys 2.0.2,
// Cargo.lock:1451-1467 — third-party code that is NOT in /root/reference: its behaviour is restated from its
// published importer, PARITY UNPINNED):
//   * every node becomes Transform::from_matrix(ObjectList[meshes of the node..., child nodes...], node matrix)
//     (assimp.rs:77-99); several scene roots hang under an identity "ROOT" node (importer), one root is the root;
//   * every mesh PRIMITIVE is one aiMesh -> one TriangleMesh with shared position / normal / uv indices
//     (assimp.rs:101-141); the importer flips v (v' = 1 - v); PostProcess::Triangulate only (assimp.rs:31-33);
//   * material (assimp.rs:143-189): emissiveFactor x KHR_materials_emissive_strength ($clr.emissive,
//     $mat.emissiveIntensity) with non-zero length -> Emissive(constant); anything else -> Glossy(baseColorFactor
//     ($clr.base, default 0.5,0.5,0.5), roughnessFactor ($mat.roughnessFactor, default 0), ior 1.5);
//     baseColor / roughness TEXTURES are read from `<dir>/<name>.<hint>` (assimp.rs:160-182), which for the embedded
//     images of a .glb does not exist (the reference panics on the unwrap): refused here with a message;
//   * an emissive mesh adds an invisible light-sampling sphere to `lights` (assimp.rs:123-129): centre = the SUM OF THE
//     TRANSLATION COLUMNS of the node chain (rotations / scales of ancestors ignored, assimp.rs:81-83,86,91), radius =
//     the smallest vertex distance from the mesh origin, the emissive material;
//   * camera (assimp.rs:41-51): aiCamera.position / look_at are NODE-LOCAL and the reference does not apply the camera
//     node's transform: for glTF that is (0,0,0) looking at (0,0,-1); aspect = aspectRatio, focal_length =
//     18 / tan(hfov / 2) with hfov = 2 atan(tan(yfov / 2) aspect) (importer). `-c=` / `-l=` override as usual.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <map>
#include <memory>

#include "host_scene.h"

namespace rthost {
namespace {

// ---- a minimal JSON reader (objects, arrays, numbers, strings, true / false / null) --------------------
struct Json {
    enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
    double num = 0;
    bool b = false;
    std::string str;
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;
    const Json* get(const std::string& k) const {
        for (auto& kv : obj) if (kv.first == k) return &kv.second;
        return nullptr;
    }
    double number(const std::string& k, double dflt) const { const Json* j = get(k); return j && j->kind == Num ? j->num : dflt; }
    int integer(const std::string& k, int dflt) const { return (int)number(k, (double)dflt); }
};

struct JsonParser {
    const char* p; const char* end; std::string err;
    void ws() { while (p < end && (*p == ' ' || *p == '\n' || *p == '\r' || *p == '\t')) p++; }
    bool fail(const char* m) { if (err.empty()) err = m; return false; }
    bool parse_string(std::string& out) {
        if (p >= end || *p != '"') return fail("JSON: expected a string");
        p++;
        while (p < end && *p != '"') {
            if (*p == '\\' && p + 1 < end) {
                p++;
                switch (*p) {
                case 'n': out += '\n'; break; case 't': out += '\t'; break; case 'r': out += '\r'; break;
                case 'b': out += '\b'; break; case 'f': out += '\f'; break;
                case 'u': out += '?'; p += std::min<ptrdiff_t>(4, end - p - 1); break; // (names only: code points are not needed)
                default: out += *p;
                }
                p++;
            } else out += *p++;
        }
        if (p >= end) return fail("JSON: unterminated string");
        p++;
        return true;
    }
    bool parse(Json& out, int depth = 0) {
        if (depth > 64) return fail("JSON: nested too deep");
        ws();
        if (p >= end) return fail("JSON: unexpected end");
        if (*p == '{') {
            out.kind = Json::Obj; p++; ws();
            if (p < end && *p == '}') { p++; return true; }
            for (;;) {
                ws();
                std::string k;
                if (!parse_string(k)) return false;
                ws();
                if (p >= end || *p != ':') return fail("JSON: expected ':'");
                p++;
                Json v;
                if (!parse(v, depth + 1)) return false;
                out.obj.emplace_back(std::move(k), std::move(v));
                ws();
                if (p < end && *p == ',') { p++; continue; }
                if (p < end && *p == '}') { p++; return true; }
                return fail("JSON: expected ',' or '}'");
            }
        }
        if (*p == '[') {
            out.kind = Json::Arr; p++; ws();
            if (p < end && *p == ']') { p++; return true; }
            for (;;) {
                Json v;
                if (!parse(v, depth + 1)) return false;
                out.arr.push_back(std::move(v));
                ws();
                if (p < end && *p == ',') { p++; continue; }
                if (p < end && *p == ']') { p++; return true; }
                return fail("JSON: expected ',' or ']'");
            }
        }
        if (*p == '"') { out.kind = Json::Str; return parse_string(out.str); }
        if (end - p >= 4 && !std::strncmp(p, "true", 4)) { out.kind = Json::Bool; out.b = true; p += 4; return true; }
        if (end - p >= 5 && !std::strncmp(p, "false", 5)) { out.kind = Json::Bool; out.b = false; p += 5; return true; }
        if (end - p >= 4 && !std::strncmp(p, "null", 4)) { out.kind = Json::Null; p += 4; return true; }
        char* e = nullptr;
        std::string tok(p, std::min<size_t>(64, (size_t)(end - p)));
        out.num = std::strtod(tok.c_str(), &e);
        if (e == tok.c_str()) return fail("JSON: unexpected character");
        out.kind = Json::Num;
        p += e - tok.c_str();
        return true;
    }
};

// ---- GLB container + accessors -------------------------------------------------------------------------
struct Glb {
    Json json;
    std::vector<uint8_t> bin;
};

uint32_t rd32(const uint8_t* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }

int read_glb(const std::string& path, Glb& g, std::string& err) {
    std::ifstream in(path, std::ios::binary);
    if (!in.good()) { err = "cannot open model file " + path; return RTGPU_ERR_IO; }
    std::vector<uint8_t> b((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
    if (b.size() < 20 || std::memcmp(b.data(), "glTF", 4) != 0) { err = "model: only binary glTF 2.0 (.glb) is supported (the reference's other formats need Assimp)"; return RTGPU_ERR_UNSUPPORTED; }
    if (rd32(&b[4]) != 2) { err = "model: glTF container version must be 2"; return RTGPU_ERR_UNSUPPORTED; }
    size_t off = 12;
    bool have_json = false;
    while (off + 8 <= b.size()) {
        const uint32_t len = rd32(&b[off]), type = rd32(&b[off + 4]);
        if (off + 8 + (size_t)len > b.size()) { err = "model: truncated GLB chunk"; return RTGPU_ERR_PARSE; }
        if (type == 0x4E4F534Au && !have_json) { // "JSON"
            JsonParser jp{(const char*)&b[off + 8], (const char*)&b[off + 8] + len, {}};
            if (!jp.parse(g.json)) { err = "model: " + jp.err; return RTGPU_ERR_PARSE; }
            have_json = true;
        } else if (type == 0x004E4942u && g.bin.empty()) g.bin.assign(b.begin() + (ptrdiff_t)off + 8, b.begin() + (ptrdiff_t)(off + 8 + len)); // "BIN\0"
        off += 8 + (size_t)len;
    }
    if (!have_json || g.json.kind != Json::Obj) { err = "model: GLB without a JSON chunk"; return RTGPU_ERR_PARSE; }
    return 0;
}

// accessor `index` as doubles, `want` components per element (VEC3 -> 3 ...); float / normalised integer data
bool read_accessor(const Glb& g, int index, int want, std::vector<double>& out, size_t& count, std::string& err) {
    const Json* accs = g.json.get("accessors"); const Json* views = g.json.get("bufferViews");
    if (!accs || index < 0 || index >= (int)accs->arr.size()) { err = "model: accessor index out of range"; return false; }
    const Json& a = accs->arr[(size_t)index];
    if (a.get("sparse")) { err = "model: sparse accessors are not supported"; return false; }
    const Json* ty = a.get("type");
    const std::string type = ty ? ty->str : "";
    const int comps = type == "SCALAR" ? 1 : type == "VEC2" ? 2 : type == "VEC3" ? 3 : type == "VEC4" ? 4 : 0;
    if (comps != want) { err = "model: accessor has the wrong element type (" + type + ")"; return false; }
    const int ct = a.integer("componentType", 0);
    const size_t csize = ct == 5126 || ct == 5125 ? 4 : ct == 5123 || ct == 5122 ? 2 : ct == 5121 || ct == 5120 ? 1 : 0;
    if (!csize) { err = "model: unknown accessor component type"; return false; }
    count = (size_t)a.number("count", 0);
    const int bv = a.integer("bufferView", -1);
    if (!views || bv < 0 || bv >= (int)views->arr.size()) { err = "model: accessor without a buffer view"; return false; }
    const Json& v = views->arr[(size_t)bv];
    if (v.integer("buffer", 0) != 0) { err = "model: only the GLB-embedded buffer 0 is supported"; return false; }
    const size_t base = (size_t)v.number("byteOffset", 0) + (size_t)a.number("byteOffset", 0);
    size_t stride = (size_t)v.number("byteStride", 0);
    if (!stride) stride = csize * (size_t)comps;
    if (count && base + (count - 1) * stride + csize * (size_t)comps > g.bin.size()) { err = "model: accessor runs past the binary chunk"; return false; }
    const bool normalized = a.get("normalized") && a.get("normalized")->b;
    out.resize(count * (size_t)comps);
    for (size_t i = 0; i < count; i++)
        for (int c = 0; c < comps; c++) {
            const uint8_t* p = &g.bin[base + i * stride + (size_t)c * csize];
            double x;
            switch (ct) {
            case 5126: { float f; std::memcpy(&f, p, 4); x = (double)f; break; }     // `v.x as f64` (assimp.rs:103)
            case 5125: x = (double)rd32(p); break;
            case 5123: x = (double)((uint32_t)p[0] | ((uint32_t)p[1] << 8)); if (normalized) x /= 65535.0; break;
            case 5122: { int16_t s; std::memcpy(&s, p, 2); x = (double)s; if (normalized) x = std::fmax(x / 32767.0, -1.0); break; }
            case 5121: x = (double)p[0]; if (normalized) x /= 255.0; break;
            default: { int8_t s; std::memcpy(&s, p, 1); x = (double)s; if (normalized) x = std::fmax(x / 127.0, -1.0); break; }
            }
            out[i * (size_t)comps + (size_t)c] = x;
        }
    return true;
}

} // namespace
} // namespace rthost

#include "gltf_scene.inl"
