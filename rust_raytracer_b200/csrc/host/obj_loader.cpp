// obj_loader.cpp — load_mesh_from_file (reference src/loaders/obj.rs:13-107), restated.
// Only `v`, `vt`, `vn`, `f` are read; lines are split on single spaces; faces must be
// triangles `v/vt/vn` or `v//vn` (the normal index is mandatory); indices may be negative;
// `vn` is normalised on load. Also hosts the proxy-mesh generator for the missing dragon.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

#include "host_scene.h"
#include <cstdlib>
#include <thread>

// host cores this process may count on: under torchrun (one process per GPU) every rank runs the same set-up at the same
// time, so the cores are divided by LOCAL_WORLD_SIZE (8 ranks x 16 threads on 32 cores made the "Ready" phase slower at N = 8)
static inline unsigned rt_host_cores() {
    unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    if (const char* e = std::getenv("LOCAL_WORLD_SIZE")) { const int w = std::atoi(e); if (w > 1) hw = std::max(2u, hw / (unsigned)w); }
    return hw;
}


namespace rthost {

static inline const char* skip_token(const char* p) { while (*p && *p != ' ') p++; return p; }

namespace {

// One contiguous run of lines, parsed independently of the others. Faces keep their raw 1-based /
// negative indices plus the number of v / vt / vn records seen in THIS chunk before the face; the
// merge step adds the counts of the preceding chunks to resolve negative (relative) indices.
struct ObjChunk {
    std::vector<double> positions, uvs, normals;
    struct Corner { long v, t, n; bool has_t; };
    struct Face { Corner c[3]; size_t nv, nt, nn; };
    std::vector<Face> faces;
    size_t lines = 0;
    size_t err_line = 0; // 1-based line inside the chunk of the first error, 0 = none
    std::string err;
};

void parse_obj_chunk(char* begin, char* end, ObjChunk& out) {
    char* pos = begin;
    while (pos < end) {
        char* eol = (char*)std::memchr(pos, '\n', (size_t)(end - pos));
        if (!eol) eol = end;
        char* lend = eol;
        if (lend > pos && lend[-1] == '\r') lend--; // BufRead::lines strips "\r\n"
        const char saved = *lend;                    // (the buffer has one spare byte past the file)
        *lend = 0;
        const char* line = pos;
        out.lines++;
        const char* p = skip_token(line);
        const size_t cmd_len = (size_t)(p - line);
        auto parse_floats = [&](double* vals, int maxn) -> int { // params.map(parse::<f64>().unwrap())
            int n = 0;
            const char* q = p;
            while (*q == ' ') {
                q++;
                const char* tok_end = skip_token(q);
                char* num_end = nullptr;
                double v = std::strtod(q, &num_end);
                if (num_end != tok_end || tok_end == q) return -1; // unwrap() panic
                if (n < maxn) vals[n] = v;
                n++;
                q = tok_end;
            }
            return n;
        };
        auto bad = [&](const char* what) { if (!out.err_line) { out.err_line = out.lines; out.err = what; } };
        if (cmd_len == 1 && line[0] == 'v') {
            double v[8]; int n = parse_floats(v, 8);
            if (n < 3) bad("vertex needs 3 numbers");
            else out.positions.insert(out.positions.end(), {v[0], v[1], v[2]});
        } else if (cmd_len == 2 && line[0] == 'v' && line[1] == 't') {
            double v[8]; int n = parse_floats(v, 8);
            if (n < 2) bad("vt needs 2 numbers");
            else out.uvs.insert(out.uvs.end(), {v[0], v[1]});
        } else if (cmd_len == 2 && line[0] == 'v' && line[1] == 'n') {
            double v[8]; int n = parse_floats(v, 8);
            if (n < 3) bad("vn needs 3 numbers");
            else {
                double len = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
                out.normals.insert(out.normals.end(), {v[0] / len, v[1] / len, v[2] / len}); // obj.rs:48 to_unit
            }
        } else if (cmd_len == 1 && line[0] == 'f') {
            ObjChunk::Face face{};
            int corner = 0;
            bool ok = true;
            const char* q = p;
            while (ok && *q == ' ') {
                q++;
                const char* tok_end = skip_token(q);
                if (corner >= 3) { bad("face is not a triangle"); ok = false; break; } // index out of bounds panic
                // split('/') -> up to three parts, "" = None
                long part[3] = {0, 0, 0}; bool present[3] = {false, false, false};
                int np = 0;
                const char* s = q;
                while (np < 3) {
                    const char* e = s;
                    while (e < tok_end && *e != '/') e++;
                    if (e > s) {
                        char* num_end = nullptr;
                        part[np] = std::strtol(s, &num_end, 10);
                        if (num_end != e) { bad("bad face index"); ok = false; break; }
                        present[np] = true;
                    }
                    np++;
                    if (e >= tok_end) break;
                    s = e + 1;
                }
                if (!ok) break;
                if (np < 3 || !present[0] || !present[2]) { bad("face corner must be v/vt/vn or v//vn"); ok = false; break; }
                face.c[corner] = {part[0], part[1], part[2], present[1]};
                corner++;
                q = tok_end;
            }
            if (ok) {
                face.nv = out.positions.size() / 3; face.nt = out.uvs.size() / 2; face.nn = out.normals.size() / 3;
                out.faces.push_back(face);
            }
        }
        *lend = saved;
        pos = eol + 1;
    }
}

} // namespace

int load_obj(const std::string& path, MeshData& out, std::string& err) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) { err = "cannot open mesh file " + path; return RTGPU_ERR_IO; }
    std::fseek(f, 0, SEEK_END);
    long size = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    std::vector<char> text((size_t)size + 1, '\0');
    if (size > 0 && std::fread(text.data(), 1, (size_t)size, f) != (size_t)size) { std::fclose(f); err = "short read on " + path; return RTGPU_ERR_IO; }
    std::fclose(f);

    // split at line boundaries; chunks are parsed in parallel and merged in file order
    size_t n_chunks = size < (1 << 20) ? 1 : std::max(1u, std::min(32u, rt_host_cores()));
    if (const char* e = std::getenv("RTGPU_OBJ_THREADS")) n_chunks = (size_t)std::max(1, std::min(64, std::atoi(e))); // tests: any split must give the same mesh
    std::vector<size_t> cut(n_chunks + 1, (size_t)size);
    cut[0] = 0;
    for (size_t k = 1; k < n_chunks; k++) {
        size_t at = std::max(cut[k - 1], (size_t)size * k / n_chunks);
        const char* nl = at < (size_t)size ? (const char*)std::memchr(text.data() + at, '\n', (size_t)size - at) : nullptr;
        cut[k] = nl ? (size_t)(nl - text.data()) + 1 : (size_t)size;
    }
    std::vector<ObjChunk> chunks(n_chunks);
    if (n_chunks == 1) parse_obj_chunk(text.data(), text.data() + size, chunks[0]);
    else {
        std::vector<std::thread> pool;
        for (size_t k = 0; k < n_chunks; k++)
            pool.emplace_back([&, k] { parse_obj_chunk(text.data() + cut[k], text.data() + cut[k + 1], chunks[k]); });
        for (auto& th : pool) th.join();
    }
    size_t line_base = 0;
    for (const ObjChunk& c : chunks) { // first error in file order
        if (c.err_line) { err = path + ":" + std::to_string(line_base + c.err_line) + ": " + c.err; return RTGPU_ERR_PARSE; }
        line_base += c.lines;
    }
    size_t np = 0, nu = 0, nn = 0, nf = 0;
    for (const ObjChunk& c : chunks) { np += c.positions.size(); nu += c.uvs.size(); nn += c.normals.size(); nf += c.faces.size(); }
    out.positions.reserve(out.positions.size() + np); out.uvs.reserve(out.uvs.size() + nu); out.normals.reserve(out.normals.size() + nn);
    out.tri_pos.reserve(out.tri_pos.size() + nf * 3); out.tri_nrm.reserve(out.tri_nrm.size() + nf * 3); out.tri_uv.reserve(out.tri_uv.size() + nf * 3);
    for (const ObjChunk& c : chunks) {
        const long bv = (long)(out.positions.size() / 3), bt = (long)(out.uvs.size() / 2), bn = (long)(out.normals.size() / 3);
        for (const ObjChunk::Face& fc : c.faces) {
            // counts of records that precede this face in the whole file (negative indices are relative to them)
            const long nv = bv + (long)fc.nv, nt = bt + (long)fc.nt, nnn = bn + (long)fc.nn;
            int iv[3], in[3], it[3] = {-1, -1, -1};
            bool has_uvs = true;
            for (int k = 0; k < 3; k++) {
                const ObjChunk::Corner& cr = fc.c[k];
                iv[k] = (int)(cr.v > 0 ? cr.v - 1 : cr.v < 0 ? nv + cr.v : 0);
                if (cr.has_t) it[k] = (int)(cr.t > 0 ? cr.t - 1 : cr.t < 0 ? nt + cr.t : 0); else has_uvs = false;
                in[k] = (int)(cr.n > 0 ? cr.n - 1 : cr.n < 0 ? nnn + cr.n : 0);
            }
            for (int k = 0; k < 3; k++) {
                out.tri_pos.push_back(iv[k]);
                out.tri_nrm.push_back(in[k]);
                out.tri_uv.push_back(has_uvs ? it[k] : -1);
            }
        }
        out.positions.insert(out.positions.end(), c.positions.begin(), c.positions.end());
        out.uvs.insert(out.uvs.end(), c.uvs.begin(), c.uvs.end());
        out.normals.insert(out.normals.end(), c.normals.begin(), c.normals.end());
    }
    // index validation (the reference would panic at render time on a bad index)
    int nv = (int)(out.positions.size() / 3), nnrm = (int)(out.normals.size() / 3), nt = (int)(out.uvs.size() / 2);
    for (size_t i = 0; i < out.tri_pos.size(); i++) {
        if (out.tri_pos[i] < 0 || out.tri_pos[i] >= nv || out.tri_nrm[i] < 0 || out.tri_nrm[i] >= nnrm ||
            out.tri_uv[i] >= nt) { err = path + ": face index out of range"; return RTGPU_ERR_PARSE; }
    }
    return 0;
}

} // namespace rthost

using namespace rthost;

// Proxy for the missing dragon mesh: midpoint subdivision (1 -> 4) with shared edge
// midpoints per attribute, then a smooth displacement along the normal.
extern "C" int rtgpu_make_proxy_mesh(const char* src_obj, const char* dst_obj, int levels, double displacement, double target_width) {
    MeshData m; std::string err;
    int rc = load_obj(src_obj, m, err);
    if (rc) { set_error(err); return rc; }
    bool uv_ok = !m.uvs.empty();
    for (size_t i = 0; i < m.tri_uv.size(); i++) if (m.tri_uv[i] < 0) uv_ok = false;
    for (int l = 0; l < levels; l++) {
        MeshData n;
        n.positions = m.positions; n.normals = m.normals; n.uvs = m.uvs;
        std::map<std::pair<int, int>, int> mp, mn, mt;
        auto mid = [&](std::map<std::pair<int, int>, int>& cache, std::vector<double>& arr, int dim, int a, int b, bool unit) {
            std::pair<int, int> key(std::min(a, b), std::max(a, b));
            auto it = cache.find(key);
            if (it != cache.end()) return it->second;
            double v[3] = {0, 0, 0}, len = 0;
            for (int k = 0; k < dim; k++) { v[k] = 0.5 * (arr[(size_t)a * dim + k] + arr[(size_t)b * dim + k]); len += v[k] * v[k]; }
            if (unit) { len = std::sqrt(len); if (len > 0) for (int k = 0; k < dim; k++) v[k] /= len; }
            int idx = (int)(arr.size() / dim);
            for (int k = 0; k < dim; k++) arr.push_back(v[k]);
            cache[key] = idx;
            return idx;
        };
        size_t nt = m.tri_pos.size() / 3;
        for (size_t t = 0; t < nt; t++) {
            int p[3], q[3], u[3], pm[3], qm[3], um[3] = {-1, -1, -1};
            for (int k = 0; k < 3; k++) { p[k] = m.tri_pos[t * 3 + k]; q[k] = m.tri_nrm[t * 3 + k]; u[k] = m.tri_uv[t * 3 + k]; }
            for (int k = 0; k < 3; k++) {
                int k2 = (k + 1) % 3;
                pm[k] = mid(mp, n.positions, 3, p[k], p[k2], false);
                qm[k] = mid(mn, n.normals, 3, q[k], q[k2], true);
                if (uv_ok) um[k] = mid(mt, n.uvs, 2, u[k], u[k2], false);
            }
            auto emit = [&](int a, int b, int c, const int* P, const int* Q, const int* U) {
                (void)a; (void)b; (void)c;
                for (int k = 0; k < 3; k++) { n.tri_pos.push_back(P[k]); n.tri_nrm.push_back(Q[k]); n.tri_uv.push_back(uv_ok ? U[k] : -1); }
            };
            int P0[3] = {p[0], pm[0], pm[2]}, Q0[3] = {q[0], qm[0], qm[2]}, U0[3] = {u[0], um[0], um[2]};
            int P1[3] = {pm[0], p[1], pm[1]}, Q1[3] = {qm[0], q[1], qm[1]}, U1[3] = {um[0], u[1], um[1]};
            int P2[3] = {pm[2], pm[1], p[2]}, Q2[3] = {qm[2], qm[1], q[2]}, U2[3] = {um[2], um[1], u[2]};
            int P3[3] = {pm[0], pm[1], pm[2]}, Q3[3] = {qm[0], qm[1], qm[2]}, U3[3] = {um[0], um[1], um[2]};
            emit(0, 0, 0, P0, Q0, U0); emit(0, 0, 0, P1, Q1, U1); emit(0, 0, 0, P2, Q2, U2); emit(0, 0, 0, P3, Q3, U3);
        }
        m = std::move(n);
    }
    // displacement: positions move along a smooth pseudo-normal (position direction) so
    // that shared vertices stay shared and the surface gains fine relief
    if (displacement != 0.0) {
        for (size_t v = 0; v + 2 < m.positions.size(); v += 3) {
            double x = m.positions[v], y = m.positions[v + 1], z = m.positions[v + 2];
            double len = std::sqrt(x * x + y * y + z * z);
            if (len == 0) continue;
            double d = displacement * (std::sin(37.0 * x) * std::sin(41.0 * y) * std::sin(43.0 * z) +
                                       0.5 * std::sin(91.0 * x + 1.3) * std::sin(97.0 * y + 0.7) * std::sin(89.0 * z + 2.1));
            m.positions[v] += d * x / len; m.positions[v + 1] += d * y / len; m.positions[v + 2] += d * z / len;
        }
    }
    if (target_width > 0.0 && !m.positions.empty()) {
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
        for (size_t v = 0; v + 2 < m.positions.size(); v += 3)
            for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], m.positions[v + k]); hi[k] = std::max(hi[k], m.positions[v + k]); }
        const double sc = target_width / (hi[0] - lo[0]);
        const double cx = 0.5 * (lo[0] + hi[0]), cz = 0.5 * (lo[2] + hi[2]);
        for (size_t v = 0; v + 2 < m.positions.size(); v += 3) {
            m.positions[v] = (m.positions[v] - cx) * sc;
            m.positions[v + 1] = (m.positions[v + 1] - lo[1]) * sc;
            m.positions[v + 2] = (m.positions[v + 2] - cz) * sc;
        }
    }
    FILE* f = std::fopen(dst_obj, "wb");
    if (!f) { set_error(std::string("cannot write ") + dst_obj); return RTGPU_ERR_IO; }
    std::fprintf(f, "# proxy mesh generated by rtgpu_make_proxy_mesh(levels=%d, displacement=%g) from %s\n", levels, displacement, src_obj);
    for (size_t v = 0; v + 2 < m.positions.size(); v += 3) std::fprintf(f, "v %.9g %.9g %.9g\n", m.positions[v], m.positions[v + 1], m.positions[v + 2]);
    for (size_t v = 0; v + 1 < m.uvs.size(); v += 2) std::fprintf(f, "vt %.9g %.9g\n", m.uvs[v], m.uvs[v + 1]);
    for (size_t v = 0; v + 2 < m.normals.size(); v += 3) std::fprintf(f, "vn %.9g %.9g %.9g\n", m.normals[v], m.normals[v + 1], m.normals[v + 2]);
    size_t nt = m.tri_pos.size() / 3;
    for (size_t t = 0; t < nt; t++) {
        std::fprintf(f, "f");
        for (int k = 0; k < 3; k++) {
            if (uv_ok) std::fprintf(f, " %d/%d/%d", m.tri_pos[t * 3 + k] + 1, m.tri_uv[t * 3 + k] + 1, m.tri_nrm[t * 3 + k] + 1);
            else std::fprintf(f, " %d//%d", m.tri_pos[t * 3 + k] + 1, m.tri_nrm[t * 3 + k] + 1);
        }
        std::fprintf(f, "\n");
    }
    std::fclose(f);
    return 0;
}
