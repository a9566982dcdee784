// obj_loader.cpp — load_mesh_from_file (reference src/loaders/obj.rs:13-107), restated.
// Only `v`, `vt`, `vn`, `f` are read; lines are split on single spaces; faces must be
// triangles `v/vt/vn` or `v//vn` (the normal index is mandatory); indices may be negative;
// `vn` is normalised on load. Also hosts the proxy-mesh generator for the missing dragon.
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <chrono>
#include <climits>
#include <cmath>
#include <memory>
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

// Decimal token [q, end) -> f64, correctly rounded WITHOUT strtod when that is provably exact (Clinger's fast path): at most
// 15 significant digits (the integer mantissa is < 2^53, exactly a double) and a power of ten up to 10^22 (exactly a double),
// so ONE IEEE multiplication or division of two exact operands gives the correctly rounded result — the same bits strtod and
// Rust's `parse::<f64>()` (obj.rs:31-58) produce. Anything else (longer mantissas, large exponents, inf / nan, malformed
// text) returns false and goes to strtod. OBJ exporters write 6-7 digits, so this is the path of every number in practice:
// strtod was ~85 % of the parse time of the 115 MB dragon file.
static inline bool fast_parse_f64(const char* q, const char* end, double& out) {
    static const double P10[23] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
    const char* p = q;
    bool neg = false;
    if (p < end && (*p == '-' || *p == '+')) { neg = *p == '-'; p++; }
    uint64_t m = 0;
    int digits = 0, sig = 0, exp10 = 0;
    bool seen_digit = false;
    for (; p < end && *p >= '0' && *p <= '9'; p++) {
        seen_digit = true;
        if (sig > 0 || *p != '0') { if (sig >= 15) return false; m = m * 10 + (uint64_t)(*p - '0'); sig++; }
        digits++;
    }
    if (p < end && *p == '.') {
        p++;
        for (; p < end && *p >= '0' && *p <= '9'; p++) {
            seen_digit = true;
            if (sig > 0 || *p != '0') { if (sig >= 15) return false; m = m * 10 + (uint64_t)(*p - '0'); sig++; }
            exp10--;
        }
    }
    if (!seen_digit) return false;
    if (p < end && (*p == 'e' || *p == 'E')) {
        p++;
        bool eneg = false;
        if (p < end && (*p == '-' || *p == '+')) { eneg = *p == '-'; p++; }
        if (!(p < end && *p >= '0' && *p <= '9')) return false;
        int e = 0;
        for (; p < end && *p >= '0' && *p <= '9'; p++) { e = e * 10 + (*p - '0'); if (e > 10000) return false; }
        exp10 += eneg ? -e : e;
    }
    if (p != end) return false;
    double v;
    if (m == 0) v = 0.0;
    else if (exp10 == 0) v = (double)m;
    else if (exp10 < 0 && exp10 >= -22) v = (double)m / P10[-exp10];
    else if (exp10 > 0 && exp10 <= 22) v = (double)m * P10[exp10];
    else return false;
    out = neg ? -v : v;
    return true;
}

namespace {

// One contiguous run of lines, parsed independently of the others. Faces keep their raw 1-based /
// negative indices plus the number of v / vt / vn records seen in THIS chunk before the face; the
// merge step adds the counts of the preceding chunks to resolve negative (relative) indices.
struct ObjChunk {
    std::vector<double> positions, uvs, normals;
    struct Corner { int32_t v, t, n; };                       // raw 1-based / negative indices; t == 0: no uv index
    struct Face { Corner c[3]; uint32_t nv, nt, nn; };        // 48 bytes (was 120: the records are written once and merged once)
    std::vector<Face> faces;
    size_t lines = 0;
    size_t err_line = 0; // 1-based line inside the chunk of the first error, 0 = none
    std::string err;
};

void parse_obj_chunk(char* begin, char* end, ObjChunk& out) {
    { // typical line lengths: ~30 bytes per v / vn, ~20 per vt, ~45 per f — reserve once instead of doubling 20 times
        const size_t bytes = (size_t)(end - begin);
        out.positions.reserve(bytes / 40); out.normals.reserve(bytes / 40); out.uvs.reserve(bytes / 80); out.faces.reserve(bytes / 90);
    }
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
                double v;
                if (!fast_parse_f64(q, tok_end, v)) {
                    char* num_end = nullptr;
                    v = std::strtod(q, &num_end);
                    if (num_end != tok_end || tok_end == q) return -1; // unwrap() panic
                }
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
                    if (e > s) { // [+-]digits, like `parse::<isize>()` (obj.rs:64-79); nine of these per face line
                        const char* d = s;
                        const bool neg = *d == '-';
                        if (*d == '-' || *d == '+') d++;
                        long val = 0;
                        bool digits_ok = d < e && e - d <= 10;
                        for (; digits_ok && d < e; d++) {
                            if (*d < '0' || *d > '9') { digits_ok = false; break; }
                            val = val * 10 + (*d - '0');
                        }
                        if (!digits_ok) { bad("bad face index"); ok = false; break; }
                        part[np] = neg ? -val : val;
                        present[np] = true;
                    }
                    np++;
                    if (e >= tok_end) break;
                    s = e + 1;
                }
                if (!ok) break;
                if (np < 3 || !present[0] || !present[2]) { bad("face corner must be v/vt/vn or v//vn"); ok = false; break; }
                if (part[0] > INT32_MAX || part[0] < INT32_MIN || part[1] > INT32_MAX || part[1] < INT32_MIN || part[2] > INT32_MAX || part[2] < INT32_MIN ||
                    (present[1] && part[1] == 0)) { bad("face index out of range"); ok = false; break; }
                face.c[corner] = {(int32_t)part[0], present[1] ? (int32_t)part[1] : 0, (int32_t)part[2]};
                corner++;
                q = tok_end;
            }
            if (ok) {
                face.nv = (uint32_t)(out.positions.size() / 3); face.nt = (uint32_t)(out.uvs.size() / 2); face.nn = (uint32_t)(out.normals.size() / 3);
                out.faces.push_back(face);
            }
        }
        *lend = saved;
        pos = eol + 1;
    }
}

} // namespace

int load_obj(const std::string& path, MeshData& out, std::string& err) {
    const bool timing = std::getenv("RTGPU_BUILD_TIMING") != nullptr;
    const auto t0 = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (timing) std::fprintf(stderr, "[rtgpu build]   obj: %-21s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    };
    const int fd = ::open(path.c_str(), O_RDONLY);
    if (fd < 0) { err = "cannot open mesh file " + path; return RTGPU_ERR_IO; }
    struct stat st;
    if (::fstat(fd, &st) != 0 || st.st_size < 0) { ::close(fd); err = "cannot stat " + path; return RTGPU_ERR_IO; }
    const long size = (long)st.st_size;
    // uninitialised buffer (a zero-filled std::vector<char> was a 115 MB memset before the read), filled by parallel preads
    std::unique_ptr<char[]> text_buf(new char[(size_t)size + 1]);
    struct { char* p; char* data() { return p; } } text{text_buf.get()};
    text.p[size] = '\0';
    {
        const size_t n_rd = size < (1 << 22) ? 1 : std::max(1u, std::min(8u, rt_host_cores()));
        std::vector<int> bad_rd(n_rd, 0);
        auto rd = [&](size_t k) {
            size_t at = (size_t)size * k / n_rd;
            const size_t stop = (size_t)size * (k + 1) / n_rd;
            while (at < stop) {
                const ssize_t got = ::pread(fd, text.p + at, stop - at, (off_t)at);
                if (got <= 0) { bad_rd[k] = 1; return; }
                at += (size_t)got;
            }
        };
        if (n_rd == 1) rd(0);
        else {
            std::vector<std::thread> pool;
            for (size_t k = 0; k < n_rd; k++) pool.emplace_back(rd, k);
            for (auto& th : pool) th.join();
        }
        ::close(fd);
        for (int b : bad_rd) if (b) { err = "short read on " + path; return RTGPU_ERR_IO; }
    }
    lap("read");

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
    lap("parse chunks");
    size_t line_base = 0;
    for (const ObjChunk& c : chunks) { // first error in file order
        if (c.err_line) { err = path + ":" + std::to_string(line_base + c.err_line) + ": " + c.err; return RTGPU_ERR_PARSE; }
        line_base += c.lines;
    }
    // merge in file order, in parallel: prefix sums give every chunk its place in the final arrays
    const size_t p0 = out.positions.size(), u0 = out.uvs.size(), n0 = out.normals.size(), f0 = out.tri_pos.size();
    std::vector<size_t> op(n_chunks + 1, p0), ou(n_chunks + 1, u0), on(n_chunks + 1, n0), of(n_chunks + 1, f0);
    for (size_t k = 0; k < n_chunks; k++) {
        op[k + 1] = op[k] + chunks[k].positions.size(); ou[k + 1] = ou[k] + chunks[k].uvs.size();
        on[k + 1] = on[k] + chunks[k].normals.size(); of[k + 1] = of[k] + chunks[k].faces.size() * 3;
    }
    out.positions.resize(op[n_chunks]); out.uvs.resize(ou[n_chunks]); out.normals.resize(on[n_chunks]);
    out.tri_pos.resize(of[n_chunks]); out.tri_nrm.resize(of[n_chunks]); out.tri_uv.resize(of[n_chunks]);
    const long nv_all = (long)(op[n_chunks] / 3), nn_all = (long)(on[n_chunks] / 3), nt_all = (long)(ou[n_chunks] / 2);
    std::vector<int> bad_index(n_chunks, 0);
    auto merge = [&](size_t k) {
        const ObjChunk& c = chunks[k];
        if (!c.positions.empty()) std::memcpy(&out.positions[op[k]], c.positions.data(), c.positions.size() * sizeof(double));
        if (!c.uvs.empty()) std::memcpy(&out.uvs[ou[k]], c.uvs.data(), c.uvs.size() * sizeof(double));
        if (!c.normals.empty()) std::memcpy(&out.normals[on[k]], c.normals.data(), c.normals.size() * sizeof(double));
        // counts of records that precede a face in the whole file (negative indices are relative to them)
        const long bv = (long)(op[k] / 3), bt = (long)(ou[k] / 2), bn = (long)(on[k] / 3);
        size_t at = of[k];
        for (const ObjChunk::Face& fc : c.faces) {
            const long nv = bv + (long)fc.nv, nt = bt + (long)fc.nt, nnn = bn + (long)fc.nn;
            long iv[3], in[3], it[3] = {-1, -1, -1};
            bool has_uvs = true;
            for (int j = 0; j < 3; j++) {
                const ObjChunk::Corner& cr = fc.c[j];
                iv[j] = cr.v > 0 ? cr.v - 1 : cr.v < 0 ? nv + cr.v : 0;
                if (cr.t != 0) it[j] = cr.t > 0 ? cr.t - 1 : nt + cr.t; else has_uvs = false;
                in[j] = cr.n > 0 ? cr.n - 1 : cr.n < 0 ? nnn + cr.n : 0;
            }
            for (int j = 0; j < 3; j++, at++) {
                // index validation (the reference would panic at render time on a bad index)
                if (iv[j] < 0 || iv[j] >= nv_all || in[j] < 0 || in[j] >= nn_all || (has_uvs && (it[j] < 0 || it[j] >= nt_all))) bad_index[k] = 1;
                out.tri_pos[at] = (int32_t)iv[j];
                out.tri_nrm[at] = (int32_t)in[j];
                out.tri_uv[at] = has_uvs ? (int32_t)it[j] : -1;
            }
        }
    };
    if (n_chunks == 1) merge(0);
    else {
        std::vector<std::thread> pool;
        for (size_t k = 0; k < n_chunks; k++) pool.emplace_back(merge, k);
        for (auto& th : pool) th.join();
    }
    for (int b : bad_index) if (b) { err = path + ": face index out of range"; return RTGPU_ERR_PARSE; }
    lap("merge + validate");
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
