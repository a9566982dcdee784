// oracle_math.h — TEST INFRASTRUCTURE (CPU oracle). Not part of the product path.
//
// f64 restatement of the reference's math layer, operation for operation:
//   Vec4   src/vec4.rs        (all four components take part in + - * /, only xyz in dot/length/cross)
//   Mat4   src/mat4.rs        (row-major)
//   Ray    src/ray.rs:19-37
//   AABB   src/aabb.rs:9-87   (Williams slab test, +-0.001 padding)
//   utils  src/utils.rs:17-36 (onb_from_vec, Schlick reflectance)
// Build with -ffp-contract=off so no FMA contraction changes a result.
#pragma once
#include <cmath>
#include <cstdint>
#include <limits>

namespace oracle {

static const double PI = 3.14159265358979323846;
static const double INF = std::numeric_limits<double>::infinity();
static const double F64_EPSILON = 2.220446049250313e-16; // f64::EPSILON

struct Vec4 {
    double v[4];
    Vec4() : v{0, 0, 0, 0} {}
    Vec4(double x, double y, double z, double w) : v{x, y, z, w} {}
    static Vec4 vec(double x, double y, double z) { return Vec4(x, y, z, 0.0); }   // vec4.rs:19
    static Vec4 point(double x, double y, double z) { return Vec4(x, y, z, 1.0); } // vec4.rs:23
    double x() const { return v[0]; }
    double y() const { return v[1]; }
    double z() const { return v[2]; }
    double w() const { return v[3]; }
    double operator[](int i) const { return v[i]; }
    double& operator[](int i) { return v[i]; }
    double length_squared() const { return v[0] * v[0] + v[1] * v[1] + v[2] * v[2]; } // vec4.rs:105
    double length() const { return std::sqrt(length_squared()); }
    double dot(const Vec4& o) const { return v[0] * o.v[0] + v[1] * o.v[1] + v[2] * o.v[2]; } // vec4.rs:109
    Vec4 cross(const Vec4& o) const { // vec4.rs:113
        return Vec4(v[1] * o.v[2] - v[2] * o.v[1], v[2] * o.v[0] - v[0] * o.v[2], v[0] * o.v[1] - v[1] * o.v[0], 0.0);
    }
    Vec4 to_unit() const; // vec4.rs:122 (divides all four components)
    Vec4 reflect(const Vec4& n) const;                   // vec4.rs:135
    Vec4 refract(const Vec4& n, double ior_ratio) const; // vec4.rs:140
};
inline Vec4 operator+(const Vec4& a, const Vec4& b) { return Vec4(a.v[0] + b.v[0], a.v[1] + b.v[1], a.v[2] + b.v[2], a.v[3] + b.v[3]); }
inline Vec4 operator-(const Vec4& a, const Vec4& b) { return Vec4(a.v[0] - b.v[0], a.v[1] - b.v[1], a.v[2] - b.v[2], a.v[3] - b.v[3]); }
inline Vec4 operator*(const Vec4& a, const Vec4& b) { return Vec4(a.v[0] * b.v[0], a.v[1] * b.v[1], a.v[2] * b.v[2], a.v[3] * b.v[3]); }
inline Vec4 operator+(const Vec4& a, double s) { return Vec4(a.v[0] + s, a.v[1] + s, a.v[2] + s, a.v[3] + s); }
inline Vec4 operator-(const Vec4& a, double s) { return Vec4(a.v[0] - s, a.v[1] - s, a.v[2] - s, a.v[3] - s); }
inline Vec4 operator*(const Vec4& a, double s) { return Vec4(a.v[0] * s, a.v[1] * s, a.v[2] * s, a.v[3] * s); }
inline Vec4 operator/(const Vec4& a, double s) { return Vec4(a.v[0] / s, a.v[1] / s, a.v[2] / s, a.v[3] / s); }
inline Vec4 operator-(const Vec4& a) { return Vec4(-a.v[0], -a.v[1], -a.v[2], -a.v[3]); }
inline Vec4& operator+=(Vec4& a, const Vec4& b) { a = a + b; return a; }
inline Vec4& operator*=(Vec4& a, double s) { a = a * s; return a; }
inline Vec4& operator/=(Vec4& a, double s) { a = a / s; return a; }
inline Vec4 Vec4::to_unit() const { return *this / length(); }
inline Vec4 Vec4::reflect(const Vec4& n) const { return *this - n * (2.0 * dot(n)); }
inline Vec4 Vec4::refract(const Vec4& n, double ior_ratio) const {
    double cos_theta = std::fmin(1.0, (-*this).dot(n));
    Vec4 perp = (*this + (n * cos_theta)) * ior_ratio;
    Vec4 parallel = n * -std::sqrt(1.0 - perp.length_squared());
    return perp + parallel;
}

struct Mat4 {
    double m[16];
    static Mat4 from_columns(const Vec4& c0, const Vec4& c1, const Vec4& c2, const Vec4& c3) { // mat4.rs:37
        Mat4 r;
        for (int i = 0; i < 4; i++) { r.m[i * 4 + 0] = c0[i]; r.m[i * 4 + 1] = c1[i]; r.m[i * 4 + 2] = c2[i]; r.m[i * 4 + 3] = c3[i]; }
        return r;
    }
    Vec4 operator*(const Vec4& r) const { // mat4.rs:345-352
        return Vec4(m[0] * r[0] + m[1] * r[1] + m[2] * r[2] + m[3] * r[3],
                    m[4] * r[0] + m[5] * r[1] + m[6] * r[2] + m[7] * r[3],
                    m[8] * r[0] + m[9] * r[1] + m[10] * r[2] + m[11] * r[3],
                    m[12] * r[0] + m[13] * r[1] + m[14] * r[2] + m[15] * r[3]);
    }
};

struct Interval { double lo, hi; }; // interval.rs

struct Ray { // ray.rs:4-33
    Vec4 origin, dir, inv_dir;
    uint8_t sign[3];
    Ray() {}
    Ray(const Vec4& o, const Vec4& d) : origin(o), dir(d) {
        inv_dir = Vec4::vec(1.0 / d[0], 1.0 / d[1], 1.0 / d[2]);
        for (int i = 0; i < 3; i++) sign[i] = inv_dir[i] < 0.0 ? 1 : 0;
    }
    Vec4 at(double t) const { return origin + (dir * t); }
};

struct AABB { Vec4 b[2]; }; // aabb.rs:6

static const Vec4 EPSILON_VEC(0.001, 0.001, 0.001, 0.0); // aabb.rs:9
static const Vec4 INFINITY_VEC(INF, INF, INF, 1.0);      // constants.rs:3

inline AABB combine_bounds(const AABB* boxes, int n) { // aabb.rs:11-27
    Vec4 lo = INFINITY_VEC, hi = -lo;
    for (int k = 0; k < n; k++)
        for (int i = 0; i < 3; i++) {
            if (boxes[k].b[0][i] < lo[i]) lo[i] = boxes[k].b[0][i];
            if (boxes[k].b[1][i] > hi[i]) hi[i] = boxes[k].b[1][i];
        }
    return AABB{{lo - EPSILON_VEC, hi + EPSILON_VEC}};
}
inline AABB get_bounding_box(const Vec4* verts, int n) { // aabb.rs:29-45
    Vec4 lo = INFINITY_VEC, hi = -lo;
    for (int k = 0; k < n; k++)
        for (int i = 0; i < 3; i++) {
            if (verts[k][i] < lo[i]) lo[i] = verts[k][i];
            if (verts[k][i] > hi[i]) hi[i] = verts[k][i];
        }
    return AABB{{lo - EPSILON_VEC, hi + EPSILON_VEC}};
}

inline bool test_bounding_box(const AABB& bounds, const Ray& ray, const Interval& t_int) { // aabb.rs:50-87
    const Vec4& inv_dir = ray.inv_dir;
    const uint8_t* sign = ray.sign;
    double t_min = (bounds.b[sign[0]][0] - ray.origin[0]) * inv_dir[0];
    double t_max = (bounds.b[1 - sign[0]][0] - ray.origin[0]) * inv_dir[0];
    double ty_min = (bounds.b[sign[1]][1] - ray.origin[1]) * inv_dir[1];
    double ty_max = (bounds.b[1 - sign[1]][1] - ray.origin[1]) * inv_dir[1];
    if ((t_min > ty_max) || (ty_min > t_max)) return false;
    if (ty_min > t_min) t_min = ty_min;
    if (ty_max < t_max) t_max = ty_max;
    double tz_min = (bounds.b[sign[2]][2] - ray.origin[2]) * inv_dir[2];
    double tz_max = (bounds.b[1 - sign[2]][2] - ray.origin[2]) * inv_dir[2];
    if ((t_min > tz_max) || (tz_min > t_max)) return false;
    if (tz_min > t_min) t_min = tz_min;
    if (tz_max < t_max) t_max = tz_max;
    return t_min < t_int.hi && t_max > t_int.lo;
}

inline Mat4 onb_from_vec(const Vec4& w) { // utils.rs:17-28
    Vec4 a = std::fabs(w.x()) > 0.9 ? Vec4::vec(0.0, 1.0, 0.0) : Vec4::vec(1.0, 0.0, 0.0);
    Vec4 v = w.cross(a).to_unit();
    Vec4 u = w.cross(v);
    return Mat4::from_columns(u, v, w, Vec4(0.0, 0.0, 0.0, 1.0));
}

inline double powi5(double a) { // f64::powi(5): compiler-rt __powidf2 square-and-multiply order
    double r = a;      // n = 5: bit0 set
    a = a * a;         // a^2
    a = a * a;         // a^4 (bit1 clear, bit2 set)
    return r * a;
}
inline double reflectance(double cos_theta, double ior_ratio) { // utils.rs:31-36
    double r0 = (1.0 - ior_ratio) / (1.0 + ior_ratio);
    r0 = r0 * r0;
    return r0 + (1.0 - r0) * powi5(1.0 - cos_theta);
}

// ---- RNG: the reference owns one Pcg64Mcg per render thread, seeded from OS entropy
// (camera.rs:208-209) => "parity unpinned"; only the draw ORDER and distributions are
// restated. Generator: published PCG-XSL-RR 128/64 MCG (rand_pcg 0.3.1 Mcg128Xsl64).
struct Rng {
    unsigned __int128 state;
    explicit Rng(uint64_t seed, uint64_t stream = 0) {
        uint64_t s = seed * 0x9E3779B97F4A7C15ULL + stream * 0xD1B54A32D192ED03ULL + 0x632BE59BD9B4E019ULL;
        uint64_t w[2];
        for (int i = 0; i < 2; i++) { // splitmix64 to fill 128 bits
            s += 0x9E3779B97F4A7C15ULL;
            uint64_t z = s;
            z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
            z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
            w[i] = z ^ (z >> 31);
        }
        state = (((unsigned __int128)w[1] << 64) | w[0]) | 1;
    }
    uint64_t next_u64() {
        const unsigned __int128 mult = ((unsigned __int128)0x2360ED051FC65DA4ULL << 64) | 0x4385DF649FCCF645ULL;
        state = state * mult;
        uint64_t hi = (uint64_t)(state >> 64), lo = (uint64_t)state;
        uint32_t rot = (uint32_t)(hi >> 58);
        uint64_t x = hi ^ lo;
        return (x >> rot) | (x << ((64 - rot) & 63));
    }
    double uniform() { return (double)(next_u64() >> 11) * (1.0 / 9007199254740992.0); } // Standard / gen_range(0.0..1.0)
    uint64_t below(uint64_t n) { // gen_range(0..n)
        uint64_t zone = (n << __builtin_clzll(n)) - 1;
        for (;;) {
            unsigned __int128 m = (unsigned __int128)next_u64() * n;
            if ((uint64_t)m <= zone) return (uint64_t)(m >> 64);
        }
    }
    double normal() { // StandardNormal (distribution only; Box-Muller, one value per call)
        double u1 = 1.0 - uniform(), u2 = uniform();
        return std::sqrt(-2.0 * std::log(u1)) * std::cos(2.0 * PI * u2);
    }
    // injected-sample mode for per-event parity tests: when `inject` is set, draws come
    // from the array instead (uniforms and normals share the stream, in draw order)
    const double* inject = nullptr;
    int inject_n = 0, inject_pos = 0;
    double draw_uniform() { return (inject && inject_pos < inject_n) ? inject[inject_pos++] : uniform(); }
    double draw_normal() { return (inject && inject_pos < inject_n) ? inject[inject_pos++] : normal(); }
};

inline Vec4 random_in_unit_disk(Rng& rng) { // vec4.rs:35-40: on the unit CIRCLE
    double x = rng.draw_normal(), y = rng.draw_normal();
    return Vec4(x, y, 0.0, 0.0).to_unit();
}
inline Vec4 random_unit(Rng& rng) { // vec4.rs:42-48
    double x = rng.draw_normal(), y = rng.draw_normal(), z = rng.draw_normal();
    return Vec4(x, y, z, 0.0).to_unit();
}
inline Vec4 random_cosine(Rng& rng) { // vec4.rs:50-61
    double r1 = rng.draw_uniform(), r2 = rng.draw_uniform();
    double phi = r1 * 2.0 * PI;
    double sqrt_r2 = std::sqrt(r2);
    double x = std::cos(phi) * sqrt_r2, y = std::sin(phi) * sqrt_r2, z = std::sqrt(1.0 - r2);
    return Vec4::vec(x, y, z);
}

} // namespace oracle
