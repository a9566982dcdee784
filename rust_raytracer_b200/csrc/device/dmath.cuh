// dmath.cuh — small vector math for the kernels. Everything is RT_HD so the same code also
// compiles for the host-simulation unit tests (tests/hostsim), which exist to debug the
// device logic on a machine without a GPU; the product never runs it on the CPU.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#include "scene_types.h"

namespace rt {

template <typename T> struct V3 {
    T x, y, z;
    RT_HD V3() {}
    RT_HD V3(T a, T b, T c) : x(a), y(b), z(c) {}
    template <typename U> RT_HD explicit V3(const V3<U>& o) : x((T)o.x), y((T)o.y), z((T)o.z) {}
};
typedef V3<float> f3;
typedef V3<double> d3;

template <typename T> RT_HD V3<T> operator+(V3<T> a, V3<T> b) { return V3<T>(a.x + b.x, a.y + b.y, a.z + b.z); }
template <typename T> RT_HD V3<T> operator-(V3<T> a, V3<T> b) { return V3<T>(a.x - b.x, a.y - b.y, a.z - b.z); }
template <typename T> RT_HD V3<T> operator*(V3<T> a, V3<T> b) { return V3<T>(a.x * b.x, a.y * b.y, a.z * b.z); }
template <typename T> RT_HD V3<T> operator*(V3<T> a, T s) { return V3<T>(a.x * s, a.y * s, a.z * s); }
template <typename T> RT_HD V3<T> operator/(V3<T> a, T s) { return V3<T>(a.x / s, a.y / s, a.z / s); }
// fp32: one reciprocal, three multiplies. Three IEEE divisions cost three times the fast path AND, for every
// zero component (axis-aligned wall normals), the out-of-line slow path: ncu counted 3 calls of it per shaded
// path in k_shade<0> (5 % of the kernel's instructions). The f64 overload above keeps the reference's divisions.
RT_HD V3<float> operator/(V3<float> a, float s) { const float r = 1.0f / s; return V3<float>(a.x * r, a.y * r, a.z * r); }
template <typename T> RT_HD V3<T> operator-(V3<T> a) { return V3<T>(-a.x, -a.y, -a.z); }
template <typename T> RT_HD T dot(V3<T> a, V3<T> b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <typename T> RT_HD V3<T> cross(V3<T> a, V3<T> b) { return V3<T>(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
// fp32 shading helpers. On the device: MUFU approximations (sqrt / rsqrt ~2 ulp, sin / cos ~1e-6 absolute on
// [0, 2 pi]) — an IEEE sqrtf / division is a ~10-instruction dependent chain with an out-of-line slow path, and
// the shade kernels are bound by exactly such chains. The f64 helpers (reference parity) are untouched.
#if defined(__CUDA_ARCH__)
RT_HD float rt_sqrtf(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
RT_HD float rt_rsqrtf(float x) { return rsqrtf(x); }
RT_HD void rt_sincosf(float x, float& s, float& c) { __sincosf(x, &s, &c); }
RT_HD float rt_divf(float a, float b) { return __fdividef(a, b); }
#else
RT_HD float rt_sqrtf(float x) { return sqrtf(x); }
RT_HD float rt_rsqrtf(float x) { return 1.0f / sqrtf(x); }
RT_HD void rt_sincosf(float x, float& s, float& c) { s = sinf(x); c = cosf(x); }
RT_HD float rt_divf(float a, float b) { return a / b; }
#endif
RT_HD float length(f3 a) { return rt_sqrtf(dot(a, a)); }
RT_HD double length(d3 a) { return sqrt(dot(a, a)); }
RT_HD f3 normalize(f3 a) { const float r = rt_rsqrtf(dot(a, a)); return f3(a.x * r, a.y * r, a.z * r); }
RT_HD d3 normalize(d3 a) { return a / length(a); }
template <typename T> RT_HD V3<T> reflect(V3<T> v, V3<T> n) { return v - n * ((T)2 * dot(v, n)); } // vec4.rs:135

RT_HD uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
RT_HD float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

// bit helpers with identical semantics on host and device
RT_HD int rt_popc(uint32_t v) {
#if defined(__CUDA_ARCH__)
    return __popc(v);
#else
    return __builtin_popcount(v);
#endif
}
RT_HD int rt_bfind(uint32_t v) { // index of the highest set bit (v != 0)
#if defined(__CUDA_ARCH__)
    return 31 - __clz(v);
#else
    return 31 - __builtin_clz(v);
#endif
}

// ---- exact f64 arithmetic: never contracted into FMAs, so results are bit-identical to the
// reference's (and the oracle's) plain IEEE operations --------------------------------------
#if defined(__CUDA_ARCH__)
RT_HD double xmul(double a, double b) { return __dmul_rn(a, b); }
RT_HD double xadd(double a, double b) { return __dadd_rn(a, b); }
RT_HD double xsub(double a, double b) { return __dsub_rn(a, b); }
RT_HD double xdiv(double a, double b) { return __ddiv_rn(a, b); }
RT_HD double xsqrt(double a) { return __dsqrt_rn(a); }
#else
// host simulation: the translation unit is built with -ffp-contract=off
RT_HD double xmul(double a, double b) { return a * b; }
RT_HD double xadd(double a, double b) { return a + b; }
RT_HD double xsub(double a, double b) { return a - b; }
RT_HD double xdiv(double a, double b) { return a / b; }
RT_HD double xsqrt(double a) { return sqrt(a); }
#endif

struct X3 { double x, y, z; }; // exact-op vector (reference Vec4 xyz)
RT_HD X3 x3(double a, double b, double c) { X3 r; r.x = a; r.y = b; r.z = c; return r; }
RT_HD X3 xsub3(X3 a, X3 b) { return x3(xsub(a.x, b.x), xsub(a.y, b.y), xsub(a.z, b.z)); }
RT_HD X3 xadd3(X3 a, X3 b) { return x3(xadd(a.x, b.x), xadd(a.y, b.y), xadd(a.z, b.z)); }
RT_HD X3 xscale3(X3 a, double s) { return x3(xmul(a.x, s), xmul(a.y, s), xmul(a.z, s)); }
RT_HD X3 xdivs3(X3 a, double s) { return x3(xdiv(a.x, s), xdiv(a.y, s), xdiv(a.z, s)); }
// vec4.rs:109  x*x' + y*y' + z*z' evaluated left to right
RT_HD double xdot3(X3 a, X3 b) { return xadd(xadd(xmul(a.x, b.x), xmul(a.y, b.y)), xmul(a.z, b.z)); }
// vec4.rs:113
RT_HD X3 xcross3(X3 a, X3 b) {
    return x3(xsub(xmul(a.y, b.z), xmul(a.z, b.y)), xsub(xmul(a.z, b.x), xmul(a.x, b.z)), xsub(xmul(a.x, b.y), xmul(a.y, b.x)));
}
RT_HD double xlen3(X3 a) { return xsqrt(xdot3(a, a)); }
// Mat4 * Vec4 rows 0..2 (mat4.rs:345-352): m0*x + m1*y + m2*z + m3*w, left to right
RT_HD X3 xmat_apply(const double* m, X3 p, double w) {
    return x3(xadd(xadd(xadd(xmul(m[0], p.x), xmul(m[1], p.y)), xmul(m[2], p.z)), xmul(m[3], w)),
              xadd(xadd(xadd(xmul(m[4], p.x), xmul(m[5], p.y)), xmul(m[6], p.z)), xmul(m[7], w)),
              xadd(xadd(xadd(xmul(m[8], p.x), xmul(m[9], p.y)), xmul(m[10], p.z)), xmul(m[11], w)));
}

#define RT_PI_F 3.14159265358979323846f
#define RT_PI_D 3.14159265358979323846

} // namespace rt
