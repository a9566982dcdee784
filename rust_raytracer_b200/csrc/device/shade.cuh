// shade.cuh — one bounce of the reference's integrator (Camera::ray_color,
// src/camera.rs:294-328) for a path whose closest hit is known: emission, Material::scatter,
// the light-biased MixPDF (src/pdf/mix.rs) over CosinePDF / UniformPDF and HittablePDF(lights),
// and the throughput weight attenuation * scattering_pdf / pdf. The RNG draw ORDER per event
// follows the reference so that injected-number parity tests line up.
#pragma once
#include "rng.cuh"
#include "surface.cuh"
#include "texture.cuh"

namespace rt {

enum : int32_t { BOUNCE_MISS = 0, BOUNCE_PDF = 1, BOUNCE_RAY = 2, BOUNCE_ABSORBED = 3, BOUNCE_EMISSIVE = 4 };

struct Bounce {
    int32_t kind;
    f3 emission;   // radiance emitted at the hit (or background on a miss)
    f3 weight;     // throughput multiplier for the continued path
    d3 origin;     // next ray
    f3 dir;
    float pdf, spdf;
};

// utils.rs:31-36
RT_HD float reflectance_f(float cos_theta, float ior_ratio) {
    float r0 = (1.0f - ior_ratio) / (1.0f + ior_ratio);
    r0 = r0 * r0;
    const float a = 1.0f - cos_theta, a2 = a * a;
    return r0 + (1.0f - r0) * (a * (a2 * a2));
}

// utils.rs:17-28: columns (u, v, w) with w NOT necessarily unit
RT_HD f3 onb_apply(f3 w, f3 local) {
    const f3 a = fabsf(w.x) > 0.9f ? f3(0.f, 1.f, 0.f) : f3(1.f, 0.f, 0.f);
    const f3 v = normalize(cross(w, a));
    const f3 u = cross(w, v);
    return u * local.x + v * local.y + w * local.z;
}

template <typename RNG> RT_HD f3 random_unit_f(RNG& rng) { // vec4.rs:42-48
    const float x = rng.normal(), y = rng.normal(), z = rng.normal();
    return normalize(f3(x, y, z));
}
template <typename RNG> RT_HD f3 random_cosine_f(RNG& rng) { // vec4.rs:50-61
    const float r1 = rng.uniform(), r2 = rng.uniform();
    const float phi = r1 * 2.0f * RT_PI_F, sq = rt_sqrtf(r2);
    float sn, cs;
    rt_sincosf(phi, sn, cs);
    return f3(cs * sq, sn * sq, rt_sqrtf(1.0f - r2));
}

// ---- HittablePDF over the flattened `lights` (pdf/hittable.rs, list.rs:80-100) -----------------
template <bool SPH = true>
RT_HD float light_pdf_value(const DeviceScene& s, const Light& l, d3 origin, f3 dir) {
    if (!SPH && l.kind == LIGHT_SPHERE) return 0.f; // (no such light in the scene)
    switch (l.kind) {
    case LIGHT_PLANE: { // plane.rs:107-118: pdf = dist^2 / (cos * area) when the ray hits the quad, else 0
        // fp32 with the one cancellation-prone step (corner - origin, coordinates ~555 apart by ~1) taken in f64: the f64
        // intersection this replaces (two dot products, a division, the hit point, two more dot products) was the longest
        // dependent chain of every diffuse bounce
        const APrim& a = s.light_aprims[l.aprim];
        const f3 n((float)a.n[0], (float)a.n[1], (float)a.n[2]);
        const f3 b((float)(a.c[0] - origin.x), (float)(a.c[1] - origin.y), (float)(a.c[2] - origin.z)); // corner - origin
        const float dn = dot(n, dir);
        const float dd = (a.flags & APRIM_BACKFACE) ? fabsf(dn) : -dn;
        if (dd < 1.1920929e-7f) return 0.f;
        const float t = rt_divf(dot(n, b), dn);
        if (!(t > 0.001f)) return 0.f;
        const f3 local = dir * t - b; // hit point - corner
        const float u = dot(local, f3((float)a.iu[0], (float)a.iu[1], (float)a.iu[2])), v = dot(local, f3((float)a.iv[0], (float)a.iv[1], (float)a.iv[2]));
        if (u < 0.f || u > 1.f || v < 0.f || v > 1.f) return 0.f;
        const float len2 = dot(dir, dir);
        const float cosine = fabsf(dn) * rt_rsqrtf(len2);
        return rt_divf(t * t * len2, cosine * (float)a.area);
    }
    case LIGHT_SPHERE: { // sphere.rs:106-121: 1 / (2 pi (1 - cos_theta_max)) when the ray hits the sphere, else 0
        const APrim& a = s.light_aprims[l.aprim];
        const f3 oc((float)(origin.x - a.c[0]), (float)(origin.y - a.c[1]), (float)(origin.z - a.c[2]));
        const float r = (float)a.c[3];
        // hit test in the centre-to-ray form of sphere_hit_fast (no cancellation between |oc|^2 and r^2)
        const float inv_a = rt_divf(1.0f, dot(dir, dir));
        const float k = dot(dir, oc) * inv_a;
        const f3 lv = oc - dir * k;
        const float dp = r * r - dot(lv, lv);
        if (dp < 0.f) return 0.f;
        const float sq = rt_sqrtf(dp * inv_a);
        if (!(-k - sq > 0.001f) && !(-k + sq > 0.001f)) return 0.f;
        // 1 - cos_theta_max = q / (1 + sqrt(1 - q)), q = r^2 / |c|^2: stable for a small, far light
        const float q = rt_divf(r * r, dot(oc, oc));
        const float one_minus_cos = rt_divf(q, 1.0f + rt_sqrtf(fmaxf(0.f, 1.0f - q)));
        return rt_divf(1.0f, 2.0f * RT_PI_F * one_minus_cos);
    }
    case LIGHT_SKY: return 1.0f / (4.0f * RT_PI_F); // sky.rs:61-63
    case LIGHT_SUN: return 1.0f;                     // sun.rs:70-72
    default: return 0.f;
    }
}

template <bool SPH = true>
RT_HD float lights_pdf_value(const DeviceScene& s, d3 origin, f3 dir) {
    float sum = 0.f;
    for (int i = 0; i < s.n_lights; i++) sum += s.lights[i].weight * light_pdf_value<SPH>(s, s.lights[i], origin, dir);
    return sum;
}

// `fixed` (RTGPU_FLAG_FIXED, `--fixed`): the two samplers whose pdf_value does not match what they sample are corrected —
// Plane::random covers the whole quad (the reference covers one quadrant, Q1), Sphere::random builds its cone around the
// NORMALISED centre direction (the reference's basis is sheared by the distance, Q2). Off by default: parity is defined
// against the reference.
template <bool SPH = true, typename RNG> RT_HD f3 lights_random(const DeviceScene& s, d3 origin, RNG& rng, bool fixed = false) {
    if (s.n_lights == 0) return f3(1.f, 0.f, 0.f);
    int idx = 0;
    if (s.n_lights > 1) { // gen_range(0..size) at each list level == one weighted pick over the leaves
        const float u = rng.uniform();
        float acc = 0.f;
        idx = s.n_lights - 1;
        for (int i = 0; i < s.n_lights; i++) { acc += s.lights[i].weight; if (u < acc) { idx = i; break; } }
    } else if (s.lights[0].kind != LIGHT_NULL || s.lights[0].weight < 1.0f) {
        (void)rng.uniform(); // a one-element list still draws its index (list.rs:97)
    }
    const Light& l = s.lights[idx];
    if (!SPH && l.kind == LIGHT_SPHERE) return f3(1.f, 0.f, 0.f); // (no such light in the scene)
    switch (l.kind) {
    case LIGHT_PLANE: { // plane.rs:120-126: one quadrant only
        const APrim& a = s.light_aprims[l.aprim];
        const double k = fixed ? 2.0 : 1.0; // a.u, a.v are HALF extents (plane.rs:41-51)
        const double ru = k * (double)rng.uniform(), rv = k * (double)rng.uniform();
        return f3((float)(a.c[0] + a.u[0] * ru + a.v[0] * rv - origin.x), (float)(a.c[1] + a.u[1] * ru + a.v[1] * rv - origin.y),
                  (float)(a.c[2] + a.u[2] * ru + a.v[2] * rv - origin.z));
    }
    case LIGHT_SPHERE: { // sphere.rs:123-145: basis from the UNNORMALISED direction
        const APrim& a = s.light_aprims[l.aprim];
        const d3 dir(a.c[0] - origin.x, a.c[1] - origin.y, a.c[2] - origin.z);
        const double cos_theta_max = sqrt(1.0 - a.c[3] * a.c[3] / dot(dir, dir));
        const float r1 = rng.uniform(), r2 = rng.uniform();
        const float phi = r1 * 2.0f * RT_PI_F;
        const float z = 1.0f + r2 * ((float)cos_theta_max - 1.0f);
        const float sxy = rt_sqrtf(fmaxf(0.f, 1.0f - z * z));
        float sn, cs;
        rt_sincosf(phi, sn, cs);
        return onb_apply(fixed ? normalize(f3(dir)) : f3(dir), f3(cs * sxy, sn * sxy, z));
    }
    case LIGHT_SKY: return random_unit_f(rng); // sky.rs:65-67
    case LIGHT_SUN: return f3((float)l.dir[0], (float)l.dir[1], (float)l.dir[2]); // sun.rs:74-76
    default: return f3(1.f, 0.f, 0.f);
    }
}

// glossy.rs:35-50 / normal_debug.rs:25-40
RT_HD f3 mapped_normal(const DeviceScene& s, int normal_map, const Surface<float>& h, f3 pos) {
    if (normal_map < 0) return h.normal;
    const F4 c = tex_eval(s, normal_map, h.u, h.v, pos);
    const f3 sm(c.x - 0.5f, c.y - 0.5f, c.z - 0.5f);
    return normalize(h.tangent * sm.x + h.bitangent * sm.y + h.normal * sm.z);
}

RT_HD float cosine_scattering_pdf(f3 normal, f3 dir) { // lambertian.rs:37-45, glossy.rs:86-95
    const float c = dot(normal, normalize(dir));
    return c < 0.f ? 0.f : c * (1.0f / RT_PI_F);
}

// The diffuse branch shared by Lambertian / Glossy / Isotropic: MixPDF generate + value +
// scattering_pdf, guarded like the oracle.
template <bool SPH = true, typename RNG>
RT_HD void scatter_with_pdf(const DeviceScene& s, const Surface<float>& h, f3 w, bool uniform_sphere, f3 attenuation,
                            float light_bias, RNG& rng, Bounce& b, bool fixed = false) {
    f3 dir;
    if (rng.uniform() < light_bias) dir = lights_random<SPH>(s, h.pos, rng, fixed);        // mix.rs:30-36
    else dir = uniform_sphere ? random_unit_f(rng) : onb_apply(w, random_cosine_f(rng));
    const float mat_pdf = uniform_sphere ? 1.0f / (4.0f * RT_PI_F) : fmaxf(dot(normalize(dir), w) * (1.0f / RT_PI_F), 0.f); // cosine.rs:26-29
    const float pdf = mat_pdf * (1.0f - light_bias) + lights_pdf_value<SPH>(s, h.pos, dir) * light_bias;               // mix.rs:23-28
    const float spdf = uniform_sphere ? 1.0f / (4.0f * RT_PI_F) : cosine_scattering_pdf(w, dir);
    b.pdf = pdf; b.spdf = spdf; b.dir = dir; b.origin = h.pos;
    const float k = rt_divf(spdf, pdf);
    if (!(pdf > 0.f) || !isfinite(k)) { b.kind = BOUNCE_EMISSIVE; return; } // path ends with its emission only
    b.kind = BOUNCE_PDF;
    b.weight = attenuation * k;
}

// Material::emit + Material::scatter for the hit `h` reached by `ray` (camera.rs:295-328).
// CLASS >= 0: the caller guarantees the material class (a k_shade<CLASS> queue), so the other
// materials' code — and its registers — is compiled out; CLASS < 0: any material (parity hook).
#define RT_ONLY(C) if (CLASS >= 0 && CLASS != (C)) break;
template <int CLASS, bool NMAP = true, bool CTEX = false, bool SPH = true, typename RNG>
RT_HD void shade_bounce(const DeviceScene& s, const RayD& ray, const Surface<float>& h, float light_bias, RNG& rng, Bounce& b, bool fixed = false) {
    b.emission = f3(0.f, 0.f, 0.f); b.weight = f3(0.f, 0.f, 0.f); b.pdf = 0.f; b.spdf = 0.f;
    b.origin = h.pos; b.dir = f3(1.f, 0.f, 0.f);
    const f3 pos((float)h.pos.x, (float)h.pos.y, (float)h.pos.z);
    const f3 rd((float)ray.d.x, (float)ray.d.y, (float)ray.d.z);
    if ((CLASS < 0 || CLASS == CLASS_EMISSIVE) && h.material < 0) { // Sky / Sun: private Emissive (emissive.rs:28-34)
        const F4 e = tex_eval<CTEX>(s, h.texture, h.u, h.v, pos);
        b.emission = h.front_face ? f3(e.x, e.y, e.z) : f3(0.f, 0.f, 0.f);
        b.kind = BOUNCE_EMISSIVE;
        return;
    }
    const Material m = s.materials[h.material];
    switch (m.kind) {
    case MAT_LAMBERTIAN: { RT_ONLY(CLASS_LAMBERTIAN) // lambertian.rs:24-35
        const F4 a = tex_eval<CTEX>(s, m.albedo, h.u, h.v, pos);
        scatter_with_pdf<SPH>(s, h, h.normal, false, f3(a.x, a.y, a.z), light_bias, rng, b, fixed);
        return;
    }
    case MAT_METAL: { RT_ONLY(CLASS_METAL) // metal.rs:27-45
        const f3 reflected = reflect(rd, h.normal);
        const f3 ru = random_unit_f(rng);
        const float rough = tex_eval<CTEX>(s, m.roughness, h.u, h.v, pos).x;
        const f3 dir = reflected + ru * rough * length(reflected);
        if (dot(dir, h.normal) > 0.f) {
            const F4 a = tex_eval<CTEX>(s, m.albedo, h.u, h.v, pos);
            b.kind = BOUNCE_RAY; b.weight = f3(a.x, a.y, a.z); b.dir = dir;
        } else b.kind = BOUNCE_ABSORBED;
        return;
    }
    case MAT_DIELECTRIC: { RT_ONLY(CLASS_DIELECTRIC) // dielectric.rs:28-54
        const float ior_ratio = h.front_face ? 1.0f / m.ior : m.ior;
        const f3 unit = normalize(rd);
        const float cos_theta = fminf(1.0f, dot(-unit, h.normal));
        const float sin_theta = sqrtf(1.0f - cos_theta * cos_theta);
        const bool tir = ior_ratio * sin_theta > 1.0f;
        const bool reflected = tir || reflectance_f(cos_theta, ior_ratio) > rng.uniform();
        f3 dir;
        if (reflected) dir = reflect(unit, h.normal);
        else { // vec4.rs:140-147
            const float ct = fminf(1.0f, dot(-unit, h.normal));
            const f3 perp = (unit + h.normal * ct) * ior_ratio;
            const f3 par = h.normal * -sqrtf(1.0f - dot(perp, perp));
            dir = perp + par;
        }
        b.kind = BOUNCE_RAY; b.weight = f3(1.f, 1.f, 1.f); b.dir = dir;
        return;
    }
    case MAT_GLOSSY: { RT_ONLY(CLASS_GLOSSY) // glossy.rs:52-84
        const f3 n = NMAP ? mapped_normal(s, m.normal_map, h, pos) : h.normal;
        const f3 unit = normalize(rd);
        const float cos_theta = fminf(1.0f, dot(-unit, n));
        const bool specular = reflectance_f(cos_theta, m.inv_ior) > rng.uniform();
        if (specular) {
            const float rough = tex_eval<CTEX>(s, m.roughness, h.u, h.v, pos).x;
            const f3 reflected = reflect(rd, n);
            const f3 dir = reflected + random_unit_f(rng) * rough * length(reflected);
            if (dot(dir, n) > 0.f) { b.kind = BOUNCE_RAY; b.weight = f3(1.f, 1.f, 1.f); b.dir = dir; }
            else b.kind = BOUNCE_ABSORBED;
        } else {
            const F4 a = tex_eval<CTEX>(s, m.albedo, h.u, h.v, pos);
            scatter_with_pdf<SPH>(s, h, n, false, f3(a.x, a.y, a.z), light_bias, rng, b, fixed);
        }
        return;
    }
    case MAT_EMISSIVE: { RT_ONLY(CLASS_EMISSIVE) // emissive.rs:24-34
        if (h.front_face) { const F4 e = tex_eval<CTEX>(s, m.albedo, h.u, h.v, pos); b.emission = f3(e.x, e.y, e.z); }
        b.kind = BOUNCE_EMISSIVE;
        return;
    }
    case MAT_ISOTROPIC: { RT_ONLY(CLASS_ISOTROPIC) // isotropic.rs:24-36
        const F4 a = tex_eval<CTEX>(s, m.albedo, h.u, h.v, pos);
        scatter_with_pdf<SPH>(s, h, h.normal, true, f3(a.x, a.y, a.z), light_bias, rng, b, fixed);
        return;
    }
    case MAT_NORMAL_DEBUG: { RT_ONLY(CLASS_EMISSIVE) // normal_debug.rs:43-49
        const f3 n = NMAP ? mapped_normal(s, m.normal_map, h, pos) : h.normal;
        b.emission = n * 0.5f + f3(0.5f, 0.5f, 0.5f);
        b.kind = BOUNCE_EMISSIVE;
        return;
    }
    default: break;
    }
    b.kind = BOUNCE_ABSORBED;
}
#undef RT_ONLY

} // namespace rt
