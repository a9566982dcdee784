// oracle_shading.cpp — TEST INFRASTRUCTURE (CPU oracle). Not part of the product path.
//
// Textures (src/texture/*.rs), Perlin noise (src/noise/perlin.rs) and materials
// (src/material/*.rs) restated in f64 with the reference's operation and RNG-draw order.
#include <string>

#include "oracle_scene.h"

namespace oracle {

// ---- noise/perlin.rs ------------------------------------------------------------------
static double trilinear_interpolation(const Vec4 c[8], const Vec4& uvw) { // perlin.rs:54-73
    double acc = 0.0;
    double u = uvw[0], v = uvw[1], w = uvw[2];
    double uu = u * u * (3.0 - 2.0 * u), vv = v * v * (3.0 - 2.0 * v), ww = w * w * (3.0 - 2.0 * w);
    for (int i = 0; i < 2; i++)
        for (int j = 0; j < 2; j++)
            for (int k = 0; k < 2; k++) {
                double fi = i, fj = j, fk = k;
                Vec4 v_weight = Vec4::vec(u - fi, v - fj, w - fk);
                acc += (fi * uu + (1.0 - fi) * (1.0 - uu)) * (fj * vv + (1.0 - fj) * (1.0 - vv)) *
                       (fk * ww + (1.0 - fk) * (1.0 - ww)) * c[(i << 2) + (j << 1) + k].dot(v_weight);
            }
    return acc;
}

static int32_t floor_as_i32(double x) { // `x.floor() as i32` saturates, NaN -> 0
    double f = std::floor(x);
    if (f != f) return 0;
    if (f >= 2147483647.0) return 2147483647;
    if (f <= -2147483648.0) return (int32_t)-2147483647 - 1;
    return (int32_t)f;
}

double PerlinNoise::sample(const Vec4& p) const { // perlin.rs:79-99
    Vec4 uvw(p[0] - std::floor(p[0]), p[1] - std::floor(p[1]), p[2] - std::floor(p[2]), p[3] - std::floor(p[3]));
    int32_t i = floor_as_i32(p.x()), j = floor_as_i32(p.y()), k = floor_as_i32(p.z());
    Vec4 c[8];
    for (int di = 0; di < 2; di++)
        for (int dj = 0; dj < 2; dj++)
            for (int dk = 0; dk < 2; dk++)
                c[(di << 2) + (dj << 1) + dk] =
                    random_vec[perm_x[(uint32_t)(di + i) & 255] ^ perm_y[(uint32_t)(dj + j) & 255] ^ perm_z[(uint32_t)(dk + k) & 255]];
    return trilinear_interpolation(c, uvw);
}

double PerlinNoise::sample_turbulence(const Vec4& p0, int samples) const { // perlin.rs:101-113
    double acc = 0.0, weight = 1.0;
    Vec4 p = p0;
    for (int s = 0; s < samples; s++) {
        acc += weight * sample(p);
        weight *= 0.5;
        p *= 2.0;
    }
    return std::fabs(acc);
}

// ---- textures ---------------------------------------------------------------------------
namespace {

struct ConstantTex : Sampler { // constant.rs
    Vec4 value;
    Vec4 sample(double, double, const Vec4&) const override { return value; }
};

static uint32_t f64_as_u32(double x) { // Rust `as u32`: saturating, NaN -> 0
    if (!(x > 0.0)) return 0;
    if (x >= 4294967295.0) return 4294967295u;
    return (uint32_t)x;
}

struct CheckerTex : Sampler { // checkerboard.rs:34-44
    const Sampler *even, *odd;
    double scale;
    Vec4 sample(double u, double v, const Vec4& p) const override {
        uint32_t iu = f64_as_u32(u * 2.0 / scale), iv = f64_as_u32(v * 2.0 / scale);
        bool is_even = ((uint64_t)iu + iv) % 2 == 0; // (debug builds would panic on u32 overflow; release wraps — parity is the same for the sum's low bit)
        return is_even ? even->sample(u, v, p) : odd->sample(u, v, p);
    }
};

struct CheckerSolidTex : Sampler { // checkerboard.rs:74-86
    const Sampler *even, *odd;
    double scale;
    Vec4 sample(double u, double v, const Vec4& p) const override {
        int32_t ix = floor_as_i32(p.x() / scale), iy = floor_as_i32(p.y() / scale), iz = floor_as_i32(p.z() / scale);
        int32_t sum = (int32_t)((uint32_t)ix + (uint32_t)iy + (uint32_t)iz); // wrapping add (release profile)
        bool is_even = sum % 2 == 0; // Rust % keeps the sign: -1 % 2 == -1 => odd
        return is_even ? even->sample(u, v, p) : odd->sample(u, v, p);
    }
};

struct ImageTex : Sampler { // image.rs:40-53 + buffer.rs:30-60
    int width = 0, height = 0;
    bool repeat = true;
    std::vector<float> rgb; // into_rgb32f: f32 texels, widened to f64 at use
    Vec4 sample(double u, double v, const Vec4&) const override {
        if (repeat) { u = u - std::floor(u); v = v - std::floor(v); }
        else { u = u < 0.0 ? 0.0 : (u > 1.0 ? 1.0 : u); v = v < 0.0 ? 0.0 : (v > 1.0 ? 1.0 : v); }
        double w = (double)width - 0.001, h = (double)height - 0.001;
        double fx = u * w, fy = v * h;
        size_t x = !(fx > 0.0) ? 0 : (size_t)fx, y = !(fy > 0.0) ? 0 : (size_t)fy; // `as usize` saturates
        if (x >= (size_t)width) x = width - 1;   // (the reference would index out of bounds and panic)
        if (y >= (size_t)height) y = height - 1;
        const float* px = &rgb[(y * width + x) * 3];
        return Vec4::vec((double)px[0], (double)px[1], (double)px[2]);
    }
};

struct LerpTex : Sampler { // interpolate.rs:29-39
    const Sampler *start, *end, *t;
    bool is_float;
    Vec4 sample(double u, double v, const Vec4& p) const override {
        double tt = t->sample(u, v, p)[0];
        if (tt == 0.0) return start->sample(u, v, p);
        if (tt == 1.0) return end->sample(u, v, p);
        return start->sample(u, v, p) * (1.0 - tt) + end->sample(u, v, p) * tt;
    }
};

struct NoiseSolidTex : Sampler { // texture/noise.rs:20-38
    const PerlinNoise* noise;
    Vec4 scale;
    int samples;
    Vec4 sample(double, double, const Vec4& p) const override {
        Vec4 p_scaled = p * scale;
        double sampled = noise->sample_turbulence(p_scaled, samples);
        double r = 0.5 * (1.0 + std::sin(p_scaled.z() + 10.0 * sampled));
        return Vec4(r, 0, 0, 0);
    }
};

struct ChannelTex : Sampler { // channel.rs
    const Sampler* color;
    int channel;
    Vec4 sample(double u, double v, const Vec4& p) const override { return Vec4(color->sample(u, v, p)[channel], 0, 0, 0); }
};

struct UvDebugTex : Sampler { // uv_debug.rs
    Vec4 sample(double u, double v, const Vec4&) const override { return Vec4::vec(u, v, 0.5); }
};

// ---- materials --------------------------------------------------------------------------
static double cosine_scattering_pdf(const Vec4& normal, const Ray& scattered) { // lambertian.rs:37-45, glossy.rs:86-95
    double cos_theta = normal.dot(scattered.dir.to_unit());
    return cos_theta < 0.0 ? 0.0 : cos_theta / PI;
}

static Vec4 mapped_normal(const Sampler* normal_map, const HitRecord& hit) { // glossy.rs:35-50, normal_debug.rs:25-40
    if (!normal_map) return hit.normal;
    Vec4 sampled = normal_map->sample(hit.u, hit.v, hit.hit_pos);
    Mat4 basis = Mat4::from_columns(hit.tangent, hit.bitangent, hit.normal, Vec4(0.0, 0.0, 0.0, 1.0));
    return (basis * (sampled - Vec4::vec(0.5, 0.5, 0.5))).to_unit();
}

struct Lambertian : Material { // lambertian.rs
    const Sampler* albedo;
    ScatterResult scatter(const Ray&, const HitRecord& hit, Rng&) const override {
        ScatterResult r; r.kind = ScatterKind::WithPDF; r.pdf = PdfKind::Cosine; r.pdf_w = hit.normal;
        r.attenuation = albedo->sample(hit.u, hit.v, hit.hit_pos);
        return r;
    }
    double scattering_pdf(const Ray&, const Ray& scattered, const HitRecord& hit) const override { return cosine_scattering_pdf(hit.normal, scattered); }
};

struct Metal : Material { // metal.rs:27-49
    const Sampler *albedo, *roughness;
    ScatterResult scatter(const Ray& ray, const HitRecord& hit, Rng& rng) const override {
        Vec4 reflected = ray.dir.reflect(hit.normal);
        Vec4 scatter_dir = reflected + random_unit(rng) * roughness->sample(hit.u, hit.v, hit.hit_pos)[0] * reflected.length();
        ScatterResult r;
        if (scatter_dir.dot(hit.normal) > 0.0) {
            r.kind = ScatterKind::WithRay; r.scattered = Ray(hit.hit_pos, scatter_dir);
            r.attenuation = albedo->sample(hit.u, hit.v, hit.hit_pos);
        } else r.kind = ScatterKind::Absorbed;
        return r;
    }
    double scattering_pdf(const Ray&, const Ray&, const HitRecord&) const override { return 1.0; }
};

struct Dielectric : Material { // dielectric.rs:28-59
    double ior;
    ScatterResult scatter(const Ray& ray, const HitRecord& hit, Rng& rng) const override {
        double ior_ratio = hit.front_face ? 1.0 / ior : ior;
        Vec4 unit_dir = ray.dir.to_unit();
        double cos_theta = std::fmin(1.0, (-unit_dir).dot(hit.normal));
        double sin_theta = std::sqrt(1.0 - cos_theta * cos_theta);
        bool tir = ior_ratio * sin_theta > 1.0;
        bool reflected = tir || reflectance(cos_theta, ior_ratio) > rng.draw_uniform(); // short-circuit: no draw on TIR
        Vec4 scatter_dir = reflected ? unit_dir.reflect(hit.normal) : unit_dir.refract(hit.normal, ior_ratio);
        ScatterResult r; r.kind = ScatterKind::WithRay; r.scattered = Ray(hit.hit_pos, scatter_dir);
        r.attenuation = Vec4::vec(1.0, 1.0, 1.0);
        return r;
    }
    double scattering_pdf(const Ray&, const Ray&, const HitRecord&) const override { return 1.0; }
};

struct Glossy : Material { // glossy.rs:52-96
    const Sampler *albedo, *roughness, *normal_map = nullptr;
    double inv_ior;
    ScatterResult scatter(const Ray& ray, const HitRecord& hit, Rng& rng) const override {
        Vec4 normal = mapped_normal(normal_map, hit);
        Vec4 unit_dir = ray.dir.to_unit();
        double cos_theta = std::fmin(1.0, (-unit_dir).dot(normal));
        bool specular = reflectance(cos_theta, inv_ior) > rng.draw_uniform();
        ScatterResult r;
        if (specular) {
            double rough = roughness->sample(hit.u, hit.v, hit.hit_pos)[0];
            Vec4 reflected = ray.dir.reflect(normal);
            Vec4 scatter_dir = reflected + random_unit(rng) * rough * reflected.length();
            if (scatter_dir.dot(normal) > 0.0) {
                r.kind = ScatterKind::WithRay; r.scattered = Ray(hit.hit_pos, scatter_dir);
                r.attenuation = Vec4::vec(1.0, 1.0, 1.0);
            } else r.kind = ScatterKind::Absorbed;
        } else {
            r.kind = ScatterKind::WithPDF; r.pdf = PdfKind::Cosine; r.pdf_w = normal;
            r.attenuation = albedo->sample(hit.u, hit.v, hit.hit_pos);
        }
        return r;
    }
    double scattering_pdf(const Ray&, const Ray& scattered, const HitRecord& hit) const override {
        return cosine_scattering_pdf(mapped_normal(normal_map, hit), scattered);
    }
};

struct Emissive : Material { // emissive.rs
    const Sampler* emission_map;
    ScatterResult scatter(const Ray&, const HitRecord&, Rng&) const override { ScatterResult r; r.kind = ScatterKind::Emissive; return r; }
    Vec4 emit(const HitRecord& hit) const override {
        return hit.front_face ? emission_map->sample(hit.u, hit.v, hit.hit_pos) : Vec4::vec(0.0, 0.0, 0.0);
    }
    double scattering_pdf(const Ray&, const Ray&, const HitRecord&) const override { return 1.0; }
};

struct Isotropic : Material { // isotropic.rs
    const Sampler* albedo;
    ScatterResult scatter(const Ray&, const HitRecord& hit, Rng&) const override {
        ScatterResult r; r.kind = ScatterKind::WithPDF; r.pdf = PdfKind::Uniform;
        r.attenuation = albedo->sample(hit.u, hit.v, hit.hit_pos);
        return r;
    }
    double scattering_pdf(const Ray&, const Ray&, const HitRecord&) const override { return 1.0 / (4.0 * PI); }
};

struct NormalDebug : Material { // normal_debug.rs
    const Sampler* normal_map = nullptr;
    ScatterResult scatter(const Ray&, const HitRecord&, Rng&) const override { ScatterResult r; r.kind = ScatterKind::Emissive; return r; }
    Vec4 emit(const HitRecord& hit) const override { return mapped_normal(normal_map, hit) * 0.5 + Vec4::vec(0.5, 0.5, 0.5); }
    double scattering_pdf(const Ray&, const Ray&, const HitRecord&) const override { return 1.0; }
};

} // namespace

// Sky / Sun own a private Emissive (sky.rs:19-24, sun.rs:22-30)
Material* make_emissive_for(const Sampler* tex) { auto* e = new Emissive(); e->emission_map = tex; return e; }

bool build_shading(Scene& s, const RtgpuSceneDesc& d) {
    for (int i = 0; i < d.n_noises; i++) {
        auto n = std::make_unique<PerlinNoise>();
        for (int k = 0; k < 256; k++) {
            n->random_vec[k] = Vec4::vec(d.noises[i].gradients[k][0], d.noises[i].gradients[k][1], d.noises[i].gradients[k][2]);
            n->perm_x[k] = d.noises[i].perm_x[k]; n->perm_y[k] = d.noises[i].perm_y[k]; n->perm_z[k] = d.noises[i].perm_z[k];
        }
        s.noises.push_back(std::move(n));
    }
    // textures reference only earlier-or-later indices; build in two passes (allocate, then link)
    s.textures.resize(d.n_textures);
    auto ok = [&](int idx) { return idx >= 0 && idx < d.n_textures; };
    for (int i = 0; i < d.n_textures; i++) {
        const RtgpuTexture& t = d.textures[i];
        switch (t.kind) {
        case RTGPU_TEX_CONSTANT: { auto* c = new ConstantTex(); c->value = t.is_float ? Vec4(t.value[0], 0, 0, 0) : Vec4::vec(t.value[0], t.value[1], t.value[2]); s.textures[i].reset(c); break; }
        case RTGPU_TEX_CHECKER: s.textures[i].reset(new CheckerTex()); break;
        case RTGPU_TEX_CHECKER_SOLID: s.textures[i].reset(new CheckerSolidTex()); break;
        case RTGPU_TEX_IMAGE: {
            if (t.image < 0 || t.image >= d.n_images) { s.error = "texture: bad image index"; return false; }
            const RtgpuImage& im = d.images[t.image];
            auto* it = new ImageTex(); it->width = im.width; it->height = im.height; it->repeat = t.repeat != 0;
            size_t n = (size_t)im.width * im.height * 3;
            it->rgb.resize(n);
            for (size_t k = 0; k < n; k++) { // image crate into_rgb32f: value as f32 / max as f32
                if (im.format == RTGPU_IMG_RGB8) it->rgb[k] = (float)((const uint8_t*)im.data)[k] / 255.0f;
                else if (im.format == RTGPU_IMG_RGB16) it->rgb[k] = (float)((const uint16_t*)im.data)[k] / 65535.0f;
                else it->rgb[k] = ((const float*)im.data)[k];
            }
            s.textures[i].reset(it); break;
        }
        case RTGPU_TEX_LERP: s.textures[i].reset(new LerpTex()); break;
        case RTGPU_TEX_NOISE_SOLID: {
            if (t.noise < 0 || t.noise >= d.n_noises) { s.error = "texture: bad noise index"; return false; }
            auto* n = new NoiseSolidTex(); n->noise = s.noises[t.noise].get(); n->scale = Vec4::vec(t.scale, t.scale, t.scale); n->samples = t.samples;
            s.textures[i].reset(n); break;
        }
        case RTGPU_TEX_CHANNEL: s.textures[i].reset(new ChannelTex()); break;
        case RTGPU_TEX_UV_DEBUG: s.textures[i].reset(new UvDebugTex()); break;
        default: s.error = "texture: unknown kind"; return false;
        }
    }
    for (int i = 0; i < d.n_textures; i++) {
        const RtgpuTexture& t = d.textures[i];
        if (t.kind == RTGPU_TEX_CHECKER || t.kind == RTGPU_TEX_CHECKER_SOLID) {
            if (!ok(t.a) || !ok(t.b)) { s.error = "checker: bad child"; return false; }
            if (t.kind == RTGPU_TEX_CHECKER) { auto* c = (CheckerTex*)s.textures[i].get(); c->even = s.textures[t.a].get(); c->odd = s.textures[t.b].get(); c->scale = t.scale; }
            else { auto* c = (CheckerSolidTex*)s.textures[i].get(); c->even = s.textures[t.a].get(); c->odd = s.textures[t.b].get(); c->scale = t.scale; }
        } else if (t.kind == RTGPU_TEX_LERP) {
            if (!ok(t.a) || !ok(t.b) || !ok(t.c)) { s.error = "lerp: bad child"; return false; }
            auto* l = (LerpTex*)s.textures[i].get(); l->start = s.textures[t.a].get(); l->end = s.textures[t.b].get(); l->t = s.textures[t.c].get(); l->is_float = t.is_float;
        } else if (t.kind == RTGPU_TEX_CHANNEL) {
            if (!ok(t.a) || t.channel < 0 || t.channel > 3) { s.error = "channel: bad child"; return false; }
            auto* c = (ChannelTex*)s.textures[i].get(); c->color = s.textures[t.a].get(); c->channel = t.channel;
        }
    }
    for (int i = 0; i < d.n_materials; i++) {
        const RtgpuMaterial& m = d.materials[i];
        auto tex = [&](int idx) -> const Sampler* { return ok(idx) ? s.textures[idx].get() : nullptr; };
        Material* out = nullptr;
        switch (m.kind) {
        case RTGPU_MAT_LAMBERTIAN: { auto* x = new Lambertian(); x->albedo = tex(m.albedo); if (!x->albedo) { delete x; break; } out = x; break; }
        case RTGPU_MAT_METAL: { auto* x = new Metal(); x->albedo = tex(m.albedo); x->roughness = tex(m.roughness); if (!x->albedo || !x->roughness) { delete x; break; } out = x; break; }
        case RTGPU_MAT_DIELECTRIC: { auto* x = new Dielectric(); x->ior = m.ior; out = x; break; }
        case RTGPU_MAT_GLOSSY: { auto* x = new Glossy(); x->albedo = tex(m.albedo); x->roughness = tex(m.roughness); x->normal_map = tex(m.normal_map); x->inv_ior = 1.0 / m.ior; if (!x->albedo || !x->roughness) { delete x; break; } out = x; break; }
        case RTGPU_MAT_EMISSIVE: { auto* x = new Emissive(); x->emission_map = tex(m.albedo); if (!x->emission_map) { delete x; break; } out = x; break; }
        case RTGPU_MAT_ISOTROPIC: { auto* x = new Isotropic(); x->albedo = tex(m.albedo); if (!x->albedo) { delete x; break; } out = x; break; }
        case RTGPU_MAT_NORMAL_DEBUG: { auto* x = new NormalDebug(); x->normal_map = tex(m.normal_map); out = x; break; }
        }
        if (!out) { s.error = "material " + std::to_string(i) + ": bad kind or texture index"; return false; }
        s.materials.emplace_back(out);
    }
    return true;
}

double tex_sample(const Scene& s, int tex, double u, double v, const Vec4& p, Vec4* color) {
    Vec4 c = s.textures[tex]->sample(u, v, p);
    if (color) *color = c;
    return c[0];
}

} // namespace oracle
