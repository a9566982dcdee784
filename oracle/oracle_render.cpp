// oracle_render.cpp — TEST INFRASTRUCTURE (CPU oracle). Not part of the product path.
//
// Camera::render / get_ray / ray_color (src/camera.rs:189-349), MixPDF / HittablePDF /
// CosinePDF / UniformPDF (src/pdf/*.rs), Writer::save maths (src/output.rs,
// src/tonemapping/*.rs) in f64, plus the C entry points of oracle.h.
#include <atomic>
#include <cstring>
#include <string>
#include <thread>

#include "oracle.h"
#include "oracle_scene.h"

using namespace oracle;

struct OracleScene { Scene scene; };

static thread_local std::string g_err;
extern "C" const char* oracle_last_error(void) { return g_err.c_str(); }

namespace {

struct Camera { // the fields Camera::init leaves behind (camera.rs:19-44)
    int width, height;
    Vec4 position, first_pixel, delta_u, delta_v, basis_u, basis_v, background;
    bool has_aperture;
    double aperture_radius;
    int sqrt_spt, max_depth;
    double inv_sqrt_spt, light_bias;
    explicit Camera(const RtgpuCameraDesc& c, int k, int depth, double bias) {
        width = c.width; height = c.height;
        position = Vec4::point(c.position[0], c.position[1], c.position[2]);
        first_pixel = Vec4::point(c.first_pixel[0], c.first_pixel[1], c.first_pixel[2]);
        delta_u = Vec4::vec(c.pixel_delta_u[0], c.pixel_delta_u[1], c.pixel_delta_u[2]);
        delta_v = Vec4::vec(c.pixel_delta_v[0], c.pixel_delta_v[1], c.pixel_delta_v[2]);
        basis_u = Vec4::vec(c.basis_u[0], c.basis_u[1], c.basis_u[2]);
        basis_v = Vec4::vec(c.basis_v[0], c.basis_v[1], c.basis_v[2]);
        background = Vec4(c.background[0], c.background[1], c.background[2], 1.0); // config.rs:28,131 (point)
        has_aperture = c.aperture_radius >= 0.0; aperture_radius = c.aperture_radius;
        sqrt_spt = k; max_depth = depth; inv_sqrt_spt = 1.0 / (double)k; light_bias = bias;
    }
    Vec4 pixel_sample_square(int sx, int sy, Rng& rng) const { // camera.rs:334-341
        double rx = rng.draw_uniform(), ry = rng.draw_uniform();
        double x = ((double)sx + rx) * inv_sqrt_spt - 0.5;
        double y = ((double)sy + ry) * inv_sqrt_spt - 0.5;
        return delta_u * x + delta_v * y;
    }
    Vec4 defocus_disk_sample(Rng& rng) const { // camera.rs:345-349
        Vec4 v = random_in_unit_disk(rng);
        return position + (basis_u * v[0] + basis_v * v[1]) * aperture_radius;
    }
    Ray get_ray(int px, int py, int sx, int sy, Rng& rng) const { // camera.rs:260-280
        Vec4 pixel_center = first_pixel + (delta_u * (double)px) + (delta_v * (double)py);
        Vec4 pixel_sample = pixel_center + pixel_sample_square(sx, sy, rng);
        Vec4 ray_origin = has_aperture ? defocus_disk_sample(rng) : position;
        return Ray(ray_origin, pixel_sample - ray_origin);
    }
};

// MixPDF over (material pdf, HittablePDF(lights)) — pdf/mix.rs:23-36, pdf/hittable.rs
struct Mix {
    const ScatterResult& sr;
    const Hittable* lights;
    Vec4 origin;
    double mix;
    double material_value(const Vec4& dir) const {
        if (sr.pdf == PdfKind::Cosine) { // pdf/cosine.rs:26-29
            double cos_theta = dir.to_unit().dot(sr.pdf_w);
            return std::fmax(cos_theta / PI, 0.0);
        }
        return 1.0 / (4.0 * PI); // pdf/uniform.rs
    }
    Vec4 material_generate(Rng& rng) const {
        if (sr.pdf == PdfKind::Cosine) return onb_from_vec(sr.pdf_w) * random_cosine(rng); // pdf/cosine.rs:17-23,31-33
        return random_unit(rng);
    }
    Vec4 generate(Rng& rng) const { // mix.rs:30-36
        if (rng.draw_uniform() < mix) return lights->random(origin, rng);
        return material_generate(rng);
    }
    double value(const Vec4& dir, Rng& rng) const { // mix.rs:23-28
        double first = material_value(dir);
        double second = lights->pdf_value(origin, dir, rng);
        return first * (1.0 - mix) + second * mix;
    }
};

struct Tracer {
    const Scene& scene;
    const Camera& cam;
    uint64_t rays = 0;

    // ray_color (camera.rs:282-332), recursive like the reference
    Vec4 ray_color(const Ray& ray, int depth, Rng& rng) {
        if (depth == 0) return Vec4::vec(0.0, 0.0, 0.0);
        HitRecord hit;
        rays++;
        if (scene.world->test(ray, Interval{0.001, INF}, rng, hit)) {
            Vec4 from_emission = hit.material->emit(hit);
            ScatterResult sr = hit.material->scatter(ray, hit, rng);
            switch (sr.kind) {
            case ScatterKind::WithPDF: {
                Mix mix{sr, scene.lights, hit.hit_pos, cam.light_bias};
                Ray scattered(hit.hit_pos, mix.generate(rng));
                double pdf = mix.value(scattered.dir, rng);
                double scattering_pdf = hit.material->scattering_pdf(ray, scattered, hit);
                // Guarded estimator: the reference divides unguarded
                // (camera.rs:312) and can produce NaN/inf pixels; oracle and GPU both
                // terminate the path when the weight is not finite.
                if (!(pdf > 0.0) || !std::isfinite(scattering_pdf / pdf)) return from_emission;
                Vec4 scatter_color = ray_color(scattered, depth - 1, rng);
                Vec4 from_scatter = (scatter_color * sr.attenuation * scattering_pdf) / pdf;
                return from_emission + from_scatter;
            }
            case ScatterKind::WithRay: {
                Vec4 scatter_color = ray_color(sr.scattered, depth - 1, rng);
                return from_emission + scatter_color * sr.attenuation;
            }
            case ScatterKind::Absorbed: return Vec4::vec(0.0, 0.0, 0.0);
            case ScatterKind::Emissive: return from_emission;
            }
        }
        return cam.background; // camera.rs:331 (background_color is always Some, config.rs:28,41)
    }
};

static void fill_hit(const HitRecord& h, bool ok, RtgpuHit& o) {
    std::memset(&o, 0, sizeof o);
    if (!ok) { o.object = -1; o.instance = -1; return; }
    o.object = h.object; o.instance = h.instance; o.primitive = h.primitive; o.front_face = h.front_face ? 1 : 0;
    o.t = h.t;
    for (int i = 0; i < 3; i++) { o.pos[i] = h.hit_pos[i]; o.normal[i] = h.normal[i]; }
    o.uv[0] = h.u; o.uv[1] = h.v;
}

// output.rs:42-50
static double linear_to_srgb(double x) { return x < 0.0031308 ? x * 12.92 : std::pow(x, 1.0 / 2.4) * 1.055 - 0.055; }
static double clamp01(double x) { return x < 0.0 ? 0.0 : (x > 1.0 ? 1.0 : x); } // f64::clamp (NaN stays NaN)
static uint8_t as_u8(double x) { if (!(x > 0.0)) return 0; if (x >= 255.0) return 255; return (uint8_t)x; } // `as u8` saturates, NaN -> 0

} // namespace

extern "C" {

int oracle_scene_create(const RtgpuSceneDesc* desc, OracleScene** out) {
    if (!desc || !out) { g_err = "null argument"; return RTGPU_ERR_INVALID; }
    auto* s = new OracleScene();
    if (!s->scene.build(*desc)) { g_err = s->scene.error; delete s; return RTGPU_ERR_INVALID; }
    *out = s;
    return 0;
}
int oracle_scene_destroy(OracleScene* s) { delete s; return 0; }

int oracle_render_rows(const OracleScene* s, const RtgpuCameraDesc* camd, const RtgpuRenderParams* p, int n_threads,
                       int row_offset, int row_step, double* out_rgba, uint64_t* out_rays);

int oracle_render(const OracleScene* s, const RtgpuCameraDesc* camd, const RtgpuRenderParams* p, int n_threads,
                  double* out_rgba, uint64_t* out_rays) {
    return oracle_render_rows(s, camd, p, n_threads, 0, 1, out_rgba, out_rays);
}

// Bounded-sample variant for the timed CPU baseline: only rows y = row_offset (mod row_step)
// are rendered (the others stay untouched); everything else is Camera::render.
int oracle_render_rows(const OracleScene* s, const RtgpuCameraDesc* camd, const RtgpuRenderParams* p, int n_threads,
                       int row_offset, int row_step, double* out_rgba, uint64_t* out_rays) {
    if (row_step < 1 || row_offset < 0) { g_err = "bad row range"; return RTGPU_ERR_INVALID; }
    if (!s || !camd || !p || !out_rgba) { g_err = "null argument"; return RTGPU_ERR_INVALID; }
    if (p->replicas < 1 || p->sqrt_spt < 1) { g_err = "replicas and sqrt_spt must be >= 1"; return RTGPU_ERR_INVALID; }
    Camera cam(*camd, p->sqrt_spt, p->max_depth, p->light_bias);
    const int W = cam.width, H = cam.height, T = p->replicas, k = p->sqrt_spt;
    const int samples_per_pixel = T * k * k; // camera.rs:50-51
    // optional shard [sample_begin, sample_end) of the global sample index s = tid*k*k + sy*k + sx
    int s_begin = p->sample_begin, s_end = p->sample_end;
    if (s_begin == 0 && s_end == 0) s_end = samples_per_pixel;
    std::vector<std::vector<double>> thread_bufs(T);
    std::atomic<int> next_replica{0};
    std::atomic<uint64_t> total_rays{0};
    auto worker = [&]() {
        for (;;) {
            int tid = next_replica.fetch_add(1);
            if (tid >= T) break;
            std::vector<double>& buf = thread_bufs[tid];
            buf.assign((size_t)W * H * 4, 0.0);
            Rng rng(p->seed, (uint64_t)tid); // camera.rs:208-209 (OS-seeded there)
            Tracer tr{s->scene, cam};
            for (int y = row_offset; y < H; y += row_step)
                for (int x = 0; x < W; x++) {
                    Vec4 color = Vec4::vec(0.0, 0.0, 0.0);
                    for (int sy = 0; sy < k; sy++)
                        for (int sx = 0; sx < k; sx++) {
                            int gs = tid * k * k + sy * k + sx;
                            if (gs < s_begin || gs >= s_end) continue;
                            Ray ray = cam.get_ray(x, y, sx, sy, rng);
                            color += tr.ray_color(ray, cam.max_depth, rng);
                        }
                    color /= (double)samples_per_pixel; // camera.rs:229
                    double* px = &buf[((size_t)y * W + x) * 4];
                    for (int c = 0; c < 4; c++) px[c] = color[c];
                }
            total_rays += tr.rays;
        }
    };
    int nt = n_threads <= 0 ? T : std::min(n_threads, T);
    std::vector<std::thread> threads;
    for (int i = 0; i < nt; i++) threads.emplace_back(worker);
    for (auto& t : threads) t.join();
    for (int tid = 0; tid < T; tid++) // camera.rs:243-255: buffers are added in thread order
        for (size_t i = 0; i < (size_t)W * H * 4; i++) out_rgba[i] = out_rgba[i] + thread_bufs[tid][i];
    if (out_rays) *out_rays = total_rays.load();
    return 0;
}

int oracle_trace_primary(const OracleScene* s, const double* rays, int64_t n, RtgpuHit* out) {
    if (!s || !rays || !out) { g_err = "null argument"; return RTGPU_ERR_INVALID; }
    unsigned nt = std::max(1u, std::thread::hardware_concurrency());
    std::vector<std::thread> threads;
    std::atomic<int64_t> next{0};
    for (unsigned t = 0; t < nt; t++)
        threads.emplace_back([&]() {
            Rng rng(0, 0);
            for (;;) {
                int64_t b = next.fetch_add(4096);
                if (b >= n) break;
                for (int64_t i = b; i < std::min(n, b + 4096); i++) {
                    const double* r = rays + i * 6;
                    Ray ray(Vec4::point(r[0], r[1], r[2]), Vec4::vec(r[3], r[4], r[5]));
                    HitRecord h;
                    bool ok = s->scene.world->test(ray, Interval{0.001, INF}, rng, h);
                    fill_hit(h, ok, out[i]);
                }
            }
        });
    for (auto& t : threads) t.join();
    return 0;
}

int oracle_generate_rays(const RtgpuCameraDesc* camd, int sqrt_spt, const int32_t* px, const int32_t* sxy,
                         const double* xi, int64_t n, double* out_rays) {
    if (!camd || !px || !sxy || !xi || !out_rays || sqrt_spt < 1) { g_err = "bad argument"; return RTGPU_ERR_INVALID; }
    Camera cam(*camd, sqrt_spt, 1, 0.0);
    for (int64_t i = 0; i < n; i++) {
        Rng rng(0, 0);
        rng.inject = xi + i * 4; rng.inject_n = 4; rng.inject_pos = 0;
        Ray r = cam.get_ray(px[i * 2], px[i * 2 + 1], sxy[i * 2], sxy[i * 2 + 1], rng);
        for (int c = 0; c < 3; c++) { out_rays[i * 6 + c] = r.origin[c]; out_rays[i * 6 + 3 + c] = r.dir[c]; }
    }
    return 0;
}

int oracle_bounce(const OracleScene* s, double light_bias, const double* rays, const double* xi, int n_xi,
                  int64_t n, OracleBounce* out) {
    if (!s || !rays || !xi || !out) { g_err = "null argument"; return RTGPU_ERR_INVALID; }
    for (int64_t i = 0; i < n; i++) {
        OracleBounce& o = out[i];
        std::memset(&o, 0, sizeof o);
        const double* r = rays + i * 6;
        Ray ray(Vec4::point(r[0], r[1], r[2]), Vec4::vec(r[3], r[4], r[5]));
        Rng rng(0, 0);
        rng.inject = xi + i * n_xi; rng.inject_n = n_xi; rng.inject_pos = 0;
        HitRecord hit;
        if (!s->scene.world->test(ray, Interval{0.001, INF}, rng, hit)) { o.kind = 0; continue; }
        Vec4 em = hit.material->emit(hit);
        for (int c = 0; c < 3; c++) o.emission[c] = em[c];
        ScatterResult sr = hit.material->scatter(ray, hit, rng);
        if (sr.kind == ScatterKind::WithPDF) {
            Mix mix{sr, s->scene.lights, hit.hit_pos, light_bias};
            Ray scattered(hit.hit_pos, mix.generate(rng));
            double pdf = mix.value(scattered.dir, rng);
            double sp = hit.material->scattering_pdf(ray, scattered, hit);
            o.kind = 1; o.pdf = pdf; o.scattering_pdf = sp;
            for (int c = 0; c < 3; c++) { o.weight[c] = sr.attenuation[c] * sp / pdf; o.next_origin[c] = scattered.origin[c]; o.next_dir[c] = scattered.dir[c]; }
        } else if (sr.kind == ScatterKind::WithRay) {
            o.kind = 2;
            for (int c = 0; c < 3; c++) { o.weight[c] = sr.attenuation[c]; o.next_origin[c] = sr.scattered.origin[c]; o.next_dir[c] = sr.scattered.dir[c]; }
        } else if (sr.kind == ScatterKind::Absorbed) { o.kind = 3; for (int c = 0; c < 3; c++) o.emission[c] = 0.0; }
        else o.kind = 4;
    }
    return 0;
}

int oracle_tonemap_srgb8(const double* rgba, int width, int height, int tonemap, uint8_t* out_rgb) {
    if (!rgba || !out_rgb) { g_err = "null argument"; return RTGPU_ERR_INVALID; }
    // tonemapping/aces.rs:5-18 (row 3 unused for xyz: w of a colour is 0 or irrelevant)
    static const double IN[9] = {0.59719, 0.35458, 0.04823, 0.07600, 0.90834, 0.01566, 0.02840, 0.13383, 0.83777};
    static const double OUT[9] = {1.60475, -0.53108, -0.07367, -0.10208, 1.10813, -0.00605, -0.00327, -0.07276, 1.07602};
    for (size_t i = 0; i < (size_t)width * height; i++) {
        const double* c = rgba + i * 4;
        double w = c[3];
        double v[3];
        if (tonemap == 1) { // tonemap_aces (aces.rs:27-33)
            double a[3], b[3];
            // Mat4 * Vec4 includes the 4th column term m[3]*w (zero in these matrices)
            for (int r = 0; r < 3; r++) a[r] = IN[r * 3] * c[0] + IN[r * 3 + 1] * c[1] + IN[r * 3 + 2] * c[2] + 0.0 * w;
            for (int r = 0; r < 3; r++) { // rrt_and_odt_fit (aces.rs:20-24)
                double x = a[r];
                double num = x * (x + 0.0245786) - 0.000090537;
                double den = x * (x * 0.983729 + 0.4329510) + 0.238081;
                b[r] = num / den;
            }
            for (int r = 0; r < 3; r++) v[r] = clamp01(OUT[r * 3] * b[0] + OUT[r * 3 + 1] * b[1] + OUT[r * 3 + 2] * b[2] + 0.0 * 0.0);
        } else for (int r = 0; r < 3; r++) v[r] = clamp01(c[r]); // tonemap_clamp
        for (int r = 0; r < 3; r++) out_rgb[i * 3 + r] = as_u8(linear_to_srgb(v[r]) * 255.999); // output.rs:25-33
    }
    return 0;
}

int oracle_aabb_test(const double lo[3], const double hi[3], const double r[6], double t_lo, double t_hi) {
    AABB b{{Vec4::point(lo[0], lo[1], lo[2]), Vec4::point(hi[0], hi[1], hi[2])}};
    Ray ray(Vec4::point(r[0], r[1], r[2]), Vec4::vec(r[3], r[4], r[5]));
    return test_bounding_box(b, ray, Interval{t_lo, t_hi}) ? 1 : 0;
}

int oracle_octree_stats(const OracleScene* s, int mesh_index, int64_t out[4]) {
    if (!s || !out) return RTGPU_ERR_INVALID;
    OctreeStats st = s->scene.octree_stats(mesh_index);
    out[0] = st.nodes; out[1] = st.leaves; out[2] = st.leaf_entries; out[3] = st.max_depth;
    return 0;
}

int oracle_texture_sample(const OracleScene* s, int texture, double u, double v, const double p[3], double out[3]) {
    if (!s || texture < 0 || texture >= (int)s->scene.textures.size()) { g_err = "bad texture index"; return RTGPU_ERR_INVALID; }
    Vec4 c;
    tex_sample(s->scene, texture, u, v, Vec4::point(p[0], p[1], p[2]), &c);
    out[0] = c[0]; out[1] = c[1]; out[2] = c[2];
    return 0;
}

double oracle_reflectance(double cos_theta, double ior_ratio) { return reflectance(cos_theta, ior_ratio); }

} // extern "C"
