// oracle_scene.h — TEST INFRASTRUCTURE (CPU oracle). Not part of the product path.
//
// f64 restatement of the reference's scene graph: the `Hit`, `Material`, `Sampler`,
// `Noise3D` and `PDF` trait objects (src/object.rs, src/material.rs, src/texture.rs,
// src/noise.rs, src/pdf.rs) rebuilt from the flat RtgpuSceneDesc tables.
#pragma once
#include <memory>
#include <vector>

#include "oracle_math.h"
#include "rtgpu.h"

namespace oracle {

struct Material;

struct HitRecord { // object.rs:32-72
    Vec4 hit_pos, normal, tangent, bitangent;
    double t = 0;
    double u = 0, v = 0;
    bool front_face = false;
    const Material* material = nullptr;
    // bookkeeping for the parity hook (not in the reference)
    int object = -1, instance = -1, primitive = 0;

    static HitRecord make(const Ray& ray, const Vec4& hit_pos, double t, double u, double v,
                          const Vec4& outward_normal, const Vec4& tangent, const Vec4& bitangent,
                          const Material* material) {
        HitRecord h;
        h.front_face = ray.dir.dot(outward_normal) < 0.0; // object.rs:55
        h.normal = h.front_face ? outward_normal : -outward_normal;
        h.hit_pos = hit_pos; h.tangent = tangent; h.bitangent = bitangent;
        h.t = t; h.u = u; h.v = v; h.material = material;
        return h;
    }
};

struct Hittable { // object.rs:107-115
    int object = -1, instance = -1;
    virtual ~Hittable() {}
    virtual bool test(const Ray& ray, Interval t, Rng& rng, HitRecord& out) const = 0;
    virtual AABB bounding_box() const = 0;
    virtual double pdf_value(const Vec4&, const Vec4&, Rng&) const { return 0.0; }
    virtual Vec4 random(const Vec4&, Rng&) const { return Vec4::vec(1.0, 0.0, 0.0); }
};

struct Sampler { // texture.rs:21-25 (colour result in xyz, float result in v[0])
    virtual ~Sampler() {}
    virtual Vec4 sample(double u, double v, const Vec4& p) const = 0;
};

struct PerlinNoise { // noise/perlin.rs
    Vec4 random_vec[256];
    int perm_x[256], perm_y[256], perm_z[256];
    double sample(const Vec4& p) const;
    double sample_turbulence(const Vec4& p, int samples) const;
};

enum class ScatterKind { WithPDF, WithRay, Absorbed, Emissive }; // material.rs:25-36
enum class PdfKind { Cosine, Uniform };
struct ScatterResult {
    ScatterKind kind = ScatterKind::Absorbed;
    Vec4 attenuation;
    Ray scattered;        // WithRay
    PdfKind pdf = PdfKind::Cosine; // WithPDF
    Vec4 pdf_w;           // CosinePDF::new(w)
};

struct Material { // material.rs:38-47
    virtual ~Material() {}
    virtual ScatterResult scatter(const Ray& ray, const HitRecord& hit, Rng& rng) const = 0;
    virtual Vec4 emit(const HitRecord&) const { return Vec4::vec(0.0, 0.0, 0.0); }
    virtual double scattering_pdf(const Ray& ray_in, const Ray& scattered, const HitRecord& hit) const = 0;
};

struct OctreeStats { long nodes = 0, leaves = 0, leaf_entries = 0, max_depth = 0; };

struct Scene {
    std::vector<std::unique_ptr<Hittable>> owned;
    std::vector<std::unique_ptr<Material>> materials;
    std::vector<std::unique_ptr<Sampler>> textures;
    std::vector<std::unique_ptr<PerlinNoise>> noises;
    std::vector<std::shared_ptr<void>> mesh_cache; // per mesh-object index: octree + arrays
    const Hittable* world = nullptr;
    const Hittable* lights = nullptr;
    int n_instances = 0;
    std::string error;

    bool build(const RtgpuSceneDesc& d);
    OctreeStats octree_stats(int mesh_object) const;
};

// exposed for known-answer tests
double tex_sample(const Scene& s, int tex, double u, double v, const Vec4& p, Vec4* color);

} // namespace oracle
