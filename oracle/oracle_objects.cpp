// oracle_objects.cpp — TEST INFRASTRUCTURE (CPU oracle). Not part of the product path.
//
// The reference's `Hit` implementors (src/object/*.rs) restated in f64, including the
// per-mesh octree (src/object/mesh/octree.rs) and its fixed-order DFS traversal
// (src/object/mesh.rs:165-197), rebuilt from the flat RtgpuSceneDesc tables.
#include <string>

#include "oracle_scene.h"

namespace oracle {

Material* make_emissive_for(const Sampler* tex);
bool build_shading(Scene& s, const RtgpuSceneDesc& d);

namespace {

// ---- object/list.rs ---------------------------------------------------------------------
struct ObjectList : Hittable {
    std::vector<const Hittable*> objects;
    AABB bounds{{INFINITY_VEC, -INFINITY_VEC}};
    bool disable_bounds_check = false;
    void add(const Hittable* o) { // list.rs:51-55
        AABB two[2] = {bounds, o->bounding_box()};
        bounds = combine_bounds(two, 2);
        objects.push_back(o);
    }
    bool test(const Ray& ray, Interval t, Rng& rng, HitRecord& out) const override { // list.rs:58-74
        if (!disable_bounds_check && !test_bounding_box(bounds, ray, t)) return false;
        bool any = false;
        double closest_t = t.hi;
        HitRecord h;
        for (const Hittable* o : objects)
            if (o->test(ray, Interval{t.lo, closest_t}, rng, h)) { closest_t = h.t; out = h; any = true; }
        return any;
    }
    AABB bounding_box() const override { return bounds; }
    double pdf_value(const Vec4& origin, const Vec4& dir, Rng& rng) const override { // list.rs:80-89
        double weight = 1.0 / (double)objects.size();
        double sum = 0.0;
        for (const Hittable* o : objects) sum += weight * o->pdf_value(origin, dir, rng);
        return sum;
    }
    Vec4 random(const Vec4& origin, Rng& rng) const override { // list.rs:91-100
        size_t size = objects.size();
        if (size == 0) return Vec4::vec(1.0, 0.0, 0.0);
        size_t idx = rng.inject ? (size_t)std::fmin((double)size - 1, std::floor(rng.draw_uniform() * size)) : (size_t)rng.below(size);
        return objects[idx]->random(origin, rng);
    }
};

// ---- object/bvh.rs, bvh/null_obj.rs -------------------------------------------------------
struct NullObject : Hittable {
    bool test(const Ray&, Interval, Rng&, HitRecord&) const override { return false; }
    AABB bounding_box() const override { return AABB{{INFINITY_VEC, -INFINITY_VEC}}; }
};
struct BvhNode : Hittable {
    const Hittable *c0 = nullptr, *c1 = nullptr;
    AABB bounds;
    bool test(const Ray& ray, Interval t, Rng& rng, HitRecord& out) const override { // bvh.rs:84-101
        if (!test_bounding_box(bounds, ray, t)) return false;
        bool any = false;
        double closest_t = t.hi;
        HitRecord h;
        if (c0->test(ray, Interval{t.lo, closest_t}, rng, h)) { closest_t = h.t; out = h; any = true; }
        if (c1->test(ray, Interval{t.lo, closest_t}, rng, h)) { out = h; any = true; }
        return any;
    }
    AABB bounding_box() const override { return bounds; }
};

// ---- object/transform.rs ------------------------------------------------------------------
struct Transform : Hittable {
    const Hittable* child = nullptr;
    Mat4 transform, inv_transform;
    AABB bounds;
    void update_bounds() { // transform.rs:98-118
        AABB cb = child->bounding_box();
        Vec4 o_min = cb.b[0], o_max = cb.b[1];
        Vec4 s = o_max - o_min;
        double sx = s.x(), sy = s.y(), sz = s.z();
        Vec4 corners[8] = {o_min, o_min + Vec4::vec(0.0, 0.0, sz), o_min + Vec4::vec(0.0, sy, 0.0), o_min + Vec4::vec(0.0, sy, sz),
                           o_min + Vec4::vec(sx, 0.0, 0.0), o_min + Vec4::vec(sx, 0.0, sz), o_min + Vec4::vec(sx, sy, 0.0), o_max};
        for (Vec4& c : corners) c = transform * c;
        bounds = get_bounding_box(corners, 8);
    }
    bool test(const Ray& ray, Interval t, Rng& rng, HitRecord& out) const override { // transform.rs:122-139
        Ray ray_obj(inv_transform * ray.origin, inv_transform * ray.dir);
        if (!child->test(ray_obj, t, rng, out)) return false;
        out.hit_pos = transform * out.hit_pos;
        out.normal = (transform * out.normal).to_unit();
        return true;
    }
    AABB bounding_box() const override { return bounds; }
};

// ---- object/sphere.rs ---------------------------------------------------------------------
struct Sphere : Hittable {
    const Material* material = nullptr;
    Vec4 center;
    double radius = 0;
    AABB bounds;
    bool test_impl(const Ray& ray, Interval t, bool skip_uvs, HitRecord& out) const { // sphere.rs:40-94
        Vec4 center_diff = ray.origin - center;
        double a = ray.dir.length_squared();
        double half_b = ray.dir.dot(center_diff);
        double c = center_diff.length_squared() - radius * radius;
        double discriminant = half_b * half_b - a * c;
        if (discriminant < 0.0) return false;
        double d_sqrt = std::sqrt(discriminant);
        double root = (-half_b - d_sqrt) / a;
        if (root <= t.lo || t.hi <= root) {
            root = (-half_b + d_sqrt) / a;
            if (root <= t.lo || t.hi <= root) return false;
        }
        Vec4 hit_pos = ray.at(root);
        Vec4 normal = (hit_pos - center) / radius;
        double u = 0.0, v = 0.0;
        Vec4 tangent = Vec4::vec(1.0, 0.0, 0.0), bitangent = Vec4::vec(1.0, 0.0, 0.0);
        if (!skip_uvs) {
            double theta = std::acos(normal[1]);
            double phi = std::atan2(-normal[2], normal[0]) + PI;
            tangent = Vec4::vec(-normal[2], 0.0, -normal[0]);
            bitangent = normal.cross(tangent);
            u = phi / (2.0 * PI); v = theta / PI;
        }
        out = HitRecord::make(ray, hit_pos, root, u, v, normal, tangent, bitangent, material);
        out.object = object; out.instance = instance; out.primitive = 0;
        return true;
    }
    bool test(const Ray& ray, Interval t, Rng&, HitRecord& out) const override { return test_impl(ray, t, false, out); }
    AABB bounding_box() const override { return bounds; }
    double pdf_value(const Vec4& origin, const Vec4& dir, Rng&) const override { // sphere.rs:106-121
        HitRecord h;
        if (!test_impl(Ray(origin, dir), Interval{0.001, INF}, true, h)) return 0.0;
        double radius_squared = radius * radius;
        double cos_theta_max = std::sqrt(1.0 - radius_squared / (center - origin).length_squared());
        double solid_angle = 2.0 * PI * (1.0 - cos_theta_max);
        return 1.0 / solid_angle;
    }
    Vec4 random(const Vec4& origin, Rng& rng) const override { // sphere.rs:123-145
        Vec4 dir = center - origin;
        Mat4 basis = onb_from_vec(dir); // NOT normalised
        double distance_squared = dir.length_squared();
        double radius_squared = radius * radius;
        double cos_theta_max = std::sqrt(1.0 - radius_squared / distance_squared);
        double r1 = rng.draw_uniform(), r2 = rng.draw_uniform();
        double phi = r1 * 2.0 * PI;
        double z = 1.0 + r2 * (cos_theta_max - 1.0);
        double x = std::cos(phi) * std::sqrt(1.0 - z * z);
        double y = std::sin(phi) * std::sqrt(1.0 - z * z);
        return basis * Vec4::vec(x, y, z);
    }
};

// ---- object/plane.rs ----------------------------------------------------------------------
struct Plane : Hittable {
    const Material* material = nullptr;
    bool render_backface = false;
    Vec4 corner, normal, u, v, inv_u, inv_v;
    double area = 0;
    AABB bounds;
    void init(const Vec4& center, const Vec4& uu, const Vec4& vv) { // plane.rs:29-62
        u = uu; v = vv;
        Vec4 u_unit = u.to_unit(), v_unit = v.to_unit();
        Vec4 n = u.cross(v);
        area = n.length() * 4.0;
        normal = n.to_unit();
        Vec4 corners[4] = {center + u + v, center + u - v, center - u + v, center - u - v};
        bounds = get_bounding_box(corners, 4);
        corner = corners[3];
        inv_u = u_unit * 0.5 / u.length();
        inv_v = v_unit * 0.5 / v.length();
    }
    bool test(const Ray& ray, Interval t, Rng&, HitRecord& out) const override { // plane.rs:66-101
        double dot_ray_normal = normal.dot(ray.dir);
        double dd = render_backface ? std::fabs(dot_ray_normal) : -dot_ray_normal;
        if (dd < F64_EPSILON) return false;
        double hit_t = normal.dot(corner - ray.origin) / dot_ray_normal;
        if (hit_t <= t.lo || t.hi <= hit_t) return false;
        Vec4 hit_pos = ray.at(hit_t);
        Vec4 local_pos = hit_pos - corner;
        double pu = local_pos.dot(inv_u), pv = local_pos.dot(inv_v);
        if (pu < 0.0 || pu > 1.0 || pv < 0.0 || pv > 1.0) return false;
        out = HitRecord::make(ray, hit_pos, hit_t, pu, pv, normal, u.to_unit(), v.to_unit(), material);
        out.object = object; out.instance = instance; out.primitive = 0;
        return true;
    }
    AABB bounding_box() const override { return bounds; }
    double pdf_value(const Vec4& origin, const Vec4& dir, Rng& rng) const override { // plane.rs:107-118
        HitRecord hit;
        if (!test(Ray(origin, dir), Interval{0.001, INF}, rng, hit)) return 0.0;
        double dist_squared = hit.t * hit.t * dir.length_squared();
        double cosine = std::fabs(dir.dot(hit.normal) / dir.length());
        return dist_squared / (cosine * area);
    }
    Vec4 random(const Vec4& origin, Rng& rng) const override { // plane.rs:120-126: one QUADRANT only
        double ru = rng.draw_uniform(), rv = rng.draw_uniform();
        Vec4 p = corner + u * ru + v * rv;
        return p - origin;
    }
};

// ---- object/sky.rs, object/sun.rs -----------------------------------------------------------
static AABB infinite_box() {
    const double MAX = std::numeric_limits<double>::max();
    return AABB{{Vec4::point(-MAX, -MAX, -MAX), Vec4::point(MAX, MAX, MAX)}};
}
struct Sky : Hittable {
    std::unique_ptr<Material> material;
    bool test(const Ray& ray, Interval t, Rng&, HitRecord& out) const override { // sky.rs:28-52
        double hit_t = INF;
        if (hit_t > t.hi) return false;
        Vec4 hit_pos = ray.at(hit_t);
        Vec4 unit_dir = ray.dir.to_unit();
        Vec4 normal = -unit_dir;
        double u = std::atan2(unit_dir.x(), unit_dir.z()) / (2.0 * PI) + 0.5;
        double v = unit_dir.dot(Vec4::vec(0.0, 1.0, 0.0)) / 2.0 + 0.5;
        out = HitRecord::make(ray, hit_pos, hit_t, u, v, normal, Vec4::vec(1.0, 0.0, 0.0), Vec4::vec(1.0, 0.0, 0.0), material.get());
        out.object = object; out.instance = instance; out.primitive = 0;
        return true;
    }
    AABB bounding_box() const override { return infinite_box(); }
    double pdf_value(const Vec4&, const Vec4&, Rng&) const override { return 1.0 / (4.0 * PI); }
    Vec4 random(const Vec4&, Rng& rng) const override { return random_unit(rng); }
};
struct Sun : Hittable {
    std::unique_ptr<Material> material;
    Vec4 direction;
    bool test(const Ray& ray, Interval t, Rng&, HitRecord& out) const override { // sun.rs:33-61
        Vec4 unit_dir = ray.dir.to_unit();
        if (std::fabs(direction.dot(unit_dir) - 1.0) > 0.001) return false;
        double hit_t = std::numeric_limits<double>::max();
        if (hit_t >= t.hi) return false;
        Vec4 hit_pos = ray.at(hit_t);
        Vec4 normal = -unit_dir;
        out = HitRecord::make(ray, hit_pos, hit_t, 0.0, 0.0, normal, Vec4::vec(1.0, 0.0, 0.0), Vec4::vec(1.0, 0.0, 0.0), material.get());
        out.object = object; out.instance = instance; out.primitive = 0;
        return true;
    }
    AABB bounding_box() const override { return infinite_box(); }
    double pdf_value(const Vec4&, const Vec4&, Rng&) const override { return 1.0; } // sun.rs:70-72
    Vec4 random(const Vec4&, Rng&) const override { return direction; }
};

// ---- object/volume.rs ---------------------------------------------------------------------
struct Volume : Hittable {
    const Hittable* boundary = nullptr;
    const Material* material = nullptr;
    double neg_inv_density = 0;
    bool test(const Ray& ray, Interval t, Rng& rng, HitRecord& out) const override { // volume.rs:33-71
        HitRecord hit_enter, hit_exit;
        if (!boundary->test(ray, Interval{-INF, INF}, rng, hit_enter)) return false;
        if (!boundary->test(ray, Interval{hit_enter.t + 0.0001, INF}, rng, hit_exit)) return false;
        double t_min = std::fmax(hit_enter.t, t.lo);
        double t_max = std::fmin(hit_exit.t, t.hi);
        if (t_min >= t_max) return false;
        t_min = std::fmax(t_min, 0.0);
        double ray_len = ray.dir.length();
        double dist_inside_boundary = (t_max - t_min) * ray_len;
        double hit_dist = neg_inv_density * std::log(rng.draw_uniform());
        if (hit_dist > dist_inside_boundary) return false;
        double tt = t_min + hit_dist / ray_len;
        Vec4 hit_pos = ray.at(tt);
        out = HitRecord::make(ray, hit_pos, tt, 0.0, 0.0, Vec4::vec(1.0, 0.0, 0.0), Vec4::vec(1.0, 0.0, 0.0), Vec4::vec(1.0, 0.0, 0.0), material);
        out.object = object; out.instance = instance; out.primitive = 0;
        return true;
    }
    AABB bounding_box() const override { return boundary->bounding_box(); }
};

// ---- object/mesh.rs + mesh/octree.rs ------------------------------------------------------
struct Triangle { int vi[3], ni[3], ti[3]; bool has_uv; };

struct OctreeNode {
    bool leaf = true;
    std::vector<int> tris;                      // Leaf(Vec<usize>)
    std::unique_ptr<OctreeNode> children[8];    // Branch([Box<OctreeNode>; 8])
    AABB bounding_box;
};

static const size_t MAX_TRIS_PER_LEAF = 50; // octree.rs:6
static const size_t MAX_DEPTH = 50;         // octree.rs:7

struct MeshData {
    std::vector<Vec4> vertices, uvs, normals;
    std::vector<Triangle> triangles;
    AABB bounds;
    std::unique_ptr<OctreeNode> octree;
    OctreeStats stats;

    std::unique_ptr<OctreeNode> build_node(const std::vector<int>* filter, const Vec4& b_min, const Vec4& b_max, size_t depth) { // octree.rs:31-210
        auto node = std::make_unique<OctreeNode>();
        node->bounding_box = AABB{{b_min, b_max}};
        size_t count = filter ? filter->size() : triangles.size();
        stats.nodes++;
        if ((long)depth > stats.max_depth) stats.max_depth = (long)depth;
        if (count <= MAX_TRIS_PER_LEAF || depth >= MAX_DEPTH) {
            node->leaf = true;
            if (filter) node->tris = *filter;
            else { node->tris.resize(count); for (size_t i = 0; i < count; i++) node->tris[i] = (int)i; }
            stats.leaves++; stats.leaf_entries += (long)node->tris.size();
            return node;
        }
        Vec4 midpoint = (b_min + b_max) / 2.0;
        std::vector<int> lists[8];
        for (size_t k = 0; k < count; k++) {
            int index = filter ? (*filter)[k] : (int)k;
            const Triangle& tri = triangles[index];
            Vec4 tv[3] = {vertices[tri.vi[0]], vertices[tri.vi[1]], vertices[tri.vi[2]]};
            AABB tb = get_bounding_box(tv, 3); // padded
            const Vec4 &t_min = tb.b[0], &t_max = tb.b[1];
            bool in[8] = {true, true, true, true, true, true, true, true};
            if (t_min.x() > midpoint.x()) in[0] = in[1] = in[2] = in[3] = false;
            if (t_max.x() < midpoint.x()) in[4] = in[5] = in[6] = in[7] = false;
            if (t_min.y() > midpoint.y()) in[0] = in[1] = in[4] = in[5] = false;
            if (t_max.y() < midpoint.y()) in[2] = in[3] = in[6] = in[7] = false;
            if (t_min.z() > midpoint.z()) in[0] = in[2] = in[4] = in[6] = false;
            if (t_max.z() < midpoint.z()) in[1] = in[3] = in[5] = in[7] = false;
            for (int i = 0; i < 8; i++) if (in[i]) lists[i].push_back(index);
        }
        double min_x = b_min.x(), min_y = b_min.y(), min_z = b_min.z();
        double max_x = b_max.x(), max_y = b_max.y(), max_z = b_max.z();
        double mid_x = midpoint.x(), mid_y = midpoint.y(), mid_z = midpoint.z();
        node->leaf = false;
        node->children[0] = build_node(&lists[0], b_min, midpoint, depth + 1);
        node->children[1] = build_node(&lists[1], Vec4::point(min_x, min_y, mid_z), Vec4::point(mid_x, mid_y, max_z), depth + 1);
        node->children[2] = build_node(&lists[2], Vec4::point(min_x, mid_y, min_z), Vec4::point(mid_x, max_y, mid_z), depth + 1);
        node->children[3] = build_node(&lists[3], Vec4::point(min_x, mid_y, mid_z), Vec4::point(mid_x, max_y, max_z), depth + 1);
        node->children[4] = build_node(&lists[4], Vec4::point(mid_x, min_y, min_z), Vec4::point(max_x, mid_y, mid_z), depth + 1);
        node->children[5] = build_node(&lists[5], Vec4::point(mid_x, min_y, mid_z), Vec4::point(max_x, mid_y, max_z), depth + 1);
        node->children[6] = build_node(&lists[6], Vec4::point(mid_x, mid_y, min_z), Vec4::point(max_x, max_y, mid_z), depth + 1);
        node->children[7] = build_node(&lists[7], midpoint, b_max, depth + 1);
        return node;
    }
};

struct TriangleMesh : Hittable {
    std::shared_ptr<MeshData> data;
    const Material* material = nullptr;
    bool flat_shading = false, hit_back_faces = false;

    bool test_tri(int tri_index, const Ray& ray, Interval t_int, HitRecord& out) const { // mesh.rs:62-163
        const Triangle& triangle = data->triangles[tri_index];
        Vec4 v0 = data->vertices[triangle.vi[0]], v1 = data->vertices[triangle.vi[1]], v2 = data->vertices[triangle.vi[2]];
        Vec4 edge1 = v1 - v0, edge2 = v2 - v0;
        Vec4 ray_x_edge2 = ray.dir.cross(edge2);
        double det = edge1.dot(ray_x_edge2);
        double dd = hit_back_faces ? std::fabs(det) : det;
        if (dd < F64_EPSILON) return false;
        double inv_det = 1.0 / det;
        Vec4 b = ray.origin - v0;
        double u = b.dot(ray_x_edge2) * inv_det;
        if (u < 0.0 || u > 1.0) return false;
        Vec4 b_x_edge1 = b.cross(edge1);
        double v = ray.dir.dot(b_x_edge1) * inv_det;
        if (v < 0.0 || u + v > 1.0) return false;
        double t = edge2.dot(b_x_edge1) * inv_det;
        if (t <= t_int.lo || t_int.hi <= t) return false;
        Vec4 hit_pos = ray.at(t);
        double w = 1.0 - u - v;
        Vec4 normal;
        if (flat_shading) normal = edge1.cross(edge2).to_unit();
        else {
            Vec4 n0 = data->normals[triangle.ni[0]], n1 = data->normals[triangle.ni[1]], n2 = data->normals[triangle.ni[2]];
            normal = n0 * w + n1 * u + n2 * v; // unnormalised blend
        }
        Vec4 tangent = Vec4::vec(1.0, 0.0, 0.0), bitangent = Vec4::vec(1.0, 0.0, 0.0);
        Vec4 tex_coords = Vec4::vec(0.0, 0.0, 0.0);
        if (triangle.has_uv) {
            Vec4 uv0 = data->uvs[triangle.ti[0]], uv1 = data->uvs[triangle.ti[1]], uv2 = data->uvs[triangle.ti[2]];
            Vec4 duv1 = uv1 - uv0, duv2 = uv2 - uv0;
            Vec4 edge1perp = normal.cross(edge1), edge2perp = edge2.cross(normal);
            tangent = edge2perp * duv1[0] + edge1perp * duv2[0];
            bitangent = edge2perp * duv1[1] + edge1perp * duv2[1];
            double inv_max = 1.0 / std::sqrt(std::fmax(tangent.length_squared(), bitangent.length_squared()));
            tangent *= -inv_max;
            bitangent *= inv_max;
            tex_coords = uv0 * w + uv1 * u + uv2 * v;
        }
        out = HitRecord::make(ray, hit_pos, t, tex_coords[0], tex_coords[1], normal, tangent, bitangent, material);
        out.object = object; out.instance = instance; out.primitive = tri_index;
        return true;
    }

    bool test_octree_node(const OctreeNode& node, const Ray& ray, Interval t, HitRecord& out) const { // mesh.rs:165-197
        if (!test_bounding_box(node.bounding_box, ray, t)) return false;
        bool any = false;
        double closest_t = t.hi;
        HitRecord h;
        if (node.leaf) {
            for (int idx : node.tris)
                if (test_tri(idx, ray, Interval{t.lo, closest_t}, h)) { closest_t = h.t; out = h; any = true; }
        } else {
            for (int i = 0; i < 8; i++)
                if (test_octree_node(*node.children[i], ray, Interval{t.lo, closest_t}, h)) { closest_t = h.t; out = h; any = true; }
        }
        return any;
    }
    bool test(const Ray& ray, Interval t, Rng&, HitRecord& out) const override { return test_octree_node(*data->octree, ray, t, out); }
    AABB bounding_box() const override { return data->bounds; }
};

struct Builder {
    Scene& s;
    const RtgpuSceneDesc& d;
    std::vector<std::shared_ptr<MeshData>> mesh_data; // per mesh index
    int next_instance = 0;
    int depth = 0;

    template <class T> T* own(T* p) { s.owned.emplace_back(p); return p; }

    std::shared_ptr<MeshData> get_mesh(int mi) {
        if (mesh_data[mi]) return mesh_data[mi];
        const RtgpuMesh& m = d.meshes[mi];
        auto md = std::make_shared<MeshData>();
        md->vertices.resize(m.n_positions);
        for (int i = 0; i < m.n_positions; i++) md->vertices[i] = Vec4::point(m.positions[i * 3], m.positions[i * 3 + 1], m.positions[i * 3 + 2]); // obj.rs:32
        md->normals.resize(m.n_normals);
        for (int i = 0; i < m.n_normals; i++) md->normals[i] = Vec4::vec(m.normals[i * 3], m.normals[i * 3 + 1], m.normals[i * 3 + 2]);
        md->uvs.resize(m.n_uvs);
        for (int i = 0; i < m.n_uvs; i++) md->uvs[i] = Vec4::vec(m.uvs[i * 2], m.uvs[i * 2 + 1], 0.0);
        md->triangles.resize(m.n_triangles);
        for (int i = 0; i < m.n_triangles; i++) {
            Triangle& t = md->triangles[i];
            t.has_uv = m.tri_uv != nullptr;
            for (int k = 0; k < 3; k++) {
                t.vi[k] = m.tri_pos[i * 3 + k]; t.ni[k] = m.tri_nrm[i * 3 + k];
                t.ti[k] = m.tri_uv ? m.tri_uv[i * 3 + k] : -1;
                if (t.ti[k] < 0) t.has_uv = false;
            }
        }
        md->bounds = get_bounding_box(md->vertices.data(), (int)md->vertices.size()); // mesh.rs:45
        md->octree = md->build_node(nullptr, md->bounds.b[0], md->bounds.b[1], 0);     // mesh.rs:46
        mesh_data[mi] = md;
        return md;
    }

    // Expands the object DAG into a tree; leaves get DFS pre-order instance ordinals.
    const Hittable* build(int idx) {
        if (idx < 0 || idx >= d.n_objects) { s.error = "object index out of range"; return nullptr; }
        if (++depth > 64) { s.error = "object graph too deep (cycle?)"; return nullptr; }
        const RtgpuObject& o = d.objects[idx];
        const int32_t* kids = d.children + o.first_child;
        auto mat = [&](int m) -> const Material* { return (m >= 0 && m < d.n_materials) ? s.materials[m].get() : nullptr; };
        const Hittable* result = nullptr;
        switch (o.kind) {
        case RTGPU_OBJ_LIST: {
            auto* l = own(new ObjectList());
            l->disable_bounds_check = (o.flags & RTGPU_OBJF_DISABLE_BOUNDS_CHECK) != 0;
            for (int k = 0; k < o.n_children; k++) { const Hittable* c = build(kids[k]); if (!c) return nullptr; l->add(c); }
            result = l; break;
        }
        case RTGPU_OBJ_BVH: {
            if (o.n_children != 2) { s.error = "bvh node needs two children"; return nullptr; }
            auto* b = own(new BvhNode());
            b->c0 = build(kids[0]); if (!b->c0) return nullptr;
            b->c1 = build(kids[1]); if (!b->c1) return nullptr;
            if (d.objects[kids[1]].kind == RTGPU_OBJ_NULL) b->bounds = b->c0->bounding_box(); // bvh.rs:49-53
            else { AABB two[2] = {b->c0->bounding_box(), b->c1->bounding_box()}; b->bounds = combine_bounds(two, 2); }
            result = b; break;
        }
        case RTGPU_OBJ_NULL: { auto* n = own(new NullObject()); n->instance = next_instance++; result = n; break; }
        case RTGPU_OBJ_TRANSFORM: {
            if (o.n_children != 1) { s.error = "transform needs one child"; return nullptr; }
            auto* t = own(new Transform());
            for (int i = 0; i < 16; i++) { t->transform.m[i] = o.transform[i]; t->inv_transform.m[i] = o.inv_transform[i]; }
            t->child = build(kids[0]); if (!t->child) return nullptr;
            t->update_bounds();
            result = t; break;
        }
        case RTGPU_OBJ_MESH: {
            if (o.mesh < 0 || o.mesh >= d.n_meshes || !mat(o.material)) { s.error = "mesh: bad mesh/material index"; return nullptr; }
            auto* m = own(new TriangleMesh());
            m->data = get_mesh(o.mesh); m->material = mat(o.material);
            m->flat_shading = (o.flags & RTGPU_OBJF_FLAT_SHADING) != 0; m->hit_back_faces = (o.flags & RTGPU_OBJF_HIT_BACK_FACES) != 0;
            m->instance = next_instance++; result = m; break;
        }
        case RTGPU_OBJ_SPHERE: {
            if (!mat(o.material)) { s.error = "sphere: bad material index"; return nullptr; }
            auto* sp = own(new Sphere());
            sp->material = mat(o.material); sp->center = Vec4(o.p0[0], o.p0[1], o.p0[2], o.p0[3]); sp->radius = o.p1[0];
            Vec4 rv = Vec4::vec(sp->radius, sp->radius, sp->radius);
            sp->bounds = AABB{{sp->center - rv, sp->center + rv}}; // sphere.rs:27-30
            sp->instance = next_instance++; result = sp; break;
        }
        case RTGPU_OBJ_PLANE: {
            if (!mat(o.material)) { s.error = "plane: bad material index"; return nullptr; }
            auto* p = own(new Plane());
            p->material = mat(o.material); p->render_backface = (o.flags & RTGPU_OBJF_BACKFACE) != 0;
            p->init(Vec4(o.p0[0], o.p0[1], o.p0[2], o.p0[3]), Vec4::vec(o.p1[0], o.p1[1], o.p1[2]), Vec4::vec(o.p2[0], o.p2[1], o.p2[2]));
            p->instance = next_instance++; result = p; break;
        }
        case RTGPU_OBJ_VOLUME: {
            if (o.n_children != 1 || !mat(o.material)) { s.error = "volume: bad boundary/material"; return nullptr; }
            auto* v = own(new Volume());
            v->material = mat(o.material); v->neg_inv_density = -1.0 / o.p1[0];
            v->instance = next_instance++;
            v->boundary = build(kids[0]); if (!v->boundary) return nullptr;
            result = v; break;
        }
        case RTGPU_OBJ_SKY:
        case RTGPU_OBJ_SUN: {
            if (o.texture < 0 || o.texture >= d.n_textures) { s.error = "sky/sun: bad texture index"; return nullptr; }
            if (o.kind == RTGPU_OBJ_SKY) { auto* k = own(new Sky()); k->material.reset(make_emissive_for(s.textures[o.texture].get())); k->instance = next_instance++; result = k; }
            else { auto* k = own(new Sun()); k->material.reset(make_emissive_for(s.textures[o.texture].get())); k->direction = Vec4(o.p0[0], o.p0[1], o.p0[2], o.p0[3]); k->instance = next_instance++; result = k; }
            break;
        }
        default: s.error = "unknown object kind"; return nullptr;
        }
        const_cast<Hittable*>(result)->object = idx;
        depth--;
        return result;
    }
};

} // namespace

bool Scene::build(const RtgpuSceneDesc& d) {
    if (d.abi_version != RTGPU_ABI_VERSION) { error = "abi version mismatch"; return false; }
    if (!build_shading(*this, d)) return false;
    Builder b{*this, d, {}, 0, 0};
    b.mesh_data.resize(d.n_meshes);
    world = b.build(d.world_root);
    if (!world) return false;
    n_instances = b.next_instance;
    b.depth = 0;
    lights = b.build(d.lights_root); // ordinals continue; only `world` ordinals are reported
    if (!lights) return false;
    mesh_cache.clear();
    for (auto& m : b.mesh_data) mesh_cache.push_back(m);
    return true;
}

OctreeStats Scene::octree_stats(int mesh_index) const {
    if (mesh_index < 0 || mesh_index >= (int)mesh_cache.size() || !mesh_cache[mesh_index]) return OctreeStats{};
    return std::static_pointer_cast<MeshData>(mesh_cache[mesh_index])->stats;
}

} // namespace oracle
