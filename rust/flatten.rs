//! src/gpu/flatten.rs — the `flatten()` visitor that turns the reference's opaque scene graph
//! (`Arc<dyn Hit>` / `Arc<dyn Material>` / `Arc<dyn Sampler>`, src/object.rs:107-115, src/material.rs:38-47,
//! src/texture.rs:21-25) into the flat POD tables of include/rtgpu.h.
//!
//! UNCOMPILED (no Rust toolchain in the build image). The C++ host of this repository produces the same tables
//! from the same inputs (rust_raytracer_b200/csrc/host/scene_loader.cpp, builtin_scenes.cpp); this file is what a
//! maintainer of the Rust crate adds so that `rust_raytracer <scene> --gpu` needs no C++ loader at all.
//!
//! The only change to EXISTING reference code is one defaulted method per trait:
//!
//!     // object.rs   trait Hit      { ...; fn flatten(&self, b: &mut SceneBuilder) -> Result<ObjectId, Unsupported>; }
//!     // material.rs trait Material { ...; fn flatten(&self, b: &mut SceneBuilder) -> Result<MaterialId, Unsupported>; }
//!     // texture.rs  trait Sampler  { ...; fn flatten(&self, b: &mut SceneBuilder) -> Result<TextureId, Unsupported>; }
//!     // noise.rs    trait Noise3D  { ...; fn flatten(&self, b: &mut SceneBuilder) -> Result<NoiseId, Unsupported>; }
//!
//! Each `impl` below goes into the file of the type it is for (the fields it reads are private to that module);
//! the file:line beside it says where the fields are declared.
#![allow(dead_code)]

use std::collections::HashMap;
use std::sync::Arc;

use super::ffi::*;
use crate::mat4::Mat4;
use crate::vec4::{Color, Vec4};

pub type ObjectId = i32;
pub type MaterialId = i32;
pub type TextureId = i32;
pub type NoiseId = i32;

#[derive(Debug)]
pub struct Unsupported(pub &'static str);

/// Owns every array the RtgpuSceneDesc points to; `desc()` borrows them for the duration of rtgpu_scene_create
/// (which copies everything).
#[derive(Default)]
pub struct SceneBuilder {
    objects: Vec<RtgpuObject>,
    children: Vec<i32>,
    meshes: Vec<RtgpuMesh>,
    mesh_data: Vec<MeshData>, // keeps the f64 / i32 arrays the RtgpuMesh pointers refer to alive
    materials: Vec<RtgpuMaterial>,
    textures: Vec<RtgpuTexture>,
    images: Vec<RtgpuImage>,
    image_data: Vec<Vec<f32>>,
    noises: Vec<RtgpuNoise>,
    // shared sub-objects (the same Arc reachable from `world` and `lights`, the same material on many objects) are
    // emitted once: keyed by the address of the Arc's payload
    seen_objects: HashMap<usize, ObjectId>,
    seen_materials: HashMap<usize, MaterialId>,
    seen_textures: HashMap<usize, TextureId>,
    seen_noises: HashMap<usize, NoiseId>,
}

struct MeshData {
    positions: Vec<f64>,
    normals: Vec<f64>,
    uvs: Vec<f64>,
    tri_pos: Vec<i32>,
    tri_nrm: Vec<i32>,
    tri_uv: Vec<i32>,
}

fn key<T: ?Sized>(p: &T) -> usize {
    p as *const T as *const u8 as usize
}

fn xyz(v: &Vec4) -> [f64; 4] {
    [v.x(), v.y(), v.z(), v.w()]
}

fn mat16(m: &Mat4) -> [f64; 16] {
    m.0 // Mat4([f64; 16]), row-major (mat4.rs:10)
}

impl SceneBuilder {
    pub fn new() -> Self {
        Self::default()
    }

    // ---- de-duplicating entry points the trait impls call -------------------------------------
    pub fn object(&mut self, o: &Arc<dyn crate::object::Hit>) -> Result<ObjectId, Unsupported> {
        let k = key(o.as_ref());
        if let Some(id) = self.seen_objects.get(&k) {
            return Ok(*id);
        }
        let id = o.flatten(self)?;
        self.seen_objects.insert(k, id);
        Ok(id)
    }

    pub fn material(&mut self, m: &Arc<dyn crate::material::Material>) -> Result<MaterialId, Unsupported> {
        let k = key(m.as_ref());
        if let Some(id) = self.seen_materials.get(&k) {
            return Ok(*id);
        }
        let id = m.flatten(self)?;
        self.seen_materials.insert(k, id);
        Ok(id)
    }

    pub fn texture<T>(&mut self, t: &Arc<dyn crate::texture::Sampler<Output = T>>) -> Result<TextureId, Unsupported>
    where
        T: Send + Sync + Copy + 'static,
    {
        let k = key(t.as_ref());
        if let Some(id) = self.seen_textures.get(&k) {
            return Ok(*id);
        }
        let id = t.flatten(self)?;
        self.seen_textures.insert(k, id);
        Ok(id)
    }

    // ---- emitters ---------------------------------------------------------------------------------
    fn push_object(&mut self, o: RtgpuObject, kids: &[ObjectId]) -> ObjectId {
        let mut o = o;
        o.first_child = self.children.len() as i32;
        o.n_children = kids.len() as i32;
        self.children.extend_from_slice(kids);
        self.objects.push(o);
        self.objects.len() as i32 - 1
    }

    pub fn push_material(&mut self, kind: i32, albedo: i32, roughness: i32, normal_map: i32, ior: f64) -> MaterialId {
        self.materials.push(RtgpuMaterial { kind, albedo, roughness, normal_map, ior });
        self.materials.len() as i32 - 1
    }

    pub fn push_texture(&mut self, t: RtgpuTexture) -> TextureId {
        self.textures.push(t);
        self.textures.len() as i32 - 1
    }

    /// Buffer (buffer.rs:8-12: Vec<Vec4> of f64) -> RGBF32 image; values are used as stored (no sRGB decode, Q6)
    pub fn push_image(&mut self, buffer: &crate::buffer::Buffer) -> i32 {
        let (w, h) = buffer.size();
        let mut texels = Vec::with_capacity(w * h * 3);
        for y in 0..h {
            for x in 0..w {
                let c = buffer.get_pixel(x, y);
                texels.extend_from_slice(&[c.x() as f32, c.y() as f32, c.z() as f32]);
            }
        }
        self.image_data.push(texels);
        let data = self.image_data.last().unwrap().as_ptr() as *const std::os::raw::c_void;
        self.images.push(RtgpuImage { width: w as i32, height: h as i32, format: IMG_RGBF32, reserved: 0, data });
        self.images.len() as i32 - 1
    }

    pub fn push_noise(&mut self, gradients: &[Vec4; 256], px: &[usize; 256], py: &[usize; 256], pz: &[usize; 256]) -> NoiseId {
        let mut n = RtgpuNoise { gradients: [[0.0; 3]; 256], perm_x: [0; 256], perm_y: [0; 256], perm_z: [0; 256] };
        for i in 0..256 {
            n.gradients[i] = [gradients[i].x(), gradients[i].y(), gradients[i].z()];
            n.perm_x[i] = px[i] as i32;
            n.perm_y[i] = py[i] as i32;
            n.perm_z[i] = pz[i] as i32;
        }
        self.noises.push(n);
        self.noises.len() as i32 - 1
    }

    /// The tables as the C ABI wants them. `self` must outlive the rtgpu_scene_create call.
    pub fn desc(&self, world: ObjectId, lights: ObjectId) -> RtgpuSceneDesc {
        RtgpuSceneDesc {
            abi_version: RTGPU_ABI_VERSION,
            world_root: world,
            lights_root: lights,
            n_objects: self.objects.len() as i32,
            n_children: self.children.len() as i32,
            n_meshes: self.meshes.len() as i32,
            n_materials: self.materials.len() as i32,
            n_textures: self.textures.len() as i32,
            n_images: self.images.len() as i32,
            n_noises: self.noises.len() as i32,
            objects: self.objects.as_ptr(),
            children: self.children.as_ptr(),
            meshes: self.meshes.as_ptr(),
            materials: self.materials.as_ptr(),
            textures: self.textures.as_ptr(),
            images: self.images.as_ptr(),
            noises: self.noises.as_ptr(),
        }
    }
}

// =====================================================================================================
// Hit implementors (10)
// =====================================================================================================

// object/list.rs:15-24 — children in LIST ORDER (ties and sky / sun precedence depend on it, Q13)
impl ObjectList {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<ObjectId, Unsupported> {
        let mut kids = Vec::with_capacity(self.objects.len());
        for o in &self.objects {
            kids.push(b.object(o)?);
        }
        let mut o = RtgpuObject::blank(OBJ_LIST);
        if self.disable_bounds_check {
            o.flags |= OBJF_DISABLE_BOUNDS_CHECK;
        }
        Ok(b.push_object(o, &kids))
    }
}

// object/bvh.rs:26-29 — (children.0, children.1), tested in that order (bvh.rs:92-98)
impl BoundingVolumeHierarchyNode {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<ObjectId, Unsupported> {
        let kids = [b.object(&self.children.0)?, b.object(&self.children.1)?];
        Ok(b.push_object(RtgpuObject::blank(OBJ_BVH), &kids))
    }
}

// object/bvh/null_obj.rs:13
impl NullObject {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<ObjectId, Unsupported> {
        Ok(b.push_object(RtgpuObject::blank(OBJ_NULL), &[]))
    }
}

// object/transform.rs:14-19 — BOTH matrices as accumulated (transform.rs:44-92), never re-inverted
impl Transform {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<ObjectId, Unsupported> {
        let kid = b.object(&self.object)?;
        let mut o = RtgpuObject::blank(OBJ_TRANSFORM);
        o.transform = mat16(&self.transform);
        o.inv_transform = mat16(&self.inv_transform);
        Ok(b.push_object(o, &[kid]))
    }
}

// object/mesh.rs:16-35 — three independently indexed attribute arrays; the octree is NOT handed over (the device
// builds its own wide BVH; first-hit ids stay comparable because triangles keep their index)
impl TriangleMesh {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<ObjectId, Unsupported> {
        let material = b.material(&self.material)?;
        let mut d = MeshData { positions: vec![], normals: vec![], uvs: vec![], tri_pos: vec![], tri_nrm: vec![], tri_uv: vec![] };
        for v in &self.vertices {
            d.positions.extend_from_slice(&[v.x(), v.y(), v.z()]);
        }
        for n in &self.vertex_normals {
            d.normals.extend_from_slice(&[n.x(), n.y(), n.z()]);
        }
        for t in &self.vertex_uvs {
            d.uvs.extend_from_slice(&[t.x(), t.y()]);
        }
        for t in &self.triangles {
            for k in 0..3 {
                d.tri_pos.push(t.vert_indices[k] as i32);
                d.tri_nrm.push(t.normal_indices[k] as i32);
                d.tri_uv.push(t.uv_indices.map_or(-1, |uv| uv[k] as i32));
            }
        }
        let m = RtgpuMesh {
            positions: d.positions.as_ptr(),
            normals: d.normals.as_ptr(),
            uvs: if d.uvs.is_empty() { std::ptr::null() } else { d.uvs.as_ptr() },
            tri_pos: d.tri_pos.as_ptr(),
            tri_nrm: d.tri_nrm.as_ptr(),
            tri_uv: d.tri_uv.as_ptr(),
            n_positions: self.vertices.len() as i32,
            n_normals: self.vertex_normals.len() as i32,
            n_uvs: self.vertex_uvs.len() as i32,
            n_triangles: self.triangles.len() as i32,
        };
        b.mesh_data.push(d); // (Vec buffers do not move when the MeshData does)
        b.meshes.push(m);
        let mut o = RtgpuObject::blank(OBJ_MESH);
        o.mesh = b.meshes.len() as i32 - 1;
        o.material = material;
        if self.flat_shading {
            o.flags |= OBJF_FLAT_SHADING;
        }
        if self.hit_back_faces {
            o.flags |= OBJF_HIT_BACK_FACES;
        }
        Ok(b.push_object(o, &[]))
    }
}

// object/sphere.rs:18-24 — radius may be negative (golden_monkey.rs:113)
impl Sphere {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<ObjectId, Unsupported> {
        let mut o = RtgpuObject::blank(OBJ_SPHERE);
        o.material = b.material(&self.material)?;
        o.p0 = xyz(&self.center);
        o.p1[0] = self.radius;
        Ok(b.push_object(o, &[]))
    }
}

// object/plane.rs:14-26 — the ABI takes what Plane::new took: centre and the two HALF extents
// (corner = centre - u - v, plane.rs:41-51)
impl Plane {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<ObjectId, Unsupported> {
        let mut o = RtgpuObject::blank(OBJ_PLANE);
        o.material = b.material(&self.material)?;
        o.p0 = xyz(&(self.corner + self.u + self.v));
        o.p1 = xyz(&self.u);
        o.p2 = xyz(&self.v);
        if self.render_backface {
            o.flags |= OBJF_BACKFACE;
        }
        Ok(b.push_object(o, &[]))
    }
}

// object/volume.rs:15-19 — density = -1 / neg_inv_density; the boundary must flatten to spheres / planes (possibly
// under lists and transforms): rtgpu_scene_create answers RTGPU_ERR_UNSUPPORTED for a mesh boundary
impl Volume {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<ObjectId, Unsupported> {
        let boundary = b.object(&self.boundary)?;
        let mut o = RtgpuObject::blank(OBJ_VOLUME);
        o.material = b.material(&self.material)?;
        o.p1[0] = -1.0 / self.neg_inv_density;
        Ok(b.push_object(o, &[boundary]))
    }
}

// object/sky.rs:16-18 — the private Emissive's emission map (emissive.rs:13-15 needs `pub(crate) emission_map`)
impl Sky {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<ObjectId, Unsupported> {
        let mut o = RtgpuObject::blank(OBJ_SKY);
        o.texture = b.texture(&self.material.emission_map)?;
        Ok(b.push_object(o, &[]))
    }
}

// object/sun.rs:17-20 — direction AFTER to_unit(), all four components (the DSL builds it with Vec4::point, so
// w = 1 / len, loaders/scene.rs:843; Q5)
impl Sun {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<ObjectId, Unsupported> {
        let mut o = RtgpuObject::blank(OBJ_SUN);
        o.texture = b.texture(&self.material.emission_map)?;
        o.p0 = xyz(&self.direction);
        Ok(b.push_object(o, &[]))
    }
}
