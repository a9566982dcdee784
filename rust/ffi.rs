//! src/gpu/ffi.rs — `#[repr(C)]` mirror of include/rtgpu.h and the `extern "C"` block that binds librtgpu.so.
//!
//! UNCOMPILED: the build image has no Rust toolchain (`cargo`, `rustc` absent). The same ABI is exercised for real by
//! the ctypes mirror rust_raytracer_b200/_abi.py (tests/test_abi.py checks struct sizes and every exported symbol).
//! Field order, types and constants follow include/rtgpu.h line by line; keep the two in step.
//!
//! build.rs of the crate:
//!     println!("cargo:rustc-link-search=native={}", std::env::var("RTGPU_LIB_DIR").unwrap());
//!     println!("cargo:rustc-link-lib=dylib=rtgpu");
#![allow(non_camel_case_types, dead_code)]

use std::ffi::CStr;
use std::os::raw::{c_char, c_int, c_void};

pub const RTGPU_ABI_VERSION: u32 = 1;

// RtgpuStatus
pub const RTGPU_OK: c_int = 0;
pub const RTGPU_ERR_INVALID: c_int = -1;
pub const RTGPU_ERR_NO_DEVICE: c_int = -2;
pub const RTGPU_ERR_CUDA: c_int = -3;
pub const RTGPU_ERR_UNSUPPORTED: c_int = -4;
pub const RTGPU_ERR_IO: c_int = -5;
pub const RTGPU_ERR_PARSE: c_int = -6;

// RtgpuObjectKind — one per `Hit` implementor (src/object.rs:107-115)
pub const OBJ_LIST: i32 = 0; // ObjectList                     src/object/list.rs
pub const OBJ_BVH: i32 = 1; // BoundingVolumeHierarchyNode    src/object/bvh.rs
pub const OBJ_NULL: i32 = 2; // NullObject                     src/object/bvh/null_obj.rs
pub const OBJ_TRANSFORM: i32 = 3; // Transform                 src/object/transform.rs
pub const OBJ_MESH: i32 = 4; // TriangleMesh                   src/object/mesh.rs
pub const OBJ_SPHERE: i32 = 5; // Sphere                       src/object/sphere.rs
pub const OBJ_PLANE: i32 = 6; // Plane                         src/object/plane.rs
pub const OBJ_VOLUME: i32 = 7; // Volume                       src/object/volume.rs
pub const OBJ_SKY: i32 = 8; // Sky                             src/object/sky.rs
pub const OBJ_SUN: i32 = 9; // Sun                             src/object/sun.rs

pub const OBJF_DISABLE_BOUNDS_CHECK: u32 = 1 << 0; // ObjectList.disable_bounds_check  list.rs:23
pub const OBJF_BACKFACE: u32 = 1 << 1; // Plane.render_backface             plane.rs:16
pub const OBJF_FLAT_SHADING: u32 = 1 << 2; // TriangleMesh.flat_shading     mesh.rs:25
pub const OBJF_HIT_BACK_FACES: u32 = 1 << 3; // TriangleMesh.hit_back_faces mesh.rs:26

// RtgpuMaterialKind — one per `Material` implementor (src/material.rs:38-47)
pub const MAT_LAMBERTIAN: i32 = 0;
pub const MAT_METAL: i32 = 1;
pub const MAT_DIELECTRIC: i32 = 2;
pub const MAT_GLOSSY: i32 = 3;
pub const MAT_EMISSIVE: i32 = 4;
pub const MAT_ISOTROPIC: i32 = 5;
pub const MAT_NORMAL_DEBUG: i32 = 6;

// RtgpuTextureKind — one per `Sampler` implementor (src/texture.rs:21-25)
pub const TEX_CONSTANT: i32 = 0;
pub const TEX_CHECKER: i32 = 1;
pub const TEX_CHECKER_SOLID: i32 = 2;
pub const TEX_IMAGE: i32 = 3;
pub const TEX_LERP: i32 = 4;
pub const TEX_NOISE_SOLID: i32 = 5;
pub const TEX_CHANNEL: i32 = 6;
pub const TEX_UV_DEBUG: i32 = 7;

pub const IMG_RGB8: i32 = 0;
pub const IMG_RGB16: i32 = 1;
pub const IMG_RGBF32: i32 = 2;

pub const FLAG_EXACT: u32 = 1 << 0;
pub const FLAG_GUARD_PDF: u32 = 1 << 1;
pub const FLAG_PROFILE: u32 = 1 << 2;

#[repr(C)]
#[derive(Clone, Copy)]
pub struct RtgpuObject {
    pub kind: i32,
    pub material: i32,
    pub texture: i32,
    pub mesh: i32,
    pub first_child: i32,
    pub n_children: i32,
    pub flags: u32,
    pub reserved: i32,
    pub p0: [f64; 4],
    pub p1: [f64; 4],
    pub p2: [f64; 4],
    pub transform: [f64; 16],
    pub inv_transform: [f64; 16],
}

impl RtgpuObject {
    pub fn blank(kind: i32) -> Self {
        let mut id = [0.0; 16];
        id[0] = 1.0;
        id[5] = 1.0;
        id[10] = 1.0;
        id[15] = 1.0;
        RtgpuObject {
            kind,
            material: -1,
            texture: -1,
            mesh: -1,
            first_child: 0,
            n_children: 0,
            flags: 0,
            reserved: 0,
            p0: [0.0; 4],
            p1: [0.0; 4],
            p2: [0.0; 4],
            transform: id,
            inv_transform: id,
        }
    }
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct RtgpuMesh {
    pub positions: *const f64,
    pub normals: *const f64,
    pub uvs: *const f64,
    pub tri_pos: *const i32,
    pub tri_nrm: *const i32,
    pub tri_uv: *const i32,
    pub n_positions: i32,
    pub n_normals: i32,
    pub n_uvs: i32,
    pub n_triangles: i32,
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct RtgpuMaterial {
    pub kind: i32,
    pub albedo: i32,
    pub roughness: i32,
    pub normal_map: i32,
    pub ior: f64,
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct RtgpuTexture {
    pub kind: i32,
    pub is_float: i32,
    pub a: i32,
    pub b: i32,
    pub c: i32,
    pub image: i32,
    pub noise: i32,
    pub channel: i32,
    pub samples: i32,
    pub repeat: i32,
    pub scale: f64,
    pub value: [f64; 4],
}

impl RtgpuTexture {
    pub fn blank(kind: i32, is_float: bool) -> Self {
        RtgpuTexture {
            kind,
            is_float: is_float as i32,
            a: -1,
            b: -1,
            c: -1,
            image: -1,
            noise: -1,
            channel: 0,
            samples: 7,
            repeat: 1,
            scale: 1.0,
            value: [0.0; 4],
        }
    }
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct RtgpuImage {
    pub width: i32,
    pub height: i32,
    pub format: i32,
    pub reserved: i32,
    pub data: *const c_void,
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct RtgpuNoise {
    pub gradients: [[f64; 3]; 256],
    pub perm_x: [i32; 256],
    pub perm_y: [i32; 256],
    pub perm_z: [i32; 256],
}

#[repr(C)]
pub struct RtgpuSceneDesc {
    pub abi_version: u32,
    pub world_root: i32,
    pub lights_root: i32,
    pub n_objects: i32,
    pub n_children: i32,
    pub n_meshes: i32,
    pub n_materials: i32,
    pub n_textures: i32,
    pub n_images: i32,
    pub n_noises: i32,
    pub objects: *const RtgpuObject,
    pub children: *const i32,
    pub meshes: *const RtgpuMesh,
    pub materials: *const RtgpuMaterial,
    pub textures: *const RtgpuTexture,
    pub images: *const RtgpuImage,
    pub noises: *const RtgpuNoise,
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct RtgpuCameraDesc {
    pub width: i32,
    pub height: i32,
    pub position: [f64; 3],
    pub first_pixel: [f64; 3],
    pub pixel_delta_u: [f64; 3],
    pub pixel_delta_v: [f64; 3],
    pub basis_u: [f64; 3],
    pub basis_v: [f64; 3],
    pub aperture_radius: f64, // < 0: aperture_radius == None (camera.rs:125-129)
    pub background: [f64; 3],
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct RtgpuRenderParams {
    pub replicas: i32,  // thread_count (config.rs:74)
    pub sqrt_spt: i32,  // sqrt_samples_per_thread (config.rs:154-155)
    pub max_depth: i32, // config.rs:76
    pub n_gpus: i32,
    pub light_bias: f64, // config.rs:77
    pub seed: u64,
    pub flags: u32,
    pub sample_begin: i32,
    pub sample_end: i32,
    pub reserved: i32,
}

#[repr(C)]
pub struct RtgpuScene {
    _opaque: [u8; 0],
}

extern "C" {
    pub fn rtgpu_init(devices: *const c_int, n_devices: c_int) -> c_int;
    /// call first thing in main(): driver + context bring-up overlaps the "Ready" phase (main.rs:28-61)
    pub fn rtgpu_init_async(devices: *const c_int, n_devices: c_int) -> c_int;
    pub fn rtgpu_shutdown() -> c_int;
    pub fn rtgpu_device_count() -> c_int;
    pub fn rtgpu_last_error() -> *const c_char;
    pub fn rtgpu_scene_create(desc: *const RtgpuSceneDesc, out_scene: *mut *mut RtgpuScene) -> c_int;
    pub fn rtgpu_scene_destroy(scene: *mut RtgpuScene) -> c_int;
    /// Camera::render (camera.rs:189-256): ADDS the per-pixel mean into out_rgba (W*H*4 f64 = Buffer.data)
    pub fn rtgpu_render(
        scene: *mut RtgpuScene,
        cam: *const RtgpuCameraDesc,
        params: *const RtgpuRenderParams,
        out_rgba: *mut f64,
    ) -> c_int;
    /// Camera::render + Writer::save maths (main.rs:74-82) without leaving the device: RGB8 out
    pub fn rtgpu_render_srgb8(
        scene: *mut RtgpuScene,
        cam: *const RtgpuCameraDesc,
        params: *const RtgpuRenderParams,
        tonemap: c_int,
        out_rgb: *mut u8,
    ) -> c_int;
    pub fn rtgpu_tonemap_srgb8(rgba: *const f64, width: c_int, height: c_int, tonemap: c_int, out_rgb: *mut u8) -> c_int;
}

/// message of the last failure on this thread
pub fn last_error() -> String {
    unsafe {
        let p = rtgpu_last_error();
        if p.is_null() {
            String::new()
        } else {
            CStr::from_ptr(p).to_string_lossy().into_owned()
        }
    }
}

#[derive(Debug)]
pub struct GpuError {
    pub code: c_int,
    pub message: String,
}

impl std::fmt::Display for GpuError {
    fn fmt(&self, f: &mut std::fmt::Formatter<'_>) -> std::fmt::Result {
        write!(f, "rtgpu error {}: {}", self.code, self.message)
    }
}

impl std::error::Error for GpuError {}

pub fn check(code: c_int) -> Result<(), GpuError> {
    if code == RTGPU_OK {
        Ok(())
    } else {
        Err(GpuError { code, message: last_error() })
    }
}
