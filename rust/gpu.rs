//! src/gpu.rs — the `--gpu` back end of rust_raytracer: owns the uploaded scene and replaces the ONE call
//! `camera.render(world, lights, &mut buf)` (main.rs:74-75). UNCOMPILED (no Rust toolchain in the build image);
//! the Python mirror of exactly this flow is rust_raytracer_b200/__main__.py, which the GPU tests run.
//!
//!     pub mod gpu;            // in lib.rs, next to the other modules (lib.rs:1-20)
//!     mod ffi; mod flatten; mod flatten_materials;   // src/gpu/{ffi,flatten,flatten_materials}.rs
//!
//! ---- the three edits to existing reference files -----------------------------------------------------------------
//!
//! 1. src/config.rs — `--gpu` is a bare flag and the flag regex (config.rs:63) only matches `-k=v`, so the loop at
//!    config.rs:81-152 falls through to `scene_name = arg`. Catch it first, and add two `-k=v` keys:
//!
//!        pub struct Config { ..., pub gpu: bool, pub n_gpus: usize, pub seed: u64 }
//!        for arg in args.skip(1) {
//!            if arg == "--gpu" { gpu = true; continue; }
//!            if let Some(groups) = arg_regex.captures(&arg) {
//!                match &groups[1] { ...,
//!                    "-gpus" => n_gpus = groups[2].parse().expect("--gpus must be a number"),
//!                    "-seed" => seed = groups[2].parse().expect("--seed must be a number"),
//!
//! 2. src/camera.rs — two getters over the private fields `init` computes (camera.rs:19-44,86-130): see
//!    `impl Camera` at the end of this file.
//!
//! 3. src/main.rs:74-75 —
//!
//!        if config_gpu { gpu::init_async(n_gpus); }          // first line of main(): overlaps "Ready" (main.rs:28-61)
//!        ...
//!        let mut buf = camera.create_buffer();
//!        if config_gpu {
//!            let scene = gpu::Scene::from_roots(&world, &lights)?;      // flatten() + rtgpu_scene_create
//!            scene.render(&camera, n_gpus, seed, &mut buf)?;            // rtgpu_render, ADDS into buf like camera.rs:247-253
//!        } else {
//!            camera.render(world, lights, &mut buf);
//!        }
//!
//!    (`config` is moved into the scene builder at main.rs:30-59: copy `gpu`, `n_gpus`, `seed` out of it first.)

use std::sync::Arc;

mod ffi;
mod flatten;
mod flatten_materials;

use ffi::*;
pub use flatten::{SceneBuilder, Unsupported};

use crate::buffer::Buffer;
use crate::camera::Camera;
use crate::object::Hit;

/// rtgpu_init_async: driver initialisation and one context per device run on a helper thread of the library
pub fn init_async(n_gpus: usize) {
    let devices: Vec<i32> = (0..n_gpus.max(1) as i32).collect();
    unsafe {
        rtgpu_init_async(devices.as_ptr(), devices.len() as i32);
    }
}

/// An uploaded scene: flat tables + wide BVH on every selected device (rtgpu_scene_create .. rtgpu_scene_destroy)
pub struct Scene {
    handle: *mut RtgpuScene,
}

// the handle is only used through &self / &mut self; the library serialises internally
unsafe impl Send for Scene {}

impl Scene {
    /// `world`, `lights`: the two roots of SceneData (scene.rs:30). A construct the GPU path does not cover
    /// (a mesh as a volume boundary, a custom noise map) is an error here, never a silent CPU fallback.
    pub fn from_roots(world: &Arc<dyn Hit>, lights: &Arc<dyn Hit>) -> Result<Self, Box<dyn std::error::Error>> {
        let mut b = SceneBuilder::new();
        let w = b.object(world).map_err(|u| format!("--gpu: unsupported scene construct: {}", u.0))?;
        let l = b.object(lights).map_err(|u| format!("--gpu: unsupported scene construct: {}", u.0))?;
        let desc = b.desc(w, l);
        let mut handle: *mut RtgpuScene = std::ptr::null_mut();
        check(unsafe { rtgpu_scene_create(&desc, &mut handle) })?; // copies every table: `b` may go after this line
        Ok(Scene { handle })
    }

    /// Camera::render (camera.rs:189-256): same spp semantics (thread_count x sqrt_spt^2), ADDS the per-pixel mean into
    /// `buf` (Buffer.data is Vec<Vec4>, Vec4 = [f64; 4]: already the W*H*4 f64 layout the ABI writes, buffer.rs:8-12)
    pub fn render(&self, camera: &Camera, n_gpus: usize, seed: u64, buf: &mut Buffer) -> Result<(), GpuError> {
        let cam = camera.gpu_desc();
        let mut params = camera.gpu_params();
        params.n_gpus = n_gpus.max(1) as i32;
        params.seed = seed;
        let out = buf.data_mut().as_mut_ptr() as *mut f64; // needs `pub fn data_mut(&mut self) -> &mut Vec<Color>` in buffer.rs
        check(unsafe { rtgpu_render(self.handle, &cam, &params, out) })
    }

    /// Camera::render + Writer::save's tonemap / sRGB / u8 (main.rs:74-82) without the f64 Buffer crossing PCIe twice
    pub fn render_srgb8(&self, camera: &Camera, n_gpus: usize, seed: u64, aces: bool) -> Result<Vec<u8>, GpuError> {
        let cam = camera.gpu_desc();
        let mut params = camera.gpu_params();
        params.n_gpus = n_gpus.max(1) as i32;
        params.seed = seed;
        let mut rgb = vec![0u8; cam.width as usize * cam.height as usize * 3];
        check(unsafe { rtgpu_render_srgb8(self.handle, &cam, &params, aces as i32, rgb.as_mut_ptr()) })?;
        Ok(rgb)
    }
}

impl Drop for Scene {
    fn drop(&mut self) {
        unsafe {
            rtgpu_scene_destroy(self.handle);
        }
    }
}

// ---- src/camera.rs: what the GPU needs of the camera after Camera::init (camera.rs:86-130) ------------------------
impl Camera {
    pub fn gpu_desc(&self) -> RtgpuCameraDesc {
        let v3 = |v: &crate::vec4::Vec4| [v.x(), v.y(), v.z()];
        let bg = self.background_color.unwrap(); // always Some after the config merge (config.rs:28,41; Q21)
        RtgpuCameraDesc {
            width: self.image_width as i32,
            height: self.image_height as i32,
            position: v3(&self.position),
            first_pixel: v3(&self.first_pixel),
            pixel_delta_u: v3(&self.pixel_delta.0),
            pixel_delta_v: v3(&self.pixel_delta.1),
            basis_u: v3(&self.basis[0]), // u = v_up x w, NOT normalised (camera.rs:101-104, Q3)
            basis_v: v3(&self.basis[1]),
            aperture_radius: self.aperture_radius.unwrap_or(-1.0),
            background: v3(&bg),
        }
    }

    pub fn gpu_params(&self) -> RtgpuRenderParams {
        RtgpuRenderParams {
            replicas: self.thread_count as i32, // -t is the replica count: spp = T x k^2 stays identical (Q18)
            sqrt_spt: self.sqrt_spt as i32,
            max_depth: self.max_depth as i32,
            n_gpus: 1,
            light_bias: self.light_bias,
            seed: 1,
            flags: 0,
            sample_begin: 0,
            sample_end: 0,
            reserved: 0,
        }
    }
}

// ---- the trait methods (one line per implementor, in the implementor's own file) ----------------------------------
//
//     impl Hit for Sphere      { ...; fn flatten(&self, b: &mut SceneBuilder) -> Result<ObjectId, Unsupported> { self.flatten_impl(b) } }
//     impl Material for Metal  { ...; fn flatten(&self, b: &mut SceneBuilder) -> Result<MaterialId, Unsupported> { self.flatten_impl(b) } }
//     impl<T> Sampler for CheckerboardTexture<T> { ...; fn flatten(...) -> Result<TextureId, Unsupported> { self.flatten_impl(b) } }
//     impl Noise3D for PerlinNoise3D { ...; fn flatten(...) -> Result<NoiseId, Unsupported> { self.flatten_impl(b) } }
//
// and the defaulted trait method for third-party implementors: `Err(Unsupported("this object has no GPU description"))`.
