//! src/gpu/flatten_materials.rs — second half of the `flatten()` visitor (see flatten.rs): the 7 `Material`
//! implementors, the 8 `Sampler` implementors and `PerlinNoise3D`. UNCOMPILED (no Rust toolchain in the build image).
//! Each `impl` goes into the file of its type; the cited lines are where the fields it reads are declared.
#![allow(dead_code)]

use super::ffi::*;
use super::flatten::{MaterialId, NoiseId, SceneBuilder, TextureId, Unsupported};
use crate::vec4::{Color, Vec4};

// =====================================================================================================
// Material implementors (7) — src/material/*.rs
// =====================================================================================================

// material/lambertian.rs:14-16
impl LambertianDiffuse {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<MaterialId, Unsupported> {
        let albedo = b.texture(&self.albedo)?;
        Ok(b.push_material(MAT_LAMBERTIAN, albedo, -1, -1, 1.0))
    }
}

// material/metal.rs:13-16
impl Metal {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<MaterialId, Unsupported> {
        let albedo = b.texture(&self.albedo)?;
        let roughness = b.texture(&self.roughness)?;
        Ok(b.push_material(MAT_METAL, albedo, roughness, -1, 1.0))
    }
}

// material/dielectric.rs:12-14
impl Dielectric {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<MaterialId, Unsupported> {
        Ok(b.push_material(MAT_DIELECTRIC, -1, -1, -1, self.ior))
    }
}

// material/glossy.rs:17-23 — the struct keeps 1 / ior; the ABI takes ior (the device recomputes inv_ior = 1 / ior)
impl Glossy {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<MaterialId, Unsupported> {
        let albedo = b.texture(&self.albedo)?;
        let roughness = b.texture(&self.roughness)?;
        let normal_map = match &self.normal_map {
            Some(t) => b.texture(t)?,
            None => -1,
        };
        Ok(b.push_material(MAT_GLOSSY, albedo, roughness, normal_map, 1.0 / self.inv_ior))
    }
}

// material/emissive.rs:13-15
impl Emissive {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<MaterialId, Unsupported> {
        let emission = b.texture(&self.emission_map)?;
        Ok(b.push_material(MAT_EMISSIVE, emission, -1, -1, 1.0))
    }
}

// material/isotropic.rs:14-16
impl Isotropic {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<MaterialId, Unsupported> {
        let albedo = b.texture(&self.albedo)?;
        Ok(b.push_material(MAT_ISOTROPIC, albedo, -1, -1, 1.0))
    }
}

// material/normal_debug.rs:14-16
impl NormalDebug {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<MaterialId, Unsupported> {
        let normal_map = match &self.normal_map {
            Some(t) => b.texture(t)?,
            None => -1,
        };
        Ok(b.push_material(MAT_NORMAL_DEBUG, -1, -1, normal_map, 1.0))
    }
}

// =====================================================================================================
// Sampler implementors (8) — src/texture/*.rs. `Output` is Color (Vec4) or f64; the table entry says which
// (is_float) because checker / lerp nodes are generic over it.
// =====================================================================================================

/// how a Sampler::Output becomes RtgpuTexture.value / is_float
pub trait TexValue: Copy {
    const IS_FLOAT: bool;
    fn value(self) -> [f64; 4];
}
impl TexValue for f64 {
    const IS_FLOAT: bool = true;
    fn value(self) -> [f64; 4] {
        [self, 0.0, 0.0, 0.0]
    }
}
impl TexValue for Color {
    const IS_FLOAT: bool = false;
    fn value(self) -> [f64; 4] {
        [self.x(), self.y(), self.z(), self.w()]
    }
}

// texture/constant.rs:8-10
impl<T: TexValue> ConstantTexture<T> {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<TextureId, Unsupported> {
        let mut t = RtgpuTexture::blank(TEX_CONSTANT, T::IS_FLOAT);
        t.value = self.value.value();
        Ok(b.push_texture(t))
    }
}

// texture/checkerboard.rs:8-12 — `(trunc(2u / scale) + trunc(2v / scale)) % 2` with saturating u32 casts (Q19)
impl<T: TexValue + Send + Sync + 'static> CheckerboardTexture<T> {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<TextureId, Unsupported> {
        let mut t = RtgpuTexture::blank(TEX_CHECKER, T::IS_FLOAT);
        t.a = b.texture(&self.even_squares)?;
        t.b = b.texture(&self.odd_squares)?;
        t.scale = self.scale;
        Ok(b.push_texture(t))
    }
}

// texture/checkerboard.rs:48-52
impl<T: TexValue + Send + Sync + 'static> CheckerboardSolidTexture<T> {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<TextureId, Unsupported> {
        let mut t = RtgpuTexture::blank(TEX_CHECKER_SOLID, T::IS_FLOAT);
        t.a = b.texture(&self.even_volumes)?;
        t.b = b.texture(&self.odd_volumes)?;
        t.scale = self.scale;
        Ok(b.push_texture(t))
    }
}

// texture/image.rs:15-18 — nearest neighbour, v not flipped, values as stored (no sRGB decode, Q6)
impl ImageTexture {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<TextureId, Unsupported> {
        let mut t = RtgpuTexture::blank(TEX_IMAGE, false);
        t.image = b.push_image(&self.buffer);
        t.repeat = match self.repeat {
            TextureRepeat::Repeat => 1,
            TextureRepeat::Clamp => 0,
        };
        Ok(b.push_texture(t))
    }
}

// texture/interpolate.rs:11-15 — exact 0 / 1 short-cuts are reproduced on the device (interpolate.rs:29-39)
impl<T: TexValue + Send + Sync + 'static> Interpolate<T> {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<TextureId, Unsupported> {
        let mut t = RtgpuTexture::blank(TEX_LERP, T::IS_FLOAT);
        t.a = b.texture(&self.start)?;
        t.b = b.texture(&self.end)?;
        t.c = b.texture(&self.t)?;
        Ok(b.push_texture(t))
    }
}

// texture/noise.rs:11-17 — `map` is a fn pointer and cannot be inspected: only the default map
// 0.5 (1 + sin(p.z + 10 turbulence)) (noise.rs:25) and a uniform scale (what the DSL sets, loaders/scene.rs:555-570)
// are representable; anything else is refused rather than rendered differently
impl NoiseSolidTexture {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<TextureId, Unsupported> {
        let s = self.scale;
        if s.x() != s.y() || s.x() != s.z() {
            return Err(Unsupported("noise_solid with a non-uniform scale"));
        }
        if self.map as usize != NoiseSolidTexture::DEFAULT_MAP as usize {
            return Err(Unsupported("noise_solid with a custom map function"));
        }
        let mut t = RtgpuTexture::blank(TEX_NOISE_SOLID, true);
        t.noise = self.noise.flatten(b)?;
        t.scale = s.x();
        t.samples = self.samples as i32;
        Ok(b.push_texture(t))
    }
}

// texture/channel.rs:8-11
impl Channel {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<TextureId, Unsupported> {
        let mut t = RtgpuTexture::blank(TEX_CHANNEL, true);
        t.a = b.texture(&self.color)?;
        t.channel = self.channel as i32;
        Ok(b.push_texture(t))
    }
}

// texture/uv_debug.rs:6
impl UvDebugTexture {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<TextureId, Unsupported> {
        Ok(b.push_texture(RtgpuTexture::blank(TEX_UV_DEBUG, false)))
    }
}

// =====================================================================================================
// Noise3D — noise/perlin.rs:11-16: 256 unit gradients + three Fisher-Yates permutations, generated on the host
// (they come from the scene RNG, perlin.rs:18-52) and only evaluated on the device
// =====================================================================================================
impl PerlinNoise3D {
    pub fn flatten_impl(&self, b: &mut SceneBuilder) -> Result<NoiseId, Unsupported> {
        Ok(b.push_noise(&self.random_vec, &self.perm_x, &self.perm_y, &self.perm_z))
    }
}
