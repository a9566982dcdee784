// texture.cuh — texture-expression evaluation on the device (reference src/texture/*.rs,
// src/noise/perlin.rs). Each texture is a postfix program over a small value stack; both
// branches of checker/lerp are evaluated and then selected, which yields the same value as
// the reference's lazy evaluation.
#pragma once
#include "dmath.cuh"

namespace rt {

RT_HD uint32_t float_as_u32_sat(float x) { // Rust `as u32`: saturating, NaN -> 0
#if defined(__CUDA_ARCH__)
    return __float2uint_rz(x);
#else
    if (!(x > 0.f)) return 0u;
    if (x >= 4294967296.f) return 0xFFFFFFFFu;
    return (uint32_t)x;
#endif
}
RT_HD int32_t float_floor_as_i32_sat(float x) { // `x.floor() as i32`
#if defined(__CUDA_ARCH__)
    return __float2int_rd(x);
#else
    float f = floorf(x);
    if (f != f) return 0;
    if (f >= 2147483648.f) return 2147483647;
    if (f <= -2147483648.f) return (int32_t)0x80000000;
    return (int32_t)f;
#endif
}

// noise/perlin.rs:79-99 + 54-73
RT_HD float perlin_sample(const DevNoise& nz, f3 p) {
    const float fx = floorf(p.x), fy = floorf(p.y), fz = floorf(p.z);
    const float u = p.x - fx, v = p.y - fy, w = p.z - fz;
    const int i = float_floor_as_i32_sat(p.x), j = float_floor_as_i32_sat(p.y), k = float_floor_as_i32_sat(p.z);
    const float uu = u * u * (3.f - 2.f * u), vv = v * v * (3.f - 2.f * v), ww = w * w * (3.f - 2.f * w);
    float acc = 0.f;
#pragma unroll
    for (int di = 0; di < 2; di++)
#pragma unroll
        for (int dj = 0; dj < 2; dj++)
#pragma unroll
            for (int dk = 0; dk < 2; dk++) {
                const uint32_t idx = (uint32_t)nz.perm_x[(uint32_t)(di + i) & 255u] ^ (uint32_t)nz.perm_y[(uint32_t)(dj + j) & 255u] ^
                                     (uint32_t)nz.perm_z[(uint32_t)(dk + k) & 255u];
                const F4 g = nz.grad[idx];
                const float fi = (float)di, fj = (float)dj, fk = (float)dk;
                acc += (fi * uu + (1.f - fi) * (1.f - uu)) * (fj * vv + (1.f - fj) * (1.f - vv)) * (fk * ww + (1.f - fk) * (1.f - ww)) *
                       (g.x * (u - fi) + g.y * (v - fj) + g.z * (w - fk));
            }
    return acc;
}

RT_HD float perlin_turbulence(const DevNoise& nz, f3 p, int samples) { // perlin.rs:101-113
    float acc = 0.f, weight = 1.f;
    for (int s = 0; s < samples; s++) {
        acc += weight * perlin_sample(nz, p);
        weight *= 0.5f;
        p = p * 2.0f;
    }
    return fabsf(acc);
}

RT_HD F4 image_fetch(const DevImage& im, float u, float v, int repeat) { // image.rs:40-53, buffer.rs:30-60
    if (repeat) { u = u - floorf(u); v = v - floorf(v); }
    else { u = fminf(fmaxf(u, 0.f), 1.f); v = fminf(fmaxf(v, 0.f), 1.f); }
    const float w = (float)im.width - 0.001f, h = (float)im.height - 0.001f;
    uint32_t x = float_as_u32_sat(u * w), y = float_as_u32_sat(v * h);
    if (x >= (uint32_t)im.width) x = (uint32_t)im.width - 1u;
    if (y >= (uint32_t)im.height) y = (uint32_t)im.height - 1u;
    const size_t idx = (size_t)y * (size_t)im.width + x;
    F4 r;
    if (im.format == 0) { // RGB8 stored as 4 bytes: into_rgb32f = byte / 255
        const uint32_t t = reinterpret_cast<const uint32_t*>(im.data)[idx];
        r.x = (float)(t & 255u) / 255.0f; r.y = (float)((t >> 8) & 255u) / 255.0f; r.z = (float)((t >> 16) & 255u) / 255.0f;
    } else if (im.format == 1) {
        const uint16_t* t = reinterpret_cast<const uint16_t*>(im.data) + idx * 4;
        r.x = (float)t[0] / 65535.0f; r.y = (float)t[1] / 65535.0f; r.z = (float)t[2] / 65535.0f;
    } else {
        const float* t = reinterpret_cast<const float*>(im.data) + idx * 4;
        r.x = t[0]; r.y = t[1]; r.z = t[2];
    }
    r.w = 0.f;
    return r;
}

// Sampler::sample(uv, p): colour in xyz, float textures in x
// CTEX: the caller guarantees that EVERY texture program of the scene is a single constant (api.cu: all_const_textures): the postfix
// interpreter below — its 128-byte value stack, the image fetch, the Perlin code — is compiled out of that kernel.
template <bool CTEX = false>
RT_HD F4 tex_eval(const DeviceScene& s, int tex, float u, float v, f3 p) {
    const TexProgram prog = s.tex_programs[tex];
    const TexOp* ops = s.tex_ops + prog.first;
    if (CTEX || (prog.count == 1 && ops[0].op == TOP_CONST)) { // by far the common case
        F4 r = {ops[0].value[0], ops[0].value[1], ops[0].value[2], ops[0].value[3]};
        return r;
    }
    F4 st[8];
    int sp = 0;
    for (int i = 0; i < prog.count; i++) {
        const TexOp op = ops[i];
        switch (op.op) {
        case TOP_CONST: st[sp].x = op.value[0]; st[sp].y = op.value[1]; st[sp].z = op.value[2]; st[sp].w = op.value[3]; sp++; break;
        case TOP_CHECKER: { // checkerboard.rs:34-44
            const uint32_t iu = float_as_u32_sat(u * 2.0f / op.value[0]), iv = float_as_u32_sat(v * 2.0f / op.value[0]);
            const bool even = ((iu + iv) & 1u) == 0u;
            sp--; // st[sp] = odd, st[sp-1] = even
            if (!even) st[sp - 1] = st[sp];
            break;
        }
        case TOP_CHECKER_SOLID: { // checkerboard.rs:74-86 (`%` keeps the sign: only an even sum is "even")
            const int32_t ix = float_floor_as_i32_sat(p.x / op.value[0]), iy = float_floor_as_i32_sat(p.y / op.value[0]),
                          iz = float_floor_as_i32_sat(p.z / op.value[0]);
            const uint32_t sum = (uint32_t)ix + (uint32_t)iy + (uint32_t)iz;
            const bool even = (sum & 1u) == 0u;
            sp--;
            if (!even) st[sp - 1] = st[sp];
            break;
        }
        case TOP_IMAGE: st[sp++] = image_fetch(s.images[op.arg], u, v, op.repeat); break;
        case TOP_LERP: { // interpolate.rs:29-39
            sp -= 2; // st[sp-1] = start, st[sp] = end, st[sp+1] = t
            const float t = st[sp + 1].x;
            const F4 a = st[sp - 1], b = st[sp];
            F4 r;
            if (t == 0.0f) r = a;
            else if (t == 1.0f) r = b;
            else { const float q = 1.0f - t; r.x = a.x * q + b.x * t; r.y = a.y * q + b.y * t; r.z = a.z * q + b.z * t; r.w = a.w * q + b.w * t; }
            st[sp - 1] = r;
            break;
        }
        case TOP_NOISE_SOLID: { // texture/noise.rs:20-38
            const f3 ps = p * op.value[0];
            const float turb = perlin_turbulence(s.noises[op.arg], ps, op.samples);
            st[sp].x = 0.5f * (1.0f + sinf(ps.z + 10.0f * turb)); st[sp].y = st[sp].z = st[sp].w = 0.f;
            sp++;
            break;
        }
        case TOP_CHANNEL: { // channel.rs
            const F4 c = st[sp - 1];
            const float val = op.arg == 0 ? c.x : op.arg == 1 ? c.y : op.arg == 2 ? c.z : c.w;
            st[sp - 1].x = val; st[sp - 1].y = st[sp - 1].z = st[sp - 1].w = 0.f;
            break;
        }
        case TOP_UV_DEBUG: st[sp].x = u; st[sp].y = v; st[sp].z = 0.5f; st[sp].w = 0.f; sp++; break;
        default: break;
        }
    }
    return st[0];
}

} // namespace rt
