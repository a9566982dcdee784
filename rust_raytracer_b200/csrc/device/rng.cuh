// rng.cuh — counter-based random numbers for the path tracer.
//
// The reference owns one Pcg64Mcg per render thread, seeded from OS entropy
// (src/camera.rs:208-209): streams are unpinned, only the DISTRIBUTIONS and the draw order
// per event are part of the contract. Here every (pixel, sample,
// bounce) owns a Philox4x32-10 stream addressed by its coordinates, so any GPU can render
// any subset of samples and the image does not depend on how work is scheduled.
#pragma once
#include "dmath.cuh"

namespace rt {

RT_HD uint32_t mulhi32(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return (uint32_t)(((uint64_t)a * b) >> 32);
#endif
}

// Philox4x32-10 (Salmon et al. 2011), the published constants.
RT_HD void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4]) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; r++) {
        const uint32_t hi0 = mulhi32(M0, c0), lo0 = M0 * c0;
        const uint32_t hi1 = mulhi32(M1, c2), lo1 = M1 * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// Draw stream of one path vertex: blocks of 4 x 32 bits generated on demand.
//   counter = (pixel, sample, bounce * 16 + block, domain), key = seed
// REG = false: the block is an array indexed by `have` — the whole stream object then lives in LOCAL memory (a dynamically
// indexed array cannot be in registers): a few LDL / STL per draw, but no register held between draws.
// REG = true: the block is four registers shifted per draw: no local memory at all (frame of k_shade<0, CTEX> 56 -> 16 bytes).
// Same words in the same order. Measured (profiles/r02_tune31, r02_tune32): the register form wins in the kernels compiled without the
// texture interpreter (-1 % on cornell_dragon) and LOSES where registers are short (general shade kernels on earth: +2 %, spills),
// so k_shade<..., CTEX> and the pinhole k_raygen use REG = true and everything else the array.
template <bool REG>
struct PhiloxStreamT {
    uint32_t pixel, sample, bounce, domain, k0, k1;
    uint32_t buf[REG ? 1 : 4];
    uint32_t b0, b1, b2, b3;
    int have, block;
    RT_HD PhiloxStreamT(uint64_t seed, uint32_t pixel_, uint32_t sample_, uint32_t bounce_, uint32_t domain_)
        : pixel(pixel_), sample(sample_), bounce(bounce_), domain(domain_), k0((uint32_t)seed), k1((uint32_t)(seed >> 32)),
          b0(0), b1(0), b2(0), b3(0), have(0), block(0) {}
    RT_HD uint32_t next_u32() {
        if (REG) {
            if (have == 0) {
                uint32_t o[4];
                philox4x32_10(pixel, sample, bounce * 16u + (uint32_t)block, domain, k0, k1, o);
                b0 = o[0]; b1 = o[1]; b2 = o[2]; b3 = o[3];
                have = 4; block++;
            }
            have--;
            const uint32_t r = b0; // words of a block are handed out in order
            b0 = b1; b1 = b2; b2 = b3;
            return r;
        }
        if (have == 0) { philox4x32_10(pixel, sample, bounce * 16u + (uint32_t)block, domain, k0, k1, buf); have = 4; block++; }
        return buf[4 - have--];
    }
    // rand `Standard` f64 is [0,1); here 24 bits -> float [0,1)
    RT_HD float uniform() { return (float)(next_u32() >> 8) * (1.0f / 16777216.0f); }
    // f64 uniform in the OPEN interval (0,1) from 32 bits: the free-flight draw of Volume::test takes its logarithm
    RT_HD double uniform_open_d() { return ((double)next_u32() + 0.5) * (1.0 / 4294967296.0); }
    // StandardNormal: distribution only (rand_distr uses a ziggurat); Box-Muller, one value per
    // call to keep the per-event draw order of the reference
    RT_HD float normal() {
        const float u1 = ((float)(next_u32() >> 8) + 0.5f) * (1.0f / 16777216.0f);
        const float u2 = (float)(next_u32() >> 8) * (1.0f / 16777216.0f);
#if defined(__CUDA_ARCH__)
        float s, c;
        __sincosf(2.0f * RT_PI_F * u2, &s, &c);
        // u1 can round to exactly 1.0f for the largest draws, and __logf is an approximation (~2^-21 absolute near 1):
        // a slightly positive logarithm must not become a NaN direction
        return sqrtf(fmaxf(0.0f, -2.0f * __logf(u1))) * c;
#else
        return sqrtf(fmaxf(0.0f, -2.0f * logf(u1))) * cosf(2.0f * RT_PI_F * u2);
#endif
    }
};
using PhiloxStream = PhiloxStreamT<false>;

// Injected draws for the per-event parity hook: uniforms and normals come from the caller's
// array in the reference's draw order (the oracle consumes the same array).
struct InjectedStream {
    const double* xi;
    int n, pos;
    RT_HD InjectedStream(const double* p, int n_) : xi(p), n(n_), pos(0) {}
    RT_HD float uniform() { return pos < n ? (float)xi[pos++] : 0.5f; }
    RT_HD float normal() { return pos < n ? (float)xi[pos++] : 0.0f; }
    RT_HD double uniform_open_d() { return pos < n ? xi[pos++] : 0.5; }
};

} // namespace rt
