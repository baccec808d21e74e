/*
 * ocean_noise.h -- seeded, order-independent Gaussian noise for the ocean spectrum.
 *
 * Replaces WaterBody.GenerateRandomNumber / GenerateRandomNoiseTexture
 * (reference Assets/Scripts/Water/WaterBody.cs:71-100).  The reference draws from Unity's
 * unseeded Random.Range, so its noise is not reproducible; here the *shape* of the generator is
 * kept (Marsaglia polar rejection loop, two normals per texel, r then g, x-outer / y-inner draw
 * order WB:90-96) but the uniform source is a counter-based hash so that every texel can be
 * produced independently (on the host or by one GPU thread) with bit-identical results.
 *
 * The same header is compiled by gcc (oracle, host) and nvcc (device kernel).  Every floating
 * point operation below is either exact, a single IEEE-754 correctly rounded op, or an explicit
 * fmaf, and the logarithm is a private polynomial -- no libm / CUDA-libm call -- which is what
 * makes host and device agree to the bit.
 */
#ifndef OCEAN_NOISE_H
#define OCEAN_NOISE_H

#include <stdint.h>

#if defined(__CUDACC__)
#define OCEAN_HD __host__ __device__ __forceinline__
#else
#define OCEAN_HD static inline
#endif

#if defined(__CUDA_ARCH__)
#define OCEAN_FMUL(a, b) __fmul_rn((a), (b))
#define OCEAN_FADD(a, b) __fadd_rn((a), (b))
#define OCEAN_FSUB(a, b) __fsub_rn((a), (b))
#define OCEAN_FDIV(a, b) __fdiv_rn((a), (b))
#define OCEAN_FSQRT(a) __fsqrt_rn((a))
#define OCEAN_FMA(a, b, c) __fmaf_rn((a), (b), (c))
#else
#include <math.h>
/* host: built with -ffp-contract=off, so plain operators are single rounded ops */
#define OCEAN_FMUL(a, b) ((float)((float)(a) * (float)(b)))
#define OCEAN_FADD(a, b) ((float)((float)(a) + (float)(b)))
#define OCEAN_FSUB(a, b) ((float)((float)(a) - (float)(b)))
#define OCEAN_FDIV(a, b) ((float)((float)(a) / (float)(b)))
#define OCEAN_FSQRT(a) sqrtf((a))
#define OCEAN_FMA(a, b, c) fmaf((a), (b), (c))
#endif

/* splitmix64 finaliser: a bijective 64-bit mixer */
OCEAN_HD uint64_t ocean_mix64(uint64_t z) {
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27; z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}

/* Uniform in [0,1) with 24 random bits, keyed on (seed, draw counter, rejection attempt, lane).  The first mixing round
 * depends on (seed, counter) only: ocean_counter_key() evaluates it once per draw and the rejection loop of ocean_gaussian
 * re-uses it for every attempt and both lanes (same bits as hashing it again each time, a third fewer integer multiplies --
 * on the GPU a warp runs the loop until its slowest lane accepts, ~3 iterations per normal). */
OCEAN_HD uint64_t ocean_counter_key(uint64_t seed, uint64_t counter) {
    return ocean_mix64(seed + 0x9E3779B97F4A7C15ull * (counter + 1));
}
OCEAN_HD float ocean_uniform_keyed(uint64_t key, uint32_t attempt, uint32_t lane) {
    const uint64_t z = ocean_mix64(key ^ (((uint64_t)attempt << 1 | lane) * 0xD6E8FEB86659FD93ull + 0x2545F4914F6CDD1Dull));
    return (float)(uint32_t)(z >> 40) * 5.9604644775390625e-8f; /* 2^-24, exact */
}
OCEAN_HD float ocean_uniform(uint64_t seed, uint64_t counter, uint32_t attempt, uint32_t lane) {
    return ocean_uniform_keyed(ocean_counter_key(seed, counter), attempt, lane);
}

OCEAN_HD uint32_t ocean_f2u(float f) { union { float f; uint32_t u; } c; c.f = f; return c.u; }
OCEAN_HD float ocean_u2f(uint32_t u) { union { float f; uint32_t u; } c; c.u = u; return c.f; }

/* Natural log for normal positive floats; ~1 ulp, identical bits on host and device. */
OCEAN_HD float ocean_logf(float x) {
    const float ln2_hi = 6.9313812256e-01f, ln2_lo = 9.0580006145e-06f;
    const float Lg1 = 0.66666662693f, Lg2 = 0.40000972152f, Lg3 = 0.28498786688f, Lg4 = 0.24279078841f;
    uint32_t ix = ocean_f2u(x);
    int32_t k = (int32_t)(ix >> 23) - 127;
    ix &= 0x007fffffu;
    uint32_t i = (ix + (0x95f64u << 3)) & 0x800000u;
    float m = ocean_u2f(ix | (i ^ 0x3f800000u));
    k += (int32_t)(i >> 23);
    float f = OCEAN_FSUB(m, 1.0f);
    float s = OCEAN_FDIV(f, OCEAN_FADD(2.0f, f));
    float dk = (float)k;
    float z = OCEAN_FMUL(s, s);
    float w = OCEAN_FMUL(z, z);
    float t1 = OCEAN_FMUL(w, OCEAN_FMA(w, Lg4, Lg2));
    float t2 = OCEAN_FMUL(z, OCEAN_FMA(w, Lg3, Lg1));
    float R = OCEAN_FADD(t2, t1);
    float hfsq = OCEAN_FMUL(OCEAN_FMUL(0.5f, f), f);
    /* k*ln2_hi - ((hfsq - (s*(hfsq+R) + k*ln2_lo)) - f) */
    float inner = OCEAN_FMA(s, OCEAN_FADD(hfsq, R), OCEAN_FMUL(dk, ln2_lo));
    return OCEAN_FSUB(OCEAN_FMUL(dk, ln2_hi), OCEAN_FSUB(OCEAN_FSUB(hfsq, inner), f));
}

/* One attempt of the WB:71-81 rejection loop for the draw keyed by `key`: (v1, s) and whether it is accepted (WB:77). */
OCEAN_HD int ocean_gaussian_attempt(uint64_t key, uint32_t attempt, float* v1, float* s) {
    *v1 = OCEAN_FSUB(OCEAN_FMUL(2.0f, ocean_uniform_keyed(key, attempt, 0u)), 1.0f);
    const float v2 = OCEAN_FSUB(OCEAN_FMUL(2.0f, ocean_uniform_keyed(key, attempt, 1u)), 1.0f);
    *s = OCEAN_FMA(*v1, *v1, OCEAN_FMUL(v2, v2));
    return !(*s >= 1.0f || *s == 0.0f);
}
/* ... and the transform of an accepted pair (WB:79-80). */
OCEAN_HD float ocean_gaussian_finish(float v1, float s) {
    s = OCEAN_FSQRT(OCEAN_FDIV(OCEAN_FMUL(-2.0f, ocean_logf(s)), s));
    return OCEAN_FMUL(v1, s);
}

/* One N(0,1) draw: the WB:71-81 rejection loop on the counter-based uniforms. */
OCEAN_HD float ocean_gaussian(uint64_t seed, uint64_t counter) {
    float v1, s;
    uint32_t attempt = 0;
    const uint64_t key = ocean_counter_key(seed, counter);
    while (!ocean_gaussian_attempt(key, attempt, &v1, &s)) ++attempt;
    return ocean_gaussian_finish(v1, s);
}

/*
 * Noise texel (x, y) of an N x N RG32F texture, channel ch (0 = r, 1 = g).
 * Draw counter follows the reference fill order WB:90-96: outer loop i = x, inner loop j = y,
 * r drawn before g  =>  counter = (x*N + y)*2 + ch.  Storage index is [y][x][ch] (SetPixel(x,y)).
 */
OCEAN_HD float ocean_noise_texel(uint64_t seed, uint32_t N, uint32_t x, uint32_t y, uint32_t ch) {
    return ocean_gaussian(seed, ((uint64_t)x * N + y) * 2u + ch);
}

#endif /* OCEAN_NOISE_H */
