// ocean_spectrum.cuh -- InitialSpectrum.compute:33-129 per texel, shared by the kernels of ocean_init.cu.
// Only for translation units built with -fmad=false: every fp32 operation rounds once, in the reference's order.
#pragma once
#include "ocean_internal.h"

#ifndef OCEAN_SPECTRUM_EXACT_POW
#define OCEAN_SPECTRUM_EXACT_POW 0
#endif

namespace ocean {

struct Texel {
    float2 h;     // h0(k)
    float4 wave;  // (kx, 1/|k|, kz, omega)
};

// Everything in IS:99-129 that depends on |k| only.  A texel and its mirror have the same |k| (kx and kz
// change sign, or stay put on the x = 0 / y = 0 Nyquist lines), so a thread evaluates this once per pair.
struct Radial {
    bool inBand;         // IS:114
    float kMag, w;       // |k|, omega(k)
    float twoTmaJonswap; // (2 * tma) * jonswap: the head of the product at IS:120
    float s, nf;         // spread exponent and its normalisation  IS:60-74
    float fade, dwdk;    // IS:95-97, IS:87-91
};

__device__ __forceinline__ Radial initial_spectrum_radial(float kx, float kz, const StateConsts& st, const CascadeConsts& cc,
                                                          const SpreadConsts& sc) {
    Radial r;
    const float kMag = sqrtf(kx * kx + kz * kz);          // length()  IS:112
    r.kMag = kMag;
    r.inBand = kMag >= cc.cutoffLow && kMag <= cc.cutoffHigh;  // IS:114
    if (!r.inBand) return r;
    const float g = st.gravity, D = st.depth, wp = st.peakOmega;
    const float w = sqrtf(g * kMag);                  // IS:33-35
    r.w = w;
    // TMACorrection IS:38-43
    const float wh = w * st.sqrtDepthOverG;
    float tma;
    if (wh <= 1.0f) tma = 0.5f * wh * wh;
    else if (wh < 2.0f) tma = 1.0f - 0.5f * (2.0f - wh) * (2.0f - wh);
    else tma = 1.0f;
    // JONSWAP IS:47-56
    const float sigma = w <= wp ? 0.07f : 0.09f;
    const float d = w - wp;
    const float rr = expf(-(d * d) / (2.0f * sigma * sigma * wp * wp));
    // pow(omega, 5), pow(omega_p / omega, 4) and pow(3.3, r) (IS:52-55) without the general powf: integer powers as multiplication
    // chains (each product rounds once: <= 2 ulp, the accuracy class of powf itself) and the constant base as one expf.  These three
    // were half of the kernel's powf calls, and the kernel is issue-bound (ncu: 88.6 % issue-active): -24 % instructions.
    // OCEAN_SPECTRUM_EXACT_POW=1 restores the three powf calls.
#if OCEAN_SPECTRUM_EXACT_POW
    const float jonswap = st.alpha * g * g / powf(w, 5.0f) * expf(-1.25f * powf(wp / w, 4.0f)) * powf(3.3f, rr);
#else
    const float w2 = w * w, w5 = w2 * w2 * w;
    const float q = wp / w, q2 = q * q, q4 = q2 * q2;
    const float jonswap = st.alpha * g * g / w5 * expf(-1.25f * q4) * expf(rr * 1.19392247f /* ln 3.3 */);
#endif
    r.twoTmaJonswap = 2.0f * tma * jonswap;
    // SpreadPower IS:60-66
    float sp;
    if (w < 1.05f * wp) sp = 6.97f * powf(fabsf(w / wp), 4.06f);
    else sp = 9.77f * powf(fabsf(w / wp), st.mu);
    // exponent of DirectionalSpread IS:78-84 with NormalizationFactor IS:69-74
    const float s = sp + 16.0f * tanhf(w / wp) * cc.swell * cc.swell;
    const float s2 = s * s, s3 = s2 * s;
    r.s = s;
    if (s <= 0.4f) r.nf = 0.09f * s3 + sc.c2 * s2 + sc.c1 * s + sc.c0;
    else r.nf = sqrtf(s) / sc.twoSqrtPi + 1.0f / (16.0f * sqrtf(sc.pi * s));
    // ShortWavesFade IS:95-97
    r.fade = expf(-cc.fade * cc.fade * kMag * kMag);
    // FrequencyDerivative IS:87-91
    const float th = tanhf(fminf(kMag * D, 20.0f));
    const float ch = coshf(kMag * D);
    r.dwdk = g * (D * kMag / ch / ch + th) / (w * 2.0f);
    return r;
}

// The direction-dependent rest of IS:99-129 for one texel (helpers IS:33-97 inlined, state scalars hoisted)
__device__ __forceinline__ Texel initial_spectrum_texel(float kx, float kz, const Radial& rd, const StateConsts& st,
                                                        const CascadeConsts& cc, float2 xi) {
    Texel r;
    if (rd.inBand) {
        const float k = cc.dk;
        const float theta = atan2f(kz, kx);               // IS:115
        const float spread = rd.nf * powf(fabsf(cosf(0.5f * (theta - st.windTheta))), 2.0f * rd.s);  // IS:78-84
        // IS:120
        const float amp = sqrtf(rd.twoTmaJonswap * spread * rd.fade * rd.dwdk / rd.kMag * k * k);
        r.h = make_float2(xi.x / 2.0f * amp, xi.y / 2.0f * amp);
        r.wave = make_float4(kx, 1.0f / rd.kMag, kz, rd.w);  // IS:123
    } else {
        r.h = make_float2(0.0f, 0.0f);                 // IS:125
        r.wave = make_float4(kx, 1.0f, kz, 0.0f);      // IS:126
    }
    return r;
}

}  // namespace ocean
