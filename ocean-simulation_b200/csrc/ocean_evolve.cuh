// ocean_evolve.cuh -- the per-texel time evolution (TimeDependentSpectrum.compute:20-47) shared by the row
// pass (ocean_frame.cu), the fused frame kernel (ocean_fused.cuh) and the spectrum-to-rows kernel
// (ocean_init.cu).  Sensitivity bit OCEAN_EXP & 32 (see ocean_frame.cu) replaces the sincos by a placeholder.
#pragma once
#include "ocean_fft.cuh"

#ifndef OCEAN_EXP
#define OCEAN_EXP 0
#endif

namespace ocean {

// sin and cos of a phase: 3-term Cody-Waite reduction to [-pi/4, pi/4] (the constants CUDA's own
// sincosf uses on its fast path, valid below 105615 rad) followed by the SFU approximations, whose
// absolute error (~4e-7) is of the order of the transform's own rounding.  Larger phases take the
// library path.
static __device__ __noinline__ float2 phase_sincos_large(float x) {
    float sn, cs;
    sincosf(x, &sn, &cs);
    return make_float2(sn, cs);
}
__device__ __forceinline__ void phase_sincos(float x, float& sn, float& cs) {
    if (fabsf(x) > 105615.0f) {
        const float2 r = phase_sincos_large(x);
        sn = r.x;
        cs = r.y;
        return;
    }
    const float magic = 12582912.0f;  // 1.5 * 2^23: adding it rounds to the nearest integer
    const float qf = __fmaf_rn(x, 0.636619772f, magic);
    const int iq = __float_as_int(qf);  // low bits = round(x * 2/pi)
    const float q = qf - magic;
    float r = __fmaf_rn(q, -1.57079601e+00f, x);
    r = __fmaf_rn(q, -3.13916473e-07f, r);
    r = __fmaf_rn(q, -5.39030253e-15f, r);
    const float s0 = __sinf(r), c0 = __cosf(r);
    const bool odd = iq & 1;
    const float sv = odd ? c0 : s0, cv = odd ? s0 : c0;
    sn = __int_as_float(__float_as_int(sv) ^ ((iq & 2) << 30));
    cs = __int_as_float(__float_as_int(cv) ^ (((iq + 1) & 2) << 30));
}

// measured on B200 at 4 x 512^2: the SFU path executes 18% fewer instructions but the frame is not
// issue-bound (27.2 us vs 26.9 us with the library sincosf), so the precise path stays the default
#ifndef OCEAN_ROWS_DUAL_FFT
#define OCEAN_ROWS_DUAL_FFT 0
#endif
#ifndef OCEAN_ROWS_FAST_SINCOS
#define OCEAN_ROWS_FAST_SINCOS 0
#endif

// CalculateTimeDependentComplexAmplitudesAndDerivatives TD:20-47 for one texel: h0 = (h0(k), conj h0(-k)),
// wv = (1/|k|, omega); returns the eight derived spectra in the packed pair layout.
__device__ __forceinline__ void evolve_texel(const float4 h0, const float2 wv, const float kx, const float kz, const float time,
                                             C2& pa, C2& pb) {
        const float ik = wv.x;
        // TD:23-26
        float sn, cs;
#if OCEAN_EXP & 32  // sensitivity: no sincos at all
        sn = wv.y * time; cs = 1.0f - sn;
#elif OCEAN_ROWS_FAST_SINCOS
        phase_sincos(wv.y * time, sn, cs);
#else
        sincosf(wv.y * time, &sn, &cs);
#endif
        const float hx = (h0.x * cs - h0.y * sn) + (h0.z * cs + h0.w * sn);
        const float hy = (h0.x * sn + h0.y * cs) + (h0.w * cs - h0.z * sn);
        const float ihx = -hy, ihy = hx;
        // TD:29-40
        const float ydx_x = ihx * kx, ydx_y = ihy * kx;
        const float ydz_x = ihx * kz, ydz_y = ihy * kz;
        const float X_x = ydx_x * ik, X_y = ydx_y * ik;
        const float Z_x = ydz_x * ik, Z_y = ydz_y * ik;
        const float ax = -hx * ik, ay = -hy * ik;
        const float xdx_x = ax * kx * kx, xdx_y = ay * kx * kx;
        const float zdz_x = ax * kz * kz, zdz_y = ay * kz * kz;
        const float zdx_x = ax * kx * kz, zdx_y = ay * kx * kz;
        // TD:42-45: A + iB packing; pa = (DxDz, DyDxz), pb = (DyxDyz, DxxDzz), each held as
        // re = (re of first, re of second), im = (im of first, im of second)
        pa = C2{pk2(X_x - Z_y, hx - zdx_y), pk2(X_y + Z_x, hy + zdx_x)};
        pb = C2{pk2(ydx_x - ydz_y, xdx_x - zdz_y), pk2(ydx_y + ydz_x, xdx_y + zdz_x)};
}

}  // namespace ocean
