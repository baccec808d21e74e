// ocean_init.cu -- one-time kernels: noise generation and the initial spectrum.
//
// Built with -fmad=false: every fp32 operation rounds once, in the reference's order, so the
// spectrum differs from the CPU transcription only through the math library (powf, expf, ...).
//
// ocean_init_spectrum_kernel fuses the reference's two dispatches
//   CalculateInitialSpectrumTextures            InitialSpectrum.compute:99-129  (WB:173)
//   CalculateConjugatedInitialSpectrumTextures  InitialSpectrum.compute:135-143 (WB:177)
// Each thread owns the texel pair {p, mirror(p)} with mirror(x, y) = ((N-x)%N, (N-y)%N): it
// evaluates h0 once at each texel and writes both (h0(p), conj h0(m)) and (h0(m), conj h0(p)),
// so the amplitude is never computed twice and the reference's in-place second pass disappears.
#include "ocean_internal.h"
#include "../../include/ocean_noise.h"

namespace ocean {

__global__ void ocean_noise_kernel(const uint64_t* __restrict__ seeds, int N, float2* __restrict__ noise) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    const int st = blockIdx.z;
    if (x >= N) return;
    const uint64_t seed = seeds[st];
    float2 v;
    v.x = ocean_noise_texel(seed, (uint32_t)N, (uint32_t)x, (uint32_t)y, 0u);
    v.y = ocean_noise_texel(seed, (uint32_t)N, (uint32_t)x, (uint32_t)y, 1u);
    noise[((size_t)st * N + y) * N + x] = v;
}

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
    const float jonswap = st.alpha * g * g / powf(w, 5.0f) * expf(-1.25f * powf(wp / w, 4.0f)) * powf(3.3f, rr);
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

__global__ void __launch_bounds__(128) ocean_init_spectrum_kernel(const InitArgs a) {
    const int N = a.N;
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;  // 0 .. N/2
    if (x >= N) return;
    const int mx = (N - x) & (N - 1), my = (N - y) & (N - 1);  // IS:139
    // rows 0 and N/2 mirror onto themselves: the pair {x, mx} is handled by the thread with x <= mx
    if (my == y && x > mx) return;
    const int layer = blockIdx.z;
    const int st = layer / a.C, ci = layer - st * a.C;
    const StateConsts sc = a.states[st];
    const CascadeConsts cc = a.casc[ci];
    const float2* noise = a.noise + (size_t)st * N * N;
    const size_t p = ((size_t)layer * N + y) * N + x;
    const size_t m = ((size_t)layer * N + my) * N + mx;
    // IS:101-103, IS:111
    const float kx = (float)(x - N / 2) * cc.dk, kz = (float)(y - N / 2) * cc.dk;
    const float kxm = (float)(mx - N / 2) * cc.dk, kzm = (float)(my - N / 2) * cc.dk;
    const Radial rd = initial_spectrum_radial(kx, kz, sc, cc, a.sc);
    const Texel tp = initial_spectrum_texel(kx, kz, rd, sc, cc, __ldg(&noise[(size_t)y * N + x]));
    if (m == p) {
        a.h0[p] = make_float4(tp.h.x, tp.h.y, tp.h.x, -tp.h.y);  // IS:141
        if (a.wave) a.wave[p] = tp.wave;
        a.waveHot[p] = make_float2(tp.wave.y, tp.wave.w);
        if (a.h0c) a.h0c[p] = tp.h;
        return;
    }
    const Texel tm = initial_spectrum_texel(kxm, kzm, rd, sc, cc, __ldg(&noise[(size_t)my * N + mx]));
    a.h0[p] = make_float4(tp.h.x, tp.h.y, tm.h.x, -tm.h.y);
    a.h0[m] = make_float4(tm.h.x, tm.h.y, tp.h.x, -tp.h.y);
    if (a.wave) {
        a.wave[p] = tp.wave;
        a.wave[m] = tm.wave;
    }
    a.waveHot[p] = make_float2(tp.wave.y, tp.wave.w);
    a.waveHot[m] = make_float2(tm.wave.y, tm.wave.w);
    if (a.h0c) {
        a.h0c[p] = tp.h;
        a.h0c[m] = tm.h;
    }
}

cudaError_t launch_noise(const uint64_t* seedsDev, int nStates, int N, float2* noise, cudaStream_t s) {
    dim3 block(N < 128 ? N : 128), grid((N + block.x - 1) / block.x, N, nStates);
    ocean_noise_kernel<<<grid, block, 0, s>>>(seedsDev, N, noise);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_init(const InitArgs& a, cudaStream_t s) {
    dim3 block(a.N < 128 ? a.N : 128), grid((a.N + block.x - 1) / block.x, a.N / 2 + 1, a.nStates * a.C);
    ocean_init_spectrum_kernel<<<grid, block, 0, s>>>(a);
    count_launch();
    return cudaGetLastError();
}

}  // namespace ocean
