/*
 * ocean_oracle.hpp -- "Oracle B": CPU restatement of the reference's spectral-ocean hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it, and
 * there only as the checker (or as the timed CPU baseline), never as the thing shipped.
 *
 * Parity pin: the reference (Unity C# + HLSL compute) ships no tests, golden vectors or
 * fixtures, so this transcription is pinned by "Oracle A" (oracle/Makefile, target `ref`): the reference's
 * own .compute shader text compiled as C++ through oracle/hlsl_compat.h and replayed in the
 * WaterBody.cs / IFFT.cs dispatch order.  tests/test_oracle.py::test_oracle_a_equals_oracle_b requires A == B bit-for-bit,
 * and tests/golden/ holds vectors generated from Oracle A.
 *
 * The transcription follows the reference op for op; citations are file:line into /root/reference
 *   IS  = Assets/Shaders/Compute Shaders/Water/InitialSpectrum.compute
 *   TD  = Assets/Shaders/Compute Shaders/Water/TimeDependentSpectrum.compute
 *   FFT = Assets/Shaders/Compute Shaders/Water/IFFT.compute
 *   RF  = Assets/Shaders/Compute Shaders/Water/ResultTexturesFiller.compute
 *   WB  = Assets/Scripts/Water/WaterBody.cs,  IFFT.cs = Assets/Scripts/Water/IFFT.cs
 *
 * Templated on the scalar type: T = float is the parity target (HLSL computes in fp32; every
 * unsuffixed HLSL literal is a float), T = double is the arbiter (same formulas, the same
 * float-rounded constants, no fp32 rounding in the operations).  Build with
 * -O2 -ffp-contract=off and without -ffast-math.
 */
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

namespace ocean_oracle {

template <class T> struct M;
template <> struct M<float> {
    static float sqrt(float x) { return ::sqrtf(x); }
    static float pow(float x, float y) { return ::powf(x, y); }
    static float exp(float x) { return ::expf(x); }
    static float log(float x) { return ::logf(x); }
    static float tanh(float x) { return ::tanhf(x); }
    static float cosh(float x) { return ::coshf(x); }
    static float atan2(float y, float x) { return ::atan2f(y, x); }
    static float cos(float x) { return ::cosf(x); }
    static float sin(float x) { return ::sinf(x); }
    static float abs(float x) { return ::fabsf(x); }
};
template <> struct M<double> {
    static double sqrt(double x) { return ::sqrt(x); }
    static double pow(double x, double y) { return ::pow(x, y); }
    static double exp(double x) { return ::exp(x); }
    static double log(double x) { return ::log(x); }
    static double tanh(double x) { return ::tanh(x); }
    static double cosh(double x) { return ::cosh(x); }
    static double atan2(double y, double x) { return ::atan2(y, x); }
    static double cos(double x) { return ::cos(x); }
    static double sin(double x) { return ::sin(x); }
    static double abs(double x) { return ::fabs(x); }
};

// Ocean-wide parameters: WaterBody serialized fields WB:10-14 -> uniforms WB:140-145.
struct OceanParams {
    float windSpeed, windDirX, windDirY, gravity, fetch, depth;
};
// Per-cascade parameters: WaterCascade WC:10-24, flattened as WB:231-251.
struct CascadeParams {
    float wavelength, cutoffLow, cutoffHigh, swell, fade;
};

template <class T> struct C2 { T x, y; };

template <class T>
class Ocean {
  public:
    using Mt = M<T>;
    int N, C, L;
    OceanParams p;
    std::vector<CascadeParams> casc;
    // "textures", flattened as index = ((layer*N + y)*N + x)*channels (WB:206, IS:120)
    std::vector<T> noise;      // N*N*2        _RandomNoiseTexture  (RG32F)
    std::vector<T> h0;         // C*N*N*4      _InitialSpectrumTextures (RGBA32F)
    std::vector<T> wave;       // C*N*N*4      _WavesDataTextures
    std::vector<T> spec[4];    // C*N*N*2      DxDz, DyDxz, DyxDyz, DxxDzz (RG32F)
    std::vector<T> ping;       // C*N*N*2      IFFT ping-pong array (IFFT.cs:51-61)
    std::vector<T> bfly;       // N*L*4        butterfly texture, width L, height N (IFFT.cs:36)
    std::vector<T> disp, deriv, turb;  // C*N*N*4
    int threads = 1;           // OpenMP threads used per dispatch

    // every HLSL literal is fp32: route constants through float first
    static T K(float v) { return (T)v; }
    T PI() const { return K(3.14159265f); }  // IS:8, FFT:7

    Ocean(int N_, int C_, const OceanParams& p_, const CascadeParams* c_, const float* noise_)
        : N(N_), C(C_), p(p_), casc(c_, c_ + C_) {
        L = 0;
        while ((1 << L) < N) ++L;  // (int)Mathf.Log(N, 2), exact for powers of two (IFFT.cs:35)
        size_t t = (size_t)N * N;
        noise.resize(t * 2);
        for (size_t i = 0; i < t * 2; ++i) noise[i] = (T)noise_[i];
        h0.assign(C * t * 4, 0); wave.assign(C * t * 4, 0);
        for (auto& s : spec) s.assign(C * t * 2, 0);
        ping.assign(C * t * 2, 0);
        bfly.assign((size_t)N * L * 4, 0);
        disp.assign(C * t * 4, 0); deriv.assign(C * t * 4, 0);
        turb.assign(C * t * 4, 0);  // fresh RenderTexture: treated as zero
        precompute_butterfly();
    }

    // ---------------- InitialSpectrum.compute helpers ----------------
    T AngularFrequency(T k) const { return Mt::sqrt(K(p.gravity) * k); }  // IS:33-35

    T TMACorrection(T w) const {  // IS:38-43
        T wh = w * Mt::sqrt(K(p.depth) / K(p.gravity));
        if (wh <= K(1.0f)) return K(0.5f) * wh * wh;
        if (wh < K(2.0f)) return K(1.0f) - K(0.5f) * (K(2.0f) - wh) * (K(2.0f) - wh);
        return K(1.0f);
    }

    T JONSWAP(T w, T wp) const {  // IS:47-56
        T U = K(p.windSpeed), F = K(p.fetch), g = K(p.gravity);
        T alpha = K(0.076f) * Mt::pow(Mt::abs(U * U / (F * g)), K(0.22f));
        T gamma = K(3.3f);
        T sigma = w <= wp ? K(0.07f) : K(0.09f);
        T d = w - wp;
        T r = Mt::exp(-(d * d) / (K(2.0f) * sigma * sigma * wp * wp));
        return alpha * g * g / Mt::pow(w, K(5.0f)) * Mt::exp(K(-1.25f) * Mt::pow(wp / w, K(4.0f))) *
               Mt::pow(Mt::abs(gamma), r);
    }

    T SpreadPower(T w, T wp) const {  // IS:60-66
        if (w < K(1.05f) * wp) return K(6.97f) * Mt::pow(Mt::abs(w / wp), K(4.06f));
        T peakSpeed = K(p.gravity) / wp;
        T mu = K(-2.33f) - K(1.45f) * (K(p.windSpeed) / peakSpeed - K(1.17f));
        return K(9.77f) * Mt::pow(Mt::abs(w / wp), mu);
    }

    T NormalizationFactor(T s) const {  // IS:69-74
        T s2 = s * s;
        T s3 = s2 * s;
        T pi = PI();
        if (s <= K(0.4f))
            return K(0.09f) * s3 + (Mt::pow(Mt::log(K(2.0f)), K(2.0f)) / pi - pi / K(12.0f)) * s2 +
                   Mt::log(K(2.0f)) / pi * s + K(1.0f) / (K(2.0f) * pi);
        return Mt::sqrt(s) / (K(2.0f) * Mt::sqrt(pi)) + K(1.0f) / (K(16.0f) * Mt::sqrt(pi * s));
    }

    T DirectionalSpread(T w, T wp, T theta, T swell) const {  // IS:78-84
        T s = SpreadPower(w, wp) + K(16.0f) * Mt::tanh(w / wp) * swell * swell;
        T wx = K(p.windDirX), wy = K(p.windDirY);
        T len = Mt::sqrt(wx * wx + wy * wy);  // normalize(v) = v / length(v)
        T nx = wx / len, ny = wy / len;
        T windTheta = Mt::atan2(ny, nx);
        return NormalizationFactor(s) * Mt::pow(Mt::abs(Mt::cos(K(0.5f) * (theta - windTheta))), K(2.0f) * s);
    }

    T FrequencyDerivative(T k, T w) const {  // IS:87-91
        T D = K(p.depth);
        T kd = k * D;
        T th = Mt::tanh(kd < K(20.0f) ? kd : K(20.0f));  // min(k*D, 20)
        T ch = Mt::cosh(k * D);                          // may be +inf: x/inf/inf = 0
        return K(p.gravity) * (D * k / ch / ch + th) / (w * K(2.0f));
    }

    T ShortWavesFade(T k, T fade) const { return Mt::exp(-fade * fade * k * k); }  // IS:95-97

    // ---------------- kernels ----------------
    size_t tex(int layer, int x, int y) const { return ((size_t)layer * N + y) * N + x; }

    // K1: CalculateInitialSpectrumTextures IS:99-129
    void CalculateInitialSpectrumTextures() {
#pragma omp parallel for num_threads(threads) schedule(static)
        for (int y = 0; y < N; ++y)
            for (int x = 0; x < N; ++x) {
                unsigned halfSize = (unsigned)N / 2;
                int nx = (int)((unsigned)x - halfSize);
                int nz = (int)((unsigned)y - halfSize);
                T g1 = noise[((size_t)y * N + x) * 2 + 0];
                T g2 = noise[((size_t)y * N + x) * 2 + 1];
                for (int i = 0; i < C; ++i) {
                    T k = K(2.0f) * PI() / K(casc[i].wavelength);
                    T kx = (T)nx * k, kz = (T)nz * k;
                    T kMag = Mt::sqrt(kx * kx + kz * kz);  // length()
                    T* H = &h0[tex(i, x, y) * 4];
                    T* W = &wave[tex(i, x, y) * 4];
                    if (kMag >= K(casc[i].cutoffLow) && kMag <= K(casc[i].cutoffHigh)) {
                        T kAngle = Mt::atan2(kz, kx);
                        T w = AngularFrequency(kMag);
                        T wp = K(22.0f) * Mt::pow(Mt::abs(K(p.gravity) * K(p.gravity) /
                                                          (K(p.windSpeed) * K(p.fetch))),
                                                  K(0.3333f));
                        T amp = Mt::sqrt(K(2.0f) * TMACorrection(w) * JONSWAP(w, wp) *
                                         DirectionalSpread(w, wp, kAngle, K(casc[i].swell)) *
                                         ShortWavesFade(kMag, K(casc[i].fade)) *
                                         FrequencyDerivative(kMag, w) / kMag * k * k);
                        H[0] = g1 / K(2.0f) * amp; H[1] = g2 / K(2.0f) * amp; H[2] = 0; H[3] = 0;
                        W[0] = kx; W[1] = K(1.0f) / kMag; W[2] = kz; W[3] = w;
                    } else {
                        H[0] = H[1] = H[2] = H[3] = 0;
                        W[0] = kx; W[1] = K(1.0f); W[2] = kz; W[3] = 0;
                    }
                }
            }
    }

    // K2: CalculateConjugatedInitialSpectrumTextures IS:135-143 (in place; only .zw changes)
    void CalculateConjugatedInitialSpectrumTextures() {
#pragma omp parallel for num_threads(threads) schedule(static)
        for (int y = 0; y < N; ++y)
            for (int x = 0; x < N; ++x)
                for (int i = 0; i < C; ++i) {
                    int mx = (N - x) % N, my = (N - y) % N;
                    T* H = &h0[tex(i, x, y) * 4];
                    const T* Hm = &h0[tex(i, mx, my) * 4];
                    H[2] = Hm[0];
                    H[3] = -Hm[1];
                }
    }

    // WB:171-178
    void init_spectrum() {
        CalculateInitialSpectrumTextures();
        CalculateConjugatedInitialSpectrumTextures();
    }

    // K3: PrecomputeTwiddleFactorsAndInputIndices FFT:37-45, dispatched (L, N/2/8, 1)x(1,8,1)
    void precompute_butterfly() {
        for (unsigned s = 0; s < (unsigned)L; ++s)
            for (unsigned y = 0; y < (unsigned)N / 2; ++y) {
                unsigned b = (unsigned)N >> (s + 1);
                T multy = K(2.0f) * PI() * K(1.0f) / (T)N;  // 2*PI*float2(0,1)/_TextureSize, .y
                unsigned i = (2 * b * (y / b) + y % b) % (unsigned)N;
                T arg = -multy * (T)((y / b) * b);
                // ComplexExp: exp(a.x)*(cos a.y, sin a.y) with a.x = -0*m = -0 -> exp = 1
                T e = Mt::exp(-(K(2.0f) * PI() * K(0.0f) / (T)N) * (T)((y / b) * b));
                T tx = e * Mt::cos(arg), ty = e * Mt::sin(arg);
                T* a = &bfly[((size_t)y * L + s) * 4];
                a[0] = tx; a[1] = ty; a[2] = (T)i; a[3] = (T)(i + b);
                T* c = &bfly[((size_t)(y + N / 2) * L + s) * 4];
                c[0] = -tx; c[1] = -ty; c[2] = (T)i; c[3] = (T)(i + b);
                (void)multy;
            }
    }

    // K4: CalculateTimeDependentComplexAmplitudesAndDerivatives TD:20-47
    void evolve(T time) {
#pragma omp parallel for num_threads(threads) schedule(static) collapse(2)
        for (int i = 0; i < C; ++i)
            for (int y = 0; y < N; ++y)
                for (int x = 0; x < N; ++x) {
                    size_t t = tex(i, x, y);
                    const T* wv = &wave[t * 4];
                    const T* H = &h0[t * 4];
                    T phase = wv[3] * time;
                    T ex = Mt::cos(phase), ey = Mt::sin(phase);
                    // ComplexMult(h0.xy, e) + ComplexMult(h0.zw, conj e)   TD:26
                    T hx = (H[0] * ex - H[1] * ey) + (H[2] * ex - H[3] * (-ey));
                    T hy = (H[0] * ey + H[1] * ex) + (H[2] * (-ey) + H[3] * ex);
                    T ihx = -hy, ihy = hx;
                    T YDx_x = ihx * wv[0], YDx_y = ihy * wv[0];
                    T YDz_x = ihx * wv[2], YDz_y = ihy * wv[2];
                    T X_x = YDx_x * wv[1], X_y = YDx_y * wv[1];
                    T Y_x = hx, Y_y = hy;
                    T Z_x = YDz_x * wv[1], Z_y = YDz_y * wv[1];
                    T ax = -hx * wv[1], ay = -hy * wv[1];
                    T XDx_x = ax * wv[0] * wv[0], XDx_y = ay * wv[0] * wv[0];
                    T ZDz_x = ax * wv[2] * wv[2], ZDz_y = ay * wv[2] * wv[2];
                    T ZDx_x = ax * wv[0] * wv[2], ZDx_y = ay * wv[0] * wv[2];
                    spec[0][t * 2] = X_x - Z_y;     spec[0][t * 2 + 1] = X_y + Z_x;      // DxDz   TD:42
                    spec[1][t * 2] = Y_x - ZDx_y;   spec[1][t * 2 + 1] = Y_y + ZDx_x;    // DyDxz  TD:43
                    spec[2][t * 2] = YDx_x - YDz_y; spec[2][t * 2 + 1] = YDx_y + YDz_x;  // DyxDyz TD:44
                    spec[3][t * 2] = XDx_x - ZDz_y; spec[3][t * 2 + 1] = XDx_y + ZDz_x;  // DxxDzz TD:45
                }
    }

    // K5/K6: one radix-2 stage along x (horizontal) or y. FFT:48-69
    void fft_step(std::vector<T>& input, bool horizontal, int step, bool pingPong) {
        std::vector<T>& src = pingPong ? ping : input;
        std::vector<T>& dst = pingPong ? input : ping;
#pragma omp parallel for num_threads(threads) schedule(static) collapse(2)
        for (int i = 0; i < C; ++i)
            for (int y = 0; y < N; ++y)
                for (int x = 0; x < N; ++x) {
                    int j = horizontal ? x : y;
                    const T* d = &bfly[((size_t)j * L + step) * 4];
                    T twx = d[0], twy = -d[1];  // float2(data.r, -data.g)
                    unsigned i0 = (unsigned)d[2], i1 = (unsigned)d[3];
                    size_t a = horizontal ? tex(i, (int)i0, y) : tex(i, x, (int)i0);
                    size_t b = horizontal ? tex(i, (int)i1, y) : tex(i, x, (int)i1);
                    T bx = src[b * 2], by = src[b * 2 + 1];
                    size_t o = tex(i, x, y);
                    dst[o * 2] = src[a * 2] + (twx * bx - twy * by);
                    dst[o * 2 + 1] = src[a * 2 + 1] + (twx * by + twy * bx);
                }
    }

    // K7: Permute FFT:73-78
    void permute(std::vector<T>& input) {
#pragma omp parallel for num_threads(threads) schedule(static) collapse(2)
        for (int i = 0; i < C; ++i)
            for (int y = 0; y < N; ++y)
                for (int x = 0; x < N; ++x) {
                    size_t o = tex(i, x, y);
                    T s = K(1.0f) - K(2.0f) * (T)(((unsigned)x + (unsigned)y) % 2);
                    input[o * 2] = input[o * 2] * s;
                    input[o * 2 + 1] = input[o * 2 + 1] * s;
                }
    }

    // IFFT.InverseFastFourierTransform IFFT.cs:66-94.  stop_after_horizontal is a test hook: it
    // returns which buffer (false = input, true = ping) holds the state after the x passes.
    bool ifft(std::vector<T>& input, bool stop_after_horizontal = false) {
        bool pingPong = false;
        for (int s = 0; s < L; ++s) { fft_step(input, true, s, pingPong); pingPong = !pingPong; }
        if (stop_after_horizontal) return pingPong;
        for (int s = 0; s < L; ++s) { fft_step(input, false, s, pingPong); pingPong = !pingPong; }
        permute(input);
        return false;
    }

    // K8: FillResultTextures RF:16-34
    void fill_results() {
#pragma omp parallel for num_threads(threads) schedule(static) collapse(2)
        for (int i = 0; i < C; ++i)
            for (int y = 0; y < N; ++y)
                for (int x = 0; x < N; ++x) {
                    size_t o = tex(i, x, y);
                    T DxDz_x = spec[0][o * 2], DxDz_y = spec[0][o * 2 + 1];
                    T DyDxz_x = spec[1][o * 2], DyDxz_y = spec[1][o * 2 + 1];
                    T DyxDyz_x = spec[2][o * 2], DyxDyz_y = spec[2][o * 2 + 1];
                    T DxxDzz_x = spec[3][o * 2], DxxDzz_y = spec[3][o * 2 + 1];
                    T* d = &disp[o * 4];
                    d[0] = DxDz_x; d[1] = DyDxz_x; d[2] = DxDz_y; d[3] = 0;  // float3 store; alpha := 0
                    T* dv = &deriv[o * 4];
                    dv[0] = DyxDyz_x; dv[1] = DyxDyz_y; dv[2] = DxxDzz_x; dv[3] = DxxDzz_y;
                    T jac = (K(1.0f) + DxxDzz_x) * (K(1.0f) + DxxDzz_y) - DyDxz_y * DyDxz_y;
                    T foam = turb[o * 4];
                    T decay = K(2.0f);
                    foam *= Mt::exp(-decay);
                    if (foam < jac) foam += jac;
                    T* tb = &turb[o * 4];
                    tb[0] = tb[1] = tb[2] = tb[3] = foam;
                }
    }

    // WaterBody.CalculateWavesTexturesAtTime WB:180-193 (GenerateMips left out: see DESIGN.md)
    void step(T time) {
        evolve(time);
        for (auto& s : spec) ifft(s);
        fill_results();
    }

    // Test hook: evolve + the L horizontal passes only.  out[k] receives the state of spectrum k.
    void step_rows_only(T time, std::vector<T> out[4]) {
        evolve(time);
        for (int k = 0; k < 4; ++k) {
            bool inPing = ifft(spec[k], true);
            out[k] = inPing ? ping : spec[k];
        }
    }

    // WaterBody.GetWaterHeight WB:195-209 evaluated on the current displacement layer 0
    T water_height(T wx, T wz) const {
        auto inverse_lerp = [](T a, T b, T v) {  // Mathf.InverseLerp = Clamp01((v-a)/(b-a))
            if (a == b) return (T)0;
            T r = (v - a) / (b - a);
            return r < 0 ? (T)0 : (r > 1 ? (T)1 : r);
        };
        int half = N / 2;
        T u = inverse_lerp((T)(-half), (T)half, wx);
        T v = inverse_lerp((T)(-half), (T)half, wz);
        int x = (int)(u * (T)N), y = (int)(v * (T)N);
        x = x < 0 ? 0 : (x > N - 1 ? N - 1 : x);
        y = y < 0 ? 0 : (y > N - 1 ? N - 1 : y);
        return disp[((size_t)y * N + x) * 4 + 1];
    }
};

}  // namespace ocean_oracle
