/*
 * ref_host.cpp -- host replay for "Oracle A": drives the reference's own compute-shader text
 * (compiled as C++ by oracle/Makefile, target `ref`) in the order WaterBody.cs / IFFT.cs dispatch it.
 * TEST INFRASTRUCTURE ONLY.  Output: oracle/_ref/libocean_ref.so (git-ignored).
 *
 * The shader translation units keep their uniforms/textures as globals (as HLSL does), so this
 * library drives ONE ocean at a time per process: every call re-binds before it dispatches,
 * which is what ComputeShader.SetTexture/SetFloat do in WB:130-169.
 */
#include <cstdint>
#include <cstring>
#include <vector>

extern "C" {
void ref_IS_bind(unsigned N, unsigned C, float* noise, float* h0, float* wave, const float* wavelengths,
                 const float* cutoffs, const float* fades, const float* swells, float U, float wx, float wy,
                 float g, float F, float D);
void ref_IS_dispatch(int kernel, unsigned N, int threads);
void ref_TD_bind(unsigned N, unsigned C, float* h0, float* wave, float* s0, float* s1, float* s2, float* s3);
void ref_TD_dispatch(unsigned N, float time, int threads);
void ref_FFT_bind(unsigned N, unsigned C, unsigned L, float* bfly, float* input, float* ping);
void ref_FFT_dispatch_precompute(unsigned N, unsigned L);
void ref_FFT_dispatch(int kernel, unsigned N, unsigned step, int pingPong, int threads);
void ref_RF_bind(unsigned N, unsigned C, float* disp3, float* deriv, float* turb, float* s0, float* s1, float* s2, float* s3);
void ref_RF_dispatch(unsigned N, int threads);
}

namespace {
struct RefOcean {
    unsigned N, C, L;
    int threads = 1;
    float params[6];
    // WB:231-251: cascade parameters flattened into four buffers; cutoffs interleaved (low, high)
    std::vector<float> wavelengths, cutoffs, swells, fades;
    std::vector<float> noise, h0, wave, spec[4], ping, bfly, disp3, deriv, turb;

    void bind_fft(std::vector<float>& input) { ref_FFT_bind(N, C, L, bfly.data(), input.data(), ping.data()); }

    // IFFT.InverseFastFourierTransform IFFT.cs:66-94
    void ifft(std::vector<float>& input) {
        bind_fft(input);
        bool pingPong = false;
        for (unsigned i = 0; i < L; ++i) { ref_FFT_dispatch(0, N, i, pingPong, threads); pingPong = !pingPong; }
        for (unsigned i = 0; i < L; ++i) { ref_FFT_dispatch(1, N, i, pingPong, threads); pingPong = !pingPong; }
        ref_FFT_dispatch(2, N, 0, pingPong, threads);
    }
};
}  // namespace

extern "C" {

void* ref_create(int N, int C, const float* params, const float* cascades, const float* noise) {
    RefOcean* o = new RefOcean;
    o->N = N; o->C = C;
    o->L = 0; while ((1u << o->L) < (unsigned)N) ++o->L;  // (int)Mathf.Log(N, 2) IFFT.cs:35
    std::memcpy(o->params, params, sizeof(o->params));
    for (int i = 0; i < C; ++i) {
        o->wavelengths.push_back(cascades[i * 5 + 0]);
        o->cutoffs.push_back(cascades[i * 5 + 1]);  // low   WB:238
        o->cutoffs.push_back(cascades[i * 5 + 2]);  // high  WB:239
        o->swells.push_back(cascades[i * 5 + 3]);
        o->fades.push_back(cascades[i * 5 + 4]);
    }
    size_t t = (size_t)N * N;
    o->noise.assign(noise, noise + t * 2);
    o->h0.assign(C * t * 4, 0.f); o->wave.assign(C * t * 4, 0.f);
    for (auto& s : o->spec) s.assign(C * t * 2, 0.f);
    o->ping.assign(C * t * 2, 0.f);
    o->bfly.assign((size_t)N * o->L * 4, 0.f);
    o->disp3.assign(C * t * 3, 0.f); o->deriv.assign(C * t * 4, 0.f); o->turb.assign(C * t * 4, 0.f);
    // IFFT ctor IFFT.cs:43-46
    o->bind_fft(o->spec[0]);
    ref_FFT_dispatch_precompute(N, o->L);
    return o;
}

void ref_destroy(void* h) { delete (RefOcean*)h; }
void ref_set_threads(void* h, int n) { ((RefOcean*)h)->threads = n < 1 ? 1 : n; }

// WB:253-254 -> WB:171-178
void ref_init_spectrum(void* h) {
    RefOcean* o = (RefOcean*)h;
    ref_IS_bind(o->N, o->C, o->noise.data(), o->h0.data(), o->wave.data(), o->wavelengths.data(), o->cutoffs.data(),
                o->fades.data(), o->swells.data(), o->params[0], o->params[1], o->params[2], o->params[3],
                o->params[4], o->params[5]);
    ref_IS_dispatch(0, o->N, o->threads);
    ref_IS_dispatch(1, o->N, o->threads);
}

// WaterBody.CalculateWavesTexturesAtTime WB:180-193 (mips left out)
void ref_step(void* h, float time) {
    RefOcean* o = (RefOcean*)h;
    ref_TD_bind(o->N, o->C, o->h0.data(), o->wave.data(), o->spec[0].data(), o->spec[1].data(), o->spec[2].data(), o->spec[3].data());
    ref_TD_dispatch(o->N, time, o->threads);
    for (auto& s : o->spec) o->ifft(s);
    ref_RF_bind(o->N, o->C, o->disp3.data(), o->deriv.data(), o->turb.data(), o->spec[0].data(), o->spec[1].data(), o->spec[2].data(), o->spec[3].data());
    ref_RF_dispatch(o->N, o->threads);
}

void ref_reset_foam(void* h) { RefOcean* o = (RefOcean*)h; std::fill(o->turb.begin(), o->turb.end(), 0.f); }

// which: same numbering as oracle_capi.cpp (0 h0, 1 wave, 2-5 spectra, 6 ping, 7 disp RGBA, 8 deriv, 9 turb, 10 bfly, 11 noise)
long long ref_size(void* h, int which) {
    RefOcean* o = (RefOcean*)h;
    size_t t = (size_t)o->N * o->N * o->C;
    switch (which) {
        case 0: case 1: case 7: case 8: case 9: return (long long)t * 4;
        case 2: case 3: case 4: case 5: case 6: return (long long)t * 2;
        case 10: return (long long)o->bfly.size();
        case 11: return (long long)o->noise.size();
        default: return -1;
    }
}

int ref_get_f32(void* h, int which, float* out) {
    RefOcean* o = (RefOcean*)h;
    const std::vector<float>* v = nullptr;
    switch (which) {
        case 0: v = &o->h0; break;
        case 1: v = &o->wave; break;
        case 2: case 3: case 4: case 5: v = &o->spec[which - 2]; break;
        case 6: v = &o->ping; break;
        case 7: {  // float3 store into an RGBA32F target: alpha reported as 0
            size_t t = (size_t)o->N * o->N * o->C;
            for (size_t i = 0; i < t; ++i) {
                out[i * 4 + 0] = o->disp3[i * 3 + 0]; out[i * 4 + 1] = o->disp3[i * 3 + 1];
                out[i * 4 + 2] = o->disp3[i * 3 + 2]; out[i * 4 + 3] = 0.f;
            }
            return 0;
        }
        case 8: v = &o->deriv; break;
        case 9: v = &o->turb; break;
        case 10: v = &o->bfly; break;
        case 11: v = &o->noise; break;
        default: return -1;
    }
    std::memcpy(out, v->data(), v->size() * sizeof(float));
    return 0;
}

}  // extern "C"
