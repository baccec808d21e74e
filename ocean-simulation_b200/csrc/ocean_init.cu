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
#include <stdlib.h>

#include "ocean_internal.h"
#include "../../include/ocean_noise.h"
#include "ocean_spectrum.cuh"
#include "ocean_evolve.cuh"
#include "ocean_tma.cuh"

namespace ocean {

__device__ __forceinline__ void noise_texel_pair(const uint64_t* __restrict__ seeds, int N, float2* __restrict__ noise, int x, int y, int st) {
    if (x >= N) return;
    const uint64_t seed = seeds[st];
    float2 v;
    v.x = ocean_noise_texel(seed, (uint32_t)N, (uint32_t)x, (uint32_t)y, 0u);
    v.y = ocean_noise_texel(seed, (uint32_t)N, (uint32_t)x, (uint32_t)y, 1u);
    noise[((size_t)st * N + y) * N + x] = v;
}
// The noise texture with the rejection loop load-balanced inside a thread.  With one draw per loop a warp iterates until its
// SLOWEST lane accepts: ~3 iterations per normal against 1.27 per lane (ncu: 592 instructions per texel, issue-bound).  Here a
// thread owns DRAWS = 8 draws -- 4 adjacent texels x (r, g) -- and runs ONE loop over all of them, moving on to its next draw
// the moment one is accepted; a warp then iterates for the slowest lane's SUM of attempts (~14 for 8 draws instead of ~24).
// Accepted (v1, s) pairs are parked in shared memory and transformed after the loop, outside the divergent part.  Attempt test
// and transform are the header's own functions (include/ocean_noise.h): the bits are those of ocean_noise_texel.
constexpr int kNoiseDraws = 8, kNoiseThreads = 128;
__global__ void __launch_bounds__(kNoiseThreads) ocean_noise_balanced_kernel(const uint64_t* __restrict__ seeds, int N, float2* __restrict__ noise,
                                                                             long long groups) {
    __shared__ float2 parked[kNoiseDraws][kNoiseThreads];
    const long long gid = (long long)blockIdx.x * kNoiseThreads + threadIdx.x;
    if (gid >= groups) return;
    const int gpr = N / 4;  // groups of 4 texels per row
    const int xg = (int)(gid % gpr), y = (int)((gid / gpr) % N), st = (int)(gid / ((long long)gpr * N));
    const uint64_t seed = seeds[st];
    const uint32_t x0 = (uint32_t)xg * 4u;
    // draw d: texel x0 + d / 2, channel d & 1; counter = (x * N + y) * 2 + ch  (ocean_noise_texel)
    auto counter = [&](int d) { return ((uint64_t)(x0 + (uint32_t)(d >> 1)) * (uint32_t)N + (uint32_t)y) * 2u + (uint32_t)(d & 1); };
    int d = 0;
    uint32_t attempt = 0;
    uint64_t key = ocean_counter_key(seed, counter(0));
    while (d < kNoiseDraws) {
        float v1, s;
        if (ocean_gaussian_attempt(key, attempt, &v1, &s)) {
            parked[d][threadIdx.x] = make_float2(v1, s);
            ++d;
            attempt = 0;
            key = ocean_counter_key(seed, counter(d));  // (d == 8: computed and never used)
        } else {
            ++attempt;
        }
    }
    float out[kNoiseDraws];
#pragma unroll
    for (int i = 0; i < kNoiseDraws; ++i) {
        const float2 p = parked[i][threadIdx.x];  // written by this thread only: no barrier
        out[i] = ocean_gaussian_finish(p.x, p.y);
    }
    float4* dst = reinterpret_cast<float4*>(noise + ((size_t)st * N + y) * N + x0);  // 4 texels = 32 B, 32-byte aligned
    dst[0] = make_float4(out[0], out[1], out[2], out[3]);
    dst[1] = make_float4(out[4], out[5], out[6], out[7]);
}

__global__ void ocean_noise_kernel(const uint64_t* __restrict__ seeds, int N, float2* __restrict__ noise) {
    noise_texel_pair(seeds, N, noise, blockIdx.x * blockDim.x + threadIdx.x, blockIdx.y, blockIdx.z);
}
// Ensembles: the same work items (gx x gy x gz CTAs' worth) walked by a grid that occupies only part of every SM, so that the
// memory-bound frame kernels of the OTHER lane find thread slots and registers beside this ALU-bound kernel instead of
// only at its tail (OceanEnsemble, OCEAN_ENSEMBLE_ALU_CTAS_PER_SM).  Same arithmetic per item: bit-identical results.
__global__ void __launch_bounds__(128) ocean_noise_persistent_kernel(const uint64_t* __restrict__ seeds, int N, float2* __restrict__ noise,
                                                                     int gx, int gy, int gz) {
    const long long total = (long long)gx * gy * gz;
    for (long long w = blockIdx.x; w < total; w += gridDim.x) {
        const int bx = (int)(w % gx), by = (int)((w / gx) % gy), bz = (int)(w / ((long long)gx * gy));
        noise_texel_pair(seeds, N, noise, bx * blockDim.x + threadIdx.x, by, bz);
    }
}

__device__ __forceinline__ void init_texel_pair(const InitArgs& a, const int x, const int y /* 0 .. N/2 */, const int layerIdx) {
    const int N = a.N;
    if (x >= N) return;
    const int mx = (N - x) & (N - 1), my = (N - y) & (N - 1);  // IS:139
    // rows 0 and N/2 mirror onto themselves: the pair {x, mx} is handled by the thread with x <= mx
    if (my == y && x > mx) return;
    const int layer = layerIdx;
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

#ifndef OCEAN_INIT_MINB
#define OCEAN_INIT_MINB 1
#endif
__global__ void __launch_bounds__(128, OCEAN_INIT_MINB) ocean_init_spectrum_kernel(const InitArgs a) {
    init_texel_pair(a, blockIdx.x * blockDim.x + threadIdx.x, blockIdx.y, blockIdx.z);
}
// (see ocean_noise_persistent_kernel)
__global__ void __launch_bounds__(128) ocean_init_spectrum_persistent_kernel(const InitArgs a, int gx, int gy, int gz) {
    const long long total = (long long)gx * gy * gz;
    for (long long w = blockIdx.x; w < total; w += gridDim.x) {
        const int bx = (int)(w % gx), by = (int)((w / gx) % gy), bz = (int)(w / ((long long)gx * gy));
        init_texel_pair(a, bx * blockDim.x + threadIdx.x, by, bz);
    }
}

// ---------------------------------------------------------------------------------------------
// Spectrum-to-rows kernel (initial spectrum fused into the first FFT pass).  For an ocean
// that is evolved once from a fresh spectrum (ensembles, BASELINE config 5; a live parameter change) h0 and
// waveData never need to exist in memory: one CTA owns the mirrored row pair (y, N - y) of one layer
// (rows 0 and N/2, which mirror onto themselves, share a CTA), evaluates IS:99-143 for it into shared
// memory -- the |k|-only factors once per mirrored texel pair, exactly as ocean_init_spectrum_kernel does, so
// h0 is bit-identical to the stand-alone kernel's -- then runs TD:20-47 and the x-direction inverse FFT of
// the two rows and writes the workspace planes the column pass reads.  Per texel-cascade it reads 8 B of
// noise per texel instead of writing 24 B and reading 24 B.  Built in this translation unit (-fmad=false),
// so the evolution arithmetic is not FMA-contracted: it differs from ocean_rows_kernel's by rounding only
// (both are within the stated tolerance of the oracle; tests/test_gpu_parity.py).
// Measured on B200 (profiles/r1_f_ensemble.txt, 3 x 512^2 per instance): the spectrum is ALU-bound, its memory
// traffic was already hidden, and the fused kernel runs the transcendental code at 24-28 warps per SM instead of
// 48: 41.6 us per instance against 38.0 us with the separate kernels.  It therefore stays opt-in
// (OCEAN_ENSEMBLE_FUSED_INIT=1); what it does buy is memory: 40 B per texel-cascade of pooled h0 / waveHot
// buffers are not allocated.
#ifndef OCEAN_RFN_MINB512
#define OCEAN_RFN_MINB512 6  // 80 registers: measured best of 1 / 5 / 6 / 7 CTAs per SM at 3 x 512^2
#endif
template <int N> struct RowsFromNoiseCfg {
    static constexpr int T = N / 8;
    static constexpr int THREADS = 2 * T;
    static constexpr int MINB = N == 512 ? OCEAN_RFN_MINB512 : 1;
    using FFT = LineFFT<N, 2, false, false, false>;
    // exchange buffer (2 rows) + h0(k) of both rows + (1/|k|, omega) of both rows
    static constexpr int SMEM_BYTES = FFT::SMEM_BYTES + 2 * N * 8 + 2 * N * 8;
};

template <int N>
__global__ void __launch_bounds__(RowsFromNoiseCfg<N>::THREADS, RowsFromNoiseCfg<N>::MINB) ocean_rows_from_noise_kernel(const InitArgs ia, const FrameArgs a) {
    using Cfg = RowsFromNoiseCfg<N>;
    using FFT = typename Cfg::FFT;
    constexpr int T = Cfg::T;
    extern __shared__ float4 smem[];
    float4* xb = smem;                             // [2][N] exchange buffer
    float2* H = (float2*)(smem + 2 * N);           // [2][N] h0(k)
    float2* Wv = H + 2 * N;                        // [2][N] (1/|k|, omega)
    const int t = threadIdx.x;
    const int layer = blockIdx.x / (N / 2), pr = blockIdx.x - layer * (N / 2);
    const bool selfPair = pr == 0;                 // rows 0 and N/2: IS:139 maps each onto itself
    const int yA = selfPair ? 0 : pr, yB = selfPair ? N / 2 : N - pr;
    const int st = layer / ia.C, ci = layer - st * ia.C;
    const StateConsts sc = ia.states[st];
    const CascadeConsts cc = ia.casc[ci];
    const float2* __restrict__ noise = ia.noise + (size_t)st * N * N;

    // ---- IS:99-143 for the two rows -> shared memory ------------------------------------------
    if (!selfPair) {
        // all 2T threads share row yA (4 texels each); a texel's mirror lies in row yB
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
            const int x = t + 2 * T * jj;
            const int mx = (N - x) & (N - 1);
            const float kx = (float)(x - N / 2) * cc.dk, kz = (float)(yA - N / 2) * cc.dk;      // IS:101-103, IS:111
            const float kxm = (float)(mx - N / 2) * cc.dk, kzm = (float)(yB - N / 2) * cc.dk;
            const Radial rd = initial_spectrum_radial(kx, kz, sc, cc, ia.sc);
            const Texel tp = initial_spectrum_texel(kx, kz, rd, sc, cc, __ldg(&noise[(size_t)yA * N + x]));
            const Texel tm = initial_spectrum_texel(kxm, kzm, rd, sc, cc, __ldg(&noise[(size_t)yB * N + mx]));
            H[x] = tp.h;
            H[N + mx] = tm.h;
            Wv[x] = make_float2(tp.wave.y, tp.wave.w);
            Wv[N + mx] = make_float2(tm.wave.y, tm.wave.w);
        }
    } else {
        // the mirror of (x, y) is (N - x, y): a thread owns the texels {x, N - x} of its row for x <= N/2
        const int r = t / T, tid = t - r * T, y = r ? N / 2 : 0;
        for (int x = tid; x <= N / 2; x += T) {
            const int mx = (N - x) & (N - 1);
            const float kx = (float)(x - N / 2) * cc.dk, kz = (float)(y - N / 2) * cc.dk;
            const float kxm = (float)(mx - N / 2) * cc.dk;
            const Radial rd = initial_spectrum_radial(kx, kz, sc, cc, ia.sc);
            const Texel tp = initial_spectrum_texel(kx, kz, rd, sc, cc, __ldg(&noise[(size_t)y * N + x]));
            H[r * N + x] = tp.h;
            Wv[r * N + x] = make_float2(tp.wave.y, tp.wave.w);
            if (mx != x) {
                const Texel tm = initial_spectrum_texel(kxm, kz, rd, sc, cc, __ldg(&noise[(size_t)y * N + mx]));
                H[r * N + mx] = tm.h;
                Wv[r * N + mx] = make_float2(tm.wave.y, tm.wave.w);
            }
        }
    }
    __syncthreads();

    // ---- TD:20-47 + the x-direction inverse FFT (FFT:48-57) of this thread's row -----------------
    const int r = t / T, tid = t - r * T;
    const int y = r ? yB : yA;
    const int mr = selfPair ? r : 1 - r;           // the row that holds this row's mirrored texels
    const float dk = a.dk[layer % a.C];
    const float kz = (float)(y - N / 2) * dk;
    C2 pa[8], pb[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int x = tid + T * j;
        const int mx = (N - x) & (N - 1);
        const float2 h = H[r * N + x], hm = H[mr * N + mx];
        const float kx = (float)(x - N / 2) * dk;
        evolve_texel(make_float4(h.x, h.y, hm.x, -hm.y), Wv[r * N + x], kx, kz, a.time, pa[j], pb[j]);  // IS:141 packing
    }
    FFT::run(pa, xb, tid, r, a.twiddle, 0);
    if (t == 0) spin_until_ge(a.colsDone + layer, a.colsTarget);  // write-after-read on the workspace (see ocean_rows_kernel)
    __syncthreads();
    {
        float4* __restrict__ out = a.rowsA + ((size_t)layer * N + y) * N;
#pragma unroll
        for (int j = 0; j < 8; ++j) c2_store(&out[tid + T * j], pa[j]);
    }
    FFT::run(pb, xb, tid, r, a.twiddle, FFT::EXCHANGES);
    {
        float4* __restrict__ out = a.rowsB + ((size_t)layer * N + y) * N;
#pragma unroll
        for (int j = 0; j < 8; ++j) c2_store(&out[tid + T * j], pb[j]);
    }
    __syncthreads();
    if (t == 0) {
        __threadfence();
        atomicAdd(a.rowsDone + layer, 2ull);
    }
}

bool rows_from_noise_available(int N) { return N == 128 || N == 256 || N == 512 || N == 1024 || N == 2048; }

template <int N> static cudaError_t rows_from_noise_launch(const InitArgs& ia, const FrameArgs& a, cudaStream_t s) {
    using Cfg = RowsFromNoiseCfg<N>;
    static PerDeviceOnce configured;
    if (configured.need()) {
        cudaError_t e = cudaFuncSetAttribute(ocean_rows_from_noise_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        if (e != cudaSuccess) return e;
        cudaFuncSetAttribute(ocean_rows_from_noise_kernel<N>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        configured.mark();
    }
    ocean_rows_from_noise_kernel<N><<<(unsigned)(a.layers * (N / 2)), Cfg::THREADS, Cfg::SMEM_BYTES, s>>>(ia, a);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_rows_from_noise(const InitArgs& ia, const FrameArgs& a, cudaStream_t s) {
    switch (a.N) {
        case 128: return rows_from_noise_launch<128>(ia, a, s);
        case 256: return rows_from_noise_launch<256>(ia, a, s);
        case 512: return rows_from_noise_launch<512>(ia, a, s);
        case 1024: return rows_from_noise_launch<1024>(ia, a, s);
        case 2048: return rows_from_noise_launch<2048>(ia, a, s);
        default: return cudaErrorNotSupported;
    }
}

cudaError_t launch_noise(const uint64_t* seedsDev, int nStates, int N, float2* noise, cudaStream_t s, int maxCtas) {
    dim3 block(N < 128 ? N : 128), grid((N + block.x - 1) / block.x, N, nStates);
    const long long total = (long long)grid.x * grid.y * grid.z;
    static const bool balanced = [] { const char* e = getenv("OCEAN_NOISE_BALANCED"); return e == nullptr || e[0] != '0'; }();
    if (maxCtas > 0 && total > maxCtas) ocean_noise_persistent_kernel<<<maxCtas, block, 0, s>>>(seedsDev, N, noise, (int)grid.x, (int)grid.y, (int)grid.z);
    else if (balanced) {
        const long long groups = (long long)nStates * N * (N / 4);
        ocean_noise_balanced_kernel<<<(unsigned)((groups + kNoiseThreads - 1) / kNoiseThreads), kNoiseThreads, 0, s>>>(seedsDev, N, noise, groups);
    } else ocean_noise_kernel<<<grid, block, 0, s>>>(seedsDev, N, noise);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_init(const InitArgs& a, cudaStream_t s, int maxCtas) {
    dim3 block(a.N < 128 ? a.N : 128), grid((a.N + block.x - 1) / block.x, a.N / 2 + 1, a.nStates * a.C);
    const long long total = (long long)grid.x * grid.y * grid.z;
    if (maxCtas > 0 && total > maxCtas) ocean_init_spectrum_persistent_kernel<<<maxCtas, block, 0, s>>>(a, (int)grid.x, (int)grid.y, (int)grid.z);
    else ocean_init_spectrum_kernel<<<grid, block, 0, s>>>(a);
    count_launch();
    return cudaGetLastError();
}

}  // namespace ocean
