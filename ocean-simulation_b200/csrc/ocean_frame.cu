// ocean_frame.cu -- the per-frame hot path: two kernels per frame instead of the reference's
// 1 + 4*(2*log2 N + 1) + 1 dispatches (WB:180-193, IFFT.cs:66-94).
//
//   ocean_rows_kernel  = CalculateTimeDependentComplexAmplitudesAndDerivatives (TD:20-47)
//                        fused into the x-direction inverse FFT of all four packed spectra
//                        (HorizontalStepIFFT x log2 N, FFT:48-57).
//   ocean_cols_kernel  = the y-direction inverse FFT (VerticalStepIFFT x log2 N, FFT:60-69)
//                        fused with Permute (FFT:73-78) and FillResultTextures (RF:16-34):
//                        (-1)^(x+y), unpack, lambda, Jacobian, foam recursion, 128-bit stores.
//
// Bytes per texel-cascade: rows read h0 16 + (1/|k|, omega) 8 and write 32 of intermediate; cols
// read that 32 (L2-resident) + turbulence 16 and write 3 x 16.  Algorithmic = 96 B.
// Bandwidth-bound fp32 FFT: tensor cores are not used (no dense contraction here).
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <utility>

#include "ocean_fft.cuh"
#include "ocean_internal.h"
#include "ocean_tma.cuh"
#include "ocean_evolve.cuh"

namespace ocean {

// L2 residency of the row->column workspace.  Measured at 4 x 512^2 (4 rotating oceans / one ocean), us per
// frame: no hints 26.6 / 26.0; 6 = output stores evict_first + workspace stores evict_last 25.7-26.0 / 25.1-25.4;
// also marking the workspace or turbulence loads (7, 14, 15) is no better; evict_first on the workspace loads 27.4.
#ifndef OCEAN_L2_HINTS
#define OCEAN_L2_HINTS 6  // bit 0: workspace TMA loads evict_last, bit 1: output TMA stores evict_first, bit 2: workspace stores evict_last, bit 3: turbulence TMA load evict_first
#endif

// Streaming (read-once / write-once) accesses keep the intermediate resident in L2.
__device__ __forceinline__ float4 ld_stream(const float4* p) { return __ldcs(p); }
__device__ __forceinline__ void st_stream(float4* p, float4 v) { __stcs(p, v); }
// 16-byte global store of a C2 with an L2 eviction policy (ocean_tma.cuh constants)
__device__ __forceinline__ void c2_store_hint(float4* p, const C2& v, uint64_t policy) {
    asm volatile("st.global.L2::cache_hint.v2.u64 [%0], {%1, %2}, %3;" ::"l"(p), "l"(v.re), "l"(v.im), "l"(policy) : "memory");
}
// read-once 8-byte load that does not allocate in L1 (the twiddle table should own what L1 is left)
__device__ __forceinline__ float2 ld_once(const float2* p) {
    float2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(p));
    return v;
}

// Ampere-style asynchronous copies (LDGSTS): global -> shared without a register round trip, so a
// thread can have its whole row in flight while holding no data registers.
// (The .L2::cache_hint form of the 16-byte copy raised "illegal instruction" on sm_100a / driver 580
// for some launch shapes, so the copies carry no explicit L2 policy.)
__device__ __forceinline__ void cp_async16(void* smemDst, const void* g) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smemDst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async8(void* smemDst, const void* g) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smemDst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async4(void* smemDst, const void* g) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smemDst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

// Programmatic dependent launch: a kernel launched with the programmatic-stream-serialization
// attribute may begin while its predecessor in the stream is still running.  pdl_trigger() lets the
// successor start early; pdl_wait() blocks until every predecessor grid has completed and flushed.
// The frame kernels do not use pdl_wait() for their mutual dependencies: those are per-layer
// release/acquire counters (FrameArgs::rowsDone / colsDone), which let the row pass of frame f+1 run
// next to the column pass of frame f on the same SMs.  Completion stays transitive along the stream
// through the counters: a column tile waits for its layer's rows, and a row CTA waits, before its first
// store, for the previous frame's columns of its layer to be finished and stored.
// (pdl_trigger / pdl_wait live in ocean_tma.cuh)

// Optional per-CTA phase trace (variants/ builds with -DOCEAN_TRACE only; never in the product build):
// thread 0 of every CTA records %globaltimer / clock64 / %smid at phase boundaries.
#ifdef OCEAN_TRACE
#define OCEAN_TRACE_CTAS 8192
#define OCEAN_TRACE_SLOTS 10
__device__ unsigned long long g_trace[2][OCEAN_TRACE_CTAS][OCEAN_TRACE_SLOTS][2];
__device__ unsigned g_trace_smid[2][OCEAN_TRACE_CTAS];
__device__ __forceinline__ void trace_mark(int kernel, int cta, int slot) {
    if (threadIdx.x == 0 && cta < OCEAN_TRACE_CTAS) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        g_trace[kernel][cta][slot][0] = t;
        g_trace[kernel][cta][slot][1] = (unsigned long long)clock64();
        if (slot == 0) {
            unsigned sm;
            asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
            g_trace_smid[kernel][cta] = sm;
        }
    }
}
#define TRACE(k, cta, slot) trace_mark(k, cta, slot)
#else
#define TRACE(k, cta, slot)
#endif

// tuning knobs for the headline size (overridable at build time for experiments)
#ifndef OCEAN_ROWS512_R
#define OCEAN_ROWS512_R 1
#endif
#ifndef OCEAN_ROWS1024_MINB
#define OCEAN_ROWS1024_MINB 7
#endif
#ifndef OCEAN_ROWS2048_MINB
#define OCEAN_ROWS2048_MINB 3
#endif
#ifndef OCEAN_ROWS512_MINB
#define OCEAN_ROWS512_MINB 14
#endif
#ifndef OCEAN_ROWS512_DB
#define OCEAN_ROWS512_DB false
#endif
#ifndef OCEAN_COLS512_W
#define OCEAN_COLS512_W 8
#endif
#ifndef OCEAN_COLS512_MINB
#define OCEAN_COLS512_MINB 2
#endif
#ifndef OCEAN_COLS512_DB
#define OCEAN_COLS512_DB false
#endif

// Permute (FFT:73-78) + FillResultTextures (RF:16-34) on the packed pair layout.
//   plane A: re = (Dx, Dy), im = (Dz, Dxz)      [DxDz = X + iZ, DyDxz = Y + i Z_x     TD:42-43]
//   plane B: re = (Dyx, Dxx), im = (Dyz, Dzz)   [DyxDyz = Y_x + i Y_z, DxxDzz = X_x + i Z_z  TD:44-45]
// sgn = (-1)^(x+y); lambda scales the horizontal displacement and its derivatives (1 = reference).
struct Epilogue {
    u64 aRe, aIm, b;
    __device__ __forceinline__ Epilogue(float sgn, float lambda)
        : aRe(pk2(sgn * lambda, sgn)), aIm(bc2(sgn * lambda)), b(pk2(sgn, sgn * lambda)) {}
    // displacement = (Dx, Dy, Dz), alpha := 0  RF:25; also returns Dxz for the Jacobian
    __device__ __forceinline__ float4 displacement(const C2& r, float& dxz) const {
        const u64 re = mul2(r.re, aRe), im = mul2(r.im, aIm);
        dxz = hi2(im);
        return make_float4(lo2(re), hi2(re), lo2(im), 0.0f);
    }
    // derivatives = (Dyx, Dyz, Dxx, Dzz)  RF:26;  jac = (1 + Dxx)(1 + Dzz) - Dxz^2  RF:27
    __device__ __forceinline__ float4 derivatives(const C2& r, float dxz, float& jac) const {
        const u64 re = mul2(r.re, b), im = mul2(r.im, b);
        const float dxx = hi2(re), dzz = hi2(im);
        jac = (1.0f + dxx) * (1.0f + dzz) - dxz * dxz;
        return make_float4(lo2(re), lo2(im), dxx, dzz);
    }
};

// Ensemble checksums fused into the column pass: 64-bit sums of the fp32 bit patterns of a state's three
// output arrays (integer addition: order-independent, so identical to a separate reduction pass).
__device__ __forceinline__ unsigned long long bits_sum(const float4 v) {
    return (unsigned long long)__float_as_uint(v.x) + __float_as_uint(v.y) + __float_as_uint(v.z) + __float_as_uint(v.w);
}
__device__ __forceinline__ void checksum_flush(unsigned long long* out, unsigned long long sd, unsigned long long sv,
                                               unsigned long long st) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sd += __shfl_xor_sync(0xffffffffu, sd, o);
        sv += __shfl_xor_sync(0xffffffffu, sv, o);
        st += __shfl_xor_sync(0xffffffffu, st, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(out + 0, sd);
        atomicAdd(out + 1, sv);
        atomicAdd(out + 2, st);
    }
}

#ifndef OCEAN_ROWS_WAVE_REGS
#define OCEAN_ROWS_WAVE_REGS 1
#endif
// Sensitivity experiments (variants/build_variants.py; results are WRONG, only the timing is of interest):
// 1: row pass reads L2-resident rows, 2: column pass stores nothing, 4: no turbulence load, 8: row pass stores nothing,
// 32: no sincos in the row pass, 64: both kernels only wait for / publish their counters (launch + hand-off floor), 128: both kernels return at once,
// 256: the column pass does not wait for its rows (variants/deferred_probe.py)
#ifndef OCEAN_EXP
#define OCEAN_EXP 0
#endif
// (1/|k|, omega) recomputed instead of loaded (see wave_hot in ocean_fused2.cuh): 8 B per texel-cascade less HBM traffic
#ifndef OCEAN_ROWS_RECOMPUTE_HOT
#define OCEAN_ROWS_RECOMPUTE_HOT 0
#endif
__device__ __forceinline__ float2 rows_wave_hot(float kx, float kz, float gravity) {
    const float kMag = __fsqrt_rn(__fadd_rn(__fmul_rn(kx, kx), __fmul_rn(kz, kz)));
    const float ik = kMag > 0.0f ? __frcp_rn(kMag) : 1.0f;
    return make_float2(ik, __fsqrt_rn(__fmul_rn(gravity, kMag)));
}
template <int N> struct RowCfg {
    static constexpr int T = N / 8;
    static constexpr int R = (N == 512) ? OCEAN_ROWS512_R : ((T >= 64) ? 1 : 64 / T);  // rows per CTA
    static constexpr int THREADS = R * T;
    // 72 registers per thread (as many warps per SM as at N = 512) for the one-row-per-CTA sizes
    static constexpr int MINB = (N == 512) ? OCEAN_ROWS512_MINB : (N == 1024 ? OCEAN_ROWS1024_MINB : (N == 2048 ? OCEAN_ROWS2048_MINB : 1));
    // the staged h0 row doubles as the (single) exchange buffer; every exchange is swizzled
    using FFT = LineFFT<N, R, false, false, false>;
#if OCEAN_ROWS_WAVE_REGS
    static constexpr int SMEM_BYTES = R * N * 16;        // h0 float4 per texel; (1/|k|, omega) travels in registers
#else
    static constexpr int SMEM_BYTES = R * N * (16 + 8);  // h0 float4 + (1/|k|, omega) float2 per texel
#endif
};

template <int N> struct ColCfg {
    static constexpr int T = N / 8;
    static constexpr int W = (N == 512) ? OCEAN_COLS512_W : ((T >= 64) ? 2 : ((T >= 32) ? 4 : 8));  // adjacent columns per CTA
    static constexpr int THREADS = W * T;
    static constexpr bool DB = (N == 512) ? OCEAN_COLS512_DB : true;
    static constexpr int MINB = (N == 512) ? OCEAN_COLS512_MINB : 1;
    using FFT = LineFFT<N, W, true, DB>;
};

// ---------------------------------------------------------------------------------------------
// workspace stores of the row pass
__device__ __forceinline__ void ws_store(float4* p, const C2& v, const uint64_t policy = L2_EVICT_LAST) {
#if OCEAN_EXP & 8
    if (v.re == 0x123456789abcdefull) c2_store(p, v);  // sensitivity: (practically) never stores
    return;
#endif
#if OCEAN_L2_HINTS & 4
    c2_store_hint(p, v, policy);
#else
    c2_store(p, v);
#endif
}
// Workspace residency per layer (FrameArgs::wsResident): the planes of the first `wsResident` layers are stored evict_last
// (they are to stay in L2 until the column pass has read them), the planes of the other layers evict_first (they do not
// fit anyway -- 4 x 1024^2 has 134 MB of planes against 126 MB of L2 -- and streaming them keeps the resident ones resident).
__device__ __forceinline__ uint64_t ws_policy(const FrameArgs& a, int layer) { return layer < a.wsResident ? L2_EVICT_LAST : L2_EVICT_FIRST; }

template <int N>
__global__ void __launch_bounds__(RowCfg<N>::THREADS, RowCfg<N>::MINB) ocean_rows_kernel(const FrameArgs a) {
    using Cfg = RowCfg<N>;
    using FFT = typename Cfg::FFT;
    constexpr int T = Cfg::T, R = Cfg::R;
    extern __shared__ float4 smem[];
    float4* hs = smem;                             // [R][N] staged h0, then the exchange buffer
#if !OCEAN_ROWS_WAVE_REGS
    float2* ws = (float2*)(smem + R * N);          // [R][N] staged (1/|k|, omega)
#endif
    const int tid = threadIdx.x & (T - 1);
    const int row = threadIdx.x / T;
    const size_t totalRows = (size_t)a.layers * N;
    size_t line = (size_t)blockIdx.x * R + row;  // = layer*N + y
    const bool active = line < totalRows;
    if (!active) line = totalRows - 1;  // keep every thread in the barriers; results are dropped
#if OCEAN_EXP & 1  // sensitivity: every CTA reads the same 8 rows (L2 hits instead of HBM)
    const float4* __restrict__ H = a.h0 + (line & 7) * N;
    const float2* __restrict__ Wd = a.waveHot + (line & 7) * N;
#else
    const float4* __restrict__ H = a.h0 + line * N;
    const float2* __restrict__ Wd = a.waveHot + line * N;
#endif
    pdl_trigger();
    TRACE(0, blockIdx.x, 0);
#if OCEAN_EXP & 128  // sensitivity: pure launch floor
    return;
#endif
#if OCEAN_EXP & 64  // sensitivity: launch / dispatch / counter floor -- publish and leave
    if (threadIdx.x == 0) {
        spin_until_ge(a.colsDone + (size_t)blockIdx.x * R / N, a.colsTarget);
        __threadfence();
        atomicAdd(a.rowsDone + (size_t)blockIdx.x * R / N, (unsigned long long)R);
    }
    return;
#endif

    // Stage the whole row: 16 asynchronous copies per thread are in flight before any arithmetic.
    // Element x lands in the slot exchange 0 will later use for it, and thread `tid` only ever
    // touches its own 8 elements (x = tid + T*j), so no barrier is needed around the staging.
#if OCEAN_ROWS_WAVE_REGS
    float2 wreg[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int x = tid + T * j;
        cp_async16(&hs[FFT::template phys<0>(x, row)], H + x);
#if !OCEAN_ROWS_RECOMPUTE_HOT
        wreg[j] = ld_once(Wd + x);
#endif
    }
#else
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int x = tid + T * j;
        cp_async16(&hs[FFT::template phys<0>(x, row)], H + x);
        cp_async8(&ws[row * N + x], Wd + x);
    }
#endif
    const int layer = (int)(line / N), y = (int)(line & (N - 1));
    const uint64_t wsPol = ws_policy(a, layer);
    const float dk = a.dk[layer % a.C];
    const float kz = (float)(y - N / 2) * dk;  // IS:111 (waveData.z)
    const float time = a.time;
    cp_async_wait_all();
    TRACE(0, blockIdx.x, 1);

    C2 pa[8], pb[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int x = tid + T * j;
        const float4 h0 = hs[FFT::template phys<0>(x, row)];
        const float kx = (float)(x - N / 2) * dk;  // IS:111 (waveData.x)
#if OCEAN_ROWS_RECOMPUTE_HOT
        const float2 wv = rows_wave_hot(kx, kz, a.states ? a.states[layer / a.C].gravity : a.gravity);
#elif OCEAN_ROWS_WAVE_REGS
        const float2 wv = wreg[j];
#else
        const float2 wv = ws[row * N + x];
#endif
        evolve_texel(h0, wv, kx, kz, time, pa[j], pb[j]);
    }

    TRACE(0, blockIdx.x, 2);
#if OCEAN_ROWS_DUAL_FFT
    FFT::run2(pa, pb, smem, tid, row, a.twiddle, 0);
    TRACE(0, blockIdx.x, 3);
    // h0 / waveHot are never written by a frame kernel, so everything above may overlap the previous
    // frame's column pass; the workspace rows it may still be reading can only be overwritten once it
    // has finished this layer.
    if (threadIdx.x == 0) spin_until_ge(a.colsDone + (size_t)blockIdx.x * R / N, a.colsTarget);
    __syncthreads();
    TRACE(0, blockIdx.x, 4);
    if (active) {
        float4* __restrict__ outA = a.rowsA + line * N;
        float4* __restrict__ outB = a.rowsB + line * N;
#pragma unroll
        for (int j = 0; j < 8; ++j) ws_store(&outA[tid + T * j], pa[j], wsPol);
#pragma unroll
        for (int j = 0; j < 8; ++j) ws_store(&outB[tid + T * j], pb[j], wsPol);
    }
    TRACE(0, blockIdx.x, 5);
    TRACE(0, blockIdx.x, 6);
#else
    FFT::run(pa, smem, tid, row, a.twiddle, 0);
    TRACE(0, blockIdx.x, 3);
    // h0 / waveHot are never written by a frame kernel, so everything above may overlap the previous
    // frame's column pass; the workspace rows it may still be reading can only be overwritten once it
    // has finished this layer.
    if (threadIdx.x == 0) spin_until_ge(a.colsDone + (size_t)blockIdx.x * R / N, a.colsTarget);
    __syncthreads();
    TRACE(0, blockIdx.x, 4);
    if (active) {
        float4* __restrict__ out = a.rowsA + line * N;
#pragma unroll
        for (int j = 0; j < 8; ++j) ws_store(&out[tid + T * j], pa[j], wsPol);
    }
    TRACE(0, blockIdx.x, 5);
    FFT::run(pb, smem, tid, row, a.twiddle, FFT::EXCHANGES);
    TRACE(0, blockIdx.x, 6);
    if (active) {
        float4* __restrict__ out = a.rowsB + line * N;
#pragma unroll
        for (int j = 0; j < 8; ++j) ws_store(&out[tid + T * j], pb[j], wsPol);
    }
#endif
    // publish: this CTA's rows of both planes are in memory (visible to the TMA loads of the column pass)
    __syncthreads();
    if (threadIdx.x == 0) {
        const size_t first = (size_t)blockIdx.x * R;  // R divides N: all rows of a CTA lie in one layer
        const size_t n = first + R <= totalRows ? R : totalRows - first;
        __threadfence();
#ifdef OCEAN_ROWS_PROXY_FENCE
        // (not needed: the column pass's producer lane executes fence.proxy.async after it acquires the
        // counter, which already puts a proxy fence on the causality chain  generic st -> release ->
        // acquire -> fence.proxy.async -> TMA load; FENCE.VIEW.ASYNC here cost 7% of the kernel's warp time)
        fence_proxy_async_global();
#endif
        atomicAdd(a.rowsDone + first / N, (unsigned long long)n);
    }
    TRACE(0, blockIdx.x, 7);
}

// ---------------------------------------------------------------------------------------------
// Row pass for N >= 256: one CTA = the two rows y and N-y of one layer (the pair (0, N/2) for the
// two self-mirrored rows), 2 * N/8 threads.
//
// h(k,t) = h0(k) e^{i w t} + conj(h0(-k)) e^{-i w t}  (TD:26) satisfies h(-k,t) = conj h(k,t): the
// reference stores conj h0(-k) next to h0(k) and evaluates every texel; here a CTA owns a texel
// AND its mirror ((N-x)%N, (N-y)%N), so it
//   * reads the compact h0(k) once per texel (8 B instead of the 16 B packing) and (1/|k|, w)
//     only for the half of the texels it evaluates  ->  12 B instead of 24 B per texel-cascade;
//   * evaluates sincos and h for half of the texels and hands conj h to the mirror through
//     shared memory;
//   * builds the eight derived spectra (TD:29-45) directly in the packed pair layout with f32x2
//     arithmetic: every component is hx * (c0, c1) + hy * (c2, c3) with k-only coefficients.
// Staging, the h exchange and the FFT exchanges share one shared-memory buffer; twiddles are
// staged in shared memory by cp.async alongside the row data.
#ifndef OCEAN_ROWPAIR_MINB
#define OCEAN_ROWPAIR_MINB 7
#endif
template <int N> struct RowPairCfg {
    static constexpr int T = N / 8;
    static constexpr int THREADS = 2 * T;
    static constexpr int MINB = (N == 512) ? OCEAN_ROWPAIR_MINB : 1;
    using FFT = LineFFT<N, 2, false, false, false, true>;
    static constexpr int TW_BYTES = FFT::TWIDDLE_ENTRIES * 8;
    static constexpr int SMEM_BYTES = 2 * N * 16 + TW_BYTES;
};

template <int N>
__global__ void __launch_bounds__(RowPairCfg<N>::THREADS, RowPairCfg<N>::MINB) ocean_rows_pair_kernel(const FrameArgs a) {
    using Cfg = RowPairCfg<N>;
    using FFT = typename Cfg::FFT;
    constexpr int T = Cfg::T, HALF = N / 2;
    extern __shared__ float4 smem[];
    float4* X = smem;                         // [2][N]: staging, then h hand-over, then the FFT exchanges
    float2* twS = (float2*)(smem + 2 * N);
    const int r = threadIdx.x / T, tid = threadIdx.x - r * T;
    const int layer = blockIdx.x / HALF, p = blockIdx.x - layer * HALF;
    // pair p >= 1: rows (p, N-p), each the mirror of the other; pair 0: rows 0 and N/2, self-mirrored
    const int y = p == 0 ? (r ? HALF : 0) : (r ? N - p : p);
    const int ym = (N - y) & (N - 1);
    const int rm = p == 0 ? r : 1 - r;        // buffer slot of the mirror row
    float4* Xr = X + r * N;
    const float4* Xm = X + rm * N;
    const size_t rowBase = ((size_t)layer * N + y) * N, mirBase = ((size_t)layer * N + ym) * N;
    pdl_trigger();
    TRACE(0, blockIdx.x, 0);

    // ---- stage (all private to this thread: no barrier) ----------------------------------------
    // evaluated texels: x = tid + T*j, j < 4 (thread 0 also takes x = N/2, its own mirror column)
    //   slot(x_j).xy = h0(x_j, y), .zw = (1/|k|, w);  slot(x_{j+4}).xy = h0 of the mirror texel
    // thread 0, j = 4: h0 -> slot(x_5).zw, (1/|k|, w) -> slot(x_6).zw, mirror h0 -> slot(x_7).zw
    for (int i = threadIdx.x; i < FFT::TWIDDLE_ENTRIES / 2; i += Cfg::THREADS) cp_async16((float4*)twS + i, (const float4*)a.twiddle + i);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int x = tid + T * j, mx = (N - x) & (N - 1);
        cp_async8(&Xr[x].x, a.h0c + rowBase + x);
        cp_async8(&Xr[x].z, a.waveHot + rowBase + x);
        cp_async8(&Xr[x + 4 * T].x, a.h0c + mirBase + mx);
    }
    if (tid == 0) {
        cp_async8(&Xr[5 * T].z, a.h0c + rowBase + HALF);
        cp_async8(&Xr[6 * T].z, a.waveHot + rowBase + HALF);
        cp_async8(&Xr[7 * T].z, a.h0c + mirBase + HALF);
    }
    const float dk = a.dk[layer % a.C];
    const float kz = (float)(y - HALF) * dk;  // IS:111 (waveData.z)
    const float time = a.time;
    cp_async_wait_all();
    TRACE(0, blockIdx.x, 1);

    // ---- h(k,t) at the evaluated texels (TD:23-26) ------------------------------------------------
    float hx[8], hy[8], ik[8];
    auto evolve = [&](float2 h0, float2 hm, float2 wv, float& ox, float& oy) {
        float sn, cs;
        phase_sincos(wv.y * time, sn, cs);
        // h0 * e + conj(h0(-k)) * conj(e),  conj(h0(-k)) = (hm.x, -hm.y)
        ox = (h0.x * cs - h0.y * sn) + (hm.x * cs - hm.y * sn);
        oy = (h0.x * sn + h0.y * cs) - (hm.y * cs + hm.x * sn);
    };
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float4 own = Xr[tid + T * j];
        const float4 mir = Xr[tid + T * (j + 4)];
        evolve(make_float2(own.x, own.y), make_float2(mir.x, mir.y), make_float2(own.z, own.w), hx[j], hy[j]);
        ik[j] = own.z;
    }
    if (tid == 0) {
        const float4 s5 = Xr[5 * T], s6 = Xr[6 * T], s7 = Xr[7 * T];
        evolve(make_float2(s5.z, s5.w), make_float2(s7.z, s7.w), make_float2(s6.z, s6.w), hx[4], hy[4]);
        ik[4] = s6.z;
    }
    // ---- hand conj h to the mirror texels -----------------------------------------------------------
    // (every slot written below was staged for, and has been read by, this thread only)
#pragma unroll
    for (int j = 0; j < 4; ++j) Xr[tid + T * j] = make_float4(hx[j], hy[j], ik[j], 0.0f);
    if (tid == 0) Xr[4 * T] = make_float4(hx[4], hy[4], ik[4], 0.0f);
    __syncthreads();
#pragma unroll
    for (int j = 4; j < 8; ++j) {
        if (j == 4 && tid == 0) continue;  // x = N/2 was evaluated above
        const int x = tid + T * j;
        const float4 v = Xm[(N - x) & (N - 1)];
        hx[j] = v.x;
        hy[j] = -v.y;
        ik[j] = v.z;
    }
    TRACE(0, blockIdx.x, 2);

    // ---- derived spectra, packed: component = hx * (c0, c1) + hy * (c2, c3)  (TD:29-45) -------------
    //   DxDz   = X + iZ     = ih kx/|k| - h kz/|k|      DyDxz  = Y + i Z_x = h - i h kx kz/|k|
    //   DyxDyz = Y_x + iY_z = ih kx - h kz              DxxDzz = X_x + iZ_z = -h kx^2/|k| - i h kz^2/|k|
    C2 v[8];
    auto plane_a = [&](int j) {
        const float kx = (float)(tid + T * j - HALF) * dk;  // IS:111 (waveData.x)
        const float ca = kx * ik[j], cb = kz * ik[j], cc = ca * kz;
        const u64 H = bc2(hx[j]), G = bc2(hy[j]);
        v[j] = C2{fma2(H, pk2(-cb, 1.0f), mul2(G, pk2(-ca, cc))), fma2(H, pk2(ca, -cc), mul2(G, pk2(-cb, 1.0f)))};
    };
    auto plane_b = [&](int j) {
        const float kx = (float)(tid + T * j - HALF) * dk;
        const float ca = kx * ik[j], cb = kz * ik[j], cd = ca * kx, ce = cb * kz;
        const u64 H = bc2(hx[j]), G = bc2(hy[j]);
        v[j] = C2{fma2(H, pk2(-kz, -cd), mul2(G, pk2(-kx, ce))), fma2(H, pk2(kx, -ce), mul2(G, pk2(-kz, -cd)))};
    };
#pragma unroll
    for (int j = 0; j < 8; ++j) plane_a(j);
    FFT::run(v, X, tid, r, twS, 1);  // prior = 1: barrier before the first exchange (the hand-over reads)
    TRACE(0, blockIdx.x, 3);
    // h0c / waveHot are never written by a frame kernel, so everything above may overlap the previous
    // frame's column pass; the workspace rows it may still be reading can only be overwritten once it
    // has finished this layer.
    if (threadIdx.x == 0) spin_until_ge(a.colsDone + layer, a.colsTarget);
    __syncthreads();
    TRACE(0, blockIdx.x, 4);
    {
        float4* __restrict__ out = a.rowsA + rowBase;
#pragma unroll
        for (int j = 0; j < 8; ++j) ws_store(&out[tid + T * j], v[j]);
    }
    TRACE(0, blockIdx.x, 5);
#pragma unroll
    for (int j = 0; j < 8; ++j) plane_b(j);
    FFT::run(v, X, tid, r, twS, 1 + FFT::EXCHANGES);
    TRACE(0, blockIdx.x, 6);
    {
        float4* __restrict__ out = a.rowsB + rowBase;
#pragma unroll
        for (int j = 0; j < 8; ++j) ws_store(&out[tid + T * j], v[j]);
    }
    // publish: both rows of both planes are in memory (visible to the TMA loads of the column pass)
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        fence_proxy_async_global();
        atomicAdd(a.rowsDone + layer, 2ull);
    }
    TRACE(0, blockIdx.x, 7);
}

// ---------------------------------------------------------------------------------------------
template <int N>
__global__ void __launch_bounds__(ColCfg<N>::THREADS, ColCfg<N>::MINB) ocean_cols_kernel(const FrameArgs a) {
    using Cfg = ColCfg<N>;
    using FFT = typename Cfg::FFT;
    constexpr int T = Cfg::T, W = Cfg::W;
    extern __shared__ float4 smem[];
    const int xw = threadIdx.x & (W - 1);
    const int tid = threadIdx.x / W;
    const int x = blockIdx.x * W + xw;
    const size_t base = (size_t)blockIdx.y * N * N + x;  // texel (x, 0) of this layer
    pdl_trigger();
    pdl_wait();  // the row pass of this frame (and through it every earlier kernel) has completed
    if (a.auxDone) {  // readers of the previous frame's outputs (mip chain) may still be at work
        if (threadIdx.x == 0) spin_until_ge(a.auxDone, a.auxTarget);
        __syncthreads();
    }
    const float lambda = a.lambda;
    // Permute FFT:73-78: (x + y) % 2 with y = tid + T*j and T even  =>  parity of x + tid
    const float sgn = ((x + tid) & 1) ? -1.0f : 1.0f;

    const Epilogue ep(sgn, lambda);
    C2 r[8];
    float dxz[8];
    {
        const float4* __restrict__ in = a.rowsA + base;
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = c2_from(__ldcg(in + (size_t)(tid + T * j) * N));
    }
    FFT::run(r, smem, tid, xw, a.twiddle, 0);
    unsigned long long sumD = 0, sumV = 0, sumT = 0;
    {
        float4* __restrict__ disp = a.disp + base;
        const bool heights = a.heightOut != nullptr && blockIdx.y == 0;  // the readback plane of WB:288 (.g only)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float4 v = ep.displacement(r[j], dxz[j]);
            st_stream(disp + (size_t)(tid + T * j) * N, v);
            if (heights) a.heightOut[(size_t)(tid + T * j) * N + x] = v.y;
            if (a.checksums) sumD += bits_sum(v);
        }
    }
    {
        const float4* __restrict__ in = a.rowsB + base;
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = c2_from(__ldcg(in + (size_t)(tid + T * j) * N));
    }
    FFT::run(r, smem, tid, xw, a.twiddle, FFT::EXCHANGES);
    {
        float4* __restrict__ deriv = a.deriv + base;
        float4* __restrict__ turb = a.turb + base;
        const float decay = a.foamDecay;
        const bool zeroFoam = a.zeroFoam != 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const size_t o = (size_t)(tid + T * j) * N;
            float jac;
            const float4 dv = ep.derivatives(r[j], dxz[j], jac);
            st_stream(deriv + o, dv);
            // RF:28-32
            float foam = zeroFoam ? 0.0f : ld_stream(turb + o).x;
            foam *= decay;
            if (foam < jac) foam += jac;
            st_stream(turb + o, make_float4(foam, foam, foam, foam));
            if (a.checksums) {
                sumV += bits_sum(dv);
                sumT += 4ull * __float_as_uint(foam);
            }
        }
    }
    if (a.checksums) checksum_flush(a.checksums + (size_t)(blockIdx.y / a.C) * 3, sumD, sumV, sumT);
    // hand the layer's workspace rows back to the next frame's row pass
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(a.colsDone + blockIdx.y, (unsigned long long)W);
    }
}

// ---------------------------------------------------------------------------------------------
// Column pass, TMA version (N = 256, 512): persistent, warp-specialised CTAs, two per SM.
//
//   tile      = W adjacent columns x N rows of one layer, handled by W*N/8 compute threads;
//   producer  = one extra warp whose lane 0 owns every bulk-asynchronous copy of the CTA:
//                 TMA tensor loads of the two pair planes (boxes of 4W floats x 256 rows) that land
//                 in natural [y][x] order in the buffer that IS exchange 0 of the FFT, and of the
//                 incoming turbulence tile; TMA tensor stores of the three output tiles;
//   compute   = wait full[A] -> transform in place -> stage the displacement texels in the same
//               buffer -> fence.proxy.async + arrive on staged[A]; then the same for pair B, which
//               stages derivatives in place and turbulence in a third buffer.
// The compute warps never issue a global store and never wait for one: the producer stores the
// staged tile, waits until the store has READ the buffer, and immediately refills it with the next
// tile's plane, so loads, stores and butterflies of neighbouring tiles (and of the second CTA on the
// SM) overlap instead of running in lockstep.  Twiddles live in shared memory.
#ifndef OCEAN_COLS_TMA_W
#define OCEAN_COLS_TMA_W 4
#endif
#ifndef OCEAN_COLS_TMA_MINB
#define OCEAN_COLS_TMA_MINB 2
#endif
// persistent grid = SMs x this many CTAs (capped by occupancy).  Fewer column CTAs per SM leave room
// for row CTAs of the NEXT frame to run in the shadow of this (HBM-bound) kernel.
#ifndef OCEAN_COLS_EARLY_PRODUCER
#define OCEAN_COLS_EARLY_PRODUCER 0  // 1: an isolated column pass starts 1.1 us sooner, back-to-back frames are 0.2 us slower
#endif
#ifndef OCEAN_COLS_TW_SMEM
#define OCEAN_COLS_TW_SMEM 1  // 0: the column pass reads its twiddles through the read-only path (7 KB less shared memory per CTA at N = 512)
#endif
#ifndef OCEAN_COLS_REVERSE_LAYERS
#define OCEAN_COLS_REVERSE_LAYERS 0
#endif
#ifndef OCEAN_COLS_TMA_CTAS_PER_SM
#define OCEAN_COLS_TMA_CTAS_PER_SM 2
#endif

template <int N> struct ColTmaCfg {
    static constexpr int T = N / 8;
    static constexpr int W = OCEAN_COLS_TMA_W;
    static constexpr int COMPUTE = W * T;                      // compute threads
    static constexpr int THREADS = COMPUTE + 32;               // + producer warp
    static constexpr int MINB = N >= 1024 ? 1 : OCEAN_COLS_TMA_MINB;  // N = 1024: 3 x 64 KiB tiles, one CTA per SM
    static constexpr int BOX_ROWS = N < 256 ? N : 256;
    static constexpr int NBOX = N / BOX_ROWS;
    static constexpr int TILE_ELEMS = N * W;                    // 128-bit elements per tile plane
    using FFT = LineFFT<N, W, true, false, true, OCEAN_COLS_TW_SMEM != 0, COMPUTE>;
    static constexpr int TW_BYTES = OCEAN_COLS_TW_SMEM ? FFT::TWIDDLE_ENTRIES * 8 : 0;
    static constexpr int SMEM_BYTES = 3 * TILE_ELEMS * 16 + TW_BYTES + 64;  // A, B, turbulence staging, twiddles, mbarriers
};

// CK: also reduce the ensemble checksums (FrameArgs::checksums) -- per-warp partial sums travel to the
// producer lane through shared memory under the `staged B` barrier, one set of atomics per tile.
template <int N, bool CK = false>
__global__ void __launch_bounds__(ColTmaCfg<N>::THREADS, ColTmaCfg<N>::MINB)
ocean_cols_tma_kernel(const __grid_constant__ ColsTmaMaps maps, const FrameArgs a, const int totalTiles) {
    using Cfg = ColTmaCfg<N>;
    using FFT = typename Cfg::FFT;
    constexpr int T = Cfg::T, W = Cfg::W, TILES_PER_LAYER = N / W;
    // Tile -> layer.  Build option (off): where the row -> column workspace does not fit L2 (4 x 1024^2: 134 MB against
    // 126 MB) take the layers in REVERSE order, on the idea that the row pass wrote layer 0 first and the last layers are the
    // ones still resident.  Measured on B200: 128.0 us per frame against 126.1 in ascending order -- no gain, so ascending
    // order (columns of layer 0 start while the rows of the last layers are still running) stays everywhere.
    constexpr bool REVERSE = OCEAN_COLS_REVERSE_LAYERS && N >= 1024;
    auto layer_of = [&](int tile) { return REVERSE ? a.layers - 1 - tile / TILES_PER_LAYER : tile / TILES_PER_LAYER; };
    extern __shared__ __align__(128) float4 smem[];
    float4* bufA = smem;
    float4* bufB = smem + Cfg::TILE_ELEMS;
    float4* bufT = smem + 2 * Cfg::TILE_ELEMS;
    float2* twSmem = (float2*)(smem + 3 * Cfg::TILE_ELEMS);
    const float2* twS = OCEAN_COLS_TW_SMEM ? twSmem : a.twiddle;  // the table the transforms read
    uint64_t* bars = (uint64_t*)((char*)twSmem + Cfg::TW_BYTES);  // [0] full A, [1] full B, [2] staged A, [3] staged B
    __shared__ unsigned long long ckPart[CK ? (Cfg::COMPUTE / 32) * 3 : 1];  // per-warp (disp, deriv, turb) sums of a tile

    pdl_trigger();
    TRACE(1, blockIdx.x, 0);
#if OCEAN_COLS_EARLY_PRODUCER
    // The producer lane initialises the barriers and goes straight to its first tile's loads; the compute
    // warps copy the (immutable) twiddle table meanwhile and meet on their own named barrier, so the table's
    // L2 round trip and the tile's TMA round trip overlap instead of adding up at the head of every CTA.
    if (threadIdx.x == Cfg::COMPUTE) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_init(&bars[2], Cfg::COMPUTE);
        mbar_init(&bars[3], Cfg::COMPUTE);
        mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x < Cfg::COMPUTE) {
        if (OCEAN_COLS_TW_SMEM) for (int i = threadIdx.x; i < FFT::TWIDDLE_ENTRIES; i += Cfg::COMPUTE) twSmem[i] = __ldg(&a.twiddle[i]);
        FFT::barrier();
    }
#else
    if (threadIdx.x == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_init(&bars[2], Cfg::COMPUTE);
        mbar_init(&bars[3], Cfg::COMPUTE);
        mbar_fence_init();
    }
    // the twiddle table is immutable: copying it may overlap the previous kernel
    if (OCEAN_COLS_TW_SMEM) for (int i = threadIdx.x; i < FFT::TWIDDLE_ENTRIES; i += Cfg::THREADS) twSmem[i] = __ldg(&a.twiddle[i]);
    __syncthreads();
#endif
    // No griddepcontrol.wait here, and none in the row kernels either: the producer gates every tile on its
    // layer's row counter, and a row CTA of this frame only stores (and then counts) once colsDone[layer] says
    // the previous column pass has finished with that layer.  The invariant that makes an `early` row pass
    // legal is on the host side: every producer of h0 / waveHot / the twiddle table (ocean_init_spectrum,
    // Awake) synchronises its stream before a frame can be enqueued, so nothing a row CTA reads is in flight.
    TRACE(1, blockIdx.x, 1);

    // tile -> TMA coordinates: c0 = first fp32 of the tile's columns, c1 = first row
    auto coords = [&](int tile, int& c0, int& c1) {
        const int layer = layer_of(tile), xt = tile % TILES_PER_LAYER;
        c0 = xt * W * 4;
        c1 = layer * N;
    };

#if OCEAN_EXP & 128
    return;
#endif
#if OCEAN_EXP & 64  // sensitivity: launch / dispatch / counter floor -- wait for the rows, publish and leave
    if (threadIdx.x == 0)
        for (int tile = blockIdx.x; tile < totalTiles; tile += gridDim.x) {
            spin_until_ge(a.rowsDone + layer_of(tile), a.rowsTarget);
            __threadfence();
            atomicAdd(a.colsDone + layer_of(tile), (unsigned long long)W);
        }
    return;
#endif
    if (threadIdx.x >= Cfg::COMPUTE) {
        // ---------------- producer warp ----------------
        if (threadIdx.x != Cfg::COMPUTE) return;
        auto load = [&](const CUtensorMap* map, float4* buf, uint64_t* bar, int tile) {
            int c0, c1;
            coords(tile, c0, c1);
            const bool isTurb = map == &maps.turb;
#pragma unroll
            for (int b = 0; b < Cfg::NBOX; ++b) {
                float4* dst = buf + b * Cfg::BOX_ROWS * W;
                const int r0 = c1 + b * Cfg::BOX_ROWS;
                if (!isTurb && layer_of(tile) >= a.wsResident) tma_load_2d_hint(dst, map, bar, c0, r0, L2_EVICT_FIRST);  // streamed layer: read once
                else if ((OCEAN_L2_HINTS & 1) && !isTurb) tma_load_2d_hint(dst, map, bar, c0, r0, L2_EVICT_LAST);       // workspace: stays put
                else if ((OCEAN_L2_HINTS & 8) && isTurb) tma_load_2d_hint(dst, map, bar, c0, r0, L2_EVICT_FIRST);  // read once
                else tma_load_2d(dst, map, bar, c0, r0);
            }
        };
        // pair B travels with the tile's incoming turbulence texels (RF:28), which land in bufT
        const bool needFoam = a.zeroFoam == 0 && !(OCEAN_EXP & 4);  // sensitivity bit 4: no turbulence load
        auto loadB = [&](int tile) {
            mbar_arrive_expect_tx(&bars[1], Cfg::TILE_ELEMS * 16 * (needFoam ? 2 : 1));
            load(&maps.b, bufB, &bars[1], tile);
            if (needFoam) load(&maps.turb, bufT, &bars[1], tile);
        };
        auto store = [&](const CUtensorMap* map, const float4* buf, int tile) {
#if OCEAN_EXP & 2  // sensitivity: no output stores
            return;
#endif
            int c0, c1;
            coords(tile, c0, c1);
#pragma unroll
            for (int b = 0; b < Cfg::NBOX; ++b) {
#if OCEAN_L2_HINTS & 2
                tma_store_2d_hint(map, buf + b * Cfg::BOX_ROWS * W, c0, c1 + b * Cfg::BOX_ROWS, L2_EVICT_FIRST);  // written once
#else
                tma_store_2d(map, buf + b * Cfg::BOX_ROWS * W, c0, c1 + b * Cfg::BOX_ROWS);
#endif
            }
        };
        int readyLayer = -1;  // layers are polled once: counters only grow
        auto await_rows = [&](int tile) {
            const int layer = layer_of(tile);
            if (layer == readyLayer) return;
#if !(OCEAN_EXP & 256)  // sensitivity bit 256: the column pass does not wait for its rows (what a deferred column pass would see)
            spin_until_ge(a.rowsDone + layer, a.rowsTarget);
#endif
            fence_proxy_async_global();  // order the acquired generic-proxy writes before the TMA reads
            readyLayer = layer;
        };
        int tile = blockIdx.x;
        if (tile < totalTiles) {
            await_rows(tile);
            mbar_arrive_expect_tx(&bars[0], Cfg::TILE_ELEMS * 16);
            load(&maps.a, bufA, &bars[0], tile);
            loadB(tile);
        }
        uint32_t parity = 0;
        bool auxIdle = a.auxDone == nullptr;
        for (; tile < totalTiles; tile += gridDim.x, parity ^= 1) {
            const int next = tile + gridDim.x;
            mbar_wait(&bars[2], parity);  // displacement staged in bufA
            if (!auxIdle) {  // readers of the previous frame's outputs (mip chain) may still be at work
                spin_until_ge(a.auxDone, a.auxTarget);
                auxIdle = true;
            }
            store(&maps.disp, bufA, tile);
            bulk_commit();
            bulk_wait_read_all();
            if (next < totalTiles) {
                await_rows(next);
                mbar_arrive_expect_tx(&bars[0], Cfg::TILE_ELEMS * 16);
                load(&maps.a, bufA, &bars[0], next);
            }
            mbar_wait(&bars[3], parity);  // derivatives staged in bufB, turbulence in bufT
            if constexpr (CK) {
                unsigned long long sd = 0, sv = 0, st = 0;
#pragma unroll
                for (int w = 0; w < Cfg::COMPUTE / 32; ++w) {
                    sd += ckPart[w * 3 + 0];
                    sv += ckPart[w * 3 + 1];
                    st += ckPart[w * 3 + 2];
                }
                unsigned long long* out = a.checksums + (size_t)(layer_of(tile) / a.C) * 3;
                atomicAdd(out + 0, sd);
                atomicAdd(out + 1, sv);
                atomicAdd(out + 2, st);
            }
            store(&maps.deriv, bufB, tile);
            store(&maps.turb, bufT, tile);
            bulk_commit();
            bulk_wait_read_all();
            if (next < totalTiles) loadB(next);
            // the tile's three outputs are in memory: hand its columns back (the next frame's row pass
            // may overwrite the workspace; the next frame's column pass may read this turbulence)
            bulk_wait_all();
            fence_proxy_async_global();
            __threadfence();
            atomicAdd(a.colsDone + layer_of(tile), (unsigned long long)W);
        }
        return;
    }

    // ---------------- compute warps ----------------
    const int xw = threadIdx.x & (W - 1);
    const int tid = threadIdx.x / W;
    const float lambda = a.lambda, decay = a.foamDecay;
    const bool zeroFoam = a.zeroFoam != 0 || (OCEAN_EXP & 4);
    uint32_t parity = 0;
    for (int tile = blockIdx.x; tile < totalTiles; tile += gridDim.x, parity ^= 1) {
        const int layer = layer_of(tile), xt = tile % TILES_PER_LAYER;
        const int x = xt * W + xw;
        // Permute FFT:73-78: (x + y) % 2 with y = tid + T*j and T even  =>  parity of x + tid
        const Epilogue ep(((x + tid) & 1) ? -1.0f : 1.0f, lambda);
        C2 r[8];
        float dxz[8];

        // ---- pair A = (DxDz, DyDxz) -> displacement -------------------------------------------
        mbar_wait(&bars[0], parity);
        TRACE(1, tile, 2);
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = c2_load(&bufA[(tid + T * j) * W + xw]);
        FFT::run(r, bufA, tid, xw, twS, 0);
        FFT::barrier();  // every thread has finished reading the last exchange
        TRACE(1, tile, 3);
        unsigned long long sumD = 0, sumV = 0, sumT = 0;
        const bool heights = a.heightOut != nullptr && layer == 0;  // the readback plane of WB:288 (.g only)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float4 v = ep.displacement(r[j], dxz[j]);
            bufA[(tid + T * j) * W + xw] = v;
            if (heights) a.heightOut[(size_t)(tid + T * j) * N + x] = v.y;
            if constexpr (CK) sumD += bits_sum(v);
        }
        fence_proxy_async();
        mbar_arrive(&bars[2]);
        TRACE(1, tile, 4);

        // ---- pair B = (DyxDyz, DxxDzz) -> derivatives, Jacobian, foam --------------------------
        mbar_wait(&bars[1], parity);
        TRACE(1, tile, 5);
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = c2_load(&bufB[(tid + T * j) * W + xw]);
        FFT::run(r, bufB, tid, xw, twS, 0);
        FFT::barrier();
        TRACE(1, tile, 6);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int o = (tid + T * j) * W + xw;
            float jac;
            const float4 dv = ep.derivatives(r[j], dxz[j], jac);                     // RF:26-27
            bufB[o] = dv;
            float f = zeroFoam ? 0.0f : bufT[o].x * decay;                           // RF:28-30 (incoming texel, TMA-staged)
            if (f < jac) f += jac;                                                   // RF:31
            bufT[o] = make_float4(f, f, f, f);                                       // RF:32
            if constexpr (CK) {
                sumV += bits_sum(dv);
                sumT += 4ull * __float_as_uint(f);
            }
        }
        if constexpr (CK) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                sumD += __shfl_xor_sync(0xffffffffu, sumD, o);
                sumV += __shfl_xor_sync(0xffffffffu, sumV, o);
                sumT += __shfl_xor_sync(0xffffffffu, sumT, o);
            }
            if ((threadIdx.x & 31) == 0) {
                unsigned long long* part = ckPart + (threadIdx.x >> 5) * 3;
                part[0] = sumD; part[1] = sumV; part[2] = sumT;
            }
        }
        fence_proxy_async();
        mbar_arrive(&bars[3]);  // (release: the producer lane reads ckPart after its wait on this barrier)
        TRACE(1, tile, 7);
    }
    TRACE(1, blockIdx.x, 8);
}

#include "ocean_fused.cuh"
#include "ocean_fused2.cuh"
#include "ocean_hetero.cuh"

// ---------------------------------------------------------------------------------------------
// WaterBody.GetWaterHeight WB:195-209 (mode 0), plus bilinear extensions at the shader's own
// mapping uv = worldXZ / wavelength (Water.shader:325) with Repeat wrap.
struct QueryArgs {
    const float4* disp;
    const float2* pts;
    float* out;
    int N, C, n, mode;
    float wavelengths[OCEAN_MAX_CASCADES];
};

__device__ __forceinline__ float bilinear_height(const float4* __restrict__ layer, int N, float u, float v) {
    // texel centres at (i + 0.5)/N, Repeat wrap
    const float fx = u * (float)N - 0.5f, fy = v * (float)N - 0.5f;
    const float x0f = floorf(fx), y0f = floorf(fy);
    const float wx = fx - x0f, wy = fy - y0f;
    const int x0 = ((int)x0f) & (N - 1), y0 = ((int)y0f) & (N - 1);
    const int x1 = (x0 + 1) & (N - 1), y1 = (y0 + 1) & (N - 1);
    const float h00 = __ldg(&layer[(size_t)y0 * N + x0]).y, h10 = __ldg(&layer[(size_t)y0 * N + x1]).y;
    const float h01 = __ldg(&layer[(size_t)y1 * N + x0]).y, h11 = __ldg(&layer[(size_t)y1 * N + x1]).y;
    const float a = h00 + (h10 - h00) * wx, b = h01 + (h11 - h01) * wx;
    return a + (b - a) * wy;
}

__global__ void ocean_query_kernel(const QueryArgs q) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= q.n) return;
    const float2 p = q.pts[i];
    const int N = q.N;
    if (q.mode == OCEAN_QUERY_NEAREST_REF) {
        // Mathf.InverseLerp(-N/2, N/2, v) = Clamp01((v - a) / (b - a))   WB:199-200
        const float lo = (float)(-(N / 2)), hi = (float)(N / 2);
        float u = __fdiv_rn(__fsub_rn(p.x, lo), __fsub_rn(hi, lo));
        float v = __fdiv_rn(__fsub_rn(p.y, lo), __fsub_rn(hi, lo));
        u = fminf(fmaxf(u, 0.0f), 1.0f);
        v = fminf(fmaxf(v, 0.0f), 1.0f);
        int x = (int)__fmul_rn(u, (float)N), y = (int)__fmul_rn(v, (float)N);  // WB:203-204
        x = min(max(x, 0), N - 1);
        y = min(max(y, 0), N - 1);
        q.out[i] = __ldg(&q.disp[(size_t)y * N + x]).y;  // .g of layer 0  WB:206-208
        return;
    }
    const int nc = (q.mode == OCEAN_QUERY_BILINEAR_ALL_CASCADES) ? q.C : 1;
    float h = 0.0f;
    for (int c = 0; c < nc; ++c) {
        const float inv = 1.0f / q.wavelengths[c];
        float u = p.x * inv, v = p.y * inv;
        u -= floorf(u);
        v -= floorf(v);
        h += bilinear_height(q.disp + (size_t)c * N * N, N, u, v);
    }
    q.out[i] = h;
}

// Height channel (.y = Dy, what GetWaterHeight WB:206-208 reads as `.g`) of one displacement layer,
// packed for the host readback: 4 B instead of 16 B per texel over PCIe.
__global__ void ocean_pack_height_kernel(const float4* __restrict__ disp, float* __restrict__ out, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = __ldcg(&disp[i]).y;
}

// 64-bit sum of the fp32 bit patterns of one state's array (order-independent => deterministic)
__global__ void ocean_checksum_kernel(const uint4* __restrict__ arr, size_t float4PerState, unsigned long long* out,
                                      int outStride) {
    const uint4* base = arr + (size_t)blockIdx.y * float4PerState;
    unsigned long long acc = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < float4PerState; i += (size_t)gridDim.x * blockDim.x) {
        const uint4 v = base[i];
        acc += (unsigned long long)v.x + v.y + v.z + v.w;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(out + (size_t)blockIdx.y * outStride, acc);
}

// ---------------------------------------------------------------------------------------------
// Launch with programmatic stream serialization (PDL) unless OCEAN_NO_PDL is set in the environment.
static bool pdl_enabled() {
    static int v = -1;
    if (v < 0) v = getenv("OCEAN_NO_PDL") ? 0 : 1;
    return v != 0;
}
template <class... KArgs, class... Args>
static cudaError_t launch_pdl(bool early, const FrameArgs& fa, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smemBytes,
                              cudaStream_t s, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smemBytes; cfg.stream = s;
    cudaLaunchAttribute attr[2];
    int n = 0;
    if (early && pdl_enabled()) {
        attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[n].val.programmaticStreamSerializationAllowed = 1;
        ++n;
    }
    if (fa.l2Window && fa.l2WindowBytes > 0 && fa.l2HitRatio > 0.0f) {
        // keep the row -> column intermediate planes in the persisting part of L2: they are rewritten
        // every frame and must never travel to HBM; everything else streams through the rest of L2
        attr[n].id = cudaLaunchAttributeAccessPolicyWindow;
        attr[n].val.accessPolicyWindow.base_ptr = const_cast<void*>(fa.l2Window);
        attr[n].val.accessPolicyWindow.num_bytes = fa.l2WindowBytes;
        attr[n].val.accessPolicyWindow.hitRatio = fa.l2HitRatio;
        attr[n].val.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        attr[n].val.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        ++n;
    }
    cfg.attrs = attr;
    cfg.numAttrs = n;
    cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
    count_launch();
    return e;
}

template <int N> static cudaError_t rows_launch(const FrameArgs& a, cudaStream_t s, bool early) {
    using Cfg = RowCfg<N>;
    const size_t totalRows = (size_t)a.layers * N;
    const unsigned grid = (unsigned)((totalRows + Cfg::R - 1) / Cfg::R);
    return launch_pdl(early, a, ocean_rows_kernel<N>, dim3(grid), dim3(Cfg::THREADS), Cfg::SMEM_BYTES, s, a);
}
template <int N> static cudaError_t rows_pair_launch(const FrameArgs& a, cudaStream_t s, bool early) {
    using Cfg = RowPairCfg<N>;
    return launch_pdl(early, a, ocean_rows_pair_kernel<N>, dim3((unsigned)(a.layers * (N / 2))), dim3(Cfg::THREADS), Cfg::SMEM_BYTES, s, a);
}
template <int N> static cudaError_t configure_pair() {
    cudaError_t e = cudaFuncSetAttribute(ocean_rows_pair_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, RowPairCfg<N>::SMEM_BYTES);
    if (e != cudaSuccess) return e;
    cudaFuncSetAttribute(ocean_rows_pair_kernel<N>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    return cudaSuccess;
}
template <int N> static cudaError_t cols_launch(const FrameArgs& a, cudaStream_t s) {
    using Cfg = ColCfg<N>;
    dim3 grid(N / Cfg::W, a.layers);
    return launch_pdl(true, a, ocean_cols_kernel<N>, grid, dim3(Cfg::THREADS), Cfg::FFT::SMEM_BYTES, s, a);
}

static int g_numSMs = 0;
static int num_sms() {
    if (g_numSMs == 0) {
        int dev = 0, n = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        g_numSMs = n > 0 ? n : 148;
    }
    return g_numSMs;
}

template <int N> static cudaError_t cols_tma_launch(const FrameArgs& a, const ColsTmaMaps& m, cudaStream_t s) {
    using Cfg = ColTmaCfg<N>;
    const int totalTiles = a.layers * (N / Cfg::W);
    static int ctasPerSM = 0;  // resident CTAs per SM for this kernel: the persistent grid size
    if (ctasPerSM == 0) {
        int n = 0, nck = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, ocean_cols_tma_kernel<N, false>, Cfg::THREADS, Cfg::SMEM_BYTES);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nck, ocean_cols_tma_kernel<N, true>, Cfg::THREADS, Cfg::SMEM_BYTES);
        if (nck < n) n = nck;
        ctasPerSM = n > 0 ? n : 1;
        if (ctasPerSM > OCEAN_COLS_TMA_CTAS_PER_SM) ctasPerSM = OCEAN_COLS_TMA_CTAS_PER_SM;
    }
    const int slots = num_sms() * ctasPerSM;
    const int grid = totalTiles < slots ? totalTiles : slots;
    if (a.checksums) return launch_pdl(true, a, ocean_cols_tma_kernel<N, true>, dim3(grid), dim3(Cfg::THREADS), Cfg::SMEM_BYTES, s, m, a, totalTiles);
    return launch_pdl(true, a, ocean_cols_tma_kernel<N, false>, dim3(grid), dim3(Cfg::THREADS), Cfg::SMEM_BYTES, s, m, a, totalTiles);
}
template <int N> static cudaError_t configure_tma() {
    cudaError_t e = cudaFuncSetAttribute(ocean_cols_tma_kernel<N, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, ColTmaCfg<N>::SMEM_BYTES);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(ocean_cols_tma_kernel<N, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, ColTmaCfg<N>::SMEM_BYTES);
    if (e != cudaSuccess) return e;
    cudaFuncSetAttribute(ocean_cols_tma_kernel<N, false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(ocean_cols_tma_kernel<N, true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    return cudaSuccess;
}
template <int N> static cudaError_t frame_launch(const FrameArgs& a, const ColsTmaMaps& m, cudaStream_t s, bool early) {
    using Cfg = FusedCfg<N>;
    static PerDeviceOnce configured;
    if (configured.need()) {
        cudaError_t e = cudaFuncSetAttribute(ocean_frame_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        if (e != cudaSuccess) return e;
        cudaFuncSetAttribute(ocean_frame_kernel<N>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        configured.mark();
    }
    static int ctasPerSM = 0;
    if (ctasPerSM == 0) {
        int n = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, ocean_frame_kernel<N>, Cfg::THREADS, Cfg::SMEM_BYTES);
        ctasPerSM = n > 0 ? n : 1;
    }
    const int totalItems = 2 * a.layers * Cfg::IPL;
    const int slots = num_sms() * ctasPerSM;
    const int grid = totalItems < slots ? totalItems : slots;
    return launch_pdl(early, a, ocean_frame_kernel<N>, dim3(grid), dim3(Cfg::THREADS), Cfg::SMEM_BYTES, s, m, a);
}
template <int N> static int frame2_ctas_per_sm() {
    using Cfg = Fused2Cfg<N>;
    static PerDeviceOnce configured;
    if (configured.need()) {
        cudaFuncSetAttribute(ocean_frame2_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        cudaFuncSetAttribute(ocean_frame2_kernel<N>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        configured.mark();
    }
    static int ctasPerSM = 0;
    if (ctasPerSM == 0) {
        int n = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, ocean_frame2_kernel<N>, Cfg::THREADS, Cfg::SMEM_BYTES);
        ctasPerSM = n > 0 ? n : 1;
    }
    return ctasPerSM;
}
template <int N> static int frame2_grid(int layers) {
    const int totalItems = 2 * layers * Fused2Cfg<N>::IPL;
    const int slots = num_sms() * frame2_ctas_per_sm<N>();
    return totalItems < slots ? totalItems : slots;
}
template <int N> static cudaError_t frame2_launch(const FrameArgs& a, const ColsTmaMaps& m, cudaStream_t s, bool early) {
    using Cfg = Fused2Cfg<N>;
    return launch_pdl(early, a, ocean_frame2_kernel<N>, dim3(frame2_grid<N>(a.layers)), dim3(Cfg::THREADS), Cfg::SMEM_BYTES, s, m, a);
}
// Probe: both roles in one persistent CTA per SM (ocean_hetero.cuh).  mode 1: in order (column team waits for its rows:
// results identical to the two-kernel frame), mode 2: column team does not wait (timing of a deferred schedule; WRONG results).
int hetero_mode(int N, const ColsTmaMaps* maps) {
    static const int mode = [] { const char* e = getenv("OCEAN_HETERO"); return e ? atoi(e) : 0; }();
    return (N == 512 && maps && maps->valid) ? mode : 0;
}
cudaError_t launch_hetero(const FrameArgs& a, const ColsTmaMaps* maps, cudaStream_t s, int mode) {
    static PerDeviceOnce configured;
    if (configured.need()) {
        cudaError_t e = cudaFuncSetAttribute(ocean_hetero_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, HeteroCfg::SMEM_BYTES);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(ocean_hetero_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, HeteroCfg::SMEM_BYTES);
        if (e != cudaSuccess) return e;
        configured.mark();
    }
    const int totalTiles = a.layers * (HeteroCfg::N / HeteroCfg::W);
    const int grid = num_sms();
    if (mode == 2) return launch_pdl(true, a, ocean_hetero_kernel<false>, dim3(grid), dim3(HeteroCfg::THREADS), HeteroCfg::SMEM_BYTES, s, *maps, a, a, totalTiles);
    return launch_pdl(true, a, ocean_hetero_kernel<true>, dim3(grid), dim3(HeteroCfg::THREADS), HeteroCfg::SMEM_BYTES, s, *maps, a, a, totalTiles);
}
static bool tma_size(int N) { return N == 256 || N == 512 || N == 1024; }
// The whole frame as one persistent kernel.  Mode 1 (static round robin) was the default at N = 256 in round 1, where the
// frame is a latency chain and one launch beats two (config 1: 9.3 vs 10.5 us, config 2: 10.0 vs 13.0 us); at N = 512 it
// lost to the two-kernel path (29.4 vs 26.5 us at 4 x 512^2: its items waited for their loads and for each other).  Mode 2
// (ocean_fused2.cuh) hands items out dynamically and prefetches.  OCEAN_FUSED=0 / 1 / 2 forces a mode for both sizes.
#ifndef OCEAN_FUSED_DEFAULT_256
#define OCEAN_FUSED_DEFAULT_256 1
#endif
#ifndef OCEAN_FUSED_DEFAULT_512
#define OCEAN_FUSED_DEFAULT_512 0
#endif
#ifndef OCEAN_FUSED_DEFAULT_1024
#define OCEAN_FUSED_DEFAULT_1024 0
#endif
int frame_fused_mode(int N, const ColsTmaMaps* maps) {
    if (!(N == 256 || N == 512 || N == 1024) || !maps || !maps->valid) return 0;
    static const int mode = [] {
        const char* e = getenv("OCEAN_FUSED");
        return e == nullptr ? -1 : (e[0] - '0');
    }();
    if (N == 1024) {  // static item order only (layer-interleaved, ocean_fused.cuh); OCEAN_FUSED_1024 overrides for this size alone
        static const int m1024 = [] { const char* e = getenv("OCEAN_FUSED_1024"); return e == nullptr ? -1 : (e[0] != '0'); }();
        if (m1024 >= 0) return m1024;
        return mode >= 0 ? (mode != 0) : OCEAN_FUSED_DEFAULT_1024;
    }
    if (mode >= 0 && mode <= 2) return mode;
    return N == 256 ? OCEAN_FUSED_DEFAULT_256 : OCEAN_FUSED_DEFAULT_512;
}
int frame_fused_grid(int N, int layers, int mode) {
    if (mode != 2) return 0;
    return N == 512 ? frame2_grid<512>(layers) : frame2_grid<256>(layers);
}
cudaError_t launch_frame(const FrameArgs& a, const ColsTmaMaps* maps, cudaStream_t s, bool early, int mode) {
    if (mode == 2) {
        if (a.N == 512) return frame2_launch<512>(a, *maps, s, early);
        if (a.N == 256) return frame2_launch<256>(a, *maps, s, early);
        return cudaErrorNotSupported;
    }
    if (a.N == 512) return frame_launch<512>(a, *maps, s, early);
    if (a.N == 256) return frame_launch<256>(a, *maps, s, early);
    if (a.N == 1024) return frame_launch<1024>(a, *maps, s, early);
    return cudaErrorNotSupported;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

cudaError_t make_cols_maps(int N, int layers, const float4* rowsA, const float4* rowsB, float4* disp, float4* deriv, float4* turb,
                           ColsTmaMaps* out) {
    out->valid = 0;
    if (!tma_size(N)) return cudaSuccess;
    static EncodeTiledFn encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
        if (e != cudaSuccess) return e;
        if (!fn || q != cudaDriverEntryPointSuccess) return cudaErrorNotSupported;
        encode = (EncodeTiledFn)fn;
    }
    const int W = OCEAN_COLS_TMA_W;
    const int boxRows = N < 256 ? N : 256;
    const cuuint64_t dims[2] = {(cuuint64_t)N * 4, (cuuint64_t)layers * N};  // fp32 elements per row, rows
    const cuuint64_t strides[1] = {(cuuint64_t)N * 16};                      // bytes between rows
    const cuuint32_t box[2] = {(cuuint32_t)W * 4, (cuuint32_t)boxRows};
    const cuuint32_t estr[2] = {1, 1};
    const void* ptrs[5] = {rowsA, rowsB, disp, deriv, turb};
    CUtensorMap* maps[5] = {&out->a, &out->b, &out->disp, &out->deriv, &out->turb};
    for (int i = 0; i < 5; ++i) {
        CUresult r = encode(maps[i], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)ptrs[i], dims, strides, box, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
    }
    out->valid = 1;
    return cudaSuccess;
}

template <int N> static cudaError_t configure() {
    cudaError_t e = cudaFuncSetAttribute(ocean_rows_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, RowCfg<N>::SMEM_BYTES);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(ocean_cols_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, ColCfg<N>::FFT::SMEM_BYTES);
    if (e != cudaSuccess) return e;
    // prefer the largest shared-memory carve-out: occupancy is set by registers and smem, not L1
    cudaFuncSetAttribute(ocean_rows_kernel<N>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(ocean_cols_kernel<N>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    return cudaSuccess;
}
template <int N> static const char* describe() {
    static char buf[768];
    if (N == 1024 || N == 512 || N == 256) {
        constexpr int M = (N == 1024 || N == 512 || N == 256) ? N : 512;
        snprintf(buf, sizeof buf,
                 "rows<N=%d>: %d rows/CTA, %d threads, %d B smem (cp.async-staged row = exchange buffer); "
                 "cols_tma<N=%d>: persistent warp-specialised, %d columns/tile, %d threads, %d B smem (TMA in, TMA out)",
                 N, RowCfg<N>::R, RowCfg<N>::THREADS, RowCfg<N>::SMEM_BYTES, N, ColTmaCfg<M>::W, ColTmaCfg<M>::THREADS,
                 ColTmaCfg<M>::SMEM_BYTES);
        if (N == 256) {
            const size_t n = strlen(buf);
            snprintf(buf + n, sizeof buf - n, "; whole frames run as one fused persistent kernel (ocean_frame_kernel) unless OCEAN_FUSED=0");
        }
        return buf;
    }
    snprintf(buf, sizeof buf,
             "rows<N=%d>: %d rows/CTA, %d threads, %d B smem, %s buffer; cols<N=%d>: %d columns/CTA, %d threads, %d B smem, %s buffer",
             N, RowCfg<N>::R, RowCfg<N>::THREADS, RowCfg<N>::SMEM_BYTES, "single", N,
             ColCfg<N>::W, ColCfg<N>::THREADS, ColCfg<N>::FFT::SMEM_BYTES, ColCfg<N>::DB ? "double" : "single");
    return buf;
}

#define OCEAN_DISPATCH_N(N_, CALL)                 \
    switch (N_) {                                  \
        case 16: return CALL(16);                  \
        case 32: return CALL(32);                  \
        case 64: return CALL(64);                  \
        case 128: return CALL(128);                \
        case 256: return CALL(256);                \
        case 512: return CALL(512);                \
        case 1024: return CALL(1024);              \
        case 2048: return CALL(2048);              \
        default: break;                            \
    }

#ifdef OCEAN_TRACE
extern "C" int ocean_debug_trace_copy(unsigned long long* times, unsigned* smids) {
    cudaDeviceSynchronize();
    if (cudaMemcpyFromSymbol(times, g_trace, sizeof(g_trace)) != cudaSuccess) return 1;
    if (cudaMemcpyFromSymbol(smids, g_trace_smid, sizeof(g_trace_smid)) != cudaSuccess) return 1;
    return 0;
}
#endif

// true when the row pass in use reads FrameArgs::h0c (the pair-row kernel): only then is it allocated and filled
bool rows_need_h0c() {
#ifdef OCEAN_ROWS_USE_PAIR
    return true;
#else
    static const bool pair = getenv("OCEAN_ROWS_PAIR") != nullptr;
    return pair;
#endif
}

cudaError_t launch_rows(const FrameArgs& a, cudaStream_t s, bool early) {
    // The pair-row kernel moves half the bytes and issues fewer instructions, but on B200 at N = 512 the
    // one-row kernel is still ~1 us/frame faster (more, smaller CTAs hide more latency): it stays the
    // default until the pair kernel wins.  OCEAN_ROWS_PAIR=1 (or -DOCEAN_ROWS_USE_PAIR) selects it.
    const bool legacy = !rows_need_h0c() || a.h0c == nullptr;
    if (!legacy) {
        if (a.N == 256) return rows_pair_launch<256>(a, s, early);
        if (a.N == 512) return rows_pair_launch<512>(a, s, early);
        if (a.N == 1024) return rows_pair_launch<1024>(a, s, early);
        if (a.N == 2048) return rows_pair_launch<2048>(a, s, early);
    }
#define CALL(n) rows_launch<n>(a, s, early)
    OCEAN_DISPATCH_N(a.N, CALL)
#undef CALL
    return cudaErrorInvalidValue;
}
cudaError_t launch_cols(const FrameArgs& a, const ColsTmaMaps* maps, cudaStream_t s) {
    if (maps && maps->valid) {
        if (a.N == 512) return cols_tma_launch<512>(a, *maps, s);
        if (a.N == 256) return cols_tma_launch<256>(a, *maps, s);
        if (a.N == 1024) return cols_tma_launch<1024>(a, *maps, s);
    }
#define CALL(n) cols_launch<n>(a, s)
    OCEAN_DISPATCH_N(a.N, CALL)
#undef CALL
    return cudaErrorInvalidValue;
}
cudaError_t configure_frame_kernels(int N) {
    if (N == 256) { cudaError_t e = configure_pair<256>(); if (e != cudaSuccess) return e; }
    if (N == 512) { cudaError_t e = configure_pair<512>(); if (e != cudaSuccess) return e; }
    if (N == 1024) { cudaError_t e = configure_pair<1024>(); if (e != cudaSuccess) return e; }
    if (N == 2048) { cudaError_t e = configure_pair<2048>(); if (e != cudaSuccess) return e; }
    if (N == 512) { cudaError_t e = configure_tma<512>(); if (e != cudaSuccess) return e; }
    if (N == 256) { cudaError_t e = configure_tma<256>(); if (e != cudaSuccess) return e; }
    if (N == 1024) { cudaError_t e = configure_tma<1024>(); if (e != cudaSuccess) return e; }
#define CALL(n) configure<n>()
    OCEAN_DISPATCH_N(N, CALL)
#undef CALL
    return cudaErrorInvalidValue;
}
const char* describe_frame_kernels(int N) {
#define CALL(n) describe<n>()
    OCEAN_DISPATCH_N(N, CALL)
#undef CALL
    return "unsupported N";
}

cudaError_t launch_query(const float4* disp, int N, int C, const float* wavelengths, const float2* pts, float* out, int n,
                         int mode, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    QueryArgs q;
    q.disp = disp; q.pts = pts; q.out = out; q.N = N; q.C = C; q.n = n; q.mode = mode;
    for (int i = 0; i < OCEAN_MAX_CASCADES; ++i) q.wavelengths[i] = i < C ? wavelengths[i] : 1.0f;
    ocean_query_kernel<<<(n + 255) / 256, 256, 0, s>>>(q);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_pack_height(const float4* disp, float* out, int n, cudaStream_t s) {
    ocean_pack_height_kernel<<<(n + 255) / 256, 256, 0, s>>>(disp, out, n);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_checksum(const float4* arr, int nStates, size_t float4PerState, unsigned long long* out, int outStride,
                            cudaStream_t s) {
    dim3 grid(64, nStates);
    ocean_checksum_kernel<<<grid, 256, 0, s>>>((const uint4*)arr, float4PerState, out, outStride);
    count_launch();
    return cudaGetLastError();
}

}  // namespace ocean
