// ocean_fft.cuh -- in-register / shared-memory line FFT used by the row and column kernels.
//
// Replaces the reference's radix-2 Stockham network (IFFT.compute:48-69 driven by IFFT.cs:66-94:
// 2*log2(N) full-texture passes through a ping-pong array) by one pass per direction: a line of
// N complex points is transformed by T = N/8 threads, 8 points per thread, as radix-8
// decimation-in-frequency stages done in registers with 128-bit shared-memory exchanges in
// between.  Net effect is the same unnormalised inverse DFT, natural order in and out
// (SURVEY fact 6); only the summation order -- hence rounding -- differs from the reference.
// The index algebra is restated and proven against numpy in ocean-simulation_b200/fft_plan.py
// (tests/test_fft_plan.py); keep the two in step.
//
// Each element is a float4 = two packed complex spectra (xy, zw), so one exchange moves a pair
// of the reference's four packed spectra (TD:42-45) and every shared-memory access is 128-bit.
#pragma once
#include <cuda_runtime.h>

namespace ocean {

__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}

#define OCEAN_SQRT1_2 0.70710678118654752440f

// u[q] = sum_j v[j] e^{+2 pi i j q / 8}, in place, natural order; both packed spectra at once.
__device__ __forceinline__ void radix8_inv(float4 (&v)[8]) {
    float4 a[4], b[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        a[j] = make_float4(v[j].x + v[j + 4].x, v[j].y + v[j + 4].y, v[j].z + v[j + 4].z, v[j].w + v[j + 4].w);
        b[j] = make_float4(v[j].x - v[j + 4].x, v[j].y - v[j + 4].y, v[j].z - v[j + 4].z, v[j].w - v[j + 4].w);
    }
    // b[j] *= e^{+i pi j / 4}
    {
        float4 t = b[1];  // * (1+i)/sqrt2
        b[1] = make_float4((t.x - t.y) * OCEAN_SQRT1_2, (t.x + t.y) * OCEAN_SQRT1_2,
                           (t.z - t.w) * OCEAN_SQRT1_2, (t.z + t.w) * OCEAN_SQRT1_2);
        t = b[2];  // * i
        b[2] = make_float4(-t.y, t.x, -t.w, t.z);
        t = b[3];  // * (-1+i)/sqrt2
        b[3] = make_float4((-t.x - t.y) * OCEAN_SQRT1_2, (t.x - t.y) * OCEAN_SQRT1_2,
                           (-t.z - t.w) * OCEAN_SQRT1_2, (t.z - t.w) * OCEAN_SQRT1_2);
    }
    // two inverse DFT-4: a -> even outputs, b -> odd outputs
#define OCEAN_DFT4(c, o0, o1, o2, o3)                                                                   \
    {                                                                                                   \
        float4 e0 = make_float4(c[0].x + c[2].x, c[0].y + c[2].y, c[0].z + c[2].z, c[0].w + c[2].w);    \
        float4 e1 = make_float4(c[1].x + c[3].x, c[1].y + c[3].y, c[1].z + c[3].z, c[1].w + c[3].w);    \
        float4 d0 = make_float4(c[0].x - c[2].x, c[0].y - c[2].y, c[0].z - c[2].z, c[0].w - c[2].w);    \
        /* d1 = i * (c1 - c3) */                                                                        \
        float4 d1 = make_float4(-(c[1].y - c[3].y), c[1].x - c[3].x, -(c[1].w - c[3].w), c[1].z - c[3].z); \
        v[o0] = make_float4(e0.x + e1.x, e0.y + e1.y, e0.z + e1.z, e0.w + e1.w);                        \
        v[o2] = make_float4(e0.x - e1.x, e0.y - e1.y, e0.z - e1.z, e0.w - e1.w);                        \
        v[o1] = make_float4(d0.x + d1.x, d0.y + d1.y, d0.z + d1.z, d0.w + d1.w);                        \
        v[o3] = make_float4(d0.x - d1.x, d0.y - d1.y, d0.z - d1.z, d0.w - d1.w);                        \
    }
    OCEAN_DFT4(a, 0, 2, 4, 6)
    OCEAN_DFT4(b, 1, 3, 5, 7)
#undef OCEAN_DFT4
}

// Inverse DFT-4 of v[base + 0..3] -> out[0..3]
__device__ __forceinline__ void radix4_inv(const float4& c0, const float4& c1, const float4& c2, const float4& c3,
                                           float4& o0, float4& o1, float4& o2, float4& o3) {
    float4 e0 = make_float4(c0.x + c2.x, c0.y + c2.y, c0.z + c2.z, c0.w + c2.w);
    float4 e1 = make_float4(c1.x + c3.x, c1.y + c3.y, c1.z + c3.z, c1.w + c3.w);
    float4 d0 = make_float4(c0.x - c2.x, c0.y - c2.y, c0.z - c2.z, c0.w - c2.w);
    float4 d1 = make_float4(-(c1.y - c3.y), c1.x - c3.x, -(c1.w - c3.w), c1.z - c3.z);
    o0 = make_float4(e0.x + e1.x, e0.y + e1.y, e0.z + e1.z, e0.w + e1.w);
    o2 = make_float4(e0.x - e1.x, e0.y - e1.y, e0.z - e1.z, e0.w - e1.w);
    o1 = make_float4(d0.x + d1.x, d0.y + d1.y, d0.z + d1.z, d0.w + d1.w);
    o3 = make_float4(d0.x - d1.x, d0.y - d1.y, d0.z - d1.z, d0.w - d1.w);
}

constexpr int ilog2(int n) { return n <= 1 ? 0 : 1 + ilog2(n >> 1); }

// N points per line, NL lines per CTA.  INTERLEAVED = lines are adjacent columns (column kernel:
// CTA lane = line + NL*tid) and share one interleaved buffer, else every line owns a region
// (row kernel: CTA lane = tid + T*line).  DOUBLE_BUF trades shared memory for one barrier per
// exchange instead of two.  DENSE0: exchange 0 is unswizzled (a TMA tensor copy staged the line in
// natural order in the same buffer); otherwise every exchange is XOR-swizzled.
//
// Twiddles: table tw[(s*7 + q-1)*T + tid] = e^{+2 pi i t q C_s / N}, t = tid mod (M_s/8), one entry
// per (stage, q, thread) so that a warp reads 32 consecutive float2 (see make_twiddles, ocean_host.cu).
template <int N, int NL, bool INTERLEAVED, bool DOUBLE_BUF, bool DENSE0 = false>
struct LineFFT {
    static constexpr int T = N / 8;
    static constexpr int LOGN = ilog2(N);
    static constexpr int S = (LOGN + 2) / 3;             // stages: 8,8,...,RLAST
    static constexpr int RLAST = N >> (3 * (S - 1));     // 2, 4 or 8
    static constexpr int EXCHANGES = S - 1;
    static constexpr int BUF_ELEMS = N * NL;             // float4 elements per buffer
    static constexpr int SMEM_BYTES = BUF_ELEMS * 16 * (DOUBLE_BUF ? 2 : 1);
    static_assert(N >= 16 && (N & (N - 1)) == 0, "N must be a power of two >= 16");

    // logical element F of line `ln` -> physical float4 index; XOR swizzle keeps every 128-bit
    // access of a quarter-warp on 8 distinct 16-byte bank groups (checked in fft_plan.py)
    template <int s>
    __device__ static __forceinline__ int phys(int F, int ln) {
        int G = INTERLEAVED ? (F * NL + ln) : (ln * N + F);
        if constexpr (DENSE0 && s == 0) return G;
        return G ^ ((G >> 3) & 7);
    }
    static constexpr int TWIDDLE_ENTRIES = (S - 1) * 7 * T;

    template <int s>
    __device__ static __forceinline__ void stage(float4 (&reg)[8], float4* __restrict__ smem, const int tid,
                                                 const int ln, const float2* __restrict__ tw, const int prior) {
        constexpr int Cs = 1 << (3 * s);
        constexpr int Ms = N / Cs;
        constexpr bool last = (s == S - 1);
        if constexpr (!last || RLAST == 8) {
            radix8_inv(reg);
            if constexpr (Ms > 8) {
#pragma unroll
                for (int q = 1; q < 8; ++q) {
                    const float2 w = __ldg(&tw[(s * 7 + q - 1) * T + tid]);
                    const float4 r = reg[q];
                    reg[q] = make_float4(r.x * w.x - r.y * w.y, r.x * w.y + r.y * w.x,
                                         r.z * w.x - r.w * w.y, r.z * w.y + r.w * w.x);
                }
            }
            if constexpr (!last) {
                constexpr int Mn = Ms / 8;
                float4* buf = smem + (DOUBLE_BUF ? ((prior + s) & 1) * BUF_ELEMS : 0);
                if constexpr (!DOUBLE_BUF) {
                    if (prior + s > 0) __syncthreads();  // previous exchange fully read
                }
#pragma unroll
                for (int q = 0; q < 8; ++q) buf[phys<s>(tid + T * q, ln)] = reg[q];
                __syncthreads();
                constexpr int rn = (s + 1 == S - 1) ? RLAST : 8;
                if constexpr (rn == 8) {
                    constexpr int Q = Mn / 8;  // threads per sub-problem in the next stage
                    const int cn = tid / Q, tn = tid & (Q - 1);
#pragma unroll
                    for (int j = 0; j < 8; ++j) reg[j] = buf[phys<s>((cn * 8 + j) * Q + tn, ln)];
                } else {
#pragma unroll
                    for (int i = 0; i < 8 / rn; ++i)
#pragma unroll
                        for (int j = 0; j < rn; ++j) reg[i * rn + j] = buf[phys<s>((tid + T * i) * rn + j, ln)];
                }
            }
        } else if constexpr (RLAST == 4) {
            // two inverse DFT-4 on reg[0..3], reg[4..7]; output slot = i + 2*q
            float4 o[8];
            radix4_inv(reg[0], reg[1], reg[2], reg[3], o[0], o[2], o[4], o[6]);
            radix4_inv(reg[4], reg[5], reg[6], reg[7], o[1], o[3], o[5], o[7]);
#pragma unroll
            for (int j = 0; j < 8; ++j) reg[j] = o[j];
        } else {
            // RLAST == 2: four butterflies on (reg[2i], reg[2i+1]); output slot = i + 4*q
            float4 o[8];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float4 p = reg[2 * i], m = reg[2 * i + 1];
                o[i] = make_float4(p.x + m.x, p.y + m.y, p.z + m.z, p.w + m.w);
                o[i + 4] = make_float4(p.x - m.x, p.y - m.y, p.z - m.z, p.w - m.w);
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) reg[j] = o[j];
        }
    }

    // reg[j] = point (tid + T*j) of line `ln` on entry and on return.
    // prior = number of exchanges this CTA has already performed (EXCHANGES per earlier transform).
    // Double buffer: exchange e uses buffer e & 1, one barrier each.  Single buffer: a leading
    // barrier protects the previous exchange's reads before the buffer is overwritten.
    __device__ static __forceinline__ void run(float4 (&reg)[8], float4* __restrict__ smem, const int tid, const int ln,
                                               const float2* __restrict__ tw, const int prior) {
        stage<0>(reg, smem, tid, ln, tw, prior);
        if constexpr (S > 1) stage<1>(reg, smem, tid, ln, tw, prior);
        if constexpr (S > 2) stage<2>(reg, smem, tid, ln, tw, prior);
        if constexpr (S > 3) stage<3>(reg, smem, tid, ln, tw, prior);
    }
};

}  // namespace ocean
