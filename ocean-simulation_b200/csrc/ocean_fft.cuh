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
// Element type: C2 = TWO complex values, one from each of two packed spectra (TD:42-45), held
// structure-of-arrays as two 64-bit register pairs  re = (re_a, re_b), im = (im_a, im_b).  Every
// butterfly, rotation and twiddle multiplication is then element-wise on a pair, which is exactly
// what Blackwell's packed fp32 instructions do (PTX add/sub/mul/fma .f32x2 -> SASS FADD2 / FMUL2 /
// FFMA2, sm_100+): one issue slot per two flops-lanes, halving the FP instruction count of the
// transform.  A twiddle (wx, wy) multiplies both spectra, so it enters as a scalar broadcast
// operand (ptxas emits the .F32 broadcast form; negations fold into operand modifiers).
// In memory a C2 is one 128-bit element (re_a, re_b, im_a, im_b).
#pragma once
#include <cuda_runtime.h>

namespace ocean {

typedef unsigned long long u64;

__device__ __forceinline__ u64 pk2(float lo, float hi) {
    u64 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ float lo2(u64 v) {
    [[maybe_unused]] float a, b;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
    return a;
}
__device__ __forceinline__ float hi2(u64 v) {
    [[maybe_unused]] float a, b;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
    return b;
}
__device__ __forceinline__ u64 bc2(float s) { return pk2(s, s); }
__device__ __forceinline__ u64 add2(u64 a, u64 b) {
    u64 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ u64 sub2(u64 a, u64 b) {
    u64 r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ u64 mul2(u64 a, u64 b) {
    u64 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) {
    u64 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ u64 neg2(u64 a) { return pk2(-lo2(a), -hi2(a)); }  // folds into operand modifiers

struct C2 {
    u64 re, im;
};
__device__ __forceinline__ C2 operator+(const C2& a, const C2& b) { return C2{add2(a.re, b.re), add2(a.im, b.im)}; }
__device__ __forceinline__ C2 operator-(const C2& a, const C2& b) { return C2{sub2(a.re, b.re), sub2(a.im, b.im)}; }
__device__ __forceinline__ C2 mul_i(const C2& a) { return C2{neg2(a.im), a.re}; }  // a * i
// both spectra times the same complex scalar w
__device__ __forceinline__ C2 cmul(const C2& a, float2 w) {
    const u64 wx = bc2(w.x), wy = bc2(w.y);
    return C2{fma2(a.re, wx, neg2(mul2(a.im, wy))), fma2(a.re, wy, mul2(a.im, wx))};
}
__device__ __forceinline__ C2 c2_from(const float4& v) { return C2{pk2(v.x, v.y), pk2(v.z, v.w)}; }
__device__ __forceinline__ float4 c2_to(const C2& v) { return make_float4(lo2(v.re), hi2(v.re), lo2(v.im), hi2(v.im)); }
__device__ __forceinline__ C2 c2_load(const float4* p) {
    const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(p);
    return C2{v.x, v.y};
}
__device__ __forceinline__ void c2_store(float4* p, const C2& v) { *reinterpret_cast<ulonglong2*>(p) = make_ulonglong2(v.re, v.im); }

#define OCEAN_SQRT1_2 0.70710678118654752440f
#ifndef OCEAN_HETERO_BASE
#define OCEAN_HETERO_BASE 288  // threads of the column role (256 compute + producer warp) in the heterogeneous frame kernel
#endif

// Inverse DFT-4: o[q] = sum_j c[j] e^{+2 pi i j q / 4}
__device__ __forceinline__ void radix4_inv(const C2& c0, const C2& c1, const C2& c2, const C2& c3, C2& o0, C2& o1, C2& o2, C2& o3) {
    const C2 e0 = c0 + c2, e1 = c1 + c3, d0 = c0 - c2, t = c1 - c3;
    o0 = e0 + e1;
    o2 = e0 - e1;
    // d1 = i * t = (-t.im, t.re)
    o1 = C2{sub2(d0.re, t.im), add2(d0.im, t.re)};
    o3 = C2{add2(d0.re, t.im), sub2(d0.im, t.re)};
}

// u[q] = sum_j v[j] e^{+2 pi i j q / 8}, in place, natural order; both packed spectra at once.
__device__ __forceinline__ void radix8_inv(C2 (&v)[8]) {
    C2 a[4], b[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        a[j] = v[j] + v[j + 4];
        b[j] = v[j] - v[j + 4];
    }
    // b[j] *= e^{+i pi j / 4}
    const u64 c = bc2(OCEAN_SQRT1_2);
    {
        C2 t = b[1];  // * (1+i)/sqrt2
        b[1] = C2{mul2(sub2(t.re, t.im), c), mul2(add2(t.re, t.im), c)};
        b[2] = mul_i(b[2]);
        t = b[3];  // * (-1+i)/sqrt2
        b[3] = C2{mul2(neg2(add2(t.re, t.im)), c), mul2(sub2(t.re, t.im), c)};
    }
    // two inverse DFT-4: a -> even outputs, b -> odd outputs
    radix4_inv(a[0], a[1], a[2], a[3], v[0], v[2], v[4], v[6]);
    radix4_inv(b[0], b[1], b[2], b[3], v[1], v[3], v[5], v[7]);
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
// The table may live in global memory (read-only path) or in shared memory (persistent kernels).
// BAR_THREADS > 0: the CTA barrier is the named barrier 1 over the first BAR_THREADS threads (warp-
// specialised kernels whose producer warp does not take part); 0 = __syncthreads().
template <int N, int NL, bool INTERLEAVED, bool DOUBLE_BUF, bool DENSE0 = false, bool TW_SMEM = false, int BAR_THREADS = 0>
struct LineFFT {
    static constexpr int T = N / 8;
    static constexpr int LOGN = ilog2(N);
    static constexpr int S = (LOGN + 2) / 3;             // stages: 8,8,...,RLAST
    static constexpr int RLAST = N >> (3 * (S - 1));     // 2, 4 or 8
    static constexpr int EXCHANGES = S - 1;
    static constexpr int BUF_ELEMS = N * NL;             // 128-bit elements per buffer
    static constexpr int SMEM_BYTES = BUF_ELEMS * 16 * (DOUBLE_BUF ? 2 : 1);
    static_assert(N >= 16 && (N & (N - 1)) == 0, "N must be a power of two >= 16");

    // logical element F of line `ln` -> physical 128-bit index; XOR swizzle keeps every 128-bit
    // access of a quarter-warp on 8 distinct 16-byte bank groups (checked in fft_plan.py)
    template <int s>
    __device__ static __forceinline__ int phys(int F, int ln) {
        int G = INTERLEAVED ? (F * NL + ln) : (ln * N + F);
        if constexpr (DENSE0 && s == 0) return G;
        return G ^ ((G >> 3) & 7);
    }
    static constexpr int TWIDDLE_ENTRIES = (S - 1) * 7 * T;

    __device__ static __forceinline__ void barrier() {
        if constexpr (BAR_THREADS > 0) asm volatile("bar.sync 1, %0;" ::"n"(BAR_THREADS) : "memory");
        else if constexpr (BAR_THREADS < 0) {
            // line groups of -BAR_THREADS threads that follow the first OCEAN_HETERO_BASE threads of the CTA, one named
            // barrier (2, 3, ...) per group (heterogeneous frame kernel, ocean_hetero.cuh)
            asm volatile("bar.sync %0, %1;" ::"r"(2 + ((int)threadIdx.x - OCEAN_HETERO_BASE) / (-BAR_THREADS)), "n"(-BAR_THREADS) : "memory");
        } else __syncthreads();
    }
    __device__ static __forceinline__ float2 twiddle(const float2* __restrict__ tw, int i) {
        if constexpr (TW_SMEM) return tw[i];
        else return __ldg(&tw[i]);
    }

    template <int s>
    __device__ static __forceinline__ void stage(C2 (&reg)[8], float4* __restrict__ smem, const int tid,
                                                 const int ln, const float2* __restrict__ tw, const int prior) {
        constexpr int Cs = 1 << (3 * s);
        constexpr int Ms = N / Cs;
        constexpr bool last = (s == S - 1);
        if constexpr (!last || RLAST == 8) {
            radix8_inv(reg);
            if constexpr (Ms > 8) {
#pragma unroll
                for (int q = 1; q < 8; ++q) reg[q] = cmul(reg[q], twiddle(tw, (s * 7 + q - 1) * T + tid));
            }
            if constexpr (!last) {
                constexpr int Mn = Ms / 8;
                float4* buf = smem + (DOUBLE_BUF ? ((prior + s) & 1) * BUF_ELEMS : 0);
                if constexpr (!DOUBLE_BUF) {
                    if (prior + s > 0) barrier();  // previous exchange fully read
                }
#pragma unroll
                for (int q = 0; q < 8; ++q) c2_store(&buf[phys<s>(tid + T * q, ln)], reg[q]);
                barrier();
                constexpr int rn = (s + 1 == S - 1) ? RLAST : 8;
                if constexpr (rn == 8) {
                    constexpr int Q = Mn / 8;  // threads per sub-problem in the next stage
                    const int cn = tid / Q, tn = tid & (Q - 1);
#pragma unroll
                    for (int j = 0; j < 8; ++j) reg[j] = c2_load(&buf[phys<s>((cn * 8 + j) * Q + tn, ln)]);
                } else {
#pragma unroll
                    for (int i = 0; i < 8 / rn; ++i)
#pragma unroll
                        for (int j = 0; j < rn; ++j) reg[i * rn + j] = c2_load(&buf[phys<s>((tid + T * i) * rn + j, ln)]);
                }
            }
        } else if constexpr (RLAST == 4) {
            // two inverse DFT-4 on reg[0..3], reg[4..7]; output slot = i + 2*q
            C2 o[8];
            radix4_inv(reg[0], reg[1], reg[2], reg[3], o[0], o[2], o[4], o[6]);
            radix4_inv(reg[4], reg[5], reg[6], reg[7], o[1], o[3], o[5], o[7]);
#pragma unroll
            for (int j = 0; j < 8; ++j) reg[j] = o[j];
        } else {
            // RLAST == 2: four butterflies on (reg[2i], reg[2i+1]); output slot = i + 4*q
            C2 o[8];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                o[i] = reg[2 * i] + reg[2 * i + 1];
                o[i + 4] = reg[2 * i] - reg[2 * i + 1];
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) reg[j] = o[j];
        }
    }

    // Two transforms of the same line geometry advanced stage by stage together (single buffer only): each
    // twiddle is fetched once for both, and the two independent butterfly sets give the scheduler twice the
    // instruction-level parallelism.  The exchanges still go through the one buffer, a after b.
    template <int s>
    __device__ static __forceinline__ void exchange(C2 (&reg)[8], float4* __restrict__ buf, const int tid, const int ln) {
        constexpr int Cs = 1 << (3 * s);
        constexpr int Ms = N / Cs;
        constexpr int Mn = Ms / 8;
#pragma unroll
        for (int q = 0; q < 8; ++q) c2_store(&buf[phys<s>(tid + T * q, ln)], reg[q]);
        barrier();
        constexpr int rn = (s + 1 == S - 1) ? RLAST : 8;
        if constexpr (rn == 8) {
            constexpr int Q = Mn / 8;  // threads per sub-problem in the next stage
            const int cn = tid / Q, tn = tid & (Q - 1);
#pragma unroll
            for (int j = 0; j < 8; ++j) reg[j] = c2_load(&buf[phys<s>((cn * 8 + j) * Q + tn, ln)]);
        } else {
#pragma unroll
            for (int i = 0; i < 8 / rn; ++i)
#pragma unroll
                for (int j = 0; j < rn; ++j) reg[i * rn + j] = c2_load(&buf[phys<s>((tid + T * i) * rn + j, ln)]);
        }
    }
    __device__ static __forceinline__ void last_small(C2 (&reg)[8]) {
        C2 o[8];
        if constexpr (RLAST == 4) {
            radix4_inv(reg[0], reg[1], reg[2], reg[3], o[0], o[2], o[4], o[6]);
            radix4_inv(reg[4], reg[5], reg[6], reg[7], o[1], o[3], o[5], o[7]);
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                o[i] = reg[2 * i] + reg[2 * i + 1];
                o[i + 4] = reg[2 * i] - reg[2 * i + 1];
            }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) reg[j] = o[j];
    }
    template <int s>
    __device__ static __forceinline__ void stage2(C2 (&ra)[8], C2 (&rb)[8], float4* __restrict__ smem, const int tid,
                                                  const int ln, const float2* __restrict__ tw, const int prior) {
        static_assert(!DOUBLE_BUF, "run2 exchanges through a single buffer");
        constexpr int Cs = 1 << (3 * s);
        constexpr int Ms = N / Cs;
        constexpr bool last = (s == S - 1);
        if constexpr (!last || RLAST == 8) {
            radix8_inv(ra);
            radix8_inv(rb);
            if constexpr (Ms > 8) {
#pragma unroll
                for (int q = 1; q < 8; ++q) {
                    const float2 w = twiddle(tw, (s * 7 + q - 1) * T + tid);
                    ra[q] = cmul(ra[q], w);
                    rb[q] = cmul(rb[q], w);
                }
            }
            if constexpr (!last) {
                if (prior + s > 0) barrier();  // previous exchange fully read
                exchange<s>(ra, smem, tid, ln);
                barrier();
                exchange<s>(rb, smem, tid, ln);
            }
        } else {
            last_small(ra);
            last_small(rb);
        }
    }
    __device__ static __forceinline__ void run2(C2 (&ra)[8], C2 (&rb)[8], float4* __restrict__ smem, const int tid,
                                                const int ln, const float2* __restrict__ tw, const int prior) {
        stage2<0>(ra, rb, smem, tid, ln, tw, prior);
        if constexpr (S > 1) stage2<1>(ra, rb, smem, tid, ln, tw, prior);
        if constexpr (S > 2) stage2<2>(ra, rb, smem, tid, ln, tw, prior);
        if constexpr (S > 3) stage2<3>(ra, rb, smem, tid, ln, tw, prior);
    }

    // reg[j] = point (tid + T*j) of line `ln` on entry and on return.
    // prior = number of exchanges this CTA has already performed (EXCHANGES per earlier transform).
    // Double buffer: exchange e uses buffer e & 1, one barrier each.  Single buffer: a leading
    // barrier protects the previous exchange's reads before the buffer is overwritten.
    __device__ static __forceinline__ void run(C2 (&reg)[8], float4* __restrict__ smem, const int tid, const int ln,
                                               const float2* __restrict__ tw, const int prior) {
        stage<0>(reg, smem, tid, ln, tw, prior);
        if constexpr (S > 1) stage<1>(reg, smem, tid, ln, tw, prior);
        if constexpr (S > 2) stage<2>(reg, smem, tid, ln, tw, prior);
        if constexpr (S > 3) stage<3>(reg, smem, tid, ln, tw, prior);
    }
};

}  // namespace ocean
