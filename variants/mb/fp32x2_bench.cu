// Microbenchmark: scalar FFMA/FADD vs packed FFMA2/FADD2 issue throughput on sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp32x2_bench fp32x2_bench.cu
#include <cuda_runtime.h>
#include <stdio.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b){ u64 r; asm("mov.b64 %0, {%1,%2};":"=l"(r):"f"(a),"f"(b)); return r;}
__device__ __forceinline__ void upk(u64 v, float&a, float&b){ asm("mov.b64 {%0,%1}, %2;":"=f"(a),"=f"(b):"l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c){ u64 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;":"=l"(r):"l"(a),"l"(b),"l"(c)); return r;}
__device__ __forceinline__ u64 add2(u64 a, u64 b){ u64 r; asm volatile("add.rn.f32x2 %0, %1, %2;":"=l"(r):"l"(a),"l"(b)); return r;}
__device__ __forceinline__ float fma1(float a, float b, float c){ float r; asm volatile("fma.rn.f32 %0, %1, %2, %3;":"=f"(r):"f"(a),"f"(b),"f"(c)); return r;}
__device__ __forceinline__ float add1(float a, float b){ float r; asm volatile("add.rn.f32 %0, %1, %2;":"=f"(r):"f"(a),"f"(b)); return r;}

template <int MODE> __global__ void __launch_bounds__(256) k(float4* p, int iters) {
    float4 v = p[threadIdx.x];
    if (MODE == 0) {  // scalar FFMA, 8 chains
        float a[8]; for (int i = 0; i < 8; ++i) a[i] = v.x + i;
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = fma1(a[i], v.y, v.z);
        float s = 0; for (int i = 0; i < 8; ++i) s += a[i];
        p[threadIdx.x + blockIdx.x * blockDim.x].x = s;
    } else if (MODE == 1) {  // FFMA2, 8 chains (16 fp32 lanes-worth)
        u64 a[8]; for (int i = 0; i < 8; ++i) a[i] = pk(v.x + i, v.w + i);
        const u64 b = pk(v.y, v.y), c = pk(v.z, v.z);
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = fma2(a[i], b, c);
        float s = 0; for (int i = 0; i < 8; ++i) { float x, y; upk(a[i], x, y); s += x + y; }
        p[threadIdx.x + blockIdx.x * blockDim.x].x = s;
    } else if (MODE == 2) {  // scalar FADD
        float a[8]; for (int i = 0; i < 8; ++i) a[i] = v.x + i;
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = add1(a[i], v.y);
        float s = 0; for (int i = 0; i < 8; ++i) s += a[i];
        p[threadIdx.x + blockIdx.x * blockDim.x].x = s;
    } else if (MODE == 3) {  // FADD2
        u64 a[8]; for (int i = 0; i < 8; ++i) a[i] = pk(v.x + i, v.w + i);
        const u64 b = pk(v.y, v.y);
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = add2(a[i], b);
        float s = 0; for (int i = 0; i < 8; ++i) { float x, y; upk(a[i], x, y); s += x + y; }
        p[threadIdx.x + blockIdx.x * blockDim.x].x = s;
    } else if (MODE == 4) {  // FFMA2 + integer ALU op interleaved 1:1 (does packing free issue slots?)
        u64 a[8]; for (int i = 0; i < 8; ++i) a[i] = pk(v.x + i, v.w + i);
        const u64 b = pk(v.y, v.y), c = pk(v.z, v.z);
        unsigned q[8]; for (int i = 0; i < 8; ++i) q[i] = threadIdx.x + i;
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int i = 0; i < 8; ++i) { a[i] = fma2(a[i], b, c); asm volatile("xor.b32 %0, %0, %1;" : "+r"(q[i]) : "r"(it)); }
        float s = 0; for (int i = 0; i < 8; ++i) { float x, y; upk(a[i], x, y); s += x + y + q[i]; }
        p[threadIdx.x + blockIdx.x * blockDim.x].x = s;
    } else if (MODE == 5) {  // scalar FFMA + integer op 1:1
        float a[8]; for (int i = 0; i < 8; ++i) a[i] = v.x + i;
        unsigned q[8]; for (int i = 0; i < 8; ++i) q[i] = threadIdx.x + i;
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int i = 0; i < 8; ++i) { a[i] = fma1(a[i], v.y, v.z); asm volatile("xor.b32 %0, %0, %1;" : "+r"(q[i]) : "r"(it)); }
        float s = 0; for (int i = 0; i < 8; ++i) s += a[i] + q[i];
        p[threadIdx.x + blockIdx.x * blockDim.x].x = s;
    }
}
template <int MODE> void run(const char* name, float4* d, int flopPerInst) {
    const int iters = 4096, blocks = 148 * 8;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<blocks, 256>>>(d, 16); cudaDeviceSynchronize();
    cudaEventRecord(e0); k<MODE><<<blocks, 256>>>(d, iters); cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double inst = (double)blocks * 256 * iters * 8;  // thread-level FP instructions
    printf("%-28s %8.3f ms  %7.2f T thread-inst/s  %7.2f TFLOP/s  (%.2f warp-inst/clk/SM @1.965GHz)\n", name, ms, inst / ms / 1e9,
           inst * flopPerInst / ms / 1e9, inst / 32 / (ms * 1e-3) / 148 / 1.965e9);
}
int main() {
    float4* d; cudaMalloc(&d, 148 * 8 * 256 * sizeof(float4)); cudaMemset(d, 0, 148 * 8 * 256 * sizeof(float4));
    run<0>("FFMA  (scalar)", d, 2); run<1>("FFMA2 (packed)", d, 4); run<2>("FADD  (scalar)", d, 1); run<3>("FADD2 (packed)", d, 2);
    run<4>("FFMA2 + LOP3 1:1 (fp inst)", d, 4); run<5>("FFMA + LOP3 1:1 (fp inst)", d, 2);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
