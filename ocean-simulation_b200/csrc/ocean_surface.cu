// ocean_surface.cu -- consumers of the three output arrays, restated as CUDA kernels:
//
//   N2  mip chain of the derivative and turbulence arrays       WaterBody.cs:191-192 (GenerateMips)
//   N3  what Water.shader computes from the arrays per surface point:
//         displaced position      Domain   Water.shader:322-327
//         normal and foam term    Fragment Water.shader:337-347
//   N1  displacement-aware height queries for buoyancy (extends WaterBody.GetWaterHeight WB:195-209)
//
// Sampling follows the GPU sampler the reference configures (WB:104-128: wrap Repeat, trilinear):
// texel centres at (i + 0.5)/size, bilinear inside a level, linear between the two nearest levels.
// Unity's mip filter is not in the reference tree: the chain is DEFINED here as the 2x2 box filter
// (parity unpinned for that one step; everything downstream of it is pinned to the oracle).
#include "ocean_internal.h"

namespace ocean {

__device__ __forceinline__ float4 f4_add(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ float4 f4_lerp(float4 a, float4 b, float t) {
    return make_float4(a.x + (b.x - a.x) * t, a.y + (b.y - a.y) * t, a.z + (b.z - a.z) * t, a.w + (b.w - a.w) * t);
}

// dst[layer][y][x] = mean of the 2x2 block of src[layer]; src is n x n, dst is n/2 x n/2
__global__ void ocean_mip_kernel(const float4* __restrict__ src, size_t srcLayerStride, float4* __restrict__ dst,
                                 size_t dstLayerStride, int n) {
    const int h = n >> 1;
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= h) return;
    const float4* s = src + (size_t)blockIdx.z * srcLayerStride + (size_t)(2 * y) * n + 2 * x;
    const float4 a = s[0], b = s[1], c = s[n], d = s[n + 1];
    dst[(size_t)blockIdx.z * dstLayerStride + (size_t)y * h + x] =
        make_float4(((a.x + b.x) + (c.x + d.x)) * 0.25f, ((a.y + b.y) + (c.y + d.y)) * 0.25f,
                    ((a.z + b.z) + (c.z + d.z)) * 0.25f, ((a.w + b.w) + (c.w + d.w)) * 0.25f);
}

// One array as the sampler sees it: level 0 is the output array itself, levels >= 1 live in a
// separate buffer, per layer packed level after level (offset of level l = sum_{j=1}^{l-1} (N>>j)^2).
struct MipArray {
    const float4* base;  // [C][N][N]
    const float4* mips;  // [C][(N*N-1)/3] or nullptr (no chain: every LOD clamps to level 0)
};

__device__ __forceinline__ size_t mip_offset(int N, int level) {
    // (N^2/4) * (1 - 4^-(level-1)) / (3/4)  =  (N^2 - (N >> (level-1))^2) / 3
    const size_t n2 = (size_t)N * N, m = (size_t)(N >> (level - 1));
    return (n2 - m * m) / 3;
}

__device__ __forceinline__ float4 sample_bilinear(const float4* __restrict__ img, int n, float u, float v) {
    // Repeat wrap, texel centres at (i + 0.5)/n
    const float fx = u * (float)n - 0.5f, fy = v * (float)n - 0.5f;
    const float x0f = floorf(fx), y0f = floorf(fy);
    const float wx = fx - x0f, wy = fy - y0f;
    const int x0 = ((int)x0f) & (n - 1), y0 = ((int)y0f) & (n - 1);
    const int x1 = (x0 + 1) & (n - 1), y1 = (y0 + 1) & (n - 1);
    const float4 t00 = __ldg(&img[(size_t)y0 * n + x0]), t10 = __ldg(&img[(size_t)y0 * n + x1]);
    const float4 t01 = __ldg(&img[(size_t)y1 * n + x0]), t11 = __ldg(&img[(size_t)y1 * n + x1]);
    return f4_lerp(f4_lerp(t00, t10, wx), f4_lerp(t01, t11, wx), wy);
}

__device__ __forceinline__ const float4* level_ptr(const MipArray& a, int N, int layer, int level) {
    if (level == 0) return a.base + (size_t)layer * N * N;
    return a.mips + (size_t)layer * (((size_t)N * N - 1) / 3) + mip_offset(N, level);
}

// SAMPLE_TEXTURE2D_ARRAY_LOD(tex, sampler, uv, layer, lod) with a trilinear sampler
__device__ __forceinline__ float4 sample_lod(const MipArray& a, int N, int levels, int layer, float u, float v, float lod) {
    u -= floorf(u);
    v -= floorf(v);
    if (!a.mips || !(lod > 0.0f)) return sample_bilinear(level_ptr(a, N, layer, 0), N, u, v);
    lod = fminf(lod, (float)levels);
    const int l0 = (int)lod;
    const float t = lod - (float)l0;
    const float4 s0 = sample_bilinear(level_ptr(a, N, layer, l0), N >> l0, u, v);
    if (t == 0.0f || l0 >= levels) return s0;
    return f4_lerp(s0, sample_bilinear(level_ptr(a, N, layer, l0 + 1), N >> (l0 + 1), u, v), t);
}

struct SurfaceArgs {
    MipArray disp, deriv, turb;
    const float2* pts;
    float* outDisp;    // n x 3 or nullptr
    float* outNormal;  // n x 3 or nullptr
    float* outFoam;    // n or nullptr
    int N, C, n, levels;
    float lod;
    float wavelength[OCEAN_MAX_CASCADES];
};

__global__ void ocean_surface_kernel(const __grid_constant__ SurfaceArgs a) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    const float2 p = a.pts[i];
    float4 d = make_float4(0.f, 0.f, 0.f, 0.f), dv = d;
    float foam = 0.0f;
    for (int c = 0; c < a.C; ++c) {
        const float u = p.x / a.wavelength[c], v = p.y / a.wavelength[c];  // worldUV / _Wavelengths[i]
        if (a.outDisp) d = f4_add(d, sample_lod(a.disp, a.N, a.levels, c, u, v, a.lod));        // WS:324-326
        if (a.outNormal) dv = f4_add(dv, sample_lod(a.deriv, a.N, a.levels, c, u, v, a.lod));   // WS:342
        if (a.outFoam) {
            const float t = sample_lod(a.turb, a.N, a.levels, c, u, v, a.lod).x;
            foam += 1.0f - fminf(fmaxf(t, 0.0f), 1.0f);                                          // WS:343
        }
    }
    if (a.outDisp) {
        a.outDisp[3 * i + 0] = d.x;
        a.outDisp[3 * i + 1] = d.y;
        a.outDisp[3 * i + 2] = d.z;
    }
    if (a.outNormal) {
        // slope = (d.x / (1 + d.z), d.y / (1 + d.w)); normal = normalize(-slope.x, 1, -slope.y)   WS:346-347
        const float sx = dv.x / (1.0f + dv.z), sz = dv.y / (1.0f + dv.w);
        const float inv = 1.0f / sqrtf(sx * sx + 1.0f + sz * sz);
        a.outNormal[3 * i + 0] = -sx * inv;
        a.outNormal[3 * i + 1] = inv;
        a.outNormal[3 * i + 2] = -sz * inv;
    }
    if (a.outFoam) a.outFoam[i] = foam;
}

// Height of the water surface ABOVE a world position (x, z).  The surface point that ends up over
// (x, z) started at q with q + Dxz(q) = (x, z) (Water.shader:327 displaces horizontally too); solve by
// fixed-point iteration q <- (x, z) - Dxz(q), then return Dy(q).  All cascades, level 0, bilinear.
struct DisplacedQueryArgs {
    MipArray disp;
    const float2* pts;
    float* out;
    int N, C, n, iterations;
    float wavelength[OCEAN_MAX_CASCADES];
};

__global__ void ocean_displaced_query_kernel(const __grid_constant__ DisplacedQueryArgs a) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    const float2 p = a.pts[i];
    float qx = p.x, qz = p.y, height = 0.0f;
    for (int it = 0; it <= a.iterations; ++it) {
        float4 d = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int c = 0; c < a.C; ++c)
            d = f4_add(d, sample_lod(a.disp, a.N, 0, c, qx / a.wavelength[c], qz / a.wavelength[c], 0.0f));
        height = d.y;
        qx = p.x - d.x;
        qz = p.y - d.z;
    }
    a.out[i] = height;
}

cudaError_t launch_mip_chain(const float4* base, float4* mips, int N, int layers, cudaStream_t s) {
    const size_t perLayer = ((size_t)N * N - 1) / 3;
    const float4* src = base;
    size_t srcStride = (size_t)N * N;
    for (int level = 1, n = N; n > 1; ++level, n >>= 1) {
        float4* dst = mips + ((size_t)N * N - (size_t)n * n) / 3;  // = mip_offset(N, level)
        const int h = n >> 1;
        dim3 block(h < 128 ? h : 128), grid((h + block.x - 1) / block.x, h, layers);
        ocean_mip_kernel<<<grid, block, 0, s>>>(src, srcStride, dst, perLayer, n);
        count_launch();
        src = dst;
        srcStride = perLayer;
    }
    return cudaGetLastError();
}

static void fill_wavelengths(float* dst, const float* wavelengths, int C) {
    for (int i = 0; i < OCEAN_MAX_CASCADES; ++i) dst[i] = i < C ? wavelengths[i] : 1.0f;
}

cudaError_t launch_surface(const float4* disp, const float4* deriv, const float4* turb, const float4* derivMips, const float4* turbMips,
                           int N, int C, const float* wavelengths, const float2* pts, int n, float lod, float* outDisp,
                           float* outNormal, float* outFoam, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    SurfaceArgs a;
    a.disp = MipArray{disp, nullptr};  // displacementsTextures has no mip chain (WB:227): every LOD reads level 0
    a.deriv = MipArray{deriv, derivMips};
    a.turb = MipArray{turb, turbMips};
    a.pts = pts; a.outDisp = outDisp; a.outNormal = outNormal; a.outFoam = outFoam;
    a.N = N; a.C = C; a.n = n; a.lod = lod;
    a.levels = 0;
    while ((N >> a.levels) > 1) ++a.levels;
    fill_wavelengths(a.wavelength, wavelengths, C);
    ocean_surface_kernel<<<(n + 127) / 128, 128, 0, s>>>(a);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_displaced_query(const float4* disp, int N, int C, const float* wavelengths, const float2* pts, float* out, int n,
                                   int iterations, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    DisplacedQueryArgs a;
    a.disp = MipArray{disp, nullptr};
    a.pts = pts; a.out = out; a.N = N; a.C = C; a.n = n; a.iterations = iterations;
    fill_wavelengths(a.wavelength, wavelengths, C);
    ocean_displaced_query_kernel<<<(n + 127) / 128, 128, 0, s>>>(a);
    count_launch();
    return cudaGetLastError();
}

}  // namespace ocean
