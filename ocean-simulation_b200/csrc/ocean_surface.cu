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
#include "ocean_tma.cuh"

namespace ocean {

__device__ __forceinline__ float4 f4_add(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ float4 f4_lerp(float4 a, float4 b, float t) {
    return make_float4(a.x + (b.x - a.x) * t, a.y + (b.y - a.y) * t, a.z + (b.z - a.z) * t, a.w + (b.w - a.w) * t);
}

// The chain in two launches for both arrays (instead of one launch per level per array):
//   head: each CTA owns a T0 x T0 block of level 0 (T0 = min(32, N)) of one layer of one array and
//         reduces it through log2(T0) levels, keeping the running level in shared memory;
//   tail: one CTA per layer per array finishes the remaining levels (N/T0 squared and smaller).
// Every output texel is ((a + b) + (c + d)) * 0.25 of the level above, whatever kernel produces it.
__device__ __forceinline__ float4 box4(float4 a, float4 b, float4 c, float4 d) {
    return make_float4(((a.x + b.x) + (c.x + d.x)) * 0.25f, ((a.y + b.y) + (c.y + d.y)) * 0.25f,
                       ((a.z + b.z) + (c.z + d.z)) * 0.25f, ((a.w + b.w) + (c.w + d.w)) * 0.25f);
}

__host__ __device__ __forceinline__ size_t mip_offset(int N, int level) {
    // texels of levels 1..level-1 = (N^2 - (N >> (level-1))^2) / 3
    const size_t n2 = (size_t)N * N, m = (size_t)(N >> (level - 1));
    return (n2 - m * m) / 3;
}

struct MipArgs {
    const float4* base[2];  // level 0 of the two arrays, [layers][N][N]
    float4* mips[2];        // their chains, [layers][(N*N-1)/3]
    int N, layers, levels;  // levels = log2(N)
    unsigned long long* done;  // += 1 per CTA of the LAST kernel of the chain, once it has stopped reading
};

// The next frame's row pass does not depend on the chain, so both kernels let their successor start at
// once (pdl_trigger); the column pass that would overwrite the arrays waits for `done` instead.
__device__ __forceinline__ void mip_signal(unsigned long long* done) {
    __syncthreads();
    if (done && threadIdx.x == 0) {
        __threadfence();
        atomicAdd(done, 1ull);
    }
}

template <int T0>
__global__ void __launch_bounds__((T0 / 2) * (T0 / 2)) ocean_mip_head_kernel(const __grid_constant__ MipArgs a, const bool last) {
    constexpr int H = T0 / 2;
    __shared__ float4 cur[H * H];
    __shared__ float4 nxt[(H / 2) * (H / 2) > 0 ? (H / 2) * (H / 2) : 1];
    const int N = a.N;
    pdl_trigger();
    pdl_wait();  // the column pass that produced level 0 has completed
    const int arr = blockIdx.z / a.layers, layer = blockIdx.z % a.layers;
    const float4* src = a.base[arr] + (size_t)layer * N * N;
    float4* chain = a.mips[arr] + (size_t)layer * (((size_t)N * N - 1) / 3);
    const int tx = threadIdx.x % H, ty = threadIdx.x / H;
    {   // level 1 from level 0
        const int x = blockIdx.x * H + tx, y = blockIdx.y * H + ty, h = N >> 1;
        const float4* s = src + (size_t)(2 * y) * N + 2 * x;
        const float4 v = box4(__ldg(s), __ldg(s + 1), __ldg(s + N), __ldg(s + N + 1));
        chain[(size_t)y * h + x] = v;
        cur[ty * H + tx] = v;
    }
    float4* from = cur;
    float4* to = nxt;
    int level = 2;
    for (int w = H / 2; w >= 1; w >>= 1, ++level) {  // w x w texels of `level` in this CTA
        __syncthreads();
        if ((int)threadIdx.x < w * w) {
            const int lx = threadIdx.x % w, ly = threadIdx.x / w, sw = 2 * w;
            const float4* s = from + (2 * ly) * sw + 2 * lx;
            const float4 v = box4(s[0], s[1], s[sw], s[sw + 1]);
            const int n = N >> level;
            chain[mip_offset(N, level) + (size_t)(blockIdx.y * w + ly) * n + blockIdx.x * w + lx] = v;
            to[ly * w + lx] = v;
        }
        float4* t = from; from = to; to = t;
    }
    if (last) mip_signal(a.done);
}

// levels first..a.levels of one layer of one array, each from the level above in global memory
__global__ void __launch_bounds__(1024) ocean_mip_tail_kernel(const __grid_constant__ MipArgs a, int first) {
    const int N = a.N;
    pdl_trigger();
    pdl_wait();  // launched as a programmatic dependent of the head kernel: wait for its levels
    const int arr = blockIdx.x / a.layers, layer = blockIdx.x % a.layers;
    float4* chain = a.mips[arr] + (size_t)layer * (((size_t)N * N - 1) / 3);
    for (int level = first; level <= a.levels; ++level) {
        const int n = N >> level, sn = 2 * n;
        const float4* s = chain + mip_offset(N, level - 1);
        float4* d = chain + mip_offset(N, level);
        for (int i = threadIdx.x; i < n * n; i += blockDim.x) {
            const int x = i % n, y = i / n;
            const float4* p = s + (size_t)(2 * y) * sn + 2 * x;
            d[i] = box4(__ldcg(p), __ldcg(p + 1), __ldcg(p + sn), __ldcg(p + sn + 1));
        }
        __syncthreads();  // the level just written is the next one's source (same CTA)
    }
    mip_signal(a.done);
}

// One array as the sampler sees it: level 0 is the output array itself, levels >= 1 live in a
// separate buffer, per layer packed level after level (offset of level l = sum_{j=1}^{l-1} (N>>j)^2).
struct MipArray {
    const float4* base;  // [C][N][N]
    const float4* mips;  // [C][(N*N-1)/3] or nullptr (no chain: every LOD clamps to level 0)
};

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

cudaError_t launch_mip_chains(const float4* baseA, float4* mipsA, const float4* baseB, float4* mipsB, int N, int layers,
                              unsigned long long* done, unsigned long long* signals, cudaStream_t s) {
    MipArgs a;
    a.base[0] = baseA; a.mips[0] = mipsA; a.base[1] = baseB; a.mips[1] = mipsB;
    a.N = N; a.layers = layers; a.levels = 0; a.done = done;
    while ((N >> a.levels) > 1) ++a.levels;
    const int arrays = baseB ? 2 : 1;
    const int headLevels = N >= 32 ? 5 : 4;
    const bool headIsLast = a.levels <= headLevels;
    // both kernels are programmatic dependents: they are scheduled while their predecessor drains and
    // wait for its completion themselves (griddepcontrol.wait)
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    dim3 grid(N >= 32 ? N / 32 : 1, N >= 32 ? N / 32 : 1, arrays * layers);
    {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = grid; cfg.blockDim = dim3(N >= 32 ? 256 : 64); cfg.stream = s;
        cfg.attrs = attr; cfg.numAttrs = 1;
        cudaError_t e = N >= 32 ? cudaLaunchKernelEx(&cfg, ocean_mip_head_kernel<32>, a, headIsLast)
                                : cudaLaunchKernelEx(&cfg, ocean_mip_head_kernel<16>, a, headIsLast);  // N = 16 (OCEAN_MIN_SIZE)
        count_launch();
        if (e != cudaSuccess) return e;
    }
    if (signals) *signals = headIsLast ? (unsigned long long)grid.x * grid.y * grid.z : (unsigned long long)(arrays * layers);
    if (!headIsLast) {
        const int side = N >> (headLevels + 1);
        int threads = side * side;
        threads = threads < 32 ? 32 : (threads > 1024 ? 1024 : threads);
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(arrays * layers); cfg.blockDim = dim3(threads); cfg.stream = s;
        cfg.attrs = attr; cfg.numAttrs = 1;
        cudaError_t e = cudaLaunchKernelEx(&cfg, ocean_mip_tail_kernel, a, headLevels + 1);
        count_launch();
        if (e != cudaSuccess) return e;
    }
    return cudaGetLastError();
}

// BuoyantObject.FixedUpdate BO:18-30 for a batch of boxes (the height comes from any of the query kernels)
__global__ void ocean_buoyancy_points_kernel(const OceanBuoyantBox* __restrict__ boxes, float2* __restrict__ pts, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) pts[i] = make_float2(boxes[i].px, boxes[i].pz);  // GetWaterHeight reads worldPosition.x / .z  WB:199-200
}
__global__ void ocean_buoyancy_kernel(const OceanBuoyantBox* __restrict__ boxes, const float* __restrict__ heights, int n,
                                      float density, float absGravityY, float4* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const OceanBuoyantBox b = boxes[i];
    const float waterHeight = heights[i];                                  // BO:18
    const float heightSubmerged = fmaxf(0.0f, waterHeight - b.py);        // BO:19
    float volume = 0.0f, accel = 0.0f;
    if (heightSubmerged > 0.0f) {                                          // BO:21
        const float fullVolume = b.sx * b.sy * b.sz;                       // BO:24
        volume = fminf(fmaxf(heightSubmerged / b.sy, 0.0f), 1.0f) * fullVolume;  // BO:25
        const float buoyantForce = density * volume;                       // BO:26
        accel = absGravityY * buoyantForce;                                // BO:29
    }
    out[i] = make_float4(waterHeight, heightSubmerged, accel, volume);
}
// The whole of BuoyantObject.FixedUpdate BO:18-30 over WaterBody.GetWaterHeight WB:195-209 in one launch, reading the
// HEIGHT PLANE of a frame (the .g channel of displacement layer 0 that the column pass writes for the readback,
// FrameArgs::heightOut) instead of the RGBA layer: what ocean_update_buoyancy runs on the owning GPU so that only the
// boxes and their answers cross PCIe.  Same fp32 operations as ocean_query_kernel (NEAREST_REF) + ocean_buoyancy_kernel.
__global__ void ocean_buoyancy_plane_kernel(const OceanBuoyantBox* __restrict__ boxes, const float* __restrict__ plane, int N, int n,
                                            float density, float absGravityY, float4* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const OceanBuoyantBox b = boxes[i];
    // Mathf.InverseLerp(-N/2, N/2, v) = Clamp01((v - a) / (b - a))   WB:199-200
    const float lo = (float)(-(N / 2)), hi = (float)(N / 2);
    float u = __fdiv_rn(__fsub_rn(b.px, lo), __fsub_rn(hi, lo));
    float v = __fdiv_rn(__fsub_rn(b.pz, lo), __fsub_rn(hi, lo));
    u = fminf(fmaxf(u, 0.0f), 1.0f);
    v = fminf(fmaxf(v, 0.0f), 1.0f);
    int x = (int)__fmul_rn(u, (float)N), y = (int)__fmul_rn(v, (float)N);  // WB:203-204
    x = min(max(x, 0), N - 1);
    y = min(max(y, 0), N - 1);
    const float waterHeight = __ldcg(&plane[(size_t)y * N + x]);            // .g of layer 0  WB:206-208, BO:18
    const float heightSubmerged = fmaxf(0.0f, waterHeight - b.py);         // BO:19
    float volume = 0.0f, accel = 0.0f;
    if (heightSubmerged > 0.0f) {                                           // BO:21
        const float fullVolume = b.sx * b.sy * b.sz;                        // BO:24
        volume = fminf(fmaxf(heightSubmerged / b.sy, 0.0f), 1.0f) * fullVolume;  // BO:25
        const float buoyantForce = density * volume;                        // BO:26
        accel = absGravityY * buoyantForce;                                 // BO:29
    }
    out[i] = make_float4(waterHeight, heightSubmerged, accel, volume);
}
cudaError_t launch_buoyancy_plane(const OceanBuoyantBox* boxes, const float* plane, int N, int n, float density, float gravityY,
                                  float4* out, cudaStream_t s) {
    ocean_buoyancy_plane_kernel<<<(n + 127) / 128, 128, 0, s>>>(boxes, plane, N, n, density, fabsf(gravityY), out);
    count_launch();
    return cudaGetLastError();
}
cudaError_t launch_buoyancy_points(const OceanBuoyantBox* boxes, float2* pts, int n, cudaStream_t s) {
    ocean_buoyancy_points_kernel<<<(n + 255) / 256, 256, 0, s>>>(boxes, pts, n);
    count_launch();
    return cudaGetLastError();
}
cudaError_t launch_buoyancy(const OceanBuoyantBox* boxes, const float* heights, int n, float density, float gravityY, float4* out,
                            cudaStream_t s) {
    ocean_buoyancy_kernel<<<(n + 255) / 256, 256, 0, s>>>(boxes, heights, n, density, fabsf(gravityY), out);
    count_launch();
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
