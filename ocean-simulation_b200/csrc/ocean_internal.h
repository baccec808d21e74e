// ocean_internal.h -- declarations shared by the translation units of liboceanfft.so.
// (ocean_init.cu is built with -fmad=false so that the one-time spectrum follows the
// reference's operation order rounding for rounding; ocean_frame.cu keeps FMA contraction.)
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/oceanfft.h"

namespace ocean {

// Scalars of one ocean state, hoisted out of IS:33-97 on the host in fp32 with the reference's
// operation order (they do not depend on the texel).
struct StateConsts {
    float windSpeed, gravity, fetch, depth;
    float sqrtDepthOverG;  // sqrt(_Depth / _Gravity)                      IS:39
    float peakOmega;       // 22 * pow(|g*g/(U*F)|, 0.3333f)               IS:118
    float alpha;           // 0.076f * pow(|U*U/(F*g)|, 0.22f)             IS:48
    float mu;              // -2.33f - 1.45f*(U/(g/peakOmega) - 1.17f)     IS:63-64
    float windTheta;       // atan2(normalize(wind).y, normalize(wind).x)  IS:80-81
    float lambda;
    float pad0, pad1;
};

struct CascadeConsts {
    float dk;  // 2.0f * PI / wavelength   IS:110
    float cutoffLow, cutoffHigh, swell, fade;
    float wavelength;
};

// Constants of NormalizationFactor IS:69-74, evaluated once on the host in fp32.
struct SpreadConsts {
    float pi;     // 3.14159265f                       IS:8
    float c2;     // pow(log(2),2)/PI - PI/12.0f
    float c1;     // log(2)/PI
    float c0;     // 1/(2*PI)
    float twoSqrtPi;  // 2.0f*sqrt(PI)
};

struct InitArgs {
    int N, C;                          // texture size, cascades per state
    int nStates;                       // oceans in this launch (layers = nStates * C)
    SpreadConsts sc;
    CascadeConsts casc[OCEAN_MAX_CASCADES];
    const StateConsts* states;         // device, nStates entries
    const float2* noise;               // device, [state][y][x]
    float4* h0;                        // device, [state*C + cascade][y][x]
    float4* wave;                      // (kx, 1/|k|, kz, omega): the reference's waveData layout (inspection / parity); nullptr: not kept (ensembles)
    float2* waveHot;                   // (1/|k|, omega): what the frame kernels read; kx, kz are recomputed as n*dk
    float2* h0c;                       // h0(k) alone: the pair-row kernel reads it at k and at -k instead of the 16-byte packing
};

struct FrameArgs {
    int N;
    int layers;                        // nStates * C
    int C;                             // cascades per state: cascade of a layer = layer % C
    float dk[OCEAN_MAX_CASCADES];      // 2*PI/wavelength per cascade (IS:110)
    const float4* h0;
    const float2* waveHot;             // (1/|k|, omega) [layer][y][x]
    const float2* h0c;                 // h0(k) [layer][y][x] (compact copy of h0.xy)
    float4* rowsA;                     // (DxDz, DyDxz) after the x pass, [layer][y][x]
    float4* rowsB;                     // (DyxDyz, DxxDzz)
    float4* disp;
    float4* deriv;
    float4* turb;
    const float2* twiddle;             // e^{+2 pi i j / N}, j < N
    float time;
    float lambda;
    float foamDecay;                   // exp(-2.0f)  RF:29-30
    int zeroFoam;                      // 1: treat the incoming turbulence as zero (fresh texture)
    // Row -> column hand-off per layer: every row CTA adds the rows it finished to rowsDone[layer]
    // (release); the TMA column kernel starts a tile of `layer` once rowsDone[layer] >= rowsTarget
    // (acquire), so columns of the first layers overlap the row pass of the last ones.
    unsigned long long* rowsDone;      // device, one counter per layer
    unsigned long long rowsTarget;     // N * (row passes issued on these counters so far)
    // Column -> row hand-back per layer (write-after-read on the workspace, and the turbulence chain):
    // the column pass adds the columns it has finished AND stored to colsDone[layer]; a row CTA of the
    // next frame may overwrite its rows of the workspace once colsDone[layer] >= colsTarget.
    unsigned long long* colsDone;      // device, one counter per layer
    unsigned long long colsTarget;     // N * (column passes issued before this frame)
    // Readers of a frame's outputs that run beside the next frame (today: the mip chains WB:191-192, which
    // bump auxDone when their reads are over) do not hold its row pass back: the column pass waits for
    // auxDone >= auxTarget before its first store instead (nullptr: no such reader was ever launched on this
    // ocean).  The readback's snapshot / pack (WB:288) is ordered by the stream or by colsDone[0], not by this.
    const unsigned long long* auxDone;
    unsigned long long auxTarget;
    // Readback plane (WB:288, HEIGHT mode): when not nullptr the column pass also writes the .g channel of
    // displacement layer 0 here, N*N floats; it is complete once colsDone[0] reaches this frame's count.
    float* heightOut;
    // Ensembles: when not nullptr the column pass adds, per state (= layer / C), the 64-bit sums of the fp32
    // bit patterns of the texels it produces to checksums[state * 3 + {0: disp, 1: deriv, 2: turb}].
    unsigned long long* checksums;
    // Item queue of the fused frame kernel with dynamic scheduling (ocean_fused2.cuh): every item of the frame is handed
    // out by one atomicAdd on *itemNext; the frame's items are the values itemBase .. itemBase + totalItems - 1, and every
    // CTA makes exactly one fetch beyond them, so the counter advances by totalItems + gridDim.x per frame.  Two counters
    // alternate between consecutive frames (their kernels overlap one deep under PDL).
    unsigned long long* itemNext;
    unsigned long long itemBase;
    // Gravity for omega = sqrt(g |k|) (IS:33-35), recomputed by that kernel instead of loading (1/|k|, omega): per state
    // (= layer / C) when `states` is set (ensembles), else the scalar.
    const StateConsts* states;
    float gravity;
    // Layers [0, wsResident) keep their workspace planes L2-resident (evict_last stores); the others are streamed
    // (evict_first stores, read-once loads).  = layers when everything fits (the usual case).
    int wsResident;
    // L2 residency of the intermediate planes (host-side launch attribute, not read by the kernels):
    // accesses inside [l2Window, +l2WindowBytes) are marked persisting with this hit ratio.
    const void* l2Window;
    size_t l2WindowBytes;
    float l2HitRatio;
};

// TMA descriptors of the two intermediate planes and the three outputs, viewed as 2-D fp32 tensors
// {4N, layers*N}; the column kernel moves tiles of W adjacent columns (4W floats x up to 256 rows per box).
struct ColsTmaMaps {
    CUtensorMap a, b;               // loads: the two intermediate planes
    CUtensorMap disp, deriv, turb;  // stores: the three output arrays (same geometry)
    int valid = 0;
};
// Encodes the maps (driver entry point fetched through the runtime: no libcuda link).  Leaves
// valid = 0 for sizes the TMA column kernel does not cover; those use the generic column kernel.
cudaError_t make_cols_maps(int N, int layers, const float4* rowsA, const float4* rowsB, float4* disp, float4* deriv, float4* turb,
                           ColsTmaMaps* out);

// launchers (return cudaGetLastError()); every launch bumps the launch counter
// maxCtas > 0: walk the same work items with a persistent grid of at most maxCtas CTAs (ensembles: leaves room on every SM
// for the other lane's frame kernels); 0: one CTA per item
cudaError_t launch_noise(const uint64_t* seedsDev, int nStates, int N, float2* noise, cudaStream_t s, int maxCtas = 0);
cudaError_t launch_init(const InitArgs& a, cudaStream_t s, int maxCtas = 0);
// Initial spectrum fused into the row pass: noise -> h0 (shared memory only) -> evolution -> x-FFT
// -> workspace planes, for oceans that are evolved once from a fresh spectrum.  InitArgs::h0 / wave / waveHot are
// not written.  Call with the target device current (the kernel is configured once per device: PerDeviceOnce).
bool rows_from_noise_available(int N);
cudaError_t launch_rows_from_noise(const InitArgs& ia, const FrameArgs& a, cudaStream_t s);
// early: the row pass may start before the previous kernel in the stream has finished (PDL).  Only
// legal when that kernel does not produce h0 / waveHot / twiddles (i.e. it is a column pass).
cudaError_t launch_rows(const FrameArgs& a, cudaStream_t s, bool early);
cudaError_t launch_cols(const FrameArgs& a, const ColsTmaMaps* maps, cudaStream_t s);
// The whole frame as one persistent kernel, where available (N = 256, 512): 0 = no (two kernels), 1 = static round robin
// (ocean_fused.cuh), 2 = dynamic queue + prefetch (ocean_fused2.cuh).  OCEAN_FUSED=0/1/2 overrides the default.
int frame_fused_mode(int N, const ColsTmaMaps* maps);
int frame_fused_grid(int N, int layers, int mode);  // persistent grid size of that kernel (FrameArgs::itemBase arithmetic)
bool rows_need_h0c();  // the selected row pass reads FrameArgs::h0c / InitArgs::h0c (else they stay nullptr)
int hetero_mode(int N, const ColsTmaMaps* maps);  // probe (OCEAN_HETERO): both roles in one persistent CTA per SM
cudaError_t launch_hetero(const FrameArgs& a, const ColsTmaMaps* maps, cudaStream_t s, int mode);
cudaError_t launch_frame(const FrameArgs& a, const ColsTmaMaps* maps, cudaStream_t s, bool early, int mode);
cudaError_t launch_query(const float4* disp, int N, int C, const float* wavelengths /*host, C*/, const float2* pts,
                         float* out, int n, int mode, cudaStream_t s);
// ocean_surface.cu: mip chain (WB:191-192), surface shading inputs (Water.shader:322-347), displaced queries
inline size_t mip_texels_per_layer(int N) { return ((size_t)N * N - 1) / 3; }  // levels 1..log2(N), packed
inline size_t mip_level_offset(int N, int level) {                            // texel offset of `level` >= 1 inside a layer
    const size_t m = (size_t)(N >> (level - 1));
    return ((size_t)N * N - m * m) / 3;
}
// Builds both chains; the last kernel adds *signals to `done` in total (release) when its reads are over.
cudaError_t launch_mip_chains(const float4* baseA, float4* mipsA, const float4* baseB /*or nullptr*/, float4* mipsB, int N, int layers,
                              unsigned long long* done, unsigned long long* signals, cudaStream_t s);
cudaError_t launch_surface(const float4* disp, const float4* deriv, const float4* turb, const float4* derivMips, const float4* turbMips,
                           int N, int C, const float* wavelengths /*host, C*/, const float2* pts, int n, float lod, float* outDisp,
                           float* outNormal, float* outFoam, cudaStream_t s);
cudaError_t launch_displaced_query(const float4* disp, int N, int C, const float* wavelengths /*host, C*/, const float2* pts,
                                   float* out, int n, int iterations, cudaStream_t s);
// BuoyantObject.FixedUpdate BO:18-30: boxes -> query points (gather), heights -> (height, submerged, acceleration, volume)
cudaError_t launch_buoyancy_points(const OceanBuoyantBox* boxes, float2* pts, int n, cudaStream_t s);
cudaError_t launch_buoyancy(const OceanBuoyantBox* boxes, const float* heights, int n, float density, float gravityY, float4* out,
                            cudaStream_t s);
// BO:18-30 over WB:195-209 on a frame's height plane (N*N floats, the .g channel of displacement layer 0), one launch
cudaError_t launch_buoyancy_plane(const OceanBuoyantBox* boxes, const float* plane, int N, int n, float density, float gravityY,
                                  float4* out, cudaStream_t s);
cudaError_t launch_pack_height(const float4* dispLayer, float* out, int n, cudaStream_t s);
cudaError_t launch_checksum(const float4* arr, int nStates, size_t float4PerState, unsigned long long* out /*[state]*/,
                            int outStride, cudaStream_t s);
cudaError_t configure_frame_kernels(int N);  // opt in to >48 KiB dynamic shared memory where needed
const char* describe_frame_kernels(int N);   // "rows<N,R> grid.. smem..; cols<N,W> ..."

void count_launch();
uint64_t launch_count();

// "Configure this kernel once" flags.  cudaFuncSetAttribute (dynamic shared memory size, carve-out) applies to the CURRENT
// DEVICE only, so a process that drives several GPUs must configure every kernel once per device: a process-wide static flag
// makes the first launch on a second device fail with "invalid argument" (found by tests/test_zz_no_alloc.py on a 2-GPU box).
struct PerDeviceOnce {
    unsigned long long done = 0;  // bit d: configured on device d (the library's entry points serialise per handle; a benign
                                  // race between threads on different devices can at worst configure a kernel twice)
    static int current() { int d = 0; return cudaGetDevice(&d) == cudaSuccess ? d : -1; }
    bool need() const { const int d = current(); return d < 0 || d >= 64 || !(done & (1ull << d)); }
    void mark() { const int d = current(); if (d >= 0 && d < 64) done |= 1ull << d; }
};

}  // namespace ocean
