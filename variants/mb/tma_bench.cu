// tma_bench.cu -- how fast does one SM move column tiles?  (variants/mb, measurement only, not part of the product)
//
// The column pass moves tiles of W adjacent texel columns x 512 rows with 2-D TMA tensor copies whose box is only
// W*16 bytes wide (64 B at W = 4).  This microbenchmark times, chip-wide (148 x CTAS persistent CTAs, one issuing
// thread each, DEPTH tiles in flight per CTA):
//   ld2d<W> : TMA 2-D loads   box {4W floats, 256 rows} x 2 per 512-row tile       (what ocean_cols_tma_kernel does)
//   st2d<W> : TMA 2-D stores  of the same boxes from shared memory
//   ld1d    : 1-D bulk loads of the same number of bytes per tile, contiguous      (a tile-major workspace layout)
//   st1d    : 1-D bulk stores
//   ldg<W>  : the same tiles fetched by 256 threads with 16-byte LDG (W*16-byte segments)  and stg<W> likewise
// over a tensor of `layers` x 512 x 512 float4 (layers = 4: 33.5 MB, L2-resident after the first pass; 32: HBM).
//
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tma_bench tma_bench.cu -lcuda && ./tma_bench
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "../../ocean-simulation_b200/csrc/ocean_tma.cuh"
using namespace ocean;

#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

constexpr int N = 512, DEPTH = 3;

template <int W, int MODE>  // MODE 0: ld2d, 1: st2d, 2: ld1d, 3: st1d
__global__ void __launch_bounds__(128) tile_kernel(const __grid_constant__ CUtensorMap map, float4* base, int totalTiles, int iters) {
    extern __shared__ __align__(128) float4 smem[];
    constexpr int TILE = N * W;  // float4 per tile
    uint64_t* bars = (uint64_t*)(smem + DEPTH * TILE);
    if (threadIdx.x == 0) {
        for (int i = 0; i < DEPTH; ++i) mbar_init(&bars[i], 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    constexpr int TPL = N / W;
    uint32_t phase[DEPTH] = {0};  // mbarrier phases run on across the passes
    for (int it = 0; it < iters; ++it) {
        int issued = 0, done = 0;
        for (int tile = blockIdx.x; ; tile += gridDim.x) {
            const bool more = tile < totalTiles;
            if (more) {
                const int slot = issued % DEPTH;
                if (issued >= DEPTH) {  // free the slot: wait for its previous use
                    if (MODE == 0 || MODE == 2) { mbar_wait(&bars[slot], phase[slot]); phase[slot] ^= 1; ++done; }
                    else asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory");  // DEPTH - 1 stores may still read
                }
                const int layer = tile / TPL, xt = tile - layer * TPL;
                float4* buf = smem + slot * TILE;
                if (MODE == 0) {
                    mbar_arrive_expect_tx(&bars[slot], TILE * 16);
                    tma_load_2d(buf, &map, &bars[slot], xt * W * 4, layer * N);
                    tma_load_2d(buf + 256 * W, &map, &bars[slot], xt * W * 4, layer * N + 256);
                } else if (MODE == 1) {
                    tma_store_2d(&map, buf, xt * W * 4, layer * N);
                    tma_store_2d(&map, buf + 256 * W, xt * W * 4, layer * N + 256);
                    bulk_commit();
                } else if (MODE == 2) {
                    mbar_arrive_expect_tx(&bars[slot], TILE * 16);
                    bulk_load_1d(buf, base + (size_t)tile * TILE, TILE * 16, &bars[slot]);
                } else {
                    bulk_store_1d(base + (size_t)tile * TILE, buf, TILE * 16);
                    bulk_commit();
                }
                ++issued;
            } else {
                if (MODE == 0 || MODE == 2) {
                    while (done < issued) { const int slot = done % DEPTH; mbar_wait(&bars[slot], phase[slot]); phase[slot] ^= 1; ++done; }
                } else bulk_wait_all();
                break;
            }
        }
    }
}

// the same tiles by plain 16-byte accesses: 256 threads, thread = (row % 64.., column in tile)
template <int W, int STORE>
__global__ void __launch_bounds__(64 * W) lsu_kernel(float4* base, int totalTiles, int iters) {
    constexpr int TPL = N / W;
    const int xw = threadIdx.x % W, r0 = threadIdx.x / W;
    float4 acc = make_float4(0, 0, 0, 0);
    for (int it = 0; it < iters; ++it)
        for (int tile = blockIdx.x; tile < totalTiles; tile += gridDim.x) {
            const int layer = tile / TPL, xt = tile - layer * TPL;
            float4* p = base + ((size_t)layer * N) * N + xt * W + xw;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float4* q = p + (size_t)(r0 + 64 * j) * N;
                if (STORE) __stcs(q, make_float4(it, tile, j, xw));
                else { const float4 v = __ldcg(q); acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w; }
            }
        }
    if (!STORE && acc.x == 123.456f) base[0] = acc;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode() {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
    return (EncodeTiledFn)fn;
}

template <class F> static float time_ms(F f) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    f();  // warm (fills L2)
    CHECK(cudaDeviceSynchronize());
    cudaEventRecord(a);
    f();
    cudaEventRecord(b);
    CHECK(cudaDeviceSynchronize());
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    return ms;
}

template <int W, int MODE> static void run_tile(const char* name, float4* base, int layers, int ctasPerSM, int iters) {
    CUtensorMap map;
    const cuuint64_t dims[2] = {(cuuint64_t)N * 4, (cuuint64_t)layers * N};
    const cuuint64_t strides[1] = {(cuuint64_t)N * 16};
    const cuuint32_t box[2] = {(cuuint32_t)W * 4, 256};
    const cuuint32_t estr[2] = {1, 1};
    if (get_encode()(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
        printf("encode failed\n");
        exit(1);
    }
    const int smemBytes = DEPTH * N * W * 16 + 64;
    CHECK(cudaFuncSetAttribute(tile_kernel<W, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smemBytes));
    const int totalTiles = layers * (N / W);
    const int grid = 148 * ctasPerSM;
    const float ms = time_ms([&] { tile_kernel<W, MODE><<<grid, 128, smemBytes>>>(map, base, totalTiles, iters); });
    CHECK(cudaGetLastError());
    const double bytes = (double)layers * N * N * 16 * iters;
    printf("%-8s W=%2d layers=%2d ctas/SM=%d depth=%d : %8.1f GB/s  %6.1f B/clk/SM (1.965 GHz)  %.2f us per 33.5 MB pass\n", name, W, layers,
           ctasPerSM, DEPTH, bytes / ms / 1e6, bytes / ms / 1e6 / 148 / 1.965, ms * 1e3 / iters / (layers / 4.0));
}
template <int W, int STORE> static void run_lsu(const char* name, float4* base, int layers, int ctasPerSM, int iters) {
    const int totalTiles = layers * (N / W);
    const float ms = time_ms([&] { lsu_kernel<W, STORE><<<148 * ctasPerSM, 64 * W>>>(base, totalTiles, iters); });
    CHECK(cudaGetLastError());
    const double bytes = (double)layers * N * N * 16 * iters;
    printf("%-8s W=%2d layers=%2d ctas/SM=%d         : %8.1f GB/s  %6.1f B/clk/SM (1.965 GHz)  %.2f us per 33.5 MB pass\n", name, W, layers,
           ctasPerSM, bytes / ms / 1e6, bytes / ms / 1e6 / 148 / 1.965, ms * 1e3 / iters / (layers / 4.0));
}

int main() {
    float4* base;
    const int maxLayers = 32;
    CHECK(cudaMalloc(&base, (size_t)maxLayers * N * N * 16));
    CHECK(cudaMemset(base, 0, (size_t)maxLayers * N * N * 16));
    for (int layers : {4, 32}) {
        const int iters = layers == 4 ? 20 : 3;
        printf("---- tensor of %d x 512 x 512 float4 = %.1f MB (%s) ----\n", layers, layers * 4.194304, layers == 4 ? "L2-resident" : "HBM");
        for (int c : {1, 2}) {
            run_tile<4, 0>("ld2d", base, layers, c, iters);
            run_tile<8, 0>("ld2d", base, layers, c, iters);
            run_tile<4, 2>("ld1d", base, layers, c, iters);
            run_tile<4, 1>("st2d", base, layers, c, iters);
            run_tile<8, 1>("st2d", base, layers, c, iters);
            run_tile<4, 3>("st1d", base, layers, c, iters);
        }
        for (int c : {2, 4, 8}) {
            run_lsu<4, 0>("ldg", base, layers, c, iters);
            run_lsu<4, 1>("stg", base, layers, c, iters);
        }
        run_lsu<8, 0>("ldg", base, layers, 4, iters);
        run_lsu<8, 1>("stg", base, layers, 4, iters);
    }
    return 0;
}
