// ocean_hetero.cuh -- PROBE (OCEAN_HETERO=1, N = 512 only): one persistent CTA per SM that hosts BOTH roles at once --
// a column team (256 compute threads + a producer warp, exactly the tile pipeline of ocean_cols_tma_kernel) and G
// independent row groups of 64 threads (exactly the per-row program of ocean_rows_kernel), every role on its own named
// barriers.  The two-kernel frame alternates the roles on every SM because a full row pass and a full column pass do not
// fit the register file together; here the split is fixed -- one column team, G row groups -- and both
// run all the time.  Included by ocean_frame.cu after the two kernels it borrows from.
#pragma once

#ifndef OCEAN_HETERO_G
#define OCEAN_HETERO_G 6
#endif

struct HeteroCfg {
    static constexpr int N = 512, T = 64, W = OCEAN_COLS_TMA_W, G = OCEAN_HETERO_G;
    using Col = ColTmaCfg<N>;
    static constexpr int COL_THREADS = Col::THREADS;            // 288 = OCEAN_HETERO_BASE
    static constexpr int THREADS = COL_THREADS + G * T;
    using RowFFT = LineFFT<N, 1, false, false, false, false, -T>;  // one line per group, barrier = the group's own
    static constexpr int ROW_SMEM = N * 16;                      // staged h0 row = exchange buffer
    static constexpr int SMEM_BYTES = Col::SMEM_BYTES + 64 + G * ROW_SMEM;
    static_assert(COL_THREADS == OCEAN_HETERO_BASE, "LineFFT's group barriers assume the column role's thread count");
};

__device__ __forceinline__ void hetero_col_barrier() { asm volatile("bar.sync 15, %0;" ::"n"(HeteroCfg::COL_THREADS) : "memory"); }
__device__ __forceinline__ void hetero_group_barrier(int g) { asm volatile("bar.sync %0, 64;" ::"r"(2 + g) : "memory"); }

// ---- row role: rows `first`, `first + stride`, ... of the frame described by `a` (program of ocean_rows_kernel<512>) ----
__device__ __forceinline__ void hetero_rows(const FrameArgs& a, float4* hs, const int g, const int tid, size_t first, const size_t stride) {
    constexpr int N = HeteroCfg::N, T = HeteroCfg::T;
    using FFT = HeteroCfg::RowFFT;
    const size_t totalRows = (size_t)a.layers * N;
    const float time = a.time;
    for (size_t line = first; line < totalRows; line += stride) {
        const float4* __restrict__ H = a.h0 + line * N;
        const float2* __restrict__ Wd = a.waveHot + line * N;
        float2 wreg[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int x = tid + T * j;
            cp_async16(&hs[FFT::template phys<0>(x, 0)], H + x);
            wreg[j] = ld_once(Wd + x);
        }
        const int layer = (int)(line / N), y = (int)(line & (N - 1));
        const float dk = a.dk[layer % a.C];
        const float kz = (float)(y - N / 2) * dk;
        cp_async_wait_all();
        C2 pa[8], pb[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int x = tid + T * j;
            evolve_texel(hs[FFT::template phys<0>(x, 0)], wreg[j], (float)(x - N / 2) * dk, kz, time, pa[j], pb[j]);
        }
        FFT::run(pa, hs, tid, 0, a.twiddle, 0);
        if (tid == 0) spin_until_ge(a.colsDone + layer, a.colsTarget);  // write-after-read on the workspace
        hetero_group_barrier(g);
        {
            float4* __restrict__ out = a.rowsA + line * N;
#pragma unroll
            for (int j = 0; j < 8; ++j) ws_store(&out[tid + T * j], pa[j]);
        }
        FFT::run(pb, hs, tid, 0, a.twiddle, FFT::EXCHANGES);
        {
            float4* __restrict__ out = a.rowsB + line * N;
#pragma unroll
            for (int j = 0; j < 8; ++j) ws_store(&out[tid + T * j], pb[j]);
        }
        hetero_group_barrier(g);  // every store of the row issued; the staged row may be overwritten
        if (tid == 0) {
            __threadfence();
            atomicAdd(a.rowsDone + layer, 1ull);
        }
    }
}

// ---- column role: tiles blockIdx.x, + gridDim.x, ... of the frame described by `a` (program of ocean_cols_tma_kernel<512, false>) ----
template <bool AWAIT>
__device__ __forceinline__ void hetero_cols(const ColsTmaMaps& maps, const FrameArgs& a, float4* smem, const int totalTiles) {
    using Cfg = HeteroCfg::Col;
    using FFT = typename Cfg::FFT;
    constexpr int N = HeteroCfg::N, T = Cfg::T, W = Cfg::W, TILES_PER_LAYER = N / W;
    float4* bufA = smem;
    float4* bufB = smem + Cfg::TILE_ELEMS;
    float4* bufT = smem + 2 * Cfg::TILE_ELEMS;
    float2* twS = (float2*)(smem + 3 * Cfg::TILE_ELEMS);
    uint64_t* bars = (uint64_t*)((char*)twS + Cfg::TW_BYTES);  // [0] full A, [1] full B, [2] staged A, [3] staged B
    if (threadIdx.x == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_init(&bars[2], Cfg::COMPUTE);
        mbar_init(&bars[3], Cfg::COMPUTE);
        mbar_fence_init();
    }
    for (int i = threadIdx.x; i < FFT::TWIDDLE_ENTRIES; i += Cfg::THREADS) twS[i] = __ldg(&a.twiddle[i]);
    hetero_col_barrier();
    auto coords = [&](int tile, int& c0, int& c1) {
        const int layer = tile / TILES_PER_LAYER, xt = tile - layer * TILES_PER_LAYER;
        c0 = xt * W * 4;
        c1 = layer * N;
    };
    if (threadIdx.x >= Cfg::COMPUTE) {
        if (threadIdx.x != Cfg::COMPUTE) return;
        auto load = [&](const CUtensorMap* map, float4* buf, uint64_t* bar, int tile) {
            int c0, c1;
            coords(tile, c0, c1);
#pragma unroll
            for (int b = 0; b < Cfg::NBOX; ++b) tma_load_2d(buf + b * Cfg::BOX_ROWS * W, map, bar, c0, c1 + b * Cfg::BOX_ROWS);
        };
        const bool needFoam = a.zeroFoam == 0;
        auto loadB = [&](int tile) {
            mbar_arrive_expect_tx(&bars[1], Cfg::TILE_ELEMS * 16 * (needFoam ? 2 : 1));
            load(&maps.b, bufB, &bars[1], tile);
            if (needFoam) load(&maps.turb, bufT, &bars[1], tile);
        };
        auto store = [&](const CUtensorMap* map, const float4* buf, int tile) {
            int c0, c1;
            coords(tile, c0, c1);
#pragma unroll
            for (int b = 0; b < Cfg::NBOX; ++b) tma_store_2d_hint(map, buf + b * Cfg::BOX_ROWS * W, c0, c1 + b * Cfg::BOX_ROWS, L2_EVICT_FIRST);
        };
        int readyLayer = -1;
        auto await_rows = [&](int tile) {
            const int layer = tile / TILES_PER_LAYER;
            if (layer == readyLayer) return;
            if (AWAIT) spin_until_ge(a.rowsDone + layer, a.rowsTarget);
            fence_proxy_async_global();
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
        for (; tile < totalTiles; tile += gridDim.x, parity ^= 1) {
            const int next = tile + gridDim.x;
            mbar_wait(&bars[2], parity);
            store(&maps.disp, bufA, tile);
            bulk_commit();
            bulk_wait_read_all();
            if (next < totalTiles) {
                await_rows(next);
                mbar_arrive_expect_tx(&bars[0], Cfg::TILE_ELEMS * 16);
                load(&maps.a, bufA, &bars[0], next);
            }
            mbar_wait(&bars[3], parity);
            store(&maps.deriv, bufB, tile);
            store(&maps.turb, bufT, tile);
            bulk_commit();
            bulk_wait_read_all();
            if (next < totalTiles) loadB(next);
            bulk_wait_all();
            fence_proxy_async_global();
            __threadfence();
            atomicAdd(a.colsDone + tile / TILES_PER_LAYER, (unsigned long long)W);
        }
        return;
    }
    const int xw = threadIdx.x & (W - 1);
    const int tid = threadIdx.x / W;
    const float lambda = a.lambda, decay = a.foamDecay;
    const bool zeroFoam = a.zeroFoam != 0;
    uint32_t parity = 0;
    for (int tile = blockIdx.x; tile < totalTiles; tile += gridDim.x, parity ^= 1) {
        const int layer = tile / TILES_PER_LAYER, xt = tile - layer * TILES_PER_LAYER;
        const int x = xt * W + xw;
        const Epilogue ep(((x + tid) & 1) ? -1.0f : 1.0f, lambda);
        C2 r[8];
        float dxz[8];
        mbar_wait(&bars[0], parity);
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = c2_load(&bufA[(tid + T * j) * W + xw]);
        FFT::run(r, bufA, tid, xw, twS, 0);
        FFT::barrier();
        const bool heights = a.heightOut != nullptr && layer == 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float4 v = ep.displacement(r[j], dxz[j]);
            bufA[(tid + T * j) * W + xw] = v;
            if (heights) a.heightOut[(size_t)(tid + T * j) * N + x] = v.y;
        }
        fence_proxy_async();
        mbar_arrive(&bars[2]);
        mbar_wait(&bars[1], parity);
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = c2_load(&bufB[(tid + T * j) * W + xw]);
        FFT::run(r, bufB, tid, xw, twS, 0);
        FFT::barrier();
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int o = (tid + T * j) * W + xw;
            float jac;
            bufB[o] = ep.derivatives(r[j], dxz[j], jac);
            float f = zeroFoam ? 0.0f : bufT[o].x * decay;
            if (f < jac) f += jac;
            bufT[o] = make_float4(f, f, f, f);
        }
        fence_proxy_async();
        mbar_arrive(&bars[3]);
    }
}

// cols: the column pass to run (frame f); rows: the row pass to run beside it (frame f + 1 in a deferred schedule; the SAME
// frame in the in-order probe, where AWAIT = true makes the column team wait for its layer's rows as the two-kernel frame does)
template <bool AWAIT>
__global__ void __launch_bounds__(HeteroCfg::THREADS, 1)
ocean_hetero_kernel(const __grid_constant__ ColsTmaMaps maps, const FrameArgs cols, const FrameArgs rows, const int totalTiles) {
    extern __shared__ __align__(128) float4 smem[];
    pdl_trigger();
    if (threadIdx.x < HeteroCfg::COL_THREADS) {
        hetero_cols<AWAIT>(maps, cols, smem, totalTiles);
        return;
    }
    const int t = threadIdx.x - HeteroCfg::COL_THREADS, g = t / HeteroCfg::T, tid = t - g * HeteroCfg::T;
    float4* hs = (float4*)((char*)smem + HeteroCfg::Col::SMEM_BYTES + 64) + g * HeteroCfg::N;
    hetero_rows(rows, hs, g, tid, (size_t)blockIdx.x * HeteroCfg::G + g, (size_t)gridDim.x * HeteroCfg::G);
}
