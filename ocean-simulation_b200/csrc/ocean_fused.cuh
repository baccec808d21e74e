// ocean_fused.cuh -- the whole frame as ONE persistent, warp-specialised kernel (N = 256, 512).
// Included by ocean_frame.cu after Epilogue / ColsTmaMaps / the PDL and TRACE helpers.
//
// Why: with one launch per pass every CTA of a pass runs the same phase at the same time (load, then
// transform, then store) and every pass ends in a tail, so on B200 the two-kernel frame is bound by
// latency, not by HBM (its time does not change when everything is L2-resident).  Here the frame is a
// queue of ITEMS that persistent CTAs (two per SM) work through:
//     row item    = W consecutive rows of one layer: time evolution TD:20-47 + x-FFT FFT:48-57
//     column item = W adjacent columns of one layer: y-FFT FFT:60-69 + Permute FFT:73-78 + RF:16-34
// handed out round-robin in an order (frame_item) that keeps a layer's rows well ahead of its columns,
// so a column item rarely waits, and CTAs in different roles share an SM, so loads, butterflies and
// stores of different items overlap.  Dependencies are the per-layer release/acquire counters of FrameArgs
// (rowsDone / colsDone); they also chain consecutive frames, which overlap through PDL.
//
// One CTA = W*N/8 compute threads + one producer warp.  The compute warps never touch global memory:
// lane 0 of the producer warp issues every copy -- 1-D bulk copies for row items (W rows of h0 /
// waveHot / a workspace plane are contiguous), TMA tensor copies for column items -- and owns every
// wait on a counter, so a dependency that is not ready yet stalls one thread, not the butterflies.
//
// Shared memory (3 buffers of W*N 128-bit elements + twiddles + 4 mbarriers):
//   row item:    bufA <- h0 rows        -> staged plane A rows   bufB <- waveHot rows -> staged plane B rows
//                bufT  = FFT exchange buffer
//   column item: bufA <- plane A tile   -> staged displacement   bufB <- plane B tile -> staged derivatives
//                bufT <- turbulence tile -> staged turbulence    (the tiles are their own exchange buffers)
#pragma once

template <int N> struct FusedCfg {
    static constexpr int T = N / 8;
    static constexpr int W = OCEAN_COLS_TMA_W;                  // lines (rows or columns) per item
    static constexpr int COMPUTE = W * T;
    static constexpr int THREADS = COMPUTE + 32;
    static constexpr int BOX_ROWS = N < 256 ? N : 256;
    static constexpr int NBOX = N / BOX_ROWS;
    static constexpr int TILE_ELEMS = N * W;
    static constexpr int IPL = N / W;                           // items per layer and role
    using ColFFT = LineFFT<N, W, true, false, true, true, COMPUTE>;
    using RowFFT = LineFFT<N, W, false, false, false, true, COMPUTE>;
    static constexpr int TW_BYTES = ColFFT::TWIDDLE_ENTRIES * 8;
    static constexpr int SMEM_BYTES = 3 * TILE_ELEMS * 16 + TW_BYTES + 64;
    static_assert(OCEAN_COLS_TMA_W == W, "the TMA descriptors are encoded for tiles of W columns");
};

struct FrameItem {
    int role;   // 0 = rows, 1 = columns, -1 = none
    int layer;
    int idx;    // item within the layer: rows idx*W.., columns idx*W..
};

// queue position -> item.  OCEAN_FUSED_ORDER 0: every row group, then every column group
// (R0 .. R(L-1) C0 .. C(L-1)); 1: rows of layer L+1 between the rows and the columns of layer L
// (R0 R1 C0 R2 C1 ... C(L-1)).  With a static round-robin over ~300 CTAs order 0 keeps a full round of
// row items between a layer's rows and its columns, so a column item rarely waits.
#ifndef OCEAN_FUSED_ORDER
#define OCEAN_FUSED_ORDER 0
#endif
__device__ __forceinline__ FrameItem frame_item(int pos, int ipl, int layers, int order = OCEAN_FUSED_ORDER) {
    FrameItem it;
    const int g = pos / ipl;
    it.idx = pos - g * ipl;
    if (order == 0) {
        it.role = g >= layers;
        it.layer = g >= layers ? g - layers : g;
    } else if (layers == 1) {
        it.role = g; it.layer = 0;
    } else {
        if (g == 0) { it.role = 0; it.layer = 0; }
        else if (g == 2 * layers - 1) { it.role = 1; it.layer = layers - 1; }
        else if (g & 1) { it.role = 0; it.layer = (g + 1) >> 1; }
        else { it.role = 1; it.layer = (g >> 1) - 1; }
    }
    return it;
}
// N = 1024: the row -> column workspace of all layers (134 MB at 4 x 1024^2) does not fit the 126 MB L2, so the items run in
// the layer-interleaved order R0 R1 C0 R2 C1 ... : at most two layers' planes (67 MB) are live between their rows and
// their columns, and a layer's columns read its rows while they are still L2-resident.
template <int N> struct FusedOrder { static constexpr int value = N >= 1024 ? 1 : OCEAN_FUSED_ORDER; };

template <int N>
__device__ __forceinline__ void fused_producer(const ColsTmaMaps& maps, const FrameArgs& a, float4* bufA, float4* bufB, float4* bufT,
                                               uint64_t* bars, const int totalItems) {
    using Cfg = FusedCfg<N>;
    constexpr int W = Cfg::W, TILE_BYTES = Cfg::TILE_ELEMS * 16;
    const bool needFoam = a.zeroFoam == 0;
    auto tile_load = [&](const CUtensorMap* map, float4* buf, uint64_t* bar, const FrameItem& it) {
#pragma unroll
        for (int b = 0; b < Cfg::NBOX; ++b)
            tma_load_2d(buf + b * Cfg::BOX_ROWS * W, map, bar, it.idx * W * 4, it.layer * N + b * Cfg::BOX_ROWS);
    };
    auto tile_store = [&](const CUtensorMap* map, const float4* buf, const FrameItem& it) {
#pragma unroll
        for (int b = 0; b < Cfg::NBOX; ++b)
            if (OCEAN_L2_HINTS & 2) tma_store_2d_hint(map, buf + b * Cfg::BOX_ROWS * W, it.idx * W * 4, it.layer * N + b * Cfg::BOX_ROWS, L2_EVICT_FIRST);
            else tma_store_2d(map, buf + b * Cfg::BOX_ROWS * W, it.idx * W * 4, it.layer * N + b * Cfg::BOX_ROWS);
    };
    auto row_offset = [&](const FrameItem& it) { return ((size_t)it.layer * N + (size_t)it.idx * W) * N; };
    // a column item needs its layer's rows of THIS frame; a row item's inputs are immutable
    auto ready = [&](const FrameItem& it) { return it.role == 0 || ld_acquire_gpu(a.rowsDone + it.layer) >= a.rowsTarget; };
    auto load_a = [&](const FrameItem& it) {
        mbar_arrive_expect_tx(&bars[0], TILE_BYTES);
        if (it.role == 0) bulk_load_1d(bufA, a.h0 + row_offset(it), TILE_BYTES, &bars[0]);
        else tile_load(&maps.a, bufA, &bars[0], it);
    };
    auto load_b = [&](const FrameItem& it) {
        if (it.role == 0) {
            mbar_arrive_expect_tx(&bars[1], TILE_BYTES / 2);
            bulk_load_1d(bufB, a.waveHot + row_offset(it), TILE_BYTES / 2, &bars[1]);
        } else {  // pair B travels with the tile's incoming turbulence texels (RF:28)
            mbar_arrive_expect_tx(&bars[1], TILE_BYTES * (needFoam ? 2 : 1));
            tile_load(&maps.b, bufB, &bars[1], it);
            if (needFoam) tile_load(&maps.turb, bufT, &bars[1], it);
        }
    };
    auto begin = [&](const FrameItem& it) {  // dependencies satisfied: start both loads
        if (it.role == 1) fence_proxy_async_global();  // acquired generic/async writes -> TMA reads
        load_a(it);
        load_b(it);
    };

    int pos = blockIdx.x;
    FrameItem cur = pos < totalItems ? frame_item(pos, Cfg::IPL, a.layers, FusedOrder<N>::value) : FrameItem{-1, 0, 0};
    if (cur.role >= 0) {
        if (cur.role == 1) spin_until_ge(a.rowsDone + cur.layer, a.rowsTarget);
        begin(cur);
    }
    uint32_t parity = 0;
    while (cur.role >= 0) {
        const int npos = pos + gridDim.x;
        const FrameItem nxt = npos < totalItems ? frame_item(npos, Cfg::IPL, a.layers, FusedOrder<N>::value) : FrameItem{-1, 0, 0};
        // The next item's loads start as soon as a buffer is free AND the item is ready.  Readiness is
        // only polled here, never waited for, before this item has been published: blocking on the next
        // item's dependency could wait for a signal this very CTA still has to send.
        bool nready = false, loadedA = false, loadedB = false;
        auto poll = [&]() {
            if (!nready && nxt.role >= 0 && ready(nxt)) {
                nready = true;
                if (nxt.role == 1) fence_proxy_async_global();  // acquired writes -> TMA reads
            }
            return nready;
        };

        mbar_wait(&bars[2], parity);  // part A staged in bufA
        if (cur.role == 0) {
            // write-after-read on the workspace: the previous frame's columns of this layer are done
            spin_until_ge(a.colsDone + cur.layer, a.colsTarget);
            if (OCEAN_L2_HINTS & 4) bulk_store_1d_hint(a.rowsA + row_offset(cur), bufA, TILE_BYTES, L2_EVICT_LAST);
            else bulk_store_1d(a.rowsA + row_offset(cur), bufA, TILE_BYTES);
        } else {
            tile_store(&maps.disp, bufA, cur);
        }
        bulk_commit();
        bulk_wait_read_all();
        if (poll()) { load_a(nxt); loadedA = true; }

        mbar_wait(&bars[3], parity);  // part B staged in bufB (and bufT)
        if (cur.role == 0) {
            if (OCEAN_L2_HINTS & 4) bulk_store_1d_hint(a.rowsB + row_offset(cur), bufB, TILE_BYTES, L2_EVICT_LAST);
            else bulk_store_1d(a.rowsB + row_offset(cur), bufB, TILE_BYTES);
        } else {
            tile_store(&maps.deriv, bufB, cur);
            tile_store(&maps.turb, bufT, cur);
        }
        bulk_commit();
        bulk_wait_read_all();
        if (poll()) {
            if (!loadedA) { load_a(nxt); loadedA = true; }
            load_b(nxt);
            loadedB = true;
        }

        // everything this item wrote is in memory: publish it
        bulk_wait_all();
        fence_proxy_async_global();
        __threadfence();
        atomicAdd((cur.role == 0 ? a.rowsDone : a.colsDone) + cur.layer, (unsigned long long)W);

        if (nxt.role >= 0 && !loadedB) {
            if (!nready) {
                if (nxt.role == 1) spin_until_ge(a.rowsDone + nxt.layer, a.rowsTarget);
                if (nxt.role == 1) fence_proxy_async_global();
            }
            if (!loadedA) load_a(nxt);
            load_b(nxt);
        }
        cur = nxt;
        pos = npos;
        parity ^= 1;
    }
}

// (evolve_texel: ocean_frame.cu)

template <int N>
__global__ void __launch_bounds__(FusedCfg<N>::THREADS, N >= 1024 ? 1 : 2)
ocean_frame_kernel(const __grid_constant__ ColsTmaMaps maps, const FrameArgs a) {
    using Cfg = FusedCfg<N>;
    using ColFFT = typename Cfg::ColFFT;
    using RowFFT = typename Cfg::RowFFT;
    constexpr int T = Cfg::T, W = Cfg::W;
    extern __shared__ __align__(128) float4 smem[];
    float4* bufA = smem;
    float4* bufB = smem + Cfg::TILE_ELEMS;
    float4* bufT = smem + 2 * Cfg::TILE_ELEMS;
    float2* twS = (float2*)(smem + 3 * Cfg::TILE_ELEMS);
    uint64_t* bars = (uint64_t*)((char*)twS + Cfg::TW_BYTES);  // [0] full A, [1] full B, [2] staged A, [3] staged B

    pdl_trigger();
    if (threadIdx.x == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_init(&bars[2], Cfg::COMPUTE);
        mbar_init(&bars[3], Cfg::COMPUTE);
        mbar_fence_init();
    }
    for (int i = threadIdx.x; i < ColFFT::TWIDDLE_ENTRIES; i += Cfg::THREADS) twS[i] = __ldg(&a.twiddle[i]);  // immutable
    __syncthreads();
    const int totalItems = 2 * a.layers * Cfg::IPL;

    if (threadIdx.x >= Cfg::COMPUTE) {
        if (threadIdx.x == Cfg::COMPUTE) fused_producer<N>(maps, a, bufA, bufB, bufT, bars, totalItems);
        return;
    }

    const float lambda = a.lambda, decay = a.foamDecay, time = a.time;
    const bool zeroFoam = a.zeroFoam != 0;
    uint32_t parity = 0;
    for (int pos = blockIdx.x; pos < totalItems; pos += gridDim.x, parity ^= 1) {
        const FrameItem it = frame_item(pos, Cfg::IPL, a.layers, FusedOrder<N>::value);
        C2 r[8];
        TRACE(0, pos, 0);
#ifdef OCEAN_TRACE
        if (threadIdx.x == 0 && pos < OCEAN_TRACE_CTAS) g_trace_smid[1][pos] = (unsigned)(it.role * 16 + it.layer);
#endif
        if (it.role == 0) {
            // ================= row item: lines are rows, thread (ln, tid) owns x = tid + T*j of row y0 + ln
            const int ln = threadIdx.x / T, tid = threadIdx.x - ln * T;
            const float dk = a.dk[it.layer % a.C];
            const float kz = (float)(it.idx * W + ln - N / 2) * dk;  // IS:111 (waveData.z)
            const float4* hs = bufA + ln * N;
            const float2* ws = (const float2*)bufB + ln * N;
            C2 pb[8];
            mbar_wait(&bars[0], parity);
            mbar_wait(&bars[1], parity);
            TRACE(0, pos, 1);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int x = tid + T * j;
                evolve_texel(hs[x], ws[x], (float)(x - N / 2) * dk, kz, time, r[j], pb[j]);
            }
            TRACE(0, pos, 2);
            RowFFT::run(r, bufT, tid, ln, twS, 0);  // (its barriers also order every staging read above before the writes below)
#pragma unroll
            for (int j = 0; j < 8; ++j) c2_store(&bufA[ln * N + tid + T * j], r[j]);
            fence_proxy_async();
            mbar_arrive(&bars[2]);
            TRACE(0, pos, 3);
            RowFFT::run(pb, bufT, tid, ln, twS, RowFFT::EXCHANGES);
#pragma unroll
            for (int j = 0; j < 8; ++j) c2_store(&bufB[ln * N + tid + T * j], pb[j]);
            fence_proxy_async();
            mbar_arrive(&bars[3]);
            TRACE(0, pos, 5);
        } else {
            // ================= column item: lines are columns, thread (tid, xw) owns y = tid + T*j of column x
            const int xw = threadIdx.x & (W - 1), tid = threadIdx.x / W;
            const int x = it.idx * W + xw;
            // Permute FFT:73-78: (x + y) % 2 with y = tid + T*j and T even  =>  parity of x + tid
            const Epilogue ep(((x + tid) & 1) ? -1.0f : 1.0f, lambda);
            float dxz[8];
            // ---- pair A = (DxDz, DyDxz) -> displacement
            mbar_wait(&bars[0], parity);
            TRACE(0, pos, 1);
#pragma unroll
            for (int j = 0; j < 8; ++j) r[j] = c2_load(&bufA[(tid + T * j) * W + xw]);
            ColFFT::run(r, bufA, tid, xw, twS, 0);
            ColFFT::barrier();  // every thread has finished reading the last exchange
            const bool heights = a.heightOut != nullptr && it.layer == 0;  // the readback plane of WB:288 (.g only)
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float4 v = ep.displacement(r[j], dxz[j]);
                bufA[(tid + T * j) * W + xw] = v;
                if (heights) a.heightOut[(size_t)(tid + T * j) * N + x] = v.y;
            }
            fence_proxy_async();
            mbar_arrive(&bars[2]);
            TRACE(0, pos, 3);
            // ---- pair B = (DyxDyz, DxxDzz) -> derivatives, Jacobian, foam
            mbar_wait(&bars[1], parity);
            TRACE(0, pos, 4);
#pragma unroll
            for (int j = 0; j < 8; ++j) r[j] = c2_load(&bufB[(tid + T * j) * W + xw]);
            ColFFT::run(r, bufB, tid, xw, twS, 0);
            ColFFT::barrier();
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int o = (tid + T * j) * W + xw;
                float jac;
                bufB[o] = ep.derivatives(r[j], dxz[j], jac);       // RF:26-27
                float f = zeroFoam ? 0.0f : bufT[o].x * decay;     // RF:28-30 (incoming texel, TMA-staged)
                if (f < jac) f += jac;                             // RF:31
                bufT[o] = make_float4(f, f, f, f);                 // RF:32
            }
            fence_proxy_async();
            mbar_arrive(&bars[3]);
            TRACE(0, pos, 5);
        }
    }
}
