// ocean_fused2.cuh -- the whole frame as ONE persistent, warp-specialised kernel with a DYNAMIC item queue and a
// one-item-deep input prefetch (N = 256, 512).  Included by ocean_frame.cu after ocean_fused.cuh (FrameItem, frame_item).
//
// Same items as ocean_fused.cuh (row item = W rows of a layer: TD:20-47 + x-FFT FFT:48-57; column item = W columns:
// y-FFT FFT:60-69 + Permute FFT:73-78 + RF:16-34) and the same per-layer release/acquire counters, but:
//
//   * items come from a queue (FrameArgs::itemNext, one atomicAdd per item) in the order R0 .. R(L-1) C0 .. C(L-1),
//     so whichever CTA is free takes the next item: every row item is under way before the first column item is
//     handed out, and a CTA that starts late (it became resident when a CTA of the previous frame left) simply takes
//     fewer items.  With the static round robin of ocean_fused.cuh half of the column items waited for the slowest
//     CTA's rows (profiles/r1_c_phase_trace_fused.txt: 4.35 us of a column item's 8.2 us).
//   * the three buffers ROTATE instead of having fixed roles: item n transforms in buffers (a, b) and its successor's
//     first input (h0 rows / the plane-A tile) lands in the third one, p, WHILE item n is being computed; then
//     (a, b, p) <- (p, a, b).  Row items transform in place in their staging buffers (exchange 0 is dense), which is
//     what frees the third buffer; a row item has ONE input: (1/|k|, omega) is recomputed from (kx, kz) with the very
//     operations InitialSpectrum used (IS:33-35, IS:112, IS:123 -- bit-identical) instead of being loaded.
//   * a column item's incoming turbulence (RF:28: only .x is read) comes through the compute threads' own loads, issued
//     before the second transform, instead of a 32 KiB staged tile; the outgoing turbulence is staged in buffer a once
//     the displacement tile has left it.
//
// Barriers (shared memory): full[3] -- buffer i holds the input that was loaded into it (or, for the second buffer of a
// row item, is merely free); stagedA / stagedB -- the compute warps have staged the first / second output set;
// freeA -- the displacement tile has been read out of buffer a (column items only).
#pragma once

template <int N> struct Fused2Cfg {
    static constexpr int T = N / 8;
    static constexpr int W = OCEAN_COLS_TMA_W;                  // lines (rows or columns) per item
    static constexpr int COMPUTE = W * T;
    static constexpr int THREADS = COMPUTE + 32;
    static constexpr int BOX_ROWS = N < 256 ? N : 256;
    static constexpr int NBOX = N / BOX_ROWS;
    static constexpr int TILE_ELEMS = N * W;
    static constexpr int TILE_BYTES = TILE_ELEMS * 16;
    static constexpr int IPL = N / W;                           // items per layer and role
    using ColFFT = LineFFT<N, W, true, false, true, true, COMPUTE>;
    using RowFFT = LineFFT<N, W, false, false, true, true, COMPUTE>;  // in place: exchange 0 dense, thread-private slots
    static constexpr int TW_BYTES = ColFFT::TWIDDLE_ENTRIES * 8;
    static constexpr int SMEM_BYTES = 3 * TILE_BYTES + TW_BYTES + 128;  // 3 buffers, twiddles, 6 mbarriers + 2 item descriptors
    static_assert(OCEAN_COLS_TMA_W == W, "the TMA descriptors are encoded for tiles of W columns");
};

struct Fused2Shared {
    uint64_t full[3];
    uint64_t stagedA, stagedB, freeA;
    int desc[2][4];  // item descriptors (role, layer, idx, queue position), double buffered by item parity
};

// (1/|k|, omega) of a texel, as InitialSpectrum stored them: |k| = length(k) IS:112, 1/|k| IS:123, omega = sqrt(g |k|)
// IS:33-35, every operation rounded once in that order.  Outside the band the reference stores (1, 0) next to h0 = 0; with
// h0 = 0 any FINITE pair gives the same (zero) spectra, so only k = 0 needs the guard.
#ifndef OCEAN_F2_EXP
#define OCEAN_F2_EXP 0  // timing experiments (results WRONG): 1 = approximate (1/|k|, omega), 2 = no incoming-turbulence loads
#endif
__device__ __forceinline__ float2 wave_hot(float kx, float kz, float gravity) {
#if OCEAN_F2_EXP & 1
    const float k2 = kx * kx + kz * kz, ik0 = rsqrtf(k2 + 1e-30f);
    return make_float2(ik0, __fsqrt_rz(gravity * k2 * ik0));
#endif
    const float kMag = __fsqrt_rn(__fadd_rn(__fmul_rn(kx, kx), __fmul_rn(kz, kz)));
    const float ik = kMag > 0.0f ? __frcp_rn(kMag) : 1.0f;
    return make_float2(ik, __fsqrt_rn(__fmul_rn(gravity, kMag)));
}

template <int N>
__device__ __forceinline__ void fused2_producer(const ColsTmaMaps& maps, const FrameArgs& a, float4* const (&buf)[3], Fused2Shared* sh,
                                                const int totalItems) {
    using Cfg = Fused2Cfg<N>;
    constexpr int W = Cfg::W, TILE_BYTES = Cfg::TILE_BYTES;
    auto tile_load = [&](const CUtensorMap* map, float4* dst, uint64_t* bar, const FrameItem& it) {
#pragma unroll
        for (int b = 0; b < Cfg::NBOX; ++b)
            tma_load_2d(dst + b * Cfg::BOX_ROWS * W, map, bar, it.idx * W * 4, it.layer * N + b * Cfg::BOX_ROWS);
    };
    auto tile_store = [&](const CUtensorMap* map, const float4* src, const FrameItem& it) {
#pragma unroll
        for (int b = 0; b < Cfg::NBOX; ++b)
            if (OCEAN_L2_HINTS & 2) tma_store_2d_hint(map, src + b * Cfg::BOX_ROWS * W, it.idx * W * 4, it.layer * N + b * Cfg::BOX_ROWS, L2_EVICT_FIRST);
            else tma_store_2d(map, src + b * Cfg::BOX_ROWS * W, it.idx * W * 4, it.layer * N + b * Cfg::BOX_ROWS);
    };
    auto row_offset = [&](const FrameItem& it) { return ((size_t)it.layer * N + (size_t)it.idx * W) * N; };
    auto fetch = [&]() {  // the next item of the frame's queue; exactly one failing fetch per CTA (FrameArgs::itemBase arithmetic)
        const unsigned long long v = atomicAdd(a.itemNext, 1ull) - a.itemBase;
        if (v >= (unsigned long long)totalItems) return FrameItem{-1, 0, 0};
        return frame_item((int)v, Cfg::IPL, a.layers);
    };
    // a column item needs its layer's rows of THIS frame; a row item's input is immutable
    auto ready = [&](const FrameItem& it) { return it.role <= 0 || ld_acquire_gpu(a.rowsDone + it.layer) >= a.rowsTarget; };
    // first input of item `it` (its n-th of this CTA) into buffer x: descriptor, then the copy; role < 0: just wake the consumers
    auto load_first = [&](const FrameItem& it, int n, int x) {
        int* d = sh->desc[n & 1];
        d[0] = it.role; d[1] = it.layer; d[2] = it.idx;
        if (it.role < 0) { mbar_arrive(&sh->full[x]); return; }
        mbar_arrive_expect_tx(&sh->full[x], TILE_BYTES);
        if (it.role == 0) bulk_load_1d(buf[x], a.h0 + row_offset(it), TILE_BYTES, &sh->full[x]);
        else { fence_proxy_async_global(); tile_load(&maps.a, buf[x], &sh->full[x], it); }  // acquired writes -> TMA reads
    };
    // second buffer of item `it`: the plane-B tile of a column item; a row item only needs it to be free
    auto load_second = [&](const FrameItem& it, int x) {
        if (it.role == 0) { mbar_arrive(&sh->full[x]); return; }
        mbar_arrive_expect_tx(&sh->full[x], TILE_BYTES);
        tile_load(&maps.b, buf[x], &sh->full[x], it);
    };

    int ba = 0, bb = 1, bp = 2, n = 0;
    FrameItem cur = fetch();
    if (cur.role == 1) spin_until_ge(a.rowsDone + cur.layer, a.rowsTarget);
    load_first(cur, 0, ba);
    if (cur.role < 0) return;
    load_second(cur, bb);
    uint32_t parity = 0;
    for (;;) {
        const FrameItem nxt = fetch();
        bool loadedNext = false;
        auto prefetch = [&]() {  // never blocks: a dependency of the NEXT item may be waiting for a signal this CTA still owes
            if (!loadedNext && ready(nxt)) { load_first(nxt, n + 1, bp); loadedNext = true; }
        };
        prefetch();
        mbar_wait(&sh->stagedA, parity);  // first output set staged in buffer a
        if (cur.role == 0) {
            // write-after-read on the workspace: the previous frame's columns of this layer are done
            spin_until_ge(a.colsDone + cur.layer, a.colsTarget);
            if (OCEAN_L2_HINTS & 4) bulk_store_1d_hint(a.rowsA + row_offset(cur), buf[ba], TILE_BYTES, L2_EVICT_LAST);
            else bulk_store_1d(a.rowsA + row_offset(cur), buf[ba], TILE_BYTES);
        } else {
            tile_store(&maps.disp, buf[ba], cur);
        }
        bulk_commit();
        bulk_wait_read_all();
        if (cur.role == 1) mbar_arrive(&sh->freeA);  // the outgoing turbulence may now be staged in buffer a
        prefetch();
        mbar_wait(&sh->stagedB, parity);  // second output set staged in buffer b (and, column items, buffer a)
        if (cur.role == 0) {
            if (OCEAN_L2_HINTS & 4) bulk_store_1d_hint(a.rowsB + row_offset(cur), buf[bb], TILE_BYTES, L2_EVICT_LAST);
            else bulk_store_1d(a.rowsB + row_offset(cur), buf[bb], TILE_BYTES);
        } else {
            tile_store(&maps.deriv, buf[bb], cur);
            tile_store(&maps.turb, buf[ba], cur);
        }
        bulk_commit();
        bulk_wait_read_all();  // buffers a and b are free
        prefetch();
        bool secondIssued = false;
        if (loadedNext && nxt.role >= 0) { load_second(nxt, ba); secondIssued = true; }
        // everything this item wrote is in memory: publish it
        bulk_wait_all();
        fence_proxy_async_global();
        __threadfence();
        atomicAdd((cur.role == 0 ? a.rowsDone : a.colsDone) + cur.layer, (unsigned long long)W);
        if (!loadedNext) {
            if (nxt.role == 1) spin_until_ge(a.rowsDone + nxt.layer, a.rowsTarget);
            load_first(nxt, n + 1, bp);
        }
        if (nxt.role < 0) return;
        if (!secondIssued) load_second(nxt, ba);
        const int t = bp; bp = bb; bb = ba; ba = t;  // (a, b, p) <- (p, a, b)
        cur = nxt;
        ++n;
        parity ^= 1;
    }
}

template <int N>
__global__ void __launch_bounds__(Fused2Cfg<N>::THREADS, 2)
ocean_frame2_kernel(const __grid_constant__ ColsTmaMaps maps, const FrameArgs a) {
    using Cfg = Fused2Cfg<N>;
    using ColFFT = typename Cfg::ColFFT;
    using RowFFT = typename Cfg::RowFFT;
    constexpr int T = Cfg::T, W = Cfg::W;
    extern __shared__ __align__(128) float4 smem[];
    float4* const buf[3] = {smem, smem + Cfg::TILE_ELEMS, smem + 2 * Cfg::TILE_ELEMS};
    float2* twS = (float2*)(smem + 3 * Cfg::TILE_ELEMS);
    Fused2Shared* sh = (Fused2Shared*)((char*)twS + Cfg::TW_BYTES);

    pdl_trigger();
    if (threadIdx.x == 0) {
        for (int i = 0; i < 3; ++i) mbar_init(&sh->full[i], 1);
        mbar_init(&sh->stagedA, Cfg::COMPUTE);
        mbar_init(&sh->stagedB, Cfg::COMPUTE);
        mbar_init(&sh->freeA, 1);
        mbar_fence_init();
    }
    for (int i = threadIdx.x; i < ColFFT::TWIDDLE_ENTRIES; i += Cfg::THREADS) twS[i] = __ldg(&a.twiddle[i]);  // immutable
    __syncthreads();
    const int totalItems = 2 * a.layers * Cfg::IPL;

    if (threadIdx.x >= Cfg::COMPUTE) {
        if (threadIdx.x == Cfg::COMPUTE) fused2_producer<N>(maps, a, buf, sh, totalItems);
        return;
    }

    const float lambda = a.lambda, decay = a.foamDecay, time = a.time;
    const bool zeroFoam = a.zeroFoam != 0;
    int ba = 0, bb = 1, bp = 2;
    uint32_t fullPhase = 0;  // bit i: phase of full[i]
    uint32_t parity = 0, freeParity = 0;
    for (int n = 0;; ++n, parity ^= 1) {
#ifdef OCEAN_TRACE
        unsigned long long traceWait;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(traceWait));
#endif
        mbar_wait(&sh->full[ba], (fullPhase >> ba) & 1);
        fullPhase ^= 1u << ba;
        const int role = sh->desc[n & 1][0], layer = sh->desc[n & 1][1], idx = sh->desc[n & 1][2];
        if (role < 0) break;
#ifdef OCEAN_TRACE
        if (threadIdx.x == 0) {
            const int pos = (role * a.layers + layer) * Cfg::IPL + idx;
            if (pos < OCEAN_TRACE_CTAS) { g_trace[0][pos][5][0] = traceWait; g_trace_smid[1][pos] = blockIdx.x; }
        }
#endif
        float4* A = buf[ba];
        float4* B = buf[bb];
        C2 r[8];
        TRACE(0, (role * a.layers + layer) * Cfg::IPL + idx, 0);
        if (role == 0) {
            // ================= row item: lines are rows, thread (ln, tid) owns x = tid + T*j of row idx*W + ln
            const int ln = threadIdx.x / T, tid = threadIdx.x - ln * T;
            const float dk = a.dk[layer % a.C];
            const float gravity = a.states ? a.states[layer / a.C].gravity : a.gravity;
            const float kz = (float)(idx * W + ln - N / 2) * dk;  // IS:111 (waveData.z)
            C2 pb[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int x = tid + T * j;
                const float kx = (float)(x - N / 2) * dk;  // IS:111 (waveData.x)
                evolve_texel(A[ln * N + x], wave_hot(kx, kz, gravity), kx, kz, time, r[j], pb[j]);
            }
            TRACE(0, (role * a.layers + layer) * Cfg::IPL + idx, 1);
            // in place: exchange 0 is dense, so every thread overwrites exactly the eight slots it has just read
            RowFFT::run(r, A, tid, ln, twS, 0);
            RowFFT::barrier();  // every thread has finished reading the last exchange
#pragma unroll
            for (int j = 0; j < 8; ++j) c2_store(&A[ln * N + tid + T * j], r[j]);
            fence_proxy_async();
            mbar_arrive(&sh->stagedA);
            TRACE(0, (role * a.layers + layer) * Cfg::IPL + idx, 2);
            mbar_wait(&sh->full[bb], (fullPhase >> bb) & 1);  // the second buffer is free
            fullPhase ^= 1u << bb;
            RowFFT::run(pb, B, tid, ln, twS, 0);
            RowFFT::barrier();
#pragma unroll
            for (int j = 0; j < 8; ++j) c2_store(&B[ln * N + tid + T * j], pb[j]);
            fence_proxy_async();
            mbar_arrive(&sh->stagedB);
            TRACE(0, (role * a.layers + layer) * Cfg::IPL + idx, 4);
        } else {
            // ================= column item: lines are columns, thread (tid, xw) owns y = tid + T*j of column x
            const int xw = threadIdx.x & (W - 1), tid = threadIdx.x / W;
            const int x = idx * W + xw;
            // Permute FFT:73-78: (x + y) % 2 with y = tid + T*j and T even  =>  parity of x + tid
            const Epilogue ep(((x + tid) & 1) ? -1.0f : 1.0f, lambda);
            float dxz[8], foam[8];
            // ---- pair A = (DxDz, DyDxz) -> displacement
#pragma unroll
            for (int j = 0; j < 8; ++j) r[j] = c2_load(&A[(tid + T * j) * W + xw]);
            ColFFT::run(r, A, tid, xw, twS, 0);
            ColFFT::barrier();  // every thread has finished reading the last exchange
            const bool heights = a.heightOut != nullptr && layer == 0;  // the readback plane of WB:288 (.g only)
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float4 v = ep.displacement(r[j], dxz[j]);
                A[(tid + T * j) * W + xw] = v;
                if (heights) a.heightOut[(size_t)(tid + T * j) * N + x] = v.y;
            }
            fence_proxy_async();
            mbar_arrive(&sh->stagedA);
            TRACE(0, (role * a.layers + layer) * Cfg::IPL + idx, 2);
            // incoming turbulence (RF:28 reads .x only): written by the previous frame's column items, which this item is
            // ordered after through colsDone -> rowsDone; L1 is not coherent with those stores, so the loads go to L2
            {
                const float4* __restrict__ tin = a.turb + (size_t)layer * N * N + x;
#pragma unroll
                for (int j = 0; j < 8; ++j) foam[j] = (zeroFoam || (OCEAN_F2_EXP & 2)) ? 0.0f : __ldcg(&tin[(size_t)(tid + T * j) * N].x);
            }
            // ---- pair B = (DyxDyz, DxxDzz) -> derivatives, Jacobian, foam
            mbar_wait(&sh->full[bb], (fullPhase >> bb) & 1);
            fullPhase ^= 1u << bb;
            TRACE(0, (role * a.layers + layer) * Cfg::IPL + idx, 3);
#pragma unroll
            for (int j = 0; j < 8; ++j) r[j] = c2_load(&B[(tid + T * j) * W + xw]);
            ColFFT::run(r, B, tid, xw, twS, 0);
            ColFFT::barrier();
            mbar_wait(&sh->freeA, freeParity);  // the displacement tile has left buffer a
            freeParity ^= 1;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int o = (tid + T * j) * W + xw;
                float jac;
                B[o] = ep.derivatives(r[j], dxz[j], jac);          // RF:26-27
                float f = foam[j] * decay;                         // RF:28-30
                if (f < jac) f += jac;                             // RF:31
                A[o] = make_float4(f, f, f, f);                    // RF:32
            }
            fence_proxy_async();
            mbar_arrive(&sh->stagedB);
            TRACE(0, (role * a.layers + layer) * Cfg::IPL + idx, 4);
        }
        const int t = bp; bp = bb; bb = ba; ba = t;  // (a, b, p) <- (p, a, b)
    }
}
