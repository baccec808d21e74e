// ocean_host.cu -- host side of liboceanfft.so: the C++ mirror of WaterBody (WB) / IFFT (IFFT.cs)
// and the C ABI declared in include/oceanfft.h.  The reference's host is C# (no toolchain in
// this image), so the host logic is C++ with the same member names and call order.
//
//   WaterBodyNative::Awake                        <- WaterBody.Awake WB:211-256
//   WaterBodyNative::CalculateInitialSpectrumTextures <- WB:171-178
//   WaterBodyNative::CalculateWavesTexturesAtTime <- WB:180-193
//   WaterBodyNative::Update                       <- WB:284-297
//   WaterBodyNative::GetWaterHeight (batched)     <- WB:195-209
//   WaterBodyNative::OnDisable                    <- WB:300-309
//
// There is no CPU implementation of any kernel here: if CUDA is unavailable every entry point
// that needs the device fails with OCEAN_ERR_CUDA.
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "ocean_internal.h"

namespace ocean {

static std::atomic<uint64_t> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
uint64_t launch_count() { return g_launches.load(std::memory_order_relaxed); }

static thread_local std::string g_error;

static int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_error = buf;
    return code;
}

#define OCEAN_CUDA(expr)                                                                             \
    do {                                                                                             \
        cudaError_t e__ = (expr);                                                                    \
        if (e__ != cudaSuccess)                                                                      \
            return fail(OCEAN_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

struct DeviceScope {
    int prev = -1;
    bool ok = false;
    bool switched = false;
    // (only switches -- and later restores -- when another device is current: three runtime calls fewer per API call in the usual
    // one-device-per-process case, which is what paces small frames: at 256^2 the loop is bound by host submission)
    explicit DeviceScope(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev == dev) { ok = true; return; }
        ok = cudaSetDevice(dev) == cudaSuccess;
        switched = ok;
    }
    ~DeviceScope() {
        if (switched && prev >= 0) cudaSetDevice(prev);
    }
};

static bool finite_pos(float v) { return std::isfinite(v) && v > 0.0f; }

static int validate_params(const OceanParams& p) {
    const int N = p.textureSize;
    if (N < OCEAN_MIN_SIZE || N > OCEAN_MAX_SIZE || (N & (N - 1)) != 0)
        return fail(OCEAN_ERR_INVALID_SIZE, "textureSize %d is not a power of two in [%d, %d]", N, OCEAN_MIN_SIZE, OCEAN_MAX_SIZE);
    if (!finite_pos(p.windSpeed) || !finite_pos(p.gravity) || !finite_pos(p.fetch) || !finite_pos(p.depth))
        return fail(OCEAN_ERR_INVALID_PARAMS, "windSpeed, gravity, fetch and depth must be finite and > 0");
    if (!std::isfinite(p.windDirX) || !std::isfinite(p.windDirY) || (p.windDirX == 0.0f && p.windDirY == 0.0f))
        return fail(OCEAN_ERR_INVALID_PARAMS, "wind direction must be a finite non-zero vector");
    if (!std::isfinite(p.lambda)) return fail(OCEAN_ERR_INVALID_PARAMS, "lambda must be finite");
    return OCEAN_OK;
}

static int validate_cascades(const OceanCascade* c, int n) {
    if (n < 1 || n > OCEAN_MAX_CASCADES) return fail(OCEAN_ERR_INVALID_CASCADES, "nCascades %d outside [1, %d]", n, OCEAN_MAX_CASCADES);
    if (!c) return fail(OCEAN_ERR_INVALID_ARGUMENT, "cascades is NULL");
    for (int i = 0; i < n; ++i) {
        if (!finite_pos(c[i].wavelength)) return fail(OCEAN_ERR_INVALID_CASCADES, "cascade %d: wavelength must be > 0", i);
        if (!finite_pos(c[i].cutoffLow)) return fail(OCEAN_ERR_INVALID_CASCADES, "cascade %d: cutoffLow must be > 0 (k = 0 gives NaN, IS:114-120)", i);
        if (!(c[i].cutoffHigh > 0.0f)) return fail(OCEAN_ERR_INVALID_CASCADES, "cascade %d: cutoffHigh must be > 0", i);
        if (!std::isfinite(c[i].swell) || !std::isfinite(c[i].fade)) return fail(OCEAN_ERR_INVALID_CASCADES, "cascade %d: swell/fade must be finite", i);
    }
    return OCEAN_OK;
}

// fp32, reference operation order (this file is compiled with -ffp-contract=off on the host)
static const float kPI = 3.14159265f;  // IS:8

static StateConsts make_state_consts(const OceanParams& p) {
    StateConsts s;
    const float U = p.windSpeed, g = p.gravity, F = p.fetch, D = p.depth;
    s.windSpeed = U; s.gravity = g; s.fetch = F; s.depth = D;
    s.sqrtDepthOverG = sqrtf(D / g);                                        // IS:39
    s.peakOmega = 22.0f * powf(fabsf(g * g / (U * F)), 0.3333f);            // IS:118
    s.alpha = 0.076f * powf(fabsf(U * U / (F * g)), 0.22f);                 // IS:48
    const float peakSpeed = g / s.peakOmega;                                // IS:63
    s.mu = -2.33f - 1.45f * (U / peakSpeed - 1.17f);                        // IS:64
    const float len = sqrtf(p.windDirX * p.windDirX + p.windDirY * p.windDirY);
    s.windTheta = atan2f(p.windDirY / len, p.windDirX / len);               // IS:80-81
    s.lambda = p.lambda;
    s.pad0 = s.pad1 = 0.0f;
    return s;
}

static SpreadConsts make_spread_consts() {
    SpreadConsts c;
    c.pi = kPI;
    c.c2 = powf(logf(2.0f), 2.0f) / kPI - kPI / 12.0f;  // IS:72
    c.c1 = logf(2.0f) / kPI;
    c.c0 = 1.0f / (2.0f * kPI);
    c.twoSqrtPi = 2.0f * sqrtf(kPI);                    // IS:73
    return c;
}

static CascadeConsts make_cascade_consts(const OceanCascade& c) {
    CascadeConsts k;
    k.dk = 2.0f * kPI / c.wavelength;  // IS:110
    k.cutoffLow = c.cutoffLow; k.cutoffHigh = c.cutoffHigh; k.swell = c.swell; k.fade = c.fade;
    k.wavelength = c.wavelength;
    return k;
}

// Twiddle table of the radix-8 stages (replaces the butterfly texture of IFFT.cs:36 / FFT:37-45):
// tw[(s*7 + q-1)*T + tid] = e^{+2 pi i * t*q*8^s / N} with T = N/8, t = tid mod (N / 8^(s+1)),
// for the S-1 twiddled stages.  One entry per (stage, q, thread): warps read consecutive float2.
// Evaluated in double, rounded once.
static std::vector<float2> make_twiddles(int N) {
    int logN = 0;
    while ((1 << logN) < N) ++logN;
    const int S = (logN + 2) / 3, T = N / 8;
    std::vector<float2> tw((size_t)(S > 1 ? S - 1 : 1) * 7 * T, make_float2(1.0f, 0.0f));
    for (int s = 0; s + 1 < S; ++s) {
        const int Cs = 1 << (3 * s), Ms = N / Cs;
        for (int q = 1; q < 8; ++q)
            for (int tid = 0; tid < T; ++tid) {
                const int t = tid & (Ms / 8 - 1);
                const double a = 2.0 * M_PI * (double)((long long)t * q * Cs % N) / (double)N;
                tw[(size_t)(s * 7 + q - 1) * T + tid] = make_float2((float)cos(a), (float)sin(a));
            }
    }
    return tw;
}

// ---------------------------------------------------------------------------------------------
// Workspace: the two intermediate planes between the row and the column pass ([layers][N][N] float4
// each, contiguous).  It is scratch: nothing in it survives a frame.  Handles that were put on the
// same caller-provided stream (ocean_set_stream) are serialised by that stream, so they share one
// workspace per (device, stream, N, layers) -- like FFT plans sharing a work area -- and the planes stay
// resident in L2 however many oceans rotate through them.  The key is the geometry, not the byte size:
// the hand-off counters are one per layer and their targets count rows of N texels, so a 512 x 1 and a
// 256 x 4 ocean (equal bytes) must not meet on one workspace.  A handle on its own stream owns a
// private workspace.
struct Workspace {
    int device = 0;
    cudaStream_t stream = nullptr;  // nullptr: private
    size_t planeBytes = 0;
    int N = 0;                      // texture size of every user (part of the sharing key)
    float4* base = nullptr;
    int refs = 0;
    // Hand-off counters live with the memory they guard.  counters[0..layers) = rows written into the
    // planes (row -> column), counters[layers..2*layers) = columns consumed and stored (column -> row).
    // rowPasses / colPasses count the passes ENQUEUED on this workspace (its users are serialised by
    // one stream, so enqueue order is execution order); a pass's target is N * its ordinal.
    // counters[2*layers + {0, 1}] = the item queues of the dynamically scheduled frame kernel (ocean_fused2.cuh), used
    // alternately by consecutive frames; queueUses[i] * (items + grid) is where queue i's next frame starts.
    int layers = 0;
    unsigned long long* counters = nullptr;
    uint64_t rowPasses = 0, colPasses = 0;
    uint64_t queueFrames = 0, queueNext[2] = {0, 0};
    float4* planeA() const { return base; }
    float4* planeB() const { return (float4*)((char*)base + planeBytes); }
};
static std::mutex g_wsMutex;
static std::vector<Workspace*> g_workspaces;

static Workspace* workspace_acquire(int device, cudaStream_t shared, int N, int layers) {
    std::lock_guard<std::mutex> lock(g_wsMutex);
    const size_t planeBytes = (size_t)layers * N * N * sizeof(float4);
    if (shared)
        for (Workspace* w : g_workspaces)
            if (w->device == device && w->stream == shared && w->N == N && w->layers == layers) { ++w->refs; return w; }
    Workspace* w = new Workspace;
    w->device = device; w->stream = shared; w->planeBytes = planeBytes; w->refs = 1;
    w->N = N; w->layers = layers;
    if (cudaMalloc(&w->base, 2 * planeBytes) != cudaSuccess) { delete w; return nullptr; }
    if (cudaMalloc(&w->counters, (2 * layers + 2) * sizeof(unsigned long long)) != cudaSuccess ||
        cudaMemset(w->counters, 0, (2 * layers + 2) * sizeof(unsigned long long)) != cudaSuccess) {
        cudaFree(w->base);
        delete w;
        return nullptr;
    }
    if (shared) g_workspaces.push_back(w);
    return w;
}
static void workspace_release(Workspace* w) {
    if (!w) return;
    std::lock_guard<std::mutex> lock(g_wsMutex);
    if (--w->refs > 0) return;
    for (size_t i = 0; i < g_workspaces.size(); ++i)
        if (g_workspaces[i] == w) { g_workspaces.erase(g_workspaces.begin() + i); break; }
    cudaFree(w->base);
    cudaFree(w->counters);
    delete w;
}
// Stamps a frame's hand-off targets.  rows / cols say which passes this launch sequence contains.
// Assigns the frame an item queue (fused frame kernel, dynamic scheduling): totalItems + grid fetches will be made.
static void workspace_stamp_queue(Workspace* w, int totalItems, int grid, FrameArgs* a) {
    const int q = (int)(w->queueFrames++ & 1);
    a->itemNext = w->counters + 2 * w->layers + q;
    a->itemBase = w->queueNext[q];
    w->queueNext[q] += (uint64_t)totalItems + (uint64_t)grid;
}
static void workspace_stamp(Workspace* w, int N, bool rows, bool cols, FrameArgs* a) {
    if (rows) ++w->rowPasses;
    a->rowsDone = w->counters;
    a->colsDone = w->counters + w->layers;
    a->rowsTarget = (unsigned long long)N * w->rowPasses;  // the column pass waits for this frame's rows
    a->colsTarget = (unsigned long long)N * w->colPasses;  // the row pass waits for every earlier column pass
    if (cols) ++w->colPasses;
}

// Optional persisting-L2 set-aside for the workspace, once per device: min(device maximum,
// OCEAN_L2_PERSIST_MB MiB).  OFF by default (0): measured on B200 the per-launch access-policy window
// made the frame slower (52 us vs 27.5 us at 4 x 512^2) -- sharing the workspace already keeps the
// planes L2-resident.  Returns the bytes set aside (0 = disabled / unsupported).
static size_t l2_set_aside(int device) {
    static std::mutex m;
    static std::vector<long long> cache(64, -1);
    std::lock_guard<std::mutex> lock(m);
    if (device < 0 || device >= (int)cache.size()) return 0;
    if (cache[device] >= 0) return (size_t)cache[device];
    long long want = 0;
    if (const char* e = getenv("OCEAN_L2_PERSIST_MB")) want = atoll(e);
    int maxPersist = 0, maxWindow = 0;
    cudaDeviceGetAttribute(&maxPersist, cudaDevAttrMaxPersistingL2CacheSize, device);
    cudaDeviceGetAttribute(&maxWindow, cudaDevAttrMaxAccessPolicyWindowSize, device);
    long long bytes = want * (1ll << 20);
    if (bytes > maxPersist) bytes = maxPersist;
    if (bytes <= 0 || maxWindow <= 0 || cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)bytes) != cudaSuccess) {
        cudaGetLastError();
        bytes = 0;
    }
    cache[device] = bytes;
    return (size_t)bytes;
}
// How many layers' workspace planes (N*N*32 B each) the frame asks L2 to keep (evict_last stores); the others are streamed
// (evict_first stores, read-once loads).  Default: all of them, also where they cannot all fit (4 x 1024^2: 134 MB of planes
// against 126 MB of L2) -- measured on B200, us per 4 x 1024^2 frame with 4 / 3 / 2 / 1 / 0 resident layers:
// 126.5 / 127.6 / 128.3 / 128.2 / 131.2 (profiles/r2_experiments.txt): partial residency buys nothing.  OCEAN_WS_RESIDENT=k forces k.
static int ws_resident_layers(int /*device*/, int /*N*/, int layers) {
    static const int forced = [] { const char* e = getenv("OCEAN_WS_RESIDENT"); return e ? atoi(e) : -1; }();
    if (forced >= 0) return forced < layers ? forced : layers;
    return layers;
}
static void l2_window(int device, const Workspace* w, FrameArgs* a) {
    a->l2Window = nullptr; a->l2WindowBytes = 0; a->l2HitRatio = 0.0f;
    const size_t aside = l2_set_aside(device);
    if (!aside || !w) return;
    int maxWindow = 0;
    cudaDeviceGetAttribute(&maxWindow, cudaDevAttrMaxAccessPolicyWindowSize, device);
    size_t bytes = 2 * w->planeBytes;
    if (bytes > (size_t)maxWindow) bytes = (size_t)maxWindow;
    a->l2Window = w->base;
    a->l2WindowBytes = bytes;
    a->l2HitRatio = aside >= bytes ? 1.0f : (float)((double)aside / (double)bytes);
}

// cuStreamWaitValue64 through the runtime's driver entry point (the library links only the CUDA runtime)
typedef int (*StreamWaitValue64Fn)(cudaStream_t, unsigned long long /*CUdeviceptr*/, unsigned long long, unsigned int);
static StreamWaitValue64Fn stream_wait_value_fn() {
    static StreamWaitValue64Fn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (!getenv("OCEAN_NO_STREAM_WAIT_VALUE") &&
            cudaGetDriverEntryPoint("cuStreamWaitValue64", &p, cudaEnableDefault, &q) == cudaSuccess && p &&
            q == cudaDriverEntryPointSuccess)
            fn = (StreamWaitValue64Fn)p;
        else
            cudaGetLastError();
    }
    return fn;
}
static bool stream_wait_value_available() { return stream_wait_value_fn() != nullptr; }
static cudaError_t stream_wait_value_geq(cudaStream_t s, const unsigned long long* devAddr, unsigned long long value) {
    const int r = stream_wait_value_fn()(s, (unsigned long long)(uintptr_t)devAddr, value, 0x0u /*CU_STREAM_WAIT_VALUE_GEQ*/);
    return r == 0 ? cudaSuccess : cudaErrorUnknown;
}

// ---------------------------------------------------------------------------------------------
// Scratch of ocean_buoyancy_noalloc (no counterpart in the reference): callers
// that may not allocate -- code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc -- get the
// buoyancy chain (gather of the query points -> height query -> BO:19-29) with its intermediate arrays in this file-scope
// device array instead of OceanHandle::bBuf.  Layout, for MAX = OCEAN_NOALLOC_MAX_BOXES boxes (52 B per box, 3.25 MiB;
// every GPU that loads the library holds it until the process exits):
//     float4 answers[MAX] | float2 points[MAX] | float heights[MAX] | OceanBuoyantBox boxes[MAX]
// (answers / boxes are only used as staging when the caller's buffers live in host memory).
// In HOST code the name `g_noallocScratch` is a host-side stand-in, NOT a device address: handing it to a kernel faults,
// or silently uses the wrong memory where the GPU can reach host memory.  noalloc_scratch() therefore asks the runtime for
// the device address -- once per device, with that device current -- and caches it.
constexpr size_t kNoallocBytesPerBox = sizeof(float4) + sizeof(float2) + sizeof(float) + sizeof(OceanBuoyantBox);
static __device__ __align__(256) unsigned char g_noallocScratch[(size_t)OCEAN_NOALLOC_MAX_BOXES * kNoallocBytesPerBox];

static std::mutex g_noallocMutex;
static void* g_noallocPtr[64] = {};

// Device address of the scratch on `device`, which must be the current device.
static cudaError_t noalloc_scratch(int device, void** out) {
    *out = nullptr;
    if (device < 0 || device >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_noallocMutex);
    if (!g_noallocPtr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noallocScratch);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        g_noallocPtr[device] = p;
    }
    *out = g_noallocPtr[device];
    return cudaSuccess;
}

}  // namespace ocean

using namespace ocean;

// ---------------------------------------------------------------------------------------------
// WaterBodyNative: one WaterBody on one device.
struct OceanHandle {
    int device = 0;
    int N = 0, C = 0;
    // `params` / `cascades` are the ACTIVE set: the one the spectrum on the device was computed from, and the
    // one every frame, query and sampler uses (dk, wavelengths, lambda).  ocean_set_params only stages an edit
    // (`nextParams` / `nextCascades`); CalculateInitialSpectrumTextures commits it, so a frame issued between
    // the two still evolves a consistent (h0, omega, 1/|k|, dk) set.
    OceanParams params{};
    OceanCascade cascades[OCEAN_MAX_CASCADES]{};
    OceanParams nextParams{};
    OceanCascade nextCascades[OCEAN_MAX_CASCADES]{};
    cudaStream_t ownStream = nullptr, stream = nullptr, copyStream = nullptr;
    // arrays of WB:221-229 (+ noise WB:34); rowsA/rowsB replace the 4 spectra + ping-pong array
    float2* noise = nullptr;
    float4 *h0 = nullptr, *wave = nullptr;
    Workspace* ws = nullptr;    // rowsA / rowsB point into it
    FrameArgs l2{};             // only the l2Window* fields: computed once per workspace binding
    float4 *rowsA = nullptr, *rowsB = nullptr;
    float4 *disp = nullptr, *deriv = nullptr, *turb = nullptr;
    float2* waveHot = nullptr;  // (1/|k|, omega): compact copy of waveData read by the frame kernels
    float2* h0c = nullptr;      // h0(k): compact copy of h0.xy read by the pair-row kernel
    ColsTmaMaps colsMaps;       // TMA descriptors over rowsA / rowsB and the outputs
    int wsResident = 0;         // layers whose workspace planes stay L2-resident (ws_resident_layers)
    bool rowsPending = false;   // ocean_step_partial: a row pass was issued and its column pass has not been
    float2* twiddle = nullptr;
    StateConsts* stateDev = nullptr;
    uint64_t* seedDev = nullptr;
    bool spectrumReady = false;
    bool foamZero = true;
    int64_t frame = 0;
    // AsyncGPUReadback of displacement layer 0 (WB:288-296), triple buffered: one slot is the
    // visible buoyancyData (WB:58), up to two are in flight
    static constexpr int RB = 3;
    float4* rbStage[RB] = {nullptr, nullptr, nullptr};
    float* rbHost[RB] = {nullptr, nullptr, nullptr};
    cudaEvent_t rbDone[RB] = {nullptr, nullptr, nullptr};
    cudaEvent_t rbSnap = nullptr;
    int64_t rbFrame[RB] = {-1, -1, -1};
    int rbMode = OCEAN_READBACK_RGBA;             // what ocean_update ships to the host
    int rbSlotMode[RB] = {0, 0, 0};               // layout of each slot's host copy
    bool rbPending[RB] = {false, false, false};
    int rbLatest = -1;  // slot of the newest completed readback
    // ocean_update_buoyancy: the slot's staging plane is consumed on the device (boxes in, answers out); no plane
    // reaches the host, so such a slot never becomes the visible buoyancyData
    bool rbHasPlane[RB] = {true, true, true};
    OceanBuoyantBox* rbBoxes[RB] = {nullptr, nullptr, nullptr};
    float4* rbAnswers[RB] = {nullptr, nullptr, nullptr};
    int rbBoxCap[RB] = {0, 0, 0};
    // HEIGHT mode: the column pass itself writes the height plane of layer 0 into the slot's staging buffer;
    // the copy stream waits (stream memory operation) for the layer-0 column counter and ships it -- the
    // compute stream carries nothing but the two frame kernels
    bool SlotDone(int i, bool wait);
    // query scratch
    float2* qPts = nullptr;
    float* qOut = nullptr;
    int qCap = 0;
    // mip chains of the derivative and turbulence arrays (WB:191-192), allocated on first use
    float4 *derivMips = nullptr, *turbMips = nullptr;
    int64_t mipsFrame = -1;  // frame index the chains were built from
    bool autoMips = false;   // rebuild them at the end of every frame, as the reference does
    // readers of a frame's outputs that run beside the next frame (the mip chains)
    unsigned long long* auxDone = nullptr;  // device counter they bump when they stop reading
    unsigned long long auxSignals = 0;      // its value once every reader launched so far has finished
    // surface-sample scratch: n x (2 in + 7 out) floats
    float* sBuf = nullptr;
    float4* bBuf = nullptr;  // ocean_buoyancy scratch: out4 | points | heights | boxes, bCap boxes
    int bCap = 0;
    int sCap = 0;

    size_t texels() const { return (size_t)N * N; }
    size_t layerBytes4() const { return texels() * sizeof(float4); }

    int Awake(const float* noiseHost, uint64_t seed);
    int CalculateInitialSpectrumTextures();
    int CalculateWavesTexturesAtTime(float time, int which = 0, float* heightOut = nullptr);
    int Update(float time);
    int UpdateBuoyancy(float time, const OceanBuoyantBox* boxes, int n, float density, float gravityY, float* out4);
    int PickSlot(int* slot);
    void PollReadbacks(bool wait);
    int GetWaterHeight(const float* worldXZ, int n, int mode, float* heights, bool ptsDev, bool outDev);
    int AttachWorkspace(cudaStream_t shared);
    int GenerateMips();
    int Buoyancy(const OceanBuoyantBox* boxes, int n, int mode, float density, float gravityY, float* out4, bool boxesDev, bool outDev);
    int BuoyancyNoAlloc(const OceanBuoyantBox* boxes, int n, int mode, float density, float gravityY, float* out4, bool boxesDev,
                        bool outDev);
    int SampleSurface(const float* worldXZ, int n, float lod, float* disp3, float* normal3, float* foam, bool onDevice);
    void OnDisable();
};

// (Re)binds the intermediate planes: a private workspace, or the one shared by every handle on `shared`.
int OceanHandle::AttachWorkspace(cudaStream_t shared) {
    Workspace* w = workspace_acquire(device, shared, N, C);
    if (!w) return fail(OCEAN_ERR_CUDA, "cudaMalloc of the intermediate planes failed");
    workspace_release(ws);
    ws = w;
    rowsA = w->planeA();
    rowsB = w->planeB();
    l2_window(device, ws, &l2);
    wsResident = ws_resident_layers(device, N, C);
    OCEAN_CUDA(make_cols_maps(N, C, rowsA, rowsB, disp, deriv, turb, &colsMaps));
    return OCEAN_OK;
}

int OceanHandle::Awake(const float* noiseHost, uint64_t seed) {
    const size_t t = texels();
    OCEAN_CUDA(cudaStreamCreateWithFlags(&ownStream, cudaStreamNonBlocking));
    OCEAN_CUDA(cudaStreamCreateWithFlags(&copyStream, cudaStreamNonBlocking));
    stream = ownStream;
    OCEAN_CUDA(cudaMalloc(&noise, t * sizeof(float2)));
    OCEAN_CUDA(cudaMalloc(&h0, C * t * sizeof(float4)));
    OCEAN_CUDA(cudaMalloc(&wave, C * t * sizeof(float4)));
    OCEAN_CUDA(cudaMalloc(&waveHot, C * t * sizeof(float2)));
    if (rows_need_h0c()) OCEAN_CUDA(cudaMalloc(&h0c, C * t * sizeof(float2)));
    OCEAN_CUDA(cudaMalloc(&disp, C * t * sizeof(float4)));
    OCEAN_CUDA(cudaMalloc(&deriv, C * t * sizeof(float4)));
    OCEAN_CUDA(cudaMalloc(&turb, C * t * sizeof(float4)));
    const std::vector<float2> tw = make_twiddles(N);
    OCEAN_CUDA(cudaMalloc(&twiddle, tw.size() * sizeof(float2)));
    OCEAN_CUDA(cudaMalloc(&stateDev, sizeof(StateConsts)));
    OCEAN_CUDA(cudaMalloc(&seedDev, sizeof(uint64_t)));
    OCEAN_CUDA(cudaMemsetAsync(disp, 0, C * t * sizeof(float4), stream));
    OCEAN_CUDA(cudaMemsetAsync(deriv, 0, C * t * sizeof(float4), stream));
    OCEAN_CUDA(cudaMemsetAsync(turb, 0, C * t * sizeof(float4), stream));  // fresh RenderTexture
    OCEAN_CUDA(cudaMemcpyAsync(twiddle, tw.data(), tw.size() * sizeof(float2), cudaMemcpyHostToDevice, stream));
    if (noiseHost) {  // GenerateRandomNoiseTexture WB:86-100: the caller supplies the texture
        OCEAN_CUDA(cudaMemcpyAsync(noise, noiseHost, t * sizeof(float2), cudaMemcpyHostToDevice, stream));
    } else {          // ... or it is generated on the device from the seed
        OCEAN_CUDA(cudaMemcpyAsync(seedDev, &seed, sizeof seed, cudaMemcpyHostToDevice, stream));
        OCEAN_CUDA(launch_noise(seedDev, 1, N, noise, stream));
    }
    OCEAN_CUDA(configure_frame_kernels(N));
    { int rc = AttachWorkspace(nullptr); if (rc != OCEAN_OK) return rc; }
    for (int i = 0; i < RB; ++i) {
        OCEAN_CUDA(cudaMalloc(&rbStage[i], layerBytes4()));
        OCEAN_CUDA(cudaMallocHost(&rbHost[i], layerBytes4()));
        OCEAN_CUDA(cudaEventCreateWithFlags(&rbDone[i], cudaEventDisableTiming));
    }
    OCEAN_CUDA(cudaMalloc(&auxDone, sizeof(unsigned long long)));
    OCEAN_CUDA(cudaMemsetAsync(auxDone, 0, sizeof(unsigned long long), stream));
    OCEAN_CUDA(cudaEventCreateWithFlags(&rbSnap, cudaEventDisableTiming));
    OCEAN_CUDA(cudaStreamSynchronize(stream));  // host staging (tw, seed) may go out of scope
    return OCEAN_OK;
}

int OceanHandle::CalculateInitialSpectrumTextures() {
    params = nextParams;  // commit a staged ocean_set_params (no-op otherwise)
    memcpy(cascades, nextCascades, sizeof cascades);
    const StateConsts sc = make_state_consts(params);
    OCEAN_CUDA(cudaMemcpyAsync(stateDev, &sc, sizeof sc, cudaMemcpyHostToDevice, stream));
    InitArgs a{};
    a.N = N; a.C = C; a.nStates = 1;
    a.sc = make_spread_consts();
    for (int i = 0; i < C; ++i) a.casc[i] = make_cascade_consts(cascades[i]);
    a.states = stateDev; a.noise = noise; a.h0 = h0; a.wave = wave; a.waveHot = waveHot; a.h0c = h0c;
    OCEAN_CUDA(launch_init(a, stream));
    OCEAN_CUDA(cudaStreamSynchronize(stream));  // `sc` is stack memory
    spectrumReady = true;
    return OCEAN_OK;
}

int OceanHandle::CalculateWavesTexturesAtTime(float time, int which, float* heightOut) {
    if (!spectrumReady) return fail(OCEAN_ERR_NOT_INITIALIZED, "ocean_step before ocean_init_spectrum");
    FrameArgs a{};
    a.N = N; a.layers = C; a.C = C;
    for (int i = 0; i < C; ++i) a.dk[i] = make_cascade_consts(cascades[i]).dk;
    a.h0 = h0; a.waveHot = waveHot; a.h0c = h0c; a.rowsA = rowsA; a.rowsB = rowsB;
    a.disp = disp; a.deriv = deriv; a.turb = turb; a.twiddle = twiddle;
    a.time = time; a.lambda = params.lambda;
    a.foamDecay = expf(-2.0f);  // RF:29-30
    a.zeroFoam = foamZero ? 1 : 0;
    a.l2Window = l2.l2Window; a.l2WindowBytes = l2.l2WindowBytes; a.l2HitRatio = l2.l2HitRatio;
    const bool doRows = which == 0 || which == 1, doCols = which == 0 || which == 2;
    if (doRows && rowsPending) return fail(OCEAN_ERR_INVALID_ARGUMENT, "row pass issued twice without a column pass");
    if (doCols && !doRows && !rowsPending) return fail(OCEAN_ERR_INVALID_ARGUMENT, "column pass without a preceding row pass");
    workspace_stamp(ws, N, doRows, doCols, &a);
    a.auxDone = auxSignals ? auxDone : nullptr;  // only once a reader has been launched on this ocean
    a.auxTarget = auxSignals;
    a.heightOut = heightOut;
    a.gravity = params.gravity;
    a.wsResident = wsResident;
    const int fusedMode = which == 0 && !auxSignals ? frame_fused_mode(N, &colsMaps) : 0;
    if (fusedMode) {
        if (fusedMode == 2) workspace_stamp_queue(ws, 2 * C * (N / 4), frame_fused_grid(N, C, 2), &a);
        OCEAN_CUDA(launch_frame(a, &colsMaps, stream, true, fusedMode));
        foamZero = false;
        ++frame;
        return autoMips ? GenerateMips() : OCEAN_OK;
    }
    if (which == 0 && !auxSignals && !autoMips && hetero_mode(N, &colsMaps)) {
        OCEAN_CUDA(launch_hetero(a, &colsMaps, stream, hetero_mode(N, &colsMaps)));
        foamZero = false;
        ++frame;
        return OCEAN_OK;
    }
    if (doRows) {
        OCEAN_CUDA(launch_rows(a, stream, true));  // every producer of h0/waveHot was synchronised
        rowsPending = true;
    }
    if (doCols) {
        OCEAN_CUDA(launch_cols(a, &colsMaps, stream));
        rowsPending = false;
        foamZero = false;
        ++frame;
        if (autoMips) return GenerateMips();
    }
    return OCEAN_OK;
}

// derivativesTextures.GenerateMips(); turbulenceTextures.GenerateMips();   WB:191-192
int OceanHandle::GenerateMips() {
    const size_t bytes = (size_t)C * mip_texels_per_layer(N) * sizeof(float4);
    if (!derivMips) OCEAN_CUDA(cudaMalloc(&derivMips, bytes));
    if (!turbMips) OCEAN_CUDA(cudaMalloc(&turbMips, bytes));
    unsigned long long signals = 0;
    OCEAN_CUDA(launch_mip_chains(deriv, derivMips, turb, turbMips, N, C, auxDone, &signals, stream));
    auxSignals += signals;
    mipsFrame = frame;
    // (Building the chains on a side stream gated by the column counters was measured and dropped: with the
    // next column pass spinning for them, back-to-back frames starved the chain kernels of SM resources and
    // deadlocked; ordering that pass after them with an event instead cost 43.0 us per frame + chains
    // against 38.7 us for this in-stream programmatic chain.)
    return OCEAN_OK;
}

// BuoyantObject.FixedUpdate BO:18-30 for n boxes: gather the query points, any device query mode, the force epilogue.
int OceanHandle::Buoyancy(const OceanBuoyantBox* boxes, int n, int mode, float density, float gravityY, float* out4, bool boxesDev,
                          bool outDev) {
    if (n == 0) return OCEAN_OK;
    if (n > bCap) {
        if (bBuf) cudaFree(bBuf);
        bBuf = nullptr; bCap = 0;
        OCEAN_CUDA(cudaMalloc(&bBuf, (size_t)n * (sizeof(float4) + sizeof(float2) + sizeof(float) + sizeof(OceanBuoyantBox))));
        bCap = n;
    }
    float4* dOut = bBuf;
    float2* dPts = (float2*)(dOut + n);
    float* dHeights = (float*)(dPts + n);
    OceanBuoyantBox* dBoxes = (OceanBuoyantBox*)(dHeights + n);
    const OceanBuoyantBox* bx = boxes;
    if (!boxesDev) {
        OCEAN_CUDA(cudaMemcpyAsync(dBoxes, boxes, (size_t)n * sizeof(OceanBuoyantBox), cudaMemcpyHostToDevice, stream));
        bx = dBoxes;
    }
    OCEAN_CUDA(launch_buoyancy_points(bx, dPts, n, stream));
    float wl[OCEAN_MAX_CASCADES];
    for (int i = 0; i < C; ++i) wl[i] = cascades[i].wavelength;
    if (mode == OCEAN_QUERY_DISPLACED) {
        OCEAN_CUDA(launch_displaced_query(disp, N, C, wl, dPts, dHeights, n, OCEAN_DISPLACED_ITERATIONS, stream));
    } else {
        OCEAN_CUDA(launch_query(disp, N, C, wl, dPts, dHeights, n, mode, stream));
    }
    float4* out = outDev ? (float4*)out4 : dOut;
    OCEAN_CUDA(launch_buoyancy(bx, dHeights, n, density, gravityY, out, stream));
    if (!outDev) {
        OCEAN_CUDA(cudaMemcpyAsync(out4, dOut, (size_t)n * sizeof(float4), cudaMemcpyDeviceToHost, stream));
        OCEAN_CUDA(cudaStreamSynchronize(stream));
    }
    return OCEAN_OK;
}

// OceanHandle::Buoyancy with every intermediate in the static device array above: no cudaMalloc / cudaFree, no allocating
// library call.  Same kernels, same launch order, same results.  Every slot of the scratch that a kernel reads was written
// earlier in the same call (points and heights for i < n; the staged boxes by the copy), so nothing has to start at zero and
// the previous call's contents are simply overwritten.  The array is shared by every handle of the process on this device:
// two calls that overlap on the GPU (different streams) race on it.
int OceanHandle::BuoyancyNoAlloc(const OceanBuoyantBox* boxes, int n, int mode, float density, float gravityY, float* out4,
                                 bool boxesDev, bool outDev) {
    if (n == 0) return OCEAN_OK;
    void* base = nullptr;
    OCEAN_CUDA(noalloc_scratch(device, &base));
    const size_t MAX = OCEAN_NOALLOC_MAX_BOXES;
    float4* dOut = (float4*)base;
    float2* dPts = (float2*)(dOut + MAX);
    float* dHeights = (float*)(dPts + MAX);
    OceanBuoyantBox* dBoxes = (OceanBuoyantBox*)(dHeights + MAX);
    const OceanBuoyantBox* bx = boxes;
    if (!boxesDev) {
        OCEAN_CUDA(cudaMemcpyAsync(dBoxes, boxes, (size_t)n * sizeof(OceanBuoyantBox), cudaMemcpyHostToDevice, stream));
        bx = dBoxes;
    }
    OCEAN_CUDA(launch_buoyancy_points(bx, dPts, n, stream));
    float wl[OCEAN_MAX_CASCADES];
    for (int i = 0; i < C; ++i) wl[i] = cascades[i].wavelength;
    if (mode == OCEAN_QUERY_DISPLACED) {
        OCEAN_CUDA(launch_displaced_query(disp, N, C, wl, dPts, dHeights, n, OCEAN_DISPLACED_ITERATIONS, stream));
    } else {
        OCEAN_CUDA(launch_query(disp, N, C, wl, dPts, dHeights, n, mode, stream));
    }
    float4* out = outDev ? (float4*)out4 : dOut;
    OCEAN_CUDA(launch_buoyancy(bx, dHeights, n, density, gravityY, out, stream));
    if (!outDev) {
        OCEAN_CUDA(cudaMemcpyAsync(out4, dOut, (size_t)n * sizeof(float4), cudaMemcpyDeviceToHost, stream));
        OCEAN_CUDA(cudaStreamSynchronize(stream));
    }
    return OCEAN_OK;
}

// Water.shader Domain (WS:322-327) + Fragment (WS:337-347) arithmetic for n world positions.
int OceanHandle::SampleSurface(const float* worldXZ, int n, float lod, float* disp3, float* normal3, float* foam, bool onDevice) {
    if (n == 0) return OCEAN_OK;
    if (lod > 0.0f && (normal3 || foam) && mipsFrame != frame) {  // trilinear reads need the current frame's chain
        int rc = GenerateMips();
        if (rc != OCEAN_OK) return rc;
    }
    const bool useMips = lod > 0.0f && mipsFrame == frame;
    float wl[OCEAN_MAX_CASCADES];
    for (int i = 0; i < C; ++i) wl[i] = cascades[i].wavelength;
    if (onDevice) {
        OCEAN_CUDA(launch_surface(disp, deriv, turb, useMips ? derivMips : nullptr, useMips ? turbMips : nullptr, N, C, wl,
                                  (const float2*)worldXZ, n, lod, disp3, normal3, foam, stream));
        return OCEAN_OK;
    }
    if (n > sCap) {
        if (sBuf) cudaFree(sBuf);
        sBuf = nullptr; sCap = 0;
        OCEAN_CUDA(cudaMalloc(&sBuf, (size_t)n * 9 * sizeof(float)));
        sCap = n;
    }
    float *dPts = sBuf, *dDisp = sBuf + (size_t)2 * n, *dNormal = dDisp + (size_t)3 * n, *dFoam = dNormal + (size_t)3 * n;
    OCEAN_CUDA(cudaMemcpyAsync(dPts, worldXZ, (size_t)n * sizeof(float2), cudaMemcpyHostToDevice, stream));
    OCEAN_CUDA(launch_surface(disp, deriv, turb, useMips ? derivMips : nullptr, useMips ? turbMips : nullptr, N, C, wl,
                              (const float2*)dPts, n, lod, disp3 ? dDisp : nullptr, normal3 ? dNormal : nullptr,
                              foam ? dFoam : nullptr, stream));
    if (disp3) OCEAN_CUDA(cudaMemcpyAsync(disp3, dDisp, (size_t)n * 3 * sizeof(float), cudaMemcpyDeviceToHost, stream));
    if (normal3) OCEAN_CUDA(cudaMemcpyAsync(normal3, dNormal, (size_t)n * 3 * sizeof(float), cudaMemcpyDeviceToHost, stream));
    if (foam) OCEAN_CUDA(cudaMemcpyAsync(foam, dFoam, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, stream));
    OCEAN_CUDA(cudaStreamSynchronize(stream));
    return OCEAN_OK;
}

// Has slot i's request completed?  wait: block until it has.
bool OceanHandle::SlotDone(int i, bool wait) {
    cudaError_t e = wait ? cudaEventSynchronize(rbDone[i]) : cudaEventQuery(rbDone[i]);
    if (e == cudaErrorNotReady) cudaGetLastError();
    return e == cudaSuccess;
}

void OceanHandle::PollReadbacks(bool wait) {
    for (int i = 0; i < RB; ++i) {
        if (!rbPending[i]) continue;
        if (SlotDone(i, wait)) {
            rbPending[i] = false;
            if (rbHasPlane[i] && (rbLatest < 0 || rbFrame[i] > rbFrame[rbLatest])) rbLatest = i;
        }
    }
}

#ifdef OCEAN_HOST_PROF
#include <chrono>
static double g_prof[8]; static long g_profN;
struct ProfT { int k; std::chrono::steady_clock::time_point t0; ProfT(int k) : k(k), t0(std::chrono::steady_clock::now()) {}
    ~ProfT() { g_prof[k] += std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count(); } };
#define PROF(k) ProfT prof_##k(k)
extern "C" void ocean_debug_host_prof() { for (int i = 0; i < 8; ++i) printf("prof[%d] = %.2f us/update\n", i, g_prof[i] / (g_profN ? g_profN : 1)); }
#else
#define PROF(k)
#endif
// A readback slot that is neither in flight nor the visible buoyancyData; blocks on the oldest request if none is free.
int OceanHandle::PickSlot(int* slot) {
    int s = -1;
    for (;;) {
        PollReadbacks(false);
        for (int i = 0; i < RB && s < 0; ++i)
            if (!rbPending[i] && i != rbLatest) s = i;
        if (s >= 0) break;
        int oldest = -1;
        for (int i = 0; i < RB; ++i)
            if (rbPending[i] && (oldest < 0 || rbFrame[i] < rbFrame[oldest])) oldest = i;
        if (!SlotDone(oldest, true)) return fail(OCEAN_ERR_CUDA, "readback of frame %lld did not complete", (long long)rbFrame[oldest]);
    }
    *slot = s;
    return OCEAN_OK;
}

// One frame + BuoyantObject.FixedUpdate (BO:18-30) for n boxes answered ON THE OWNING GPU: WaterBody.Update ships layer 0
// to the host every frame so that FixedUpdate can look heights up there (WB:288-296, BO:18); here the boxes travel to the
// device (24 B each), the lookup + force arithmetic runs against the frame's height plane, and 16 B per box travel back.
// Everything but the frame itself is on the copy stream, released by the layer-0 column counter like the HEIGHT readback,
// so the compute stream still carries nothing but the two frame kernels.
int OceanHandle::UpdateBuoyancy(float time, const OceanBuoyantBox* boxes, int n, float density, float gravityY, float* out4) {
    if (!spectrumReady) return fail(OCEAN_ERR_NOT_INITIALIZED, "ocean_update_buoyancy before ocean_init_spectrum");
    int s = -1;
    { int rc = PickSlot(&s); if (rc != OCEAN_OK) return rc; }
    if (n > rbBoxCap[s]) {
        if (rbBoxes[s]) cudaFree(rbBoxes[s]);
        if (rbAnswers[s]) cudaFree(rbAnswers[s]);
        rbBoxes[s] = nullptr; rbAnswers[s] = nullptr; rbBoxCap[s] = 0;
        OCEAN_CUDA(cudaMalloc(&rbBoxes[s], (size_t)n * sizeof(OceanBuoyantBox)));
        OCEAN_CUDA(cudaMalloc(&rbAnswers[s], (size_t)n * sizeof(float4)));
        rbBoxCap[s] = n;
    }
    float* plane = (float*)rbStage[s];
    if (n > 0) OCEAN_CUDA(cudaMemcpyAsync(rbBoxes[s], boxes, (size_t)n * sizeof(OceanBuoyantBox), cudaMemcpyHostToDevice, copyStream));
    if (stream_wait_value_available()) {
        int rc = CalculateWavesTexturesAtTime(time, 0, plane);
        if (rc != OCEAN_OK) return rc;
        const unsigned long long* colsDone0 = ws->counters + ws->layers;  // layer 0 of this ocean
        OCEAN_CUDA(stream_wait_value_geq(copyStream, colsDone0, (unsigned long long)N * ws->colPasses));
    } else {
        int rc = CalculateWavesTexturesAtTime(time);
        if (rc != OCEAN_OK) return rc;
        OCEAN_CUDA(launch_pack_height(disp, plane, (int)texels(), stream));
        OCEAN_CUDA(cudaEventRecord(rbSnap, stream));
        OCEAN_CUDA(cudaStreamWaitEvent(copyStream, rbSnap, 0));
    }
    if (n > 0) {
        OCEAN_CUDA(launch_buoyancy_plane(rbBoxes[s], plane, N, n, density, gravityY, rbAnswers[s], copyStream));
        OCEAN_CUDA(cudaMemcpyAsync(out4, rbAnswers[s], (size_t)n * sizeof(float4), cudaMemcpyDeviceToHost, copyStream));
    }
    OCEAN_CUDA(cudaEventRecord(rbDone[s], copyStream));
    rbSlotMode[s] = OCEAN_READBACK_HEIGHT;
    rbHasPlane[s] = false;
    rbFrame[s] = frame - 1;
    rbPending[s] = true;
    return OCEAN_OK;
}

int OceanHandle::Update(float time) {
#ifdef OCEAN_HOST_PROF
    ++g_profN;
#endif
    PROF(0);
    if (!spectrumReady) return fail(OCEAN_ERR_NOT_INITIALIZED, "ocean_update before ocean_init_spectrum");
    int s = -1;
    { PROF(2); int rc = PickSlot(&s); if (rc != OCEAN_OK) return rc; }
    if (rbMode == OCEAN_READBACK_HEIGHT && stream_wait_value_available()) {
        // Only the channel GetWaterHeight reads (WB:206-208): the column pass writes it into the slot's staging
        // plane; the copy stream starts as soon as the columns of layer 0 are done.
        int rc;
        { PROF(1); rc = CalculateWavesTexturesAtTime(time, 0, (float*)rbStage[s]); }
        if (rc != OCEAN_OK) return rc;
        const unsigned long long* colsDone0 = ws->counters + ws->layers;  // layer 0 of this ocean
        { PROF(4); OCEAN_CUDA(stream_wait_value_geq(copyStream, colsDone0, (unsigned long long)N * ws->colPasses)); }
        { PROF(5); OCEAN_CUDA(cudaMemcpyAsync(rbHost[s], rbStage[s], texels() * sizeof(float), cudaMemcpyDeviceToHost, copyStream)); }
        { PROF(6); OCEAN_CUDA(cudaEventRecord(rbDone[s], copyStream)); }
    } else {
        int rc = CalculateWavesTexturesAtTime(time);
        if (rc != OCEAN_OK) return rc;
        // snapshot layer 0 on the compute stream, ship it to the host on the copy stream (WB:288: the whole layer)
        size_t bytes = layerBytes4();
        if (rbMode == OCEAN_READBACK_HEIGHT) {
            bytes = texels() * sizeof(float);
            OCEAN_CUDA(launch_pack_height(disp, (float*)rbStage[s], (int)texels(), stream));
        } else {
            OCEAN_CUDA(cudaMemcpyAsync(rbStage[s], disp, bytes, cudaMemcpyDeviceToDevice, stream));
        }
        OCEAN_CUDA(cudaEventRecord(rbSnap, stream));
        OCEAN_CUDA(cudaStreamWaitEvent(copyStream, rbSnap, 0));
        OCEAN_CUDA(cudaMemcpyAsync(rbHost[s], rbStage[s], bytes, cudaMemcpyDeviceToHost, copyStream));
        OCEAN_CUDA(cudaEventRecord(rbDone[s], copyStream));
    }
    rbSlotMode[s] = rbMode;
    rbHasPlane[s] = true;
    rbFrame[s] = frame - 1;
    rbPending[s] = true;
    return OCEAN_OK;
}

int OceanHandle::GetWaterHeight(const float* worldXZ, int n, int mode, float* heights, bool ptsDev, bool outDev) {
    if (n == 0) return OCEAN_OK;
    if (mode == OCEAN_QUERY_READBACK_REF) {
        // WB:195-209 verbatim semantics: a host lookup into buoyancyData, the last completed
        // AsyncGPUReadback; returns 0 until one exists (WB:197).  Host buffers only.
        if (ptsDev || outDev) return fail(OCEAN_ERR_INVALID_ARGUMENT, "READBACK_REF works on host buffers");
        PollReadbacks(false);
        const float* data = rbLatest >= 0 ? rbHost[rbLatest] : nullptr;
        const bool packed = rbLatest >= 0 && rbSlotMode[rbLatest] == OCEAN_READBACK_HEIGHT;
        for (int i = 0; i < n; ++i) {
            if (!data) { heights[i] = 0.0f; continue; }
            const float lo = (float)(-(N / 2)), hi = (float)(N / 2);
            float u = (worldXZ[2 * i] - lo) / (hi - lo), v = (worldXZ[2 * i + 1] - lo) / (hi - lo);
            u = u < 0.0f ? 0.0f : (u > 1.0f ? 1.0f : u);
            v = v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);
            int x = (int)(u * (float)N), y = (int)(v * (float)N);
            x = x < 0 ? 0 : (x > N - 1 ? N - 1 : x);
            y = y < 0 ? 0 : (y > N - 1 ? N - 1 : y);
            heights[i] = packed ? data[(size_t)y * N + x] : data[((size_t)y * N + x) * 4 + 1];  // .g  WB:206-208
        }
        return OCEAN_OK;
    }
    if (n > qCap && (!ptsDev || !outDev)) {
        if (qPts) cudaFree(qPts);
        if (qOut) cudaFree(qOut);
        qPts = nullptr; qOut = nullptr; qCap = 0;
        OCEAN_CUDA(cudaMalloc(&qPts, (size_t)n * sizeof(float2)));
        OCEAN_CUDA(cudaMalloc(&qOut, (size_t)n * sizeof(float)));
        qCap = n;
    }
    const float2* pts = (const float2*)worldXZ;
    if (!ptsDev) {
        OCEAN_CUDA(cudaMemcpyAsync(qPts, worldXZ, (size_t)n * sizeof(float2), cudaMemcpyHostToDevice, stream));
        pts = qPts;
    }
    float* out = outDev ? heights : qOut;
    float wl[OCEAN_MAX_CASCADES];
    for (int i = 0; i < C; ++i) wl[i] = cascades[i].wavelength;
    if (mode == OCEAN_QUERY_DISPLACED) {
        OCEAN_CUDA(launch_displaced_query(disp, N, C, wl, pts, out, n, OCEAN_DISPLACED_ITERATIONS, stream));
    } else {
        OCEAN_CUDA(launch_query(disp, N, C, wl, pts, out, n, mode, stream));
    }
    if (!outDev) {
        OCEAN_CUDA(cudaMemcpyAsync(heights, qOut, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, stream));
        OCEAN_CUDA(cudaStreamSynchronize(stream));
    } else if (!ptsDev) {
        OCEAN_CUDA(cudaStreamSynchronize(stream));  // the caller may reuse its host points
    }
    return OCEAN_OK;
}

void OceanHandle::OnDisable() {
    if (stream) cudaStreamSynchronize(stream);
    if (copyStream) cudaStreamSynchronize(copyStream);
    workspace_release(ws);
    ws = nullptr;
    void* dev[] = {noise, h0, wave, waveHot, h0c, disp, deriv, turb, twiddle, stateDev, seedDev, rbStage[0], rbStage[1], rbStage[2], qPts, qOut, derivMips, turbMips, sBuf, bBuf, auxDone,
                   rbBoxes[0], rbBoxes[1], rbBoxes[2], rbAnswers[0], rbAnswers[1], rbAnswers[2]};
    for (void* p : dev)
        if (p) cudaFree(p);
    for (int i = 0; i < RB; ++i) {
        if (rbHost[i]) cudaFreeHost(rbHost[i]);
        if (rbDone[i]) cudaEventDestroy(rbDone[i]);
    }
    if (rbSnap) cudaEventDestroy(rbSnap);
    if (ownStream) cudaStreamDestroy(ownStream);
    if (copyStream) cudaStreamDestroy(copyStream);
}

// ---------------------------------------------------------------------------------------------
// Ensemble runner: waves of `batch` independent oceans through pooled buffers.
//
// Two lanes.  A wave is noise -> initial spectrum (ALU-bound: ~17 transcendentals per texel-cascade)
// -> row pass -> column pass (memory-bound).  Each lane owns a stream and a full set of pooled buffers;
// consecutive waves alternate between the lanes, so the ALU-bound kernels of one wave share the SMs with
// the memory-bound kernels of the other instead of running after them (OCEAN_ENSEMBLE_LANES=1: one lane).
#ifndef OCEAN_ENSEMBLE_ALU_CTAS_PER_SM_DEFAULT
#define OCEAN_ENSEMBLE_ALU_CTAS_PER_SM_DEFAULT 0
#endif
struct EnsembleLane {
    cudaStream_t stream = nullptr;
    float2* noise = nullptr;
    float4 *h0 = nullptr, *disp = nullptr, *deriv = nullptr, *turb = nullptr;
    float2* waveHot = nullptr;
    float2* h0c = nullptr;
    Workspace* ws = nullptr;
    ColsTmaMaps colsMaps;
    // parameter / result staging of the wave in flight on this lane
    StateConsts* stateDev = nullptr;
    uint64_t* seedDev = nullptr;
    unsigned long long* sumDev = nullptr;
    StateConsts* stateHost = nullptr;
    uint64_t* seedHost = nullptr;
    unsigned long long* sumHost = nullptr;
    cudaEvent_t done = nullptr;
};

struct OceanEnsemble {
    static constexpr int MAX_LANES = 2;
    int device = 0, N = 0, C = 0, batch = 0;
    int lanes = MAX_LANES;
    // OCEAN_ENSEMBLE_FUSED_INIT=1: initial spectrum fused into the row pass (ocean_rows_from_noise_kernel, SURVEY 8f
    // N4); h0 / waveHot are never materialised (40 B per texel-cascade less pooled memory).  Off by default: the
    // spectrum is ALU-bound and the separate kernels are ~9 % faster on B200 (ocean_init.cu).
    bool fusedInit = false;
    // grid cap of the ALU-bound noise / spectrum kernels (OCEAN_ENSEMBLE_ALU_CTAS_PER_SM x SMs; 0 = one CTA per work item)
    int aluCtas = 0;
    OceanCascade cascades[OCEAN_MAX_CASCADES]{};
    EnsembleLane lane[MAX_LANES];
    float2* twiddle = nullptr;
    int lastLane = 0;   // the lane that ran the last wave (ocean_ensemble_get_output)
    int lastCount = 0;

    size_t stateFloat4() const { return (size_t)C * N * N; }
    int create();
    int run(const OceanParams* states, const uint64_t* seeds, int nStates, float time, uint64_t* checksums);
    void destroy();
};

int OceanEnsemble::create() {
    const size_t t = (size_t)N * N;
    if (const char* e = getenv("OCEAN_ENSEMBLE_LANES")) lanes = atoi(e) <= 1 ? 1 : MAX_LANES;
    if (const char* e = getenv("OCEAN_ENSEMBLE_FUSED_INIT")) fusedInit = rows_from_noise_available(N) && atoi(e) != 0;
    {
        int perSM = OCEAN_ENSEMBLE_ALU_CTAS_PER_SM_DEFAULT, sms = 0;
        if (const char* e = getenv("OCEAN_ENSEMBLE_ALU_CTAS_PER_SM")) perSM = atoi(e);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        aluCtas = (lanes > 1 && perSM > 0 && sms > 0) ? perSM * sms : 0;
    }
    const std::vector<float2> tw = make_twiddles(N);
    OCEAN_CUDA(cudaMalloc(&twiddle, tw.size() * sizeof(float2)));
    OCEAN_CUDA(cudaMemcpy(twiddle, tw.data(), tw.size() * sizeof(float2), cudaMemcpyHostToDevice));
    OCEAN_CUDA(configure_frame_kernels(N));
    for (int l = 0; l < lanes; ++l) {
        EnsembleLane& L = lane[l];
        OCEAN_CUDA(cudaStreamCreateWithFlags(&L.stream, cudaStreamNonBlocking));
        OCEAN_CUDA(cudaMalloc(&L.noise, batch * t * sizeof(float2)));
        float4** arrs[] = {&L.disp, &L.deriv, &L.turb};
        for (float4** a : arrs) OCEAN_CUDA(cudaMalloc(a, (size_t)batch * C * t * sizeof(float4)));
        if (!fusedInit) OCEAN_CUDA(cudaMalloc(&L.h0, (size_t)batch * C * t * sizeof(float4)));
        L.ws = workspace_acquire(device, nullptr, N, batch * C);
        if (!L.ws) return fail(OCEAN_ERR_CUDA, "cudaMalloc of the intermediate planes failed");
        if (!fusedInit) OCEAN_CUDA(cudaMalloc(&L.waveHot, (size_t)batch * C * t * sizeof(float2)));
        if (!fusedInit && rows_need_h0c()) OCEAN_CUDA(cudaMalloc(&L.h0c, (size_t)batch * C * t * sizeof(float2)));
        OCEAN_CUDA(cudaMalloc(&L.stateDev, batch * sizeof(StateConsts)));
        OCEAN_CUDA(cudaMalloc(&L.seedDev, batch * sizeof(uint64_t)));
        OCEAN_CUDA(cudaMalloc(&L.sumDev, (size_t)batch * 3 * sizeof(unsigned long long)));
        OCEAN_CUDA(cudaMallocHost(&L.stateHost, batch * sizeof(StateConsts)));
        OCEAN_CUDA(cudaMallocHost(&L.seedHost, batch * sizeof(uint64_t)));
        OCEAN_CUDA(cudaMallocHost(&L.sumHost, (size_t)batch * 3 * sizeof(unsigned long long)));
        OCEAN_CUDA(cudaEventCreateWithFlags(&L.done, cudaEventDisableTiming));
        OCEAN_CUDA(make_cols_maps(N, batch * C, L.ws->planeA(), L.ws->planeB(), L.disp, L.deriv, L.turb, &L.colsMaps));
    }
    return OCEAN_OK;
}

int OceanEnsemble::run(const OceanParams* states, const uint64_t* seeds, int nStates, float time, uint64_t* checksums) {
    InitArgs ia{};
    ia.N = N; ia.C = C;
    ia.sc = make_spread_consts();
    for (int i = 0; i < C; ++i) ia.casc[i] = make_cascade_consts(cascades[i]);
    int waveIdx = 0;
    int pendingBase[MAX_LANES] = {-1, -1}, pendingCount[MAX_LANES] = {0, 0};
    auto drain = [&](int b) {
        if (pendingBase[b] < 0) return cudaSuccess;
        cudaError_t e = cudaEventSynchronize(lane[b].done);
        if (e != cudaSuccess) return e;
        if (checksums)
            memcpy(checksums + (size_t)pendingBase[b] * 3, lane[b].sumHost, (size_t)pendingCount[b] * 3 * sizeof(uint64_t));
        pendingBase[b] = -1;
        return cudaSuccess;
    };
    for (int base = 0; base < nStates; base += batch, ++waveIdx) {
        const int nb = nStates - base < batch ? nStates - base : batch;
        const int b = waveIdx % lanes;
        EnsembleLane& L = lane[b];
        cudaStream_t stream = L.stream;
        OCEAN_CUDA(drain(b));
        for (int i = 0; i < nb; ++i) {
            OceanParams p = states[base + i];
            p.textureSize = N;
            int rc = validate_params(p);
            if (rc != OCEAN_OK) return rc;
            // the frame kernels take lambda as one scalar per launch (= per wave): refuse to apply the first
            // state's value silently to the others
            if (p.lambda != states[base].lambda)
                return fail(OCEAN_ERR_INVALID_PARAMS, "state %d has lambda %g but its wave (states %d..%d) starts with %g: "
                            "one ensemble wave shares one lambda (group states by lambda, or run them with batch = 1)",
                            base + i, (double)p.lambda, base, base + nb - 1, (double)states[base].lambda);
            L.stateHost[i] = make_state_consts(p);
            L.seedHost[i] = seeds ? seeds[base + i] : (uint64_t)(base + i);
        }
        OCEAN_CUDA(cudaMemcpyAsync(L.stateDev, L.stateHost, nb * sizeof(StateConsts), cudaMemcpyHostToDevice, stream));
        OCEAN_CUDA(cudaMemcpyAsync(L.seedDev, L.seedHost, nb * sizeof(uint64_t), cudaMemcpyHostToDevice, stream));
        OCEAN_CUDA(cudaMemsetAsync(L.sumDev, 0, (size_t)nb * 3 * sizeof(unsigned long long), stream));
        OCEAN_CUDA(launch_noise(L.seedDev, nb, N, L.noise, stream, aluCtas));
        // waveData (the reference's inspection layout) is not kept by ensembles: the frame reads waveHot
        ia.noise = L.noise; ia.h0 = L.h0; ia.wave = nullptr; ia.waveHot = L.waveHot; ia.h0c = L.h0c;
        ia.nStates = nb; ia.states = L.stateDev;
        if (!fusedInit) OCEAN_CUDA(launch_init(ia, stream, aluCtas));
        FrameArgs fa{};
        fa.N = N; fa.layers = nb * C; fa.C = C;
        for (int i = 0; i < C; ++i) fa.dk[i] = ia.casc[i].dk;
        fa.h0 = L.h0; fa.waveHot = L.waveHot; fa.h0c = L.h0c; fa.rowsA = L.ws->planeA(); fa.rowsB = L.ws->planeB();
        fa.disp = L.disp; fa.deriv = L.deriv; fa.turb = L.turb;
        fa.twiddle = twiddle; fa.time = time;
        fa.lambda = states[base].lambda;  // one lambda per wave: the frame kernels take it as a scalar
        fa.foamDecay = expf(-2.0f);
        fa.zeroFoam = 1;
        fa.wsResident = fa.layers;  // (ensemble waves keep round 1's policy: every plane evict_last)
        // a short last wave leaves the upper layers' counters behind: restart the counting per wave
        L.ws->rowPasses = L.ws->colPasses = 0;
        L.ws->queueFrames = 0; L.ws->queueNext[0] = L.ws->queueNext[1] = 0;
        OCEAN_CUDA(cudaMemsetAsync(L.ws->counters, 0, (size_t)(2 * L.ws->layers + 2) * sizeof(unsigned long long), stream));
        workspace_stamp(L.ws, N, true, true, &fa);
        l2_window(device, L.ws, &fa);
        if (fusedInit) {
            fa.checksums = L.sumDev;  // reduced inside the column pass: the outputs are not read back
            OCEAN_CUDA(launch_rows_from_noise(ia, fa, stream));
            OCEAN_CUDA(launch_cols(fa, &L.colsMaps, stream));
        } else if (const int fusedMode = frame_fused_mode(N, &L.colsMaps)) {
            if (fusedMode == 2) {
                fa.states = L.stateDev;  // gravity per state
                workspace_stamp_queue(L.ws, 2 * nb * C * (N / 4), frame_fused_grid(N, nb * C, 2), &fa);
            }
            OCEAN_CUDA(launch_frame(fa, &L.colsMaps, stream, false, fusedMode));  // follows the init kernel: no early start
            OCEAN_CUDA(launch_checksum(L.disp, nb, stateFloat4(), L.sumDev + 0, 3, stream));
            OCEAN_CUDA(launch_checksum(L.deriv, nb, stateFloat4(), L.sumDev + 1, 3, stream));
            OCEAN_CUDA(launch_checksum(L.turb, nb, stateFloat4(), L.sumDev + 2, 3, stream));
        } else {
            fa.checksums = L.sumDev;  // reduced inside the column pass: the outputs are not read back
            OCEAN_CUDA(launch_rows(fa, stream, false));
            OCEAN_CUDA(launch_cols(fa, &L.colsMaps, stream));
        }
        OCEAN_CUDA(cudaMemcpyAsync(L.sumHost, L.sumDev, (size_t)nb * 3 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream));
        OCEAN_CUDA(cudaEventRecord(L.done, stream));
        pendingBase[b] = base; pendingCount[b] = nb;
        lastCount = nb;
        lastLane = b;
    }
    for (int b = 0; b < lanes; ++b) OCEAN_CUDA(drain(b));
    for (int b = 0; b < lanes; ++b) OCEAN_CUDA(cudaStreamSynchronize(lane[b].stream));
    return OCEAN_OK;
}

void OceanEnsemble::destroy() {
    for (int l = 0; l < MAX_LANES; ++l) {
        EnsembleLane& L = lane[l];
        if (L.stream) cudaStreamSynchronize(L.stream);
        workspace_release(L.ws);
        L.ws = nullptr;
        void* dev[] = {L.noise, L.h0, L.waveHot, L.h0c, L.disp, L.deriv, L.turb, L.stateDev, L.seedDev, L.sumDev};
        for (void* p : dev)
            if (p) cudaFree(p);
        if (L.stateHost) cudaFreeHost(L.stateHost);
        if (L.seedHost) cudaFreeHost(L.seedHost);
        if (L.sumHost) cudaFreeHost(L.sumHost);
        if (L.done) cudaEventDestroy(L.done);
        if (L.stream) cudaStreamDestroy(L.stream);
        L = EnsembleLane{};
    }
    if (twiddle) cudaFree(twiddle);
    twiddle = nullptr;
}

// ---------------------------------------------------------------------------------------------
// C ABI
extern "C" {

int ocean_abi_version(void) { return OCEAN_ABI_VERSION; }
const char* ocean_last_error(void) { return g_error.c_str(); }
uint64_t ocean_launch_count(void) { return launch_count(); }

int ocean_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int ocean_create(const OceanParams* params, const OceanCascade* cascades, int nCascades, const float* noiseRG,
                 uint64_t seed, int device, OceanHandle** out) {
    if (!out) return fail(OCEAN_ERR_INVALID_ARGUMENT, "out is NULL");
    *out = nullptr;
    if (!params) return fail(OCEAN_ERR_INVALID_ARGUMENT, "params is NULL");
    int rc = validate_params(*params);
    if (rc != OCEAN_OK) return rc;
    rc = validate_cascades(cascades, nCascades);
    if (rc != OCEAN_OK) return rc;
    const int nDev = ocean_device_count();
    if (nDev <= 0) return fail(OCEAN_ERR_CUDA, "no CUDA device available (this library has no CPU path)");
    if (device < 0 || device >= nDev) return fail(OCEAN_ERR_INVALID_ARGUMENT, "device %d outside [0, %d)", device, nDev);
    DeviceScope scope(device);
    if (!scope.ok) return fail(OCEAN_ERR_CUDA, "cudaSetDevice(%d) failed", device);
    OceanHandle* h = new OceanHandle;
    h->device = device;
    h->N = params->textureSize;
    h->C = nCascades;
    h->params = h->nextParams = *params;
    memcpy(h->cascades, cascades, nCascades * sizeof(OceanCascade));
    memcpy(h->nextCascades, cascades, nCascades * sizeof(OceanCascade));
    rc = h->Awake(noiseRG, seed);
    if (rc != OCEAN_OK) {
        h->OnDisable();
        delete h;
        return rc;
    }
    *out = h;
    return OCEAN_OK;
}

void ocean_destroy(OceanHandle* h) {
    if (!h) return;
    DeviceScope scope(h->device);
    h->OnDisable();
    delete h;
}

int ocean_set_params(OceanHandle* h, const OceanParams* params, const OceanCascade* cascades, int nCascades) {
    if (!h || !params) return fail(OCEAN_ERR_INVALID_ARGUMENT, "NULL argument");
    int rc = validate_params(*params);
    if (rc != OCEAN_OK) return rc;
    if (params->textureSize != h->N) return fail(OCEAN_ERR_INVALID_SIZE, "textureSize cannot change after create (%d -> %d)", h->N, params->textureSize);
    if (cascades) {
        if (nCascades != h->C) return fail(OCEAN_ERR_INVALID_CASCADES, "cascade count cannot change after create (%d -> %d)", h->C, nCascades);
        rc = validate_cascades(cascades, nCascades);
        if (rc != OCEAN_OK) return rc;
        memcpy(h->nextCascades, cascades, nCascades * sizeof(OceanCascade));
    }
    h->nextParams = *params;  // staged: takes effect at the next ocean_init_spectrum
    return OCEAN_OK;
}

int ocean_init_spectrum(OceanHandle* h) {
    if (!h) return fail(OCEAN_ERR_INVALID_ARGUMENT, "handle is NULL");
    DeviceScope scope(h->device);
    return h->CalculateInitialSpectrumTextures();
}

int ocean_step(OceanHandle* h, float time) {
    if (!h) return fail(OCEAN_ERR_INVALID_ARGUMENT, "handle is NULL");
    DeviceScope scope(h->device);
    return h->CalculateWavesTexturesAtTime(time);
}

int ocean_step_partial(OceanHandle* h, float time, int which) {
    if (!h) return fail(OCEAN_ERR_INVALID_ARGUMENT, "handle is NULL");
    if (which < 0 || which > 2) return fail(OCEAN_ERR_INVALID_ARGUMENT, "which must be 0, 1 or 2");
    DeviceScope scope(h->device);
    return h->CalculateWavesTexturesAtTime(time, which);
}

int ocean_update(OceanHandle* h, float time) {
    if (!h) return fail(OCEAN_ERR_INVALID_ARGUMENT, "handle is NULL");
    DeviceScope scope(h->device);
    return h->Update(time);
}

int ocean_update_buoyancy(OceanHandle* h, float time, const OceanBuoyantBox* boxes, int n, float density, float gravityY, float* out4) {
    if (!h || (n > 0 && (!boxes || !out4))) return fail(OCEAN_ERR_INVALID_ARGUMENT, "NULL argument");
    if (n < 0) return fail(OCEAN_ERR_INVALID_ARGUMENT, "n < 0");
    DeviceScope scope(h->device);
    return h->UpdateBuoyancy(time, boxes, n, density, gravityY, out4);
}

int ocean_sync(OceanHandle* h) {
    if (!h) return fail(OCEAN_ERR_INVALID_ARGUMENT, "handle is NULL");
    DeviceScope scope(h->device);
    OCEAN_CUDA(cudaStreamSynchronize(h->stream));
    OCEAN_CUDA(cudaStreamSynchronize(h->copyStream));
    h->PollReadbacks(false);
    return OCEAN_OK;
}

int ocean_readback_mode_of_latest(OceanHandle* h) { return h && h->rbLatest >= 0 ? h->rbSlotMode[h->rbLatest] : -1; }

const float* ocean_readback_ptr(OceanHandle* h, int wait, int64_t* frame) {
    if (frame) *frame = -1;
    if (!h) return nullptr;
    DeviceScope scope(h->device);
    h->PollReadbacks(wait != 0);
    if (h->rbLatest < 0) return nullptr;
    if (frame) *frame = h->rbFrame[h->rbLatest];
    return h->rbHost[h->rbLatest];
}

int ocean_set_readback_mode(OceanHandle* h, int mode) {
    if (!h) return fail(OCEAN_ERR_INVALID_ARGUMENT, "handle is NULL");
    if (mode != OCEAN_READBACK_RGBA && mode != OCEAN_READBACK_HEIGHT) return fail(OCEAN_ERR_INVALID_ARGUMENT, "unknown readback mode %d", mode);
    h->rbMode = mode;
    return OCEAN_OK;
}

static int pick_array(OceanHandle* h, int which, void** p, size_t* bytes) {
    const size_t t = h->texels();
    switch (which) {
        case OCEAN_DISPLACEMENT: *p = h->disp; *bytes = h->C * t * sizeof(float4); break;
        case OCEAN_DERIVATIVES: *p = h->deriv; *bytes = h->C * t * sizeof(float4); break;
        case OCEAN_TURBULENCE: *p = h->turb; *bytes = h->C * t * sizeof(float4); break;
        case OCEAN_INITIAL_SPECTRUM: *p = h->h0; *bytes = h->C * t * sizeof(float4); break;
        case OCEAN_WAVES_DATA: *p = h->wave; *bytes = h->C * t * sizeof(float4); break;
        case OCEAN_NOISE: *p = h->noise; *bytes = t * sizeof(float2); break;
        case OCEAN_ROW_SPECTRA: *p = h->rowsA; *bytes = h->C * t * sizeof(float4); break;  // plane 0; plane 1 via copy
        default: return fail(OCEAN_ERR_INVALID_ARGUMENT, "unknown array id %d", which);
    }
    return OCEAN_OK;
}

int ocean_get_output(OceanHandle* h, int which, void** devPtr, size_t* bytes) {
    if (!h || !devPtr || !bytes) return fail(OCEAN_ERR_INVALID_ARGUMENT, "NULL argument");
    return pick_array(h, which, devPtr, bytes);
}

int ocean_copy_output(OceanHandle* h, int which, int cascade, float* host) {
    if (!h || !host) return fail(OCEAN_ERR_INVALID_ARGUMENT, "NULL argument");
    DeviceScope scope(h->device);
    void* p; size_t bytes;
    int rc = pick_array(h, which, &p, &bytes);
    if (rc != OCEAN_OK) return rc;
    OCEAN_CUDA(cudaStreamSynchronize(h->stream));
    if (which == OCEAN_NOISE) {
        OCEAN_CUDA(cudaMemcpy(host, p, bytes, cudaMemcpyDeviceToHost));
        return OCEAN_OK;
    }
    if (cascade < -1 || cascade >= h->C) return fail(OCEAN_ERR_INVALID_ARGUMENT, "cascade %d outside [-1, %d)", cascade, h->C);
    const size_t layer = h->layerBytes4();
    const int planes = which == OCEAN_ROW_SPECTRA ? 2 : 1;
    for (int pl = 0; pl < planes; ++pl) {
        const char* src = (const char*)(pl == 0 ? p : (void*)h->rowsB);
        char* dst = (char*)host + (size_t)pl * (cascade < 0 ? bytes : layer);
        if (cascade < 0) OCEAN_CUDA(cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost));
        else OCEAN_CUDA(cudaMemcpy(dst, src + (size_t)cascade * layer, layer, cudaMemcpyDeviceToHost));
    }
    return OCEAN_OK;
}

int ocean_get_wavelengths(OceanHandle* h, float* out, int* nCascades) {
    if (!h || !out) return fail(OCEAN_ERR_INVALID_ARGUMENT, "NULL argument");
    for (int i = 0; i < h->C; ++i) out[i] = h->cascades[i].wavelength;
    if (nCascades) *nCascades = h->C;
    return OCEAN_OK;
}

int ocean_set_turbulence(OceanHandle* h, const float* host) {
    if (!h || !host) return fail(OCEAN_ERR_INVALID_ARGUMENT, "NULL argument");
    DeviceScope scope(h->device);
    OCEAN_CUDA(cudaStreamSynchronize(h->stream));
    OCEAN_CUDA(cudaMemcpy(h->turb, host, h->C * h->layerBytes4(), cudaMemcpyHostToDevice));
    h->foamZero = false;
    return OCEAN_OK;
}

int ocean_reset_foam(OceanHandle* h) {
    if (!h) return fail(OCEAN_ERR_INVALID_ARGUMENT, "handle is NULL");
    DeviceScope scope(h->device);
    OCEAN_CUDA(cudaMemsetAsync(h->turb, 0, h->C * h->layerBytes4(), h->stream));
    h->foamZero = true;
    h->frame = 0;
    // the frame index restarts, so the readback ring (ordered by frame index) restarts with it: without this
    // READBACK_REF / ocean_readback_ptr would keep answering from the newest pre-reset layer until the new
    // index passed the old one, and that slot would be excluded from reuse for as long
    OCEAN_CUDA(cudaStreamSynchronize(h->copyStream));
    for (int i = 0; i < OceanHandle::RB; ++i) { h->rbPending[i] = false; h->rbFrame[i] = -1; }
    h->rbLatest = -1;
    return OCEAN_OK;
}

int64_t ocean_frame_index(OceanHandle* h) { return h ? h->frame : -1; }

int ocean_query_heights(OceanHandle* h, const float* worldXZ, int n, int mode, float* heights, int pointsOnDevice,
                        int heightsOnDevice) {
    if (!h || (n > 0 && (!worldXZ || !heights))) return fail(OCEAN_ERR_INVALID_ARGUMENT, "NULL argument");
    if (n < 0) return fail(OCEAN_ERR_INVALID_ARGUMENT, "n < 0");
    if (mode < OCEAN_QUERY_NEAREST_REF || mode > OCEAN_QUERY_DISPLACED) return fail(OCEAN_ERR_INVALID_ARGUMENT, "unknown query mode %d", mode);
    DeviceScope scope(h->device);
    return h->GetWaterHeight(worldXZ, n, mode, heights, pointsOnDevice != 0, heightsOnDevice != 0);
}

int ocean_buoyancy(OceanHandle* h, const OceanBuoyantBox* boxes, int n, int mode, float density, float gravityY, float* out4,
                   int boxesOnDevice, int outOnDevice) {
    if (!h || (n > 0 && (!boxes || !out4))) return fail(OCEAN_ERR_INVALID_ARGUMENT, "NULL argument");
    if (n < 0) return fail(OCEAN_ERR_INVALID_ARGUMENT, "n < 0");
    if (mode < OCEAN_QUERY_NEAREST_REF || mode > OCEAN_QUERY_DISPLACED || mode == OCEAN_QUERY_READBACK_REF)
        return fail(OCEAN_ERR_INVALID_ARGUMENT, "ocean_buoyancy takes a device query mode (not %d; READBACK_REF is a host lookup)", mode);
    if (!h->spectrumReady) return fail(OCEAN_ERR_NOT_INITIALIZED, "ocean_buoyancy before ocean_init_spectrum");
    DeviceScope scope(h->device);
    return h->Buoyancy(boxes, n, mode, density, gravityY, out4, boxesOnDevice != 0, outOnDevice != 0);
}

int ocean_buoyancy_noalloc(OceanHandle* h, const OceanBuoyantBox* boxes, int n, int mode, float density, float gravityY, float* out4,
                           int boxesOnDevice, int outOnDevice) {
    if (!h || (n > 0 && (!boxes || !out4))) return fail(OCEAN_ERR_INVALID_ARGUMENT, "NULL argument");
    if (n < 0) return fail(OCEAN_ERR_INVALID_ARGUMENT, "n < 0");
    if (n > OCEAN_NOALLOC_MAX_BOXES)
        return fail(OCEAN_ERR_INVALID_ARGUMENT, "ocean_buoyancy_noalloc: n = %d exceeds OCEAN_NOALLOC_MAX_BOXES = %d (the fixed scratch of the "
                    "allocation-free path); split the batch or use ocean_buoyancy", n, OCEAN_NOALLOC_MAX_BOXES);
    if (mode < OCEAN_QUERY_NEAREST_REF || mode > OCEAN_QUERY_DISPLACED || mode == OCEAN_QUERY_READBACK_REF)
        return fail(OCEAN_ERR_INVALID_ARGUMENT, "ocean_buoyancy_noalloc takes a device query mode (not %d; READBACK_REF is a host lookup)", mode);
    if (!h->spectrumReady) return fail(OCEAN_ERR_NOT_INITIALIZED, "ocean_buoyancy_noalloc before ocean_init_spectrum");
    DeviceScope scope(h->device);
    return h->BuoyancyNoAlloc(boxes, n, mode, density, gravityY, out4, boxesOnDevice != 0, outOnDevice != 0);
}

int ocean_generate_mips(OceanHandle* h) {
    if (!h) return fail(OCEAN_ERR_INVALID_ARGUMENT, "handle is NULL");
    DeviceScope scope(h->device);
    return h->GenerateMips();
}

int ocean_set_auto_mips(OceanHandle* h, int enabled) {
    if (!h) return fail(OCEAN_ERR_INVALID_ARGUMENT, "handle is NULL");
    h->autoMips = enabled != 0;
    return OCEAN_OK;
}

int ocean_mip_levels(OceanHandle* h) {
    if (!h) return -1;
    int levels = 0;
    while ((h->N >> levels) > 1) ++levels;
    return levels;
}

static int mip_pick(OceanHandle* h, int which, float4** base, float4** mips) {
    if (which == OCEAN_DERIVATIVES) { *base = h->deriv; *mips = h->derivMips; return OCEAN_OK; }
    if (which == OCEAN_TURBULENCE) { *base = h->turb; *mips = h->turbMips; return OCEAN_OK; }
    return fail(OCEAN_ERR_INVALID_ARGUMENT, "array %d has no mip chain (only OCEAN_DERIVATIVES and OCEAN_TURBULENCE do, WB:191-192)", which);
}

int ocean_get_mips(OceanHandle* h, int which, void** devPtr, size_t* bytes) {
    if (!h || !devPtr || !bytes) return fail(OCEAN_ERR_INVALID_ARGUMENT, "NULL argument");
    float4 *base, *mips;
    int rc = mip_pick(h, which, &base, &mips);
    if (rc != OCEAN_OK) return rc;
    if (!mips) return fail(OCEAN_ERR_NOT_INITIALIZED, "no mip chain yet: call ocean_generate_mips first");
    *devPtr = mips;
    *bytes = (size_t)h->C * mip_texels_per_layer(h->N) * sizeof(float4);
    return OCEAN_OK;
}

int ocean_copy_mip(OceanHandle* h, int which, int level, int cascade, float* host) {
    if (!h || !host) return fail(OCEAN_ERR_INVALID_ARGUMENT, "NULL argument");
    if (cascade < 0 || cascade >= h->C) return fail(OCEAN_ERR_INVALID_ARGUMENT, "cascade %d outside [0, %d)", cascade, h->C);
    const int levels = ocean_mip_levels(h);
    if (level < 0 || level > levels) return fail(OCEAN_ERR_INVALID_ARGUMENT, "level %d outside [0, %d]", level, levels);
    float4 *base, *mips;
    int rc = mip_pick(h, which, &base, &mips);
    if (rc != OCEAN_OK) return rc;
    if (level > 0 && !mips) return fail(OCEAN_ERR_NOT_INITIALIZED, "no mip chain yet: call ocean_generate_mips first");
    DeviceScope scope(h->device);
    const size_t n = (size_t)(h->N >> level);
    const float4* src = level == 0 ? base + (size_t)cascade * h->texels()
                                   : mips + (size_t)cascade * mip_texels_per_layer(h->N) + mip_level_offset(h->N, level);
    OCEAN_CUDA(cudaStreamSynchronize(h->stream));
    OCEAN_CUDA(cudaMemcpy(host, src, n * n * sizeof(float4), cudaMemcpyDeviceToHost));
    return OCEAN_OK;
}

int ocean_sample_surface(OceanHandle* h, const float* worldXZ, int n, float lod, float* displacement, float* normal, float* foam,
                         int onDevice) {
    if (!h || (n > 0 && !worldXZ)) return fail(OCEAN_ERR_INVALID_ARGUMENT, "NULL argument");
    if (n < 0) return fail(OCEAN_ERR_INVALID_ARGUMENT, "n < 0");
    if (!(lod >= 0.0f)) return fail(OCEAN_ERR_INVALID_ARGUMENT, "lod must be >= 0");
    if (n > 0 && !displacement && !normal && !foam) return fail(OCEAN_ERR_INVALID_ARGUMENT, "no output requested");
    DeviceScope scope(h->device);
    return h->SampleSurface(worldXZ, n, lod, displacement, normal, foam, onDevice != 0);
}

int ocean_get_stream(OceanHandle* h, void** s) {
    if (!h || !s) return fail(OCEAN_ERR_INVALID_ARGUMENT, "NULL argument");
    *s = (void*)h->stream;
    return OCEAN_OK;
}

int ocean_get_copy_stream(OceanHandle* h, void** s) {
    if (!h || !s) return fail(OCEAN_ERR_INVALID_ARGUMENT, "NULL argument");
    *s = (void*)h->copyStream;
    return OCEAN_OK;
}

int ocean_set_stream(OceanHandle* h, void* s) {
    if (!h) return fail(OCEAN_ERR_INVALID_ARGUMENT, "handle is NULL");
    DeviceScope scope(h->device);
    OCEAN_CUDA(cudaStreamSynchronize(h->stream));
    OCEAN_CUDA(cudaStreamSynchronize(h->copyStream));  // a pending readback waits on the old workspace's counters
    h->stream = s ? (cudaStream_t)s : h->ownStream;
    return h->AttachWorkspace(s ? (cudaStream_t)s : nullptr);
}

int ocean_describe(OceanHandle* h, char* buf, size_t bufBytes) {
    if (!h || !buf || bufBytes == 0) return fail(OCEAN_ERR_INVALID_ARGUMENT, "NULL argument");
    snprintf(buf, bufBytes, "oceanfft sm_100a: N=%d C=%d device=%d; %s", h->N, h->C, h->device, describe_frame_kernels(h->N));
    return OCEAN_OK;
}

int ocean_ensemble_create(int textureSize, const OceanCascade* cascades, int nCascades, int batch, int device, OceanEnsemble** out) {
    if (!out) return fail(OCEAN_ERR_INVALID_ARGUMENT, "out is NULL");
    *out = nullptr;
    const int N = textureSize;
    if (N < OCEAN_MIN_SIZE || N > OCEAN_MAX_SIZE || (N & (N - 1)) != 0)
        return fail(OCEAN_ERR_INVALID_SIZE, "textureSize %d is not a power of two in [%d, %d]", N, OCEAN_MIN_SIZE, OCEAN_MAX_SIZE);
    int rc = validate_cascades(cascades, nCascades);
    if (rc != OCEAN_OK) return rc;
    if (batch < 1 || batch > 4096) return fail(OCEAN_ERR_INVALID_ARGUMENT, "batch %d outside [1, 4096]", batch);
    const int nDev = ocean_device_count();
    if (nDev <= 0) return fail(OCEAN_ERR_CUDA, "no CUDA device available (this library has no CPU path)");
    if (device < 0 || device >= nDev) return fail(OCEAN_ERR_INVALID_ARGUMENT, "device %d outside [0, %d)", device, nDev);
    DeviceScope scope(device);
    OceanEnsemble* e = new OceanEnsemble;
    e->device = device; e->N = N; e->C = nCascades; e->batch = batch;
    memcpy(e->cascades, cascades, nCascades * sizeof(OceanCascade));
    rc = e->create();
    if (rc != OCEAN_OK) {
        e->destroy();
        delete e;
        return rc;
    }
    *out = e;
    return OCEAN_OK;
}

void ocean_ensemble_destroy(OceanEnsemble* e) {
    if (!e) return;
    DeviceScope scope(e->device);
    e->destroy();
    delete e;
}

int ocean_ensemble_run(OceanEnsemble* e, const OceanParams* states, const uint64_t* seeds, int nStates, float time, uint64_t* checksums) {
    if (!e || !states) return fail(OCEAN_ERR_INVALID_ARGUMENT, "NULL argument");
    if (nStates < 0) return fail(OCEAN_ERR_INVALID_ARGUMENT, "nStates < 0");
    DeviceScope scope(e->device);
    return e->run(states, seeds, nStates, time, checksums);
}

static int ens_pick(OceanEnsemble* e, int which, void** p, size_t* bytes) {
    const size_t all = (size_t)e->batch * e->stateFloat4() * sizeof(float4);
    const EnsembleLane& L = e->lane[e->lastLane];
    switch (which) {
        case OCEAN_DISPLACEMENT: *p = L.disp; break;
        case OCEAN_DERIVATIVES: *p = L.deriv; break;
        case OCEAN_TURBULENCE: *p = L.turb; break;
        case OCEAN_INITIAL_SPECTRUM:
            if (!L.h0) return fail(OCEAN_ERR_INVALID_ARGUMENT, "this ensemble fuses the spectrum into the row pass (OCEAN_ENSEMBLE_FUSED_INIT=1) and keeps no h0");
            *p = L.h0;
            break;
        case OCEAN_WAVES_DATA: return fail(OCEAN_ERR_INVALID_ARGUMENT, "ensembles do not keep waveData (inspect it on a stand-alone ocean)");
        default: return fail(OCEAN_ERR_INVALID_ARGUMENT, "unknown array id %d", which);
    }
    *bytes = all;
    return OCEAN_OK;
}

int ocean_ensemble_get_output(OceanEnsemble* e, int which, void** devPtr, size_t* bytes) {
    if (!e || !devPtr || !bytes) return fail(OCEAN_ERR_INVALID_ARGUMENT, "NULL argument");
    return ens_pick(e, which, devPtr, bytes);
}

int ocean_ensemble_copy_output(OceanEnsemble* e, int which, int slot, float* host) {
    if (!e || !host) return fail(OCEAN_ERR_INVALID_ARGUMENT, "NULL argument");
    if (slot < 0 || slot >= e->batch) return fail(OCEAN_ERR_INVALID_ARGUMENT, "slot %d outside [0, %d)", slot, e->batch);
    DeviceScope scope(e->device);
    void* p; size_t bytes;
    int rc = ens_pick(e, which, &p, &bytes);
    if (rc != OCEAN_OK) return rc;
    const size_t per = e->stateFloat4() * sizeof(float4);
    OCEAN_CUDA(cudaStreamSynchronize(e->lane[e->lastLane].stream));
    OCEAN_CUDA(cudaMemcpy(host, (const char*)p + (size_t)slot * per, per, cudaMemcpyDeviceToHost));
    return OCEAN_OK;
}

}  // extern "C"
