/*
 * oceanfft.h -- C ABI of liboceanfft.so, the B200 (sm_100a) native plugin that replaces the
 * spectral-ocean compute path of Mozobo/Ocean-Simulation behind WaterBody's surface.
 *
 * The reference has no FFI of its own: the path sits behind the C# class WaterBody and Unity's
 * ComputeShader API.  Each entry point below therefore cites the WaterBody / IFFT member it
 * stands in for (paths relative to the reference root):
 *   WB      = Assets/Scripts/Water/WaterBody.cs
 *   WC      = Assets/Scripts/Water/WaterCascade.cs
 *   IFFT.cs = Assets/Scripts/Water/IFFT.cs
 *   BO      = Assets/Scripts/Water/BuoyantObject.cs
 *   IS/TD/FFT/RF = Assets/Shaders/Compute Shaders/Water/{InitialSpectrum,TimeDependentSpectrum,
 *                  IFFT,ResultTexturesFiller}.compute
 * INTEGRATION.md shows the C# [DllImport] shim (WaterBodyNative.cs) that binds them.
 *
 * Conventions
 *   - plain C, POD structs, raw pointers and sizes; no CUDA or torch types cross the boundary
 *     (device pointers and streams travel as void*).
 *   - every function returns an OceanStatus (0 = OK) unless documented otherwise; the message
 *     for the last failure on the calling thread is ocean_last_error().  Nothing throws.
 *   - the library owns all device memory; callers own the host buffers they pass in.
 *   - one CUDA stream per handle: calls on one handle must be externally serialised and take
 *     effect in call order (Unity's in-order graphics queue, WB:180-193); different handles may
 *     be driven from different threads.
 *   - there is NO CPU fallback: without a CUDA device ocean_create fails with OCEAN_ERR_CUDA.
 *
 * Array layout (all outputs and all state): the reference's Texture2DArray layout flattened,
 *   index = ((layer * N + y) * N + x) * channels, x fastest   (WB:206, IS:120 `uint3(id.xy, i)`).
 */
#ifndef OCEANFFT_H
#define OCEANFFT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OCEAN_ABI_VERSION 1
#define OCEAN_MAX_CASCADES 5 /* Water.shader:139 sizes its per-cascade arrays at 5 */
#define OCEAN_MIN_SIZE 16
#define OCEAN_MAX_SIZE 2048

typedef enum OceanStatus {
    OCEAN_OK = 0,
    OCEAN_ERR_INVALID_ARGUMENT = 1, /* NULL pointer, bad enum, bad cascade index ... */
    OCEAN_ERR_INVALID_SIZE = 2,     /* textureSize not a power of two in [16, 2048] */
    OCEAN_ERR_INVALID_CASCADES = 3, /* nCascades outside [1, 5], wavelength <= 0, cutoffLow <= 0 ... */
    OCEAN_ERR_INVALID_PARAMS = 4,   /* non-positive windSpeed / gravity / fetch / depth, zero wind vector */
    OCEAN_ERR_CUDA = 5,             /* CUDA runtime failure (no device, out of memory, launch error) */
    OCEAN_ERR_NOT_INITIALIZED = 6   /* ocean_step before ocean_init_spectrum */
} OceanStatus;

/* WaterBody's serialized ocean fields WB:10-14 + texturesSize WB:29 (uniforms WB:131-145).
 * lambda is the horizontal-displacement (choppiness) scale.  The reference has none (it is
 * hard-wired to 1: Water.shader:325-327, RF:27); parity is defined at lambda = 1. */
typedef struct OceanParams {
    float windSpeed; /* U10, m/s            WB:10 */
    float windDirX;  /* unnormalised         WB:11 */
    float windDirY;
    float gravity;   /* m/s^2               WB:12 */
    float fetch;     /* m                   WB:13 */
    float depth;     /* m                   WB:14 */
    int32_t textureSize; /* N               WB:29 */
    float lambda;    /* 1.0 = reference */
} OceanParams;

/* WaterCascade WC:10-24 (flattened into 4 GPU buffers by WB:231-251). */
typedef struct OceanCascade {
    float wavelength; /* tile length L_i, m; dk = 2*PI/L_i   WC:10 */
    float cutoffLow;  /* |k| >= cutoffLow                    WC:20 */
    float cutoffHigh; /* |k| <= cutoffHigh                   WC:15 */
    float swell;      /* WC:22 */
    float fade;       /* WC:24 */
} OceanCascade;

typedef enum OceanArray {
    OCEAN_DISPLACEMENT = 0, /* RGBA32F (Dx, Dy, Dz, 0)                  _DisplacementsTextures WB:278, RF:25 */
    OCEAN_DERIVATIVES = 1,  /* RGBA32F (dDy/dx, dDy/dz, dDx/dx, dDz/dz) _DerivativesTextures   WB:279, RF:26 */
    OCEAN_TURBULENCE = 2,   /* RGBA32F (foam x4)                        _TurbulenceTextures    WB:280, RF:28-32 */
    OCEAN_INITIAL_SPECTRUM = 3, /* RGBA32F (h0(k), conj h0(-k))  initialSpectrumTextures WB:37, IS:120,141 */
    OCEAN_WAVES_DATA = 4,       /* RGBA32F (kx, 1/|k|, kz, omega) wavesDataTextures      WB:38, IS:123 */
    OCEAN_NOISE = 5,            /* RG32F, one layer, shared by all cascades      randomNoiseTexture WB:34 */
    OCEAN_ROW_SPECTRA = 6       /* debug: state between the x and y passes, 2 planes of RGBA32F, each [layer][y][x]:
                                   plane 0 = spectra (DxDz, DyDxz), plane 1 = (DyxDyz, DxxDzz), a texel holding
                                   (re of first, re of second, im of first, im of second) */
} OceanArray;

typedef enum OceanQueryMode {
    OCEAN_QUERY_NEAREST_REF = 0,          /* WaterBody.GetWaterHeight WB:195-209, bit-exact, on the current frame */
    OCEAN_QUERY_READBACK_REF = 1,         /* same lookup, but in the last completed readback: 0 until one exists (WB:197) */
    OCEAN_QUERY_BILINEAR = 2,             /* extension: cascade 0, uv = world/wavelength[0] (Water.shader:325), bilinear, Repeat wrap */
    OCEAN_QUERY_BILINEAR_ALL_CASCADES = 3, /* extension: sum over cascades, as the surface shader displaces vertices */
    OCEAN_QUERY_DISPLACED = 4             /* extension: height of the surface ABOVE (x, z).  Water.shader:327 moves a vertex
                                             horizontally too, so the point over (x, z) started at q with q + Dxz(q) = (x, z);
                                             solved by OCEAN_DISPLACED_ITERATIONS fixed-point steps q <- (x, z) - Dxz(q), all
                                             cascades, bilinear; returns Dy(q) */
} OceanQueryMode;
#define OCEAN_DISPLACED_ITERATIONS 4

typedef struct OceanHandle OceanHandle;     /* one WaterBody: C cascades of N x N on one device */
typedef struct OceanEnsemble OceanEnsemble; /* pooled batch runner for many independent oceans */

int ocean_abi_version(void);
const char* ocean_last_error(void);
int ocean_device_count(void); /* number of CUDA devices, 0 on failure */

/* ---- lifetime: WaterBody.Awake WB:211-256 / OnDisable WB:300-309 ------------------------ */

/* Allocates every array of WB:221-229 + IFFT.cs:36,51 on `device`, uploads or generates the
 * noise texture (WB:86-100).  noiseRG: N*N*2 floats [y][x][2] on the host, or NULL to generate
 * it on the device from `seed` (include/ocean_noise.h).  Does not compute the spectrum. */
int ocean_create(const OceanParams* params, const OceanCascade* cascades, int nCascades,
                 const float* noiseRG, uint64_t seed, int device, OceanHandle** out);
void ocean_destroy(OceanHandle* h);

/* Live parameter edit (the commented-out OnValidate WB:324-337).  textureSize must not change;
 * cascades may be NULL to keep the current ones, otherwise nCascades must equal the handle's.
 * The edit is staged: frames, queries and samplers keep using the parameter set the spectrum on the device
 * was computed from (wavelengths, dk, lambda) until the next ocean_init_spectrum commits it. */
int ocean_set_params(OceanHandle* h, const OceanParams* params, const OceanCascade* cascades, int nCascades);

/* CalculateInitialSpectrumTextures WB:171-178: IS:99-129 then IS:135-143, one fused kernel. */
int ocean_init_spectrum(OceanHandle* h);

/* ---- per frame: WaterBody.CalculateWavesTexturesAtTime(float time) WB:180-193 ----------- */

/* Enqueues one frame (TD:20-47 -> 4 x IFFT.cs:66-94 -> RF:16-34) on the handle's stream and
 * advances the foam recursion by exactly one frame.  Asynchronous. */
int ocean_step(OceanHandle* h, float time);
int ocean_sync(OceanHandle* h); /* block until everything enqueued on the handle has finished */
/* Measurement hook: enqueue only one half of a frame so each kernel can be bracketed by events.
 * which = 1: row pass (TD:20-47 + FFT:48-57), which = 2: column pass + assembly (FFT:60-78 + RF:16-34,
 * advances foam and the frame index), which = 0: both (== ocean_step). */
int ocean_step_partial(OceanHandle* h, float time, int which);

/* WaterBody.Update WB:284-297: ocean_step(time) + AsyncGPUReadback.Request of displacement
 * layer 0 (WB:288) into library-owned pinned memory on a second stream.  The readback becomes
 * visible to OCEAN_QUERY_READBACK_REF / ocean_readback_ptr once it has completed. */
int ocean_update(OceanHandle* h, float time);
/* What ocean_update ships to the host.  RGBA = the whole displacement layer 0, N*N*16 B, exactly the
 * reference's AsyncGPUReadback request (WB:288).  HEIGHT = only the channel GetWaterHeight consumes
 * (`buoyancyData[index].g`, WB:206-208), packed on the device: N*N*4 B over PCIe.  Takes effect at the
 * next ocean_update; OCEAN_QUERY_READBACK_REF answers identically in both modes. */
typedef enum OceanReadbackMode { OCEAN_READBACK_RGBA = 0, OCEAN_READBACK_HEIGHT = 1 } OceanReadbackMode;
int ocean_set_readback_mode(OceanHandle* h, int mode /*OceanReadbackMode*/);
int ocean_readback_mode_of_latest(OceanHandle* h); /* mode the latest completed readback was made in, -1 if none */
/* Latest completed readback or NULL: N*N*4 floats (RGBA of displacement layer 0) in RGBA mode, N*N
 * floats (heights) in HEIGHT mode; *frame gets the index of the frame it belongs to (-1 if none).
 * wait != 0 blocks for the newest request. */
const float* ocean_readback_ptr(OceanHandle* h, int wait, int64_t* frame);

/* ---- outputs: what WB:277-281 publishes to Water.shader ------------------------------------ */

/* Device pointer + byte size of an array (owned by the handle, valid until ocean_destroy). */
int ocean_get_output(OceanHandle* h, int which /*OceanArray*/, void** devPtr, size_t* bytes);
/* Synchronous copy of one layer (cascade) to host memory; cascade = -1 copies all layers. */
int ocean_copy_output(OceanHandle* h, int which /*OceanArray*/, int cascade, float* host);
int ocean_get_wavelengths(OceanHandle* h, float* out /*[nCascades]*/, int* nCascades); /* _Wavelengths WB:281 */

/* ---- consumers of the arrays: mip chain WB:191-192, Water.shader Domain/Fragment arithmetic ---- */

/* derivativesTextures.GenerateMips(); turbulenceTextures.GenerateMips() (WB:191-192) for the current frame.
 * The displacement array has no chain (WB:227).  Levels 1..log2(N), 2x2 box filter (Unity's filter is not
 * part of the reference tree; the box filter is this library's definition).  Off the frame's hot path:
 * explicit, or every frame after ocean_set_auto_mips(h, 1), or on demand by ocean_sample_surface(lod > 0). */
int ocean_generate_mips(OceanHandle* h);
int ocean_set_auto_mips(OceanHandle* h, int enabled);
int ocean_mip_levels(OceanHandle* h); /* log2(N) */
/* Device pointer of a chain: [layer][levels 1..log2(N) packed one after another], (N*N-1)/3 texels per layer. */
int ocean_get_mips(OceanHandle* h, int which /*OCEAN_DERIVATIVES | OCEAN_TURBULENCE*/, void** devPtr, size_t* bytes);
/* Synchronous copy of one level of one layer, (N>>level)^2 * 4 floats; level 0 is the array itself. */
int ocean_copy_mip(OceanHandle* h, int which, int level, int cascade, float* host);

/* What Water.shader derives from the arrays at n world positions (x, z), uv_i = worldXZ / wavelength_i,
 * sampler = Repeat + trilinear (WB:104-128), `lod` as the shader's lodLevel:
 *   displacement[3n] = sum_i disp_i(uv_i)                                   Domain   WS:322-327
 *   normal[3n]       = normalize(-sx, 1, -sz), (sx, sz) = (d.x/(1+d.z), d.y/(1+d.w)),
 *                      d = sum_i deriv_i(uv_i, lod)                         Fragment WS:337-347
 *   foam[n]          = sum_i (1 - saturate(turb_i(uv_i, lod).x))            Fragment WS:343 (compared with
 *                      _FoamThreshold at WS:371)
 * Any output pointer may be NULL.  onDevice != 0: every pointer is a device pointer and the call is
 * asynchronous on the handle's stream; otherwise host buffers and the call returns when they are filled. */
int ocean_sample_surface(OceanHandle* h, const float* worldXZ, int n, float lod, float* displacement, float* normal,
                         float* foam, int onDevice);

/* ---- buoyancy: BuoyantObject.FixedUpdate BO:18-30 for a batch of boxes ----------- */

/* transform.position and transform.localScale of one BuoyantObject */
typedef struct OceanBuoyantBox {
    float px, py, pz;
    float sx, sy, sz;
} OceanBuoyantBox;

/* For each of the n boxes, in the reference's fp32 operation order:
 *   waterHeight     = GetWaterHeight(position)            BO:18  (any OceanQueryMode but READBACK_REF; NEAREST_REF is
 *                                                                 the reference's lookup on the current frame)
 *   heightSubmerged = max(0, waterHeight - py)             BO:19
 *   submergedVolume = clamp(heightSubmerged / sy, 0, 1) * (sx * sy * sz)      BO:24-25 (0 when not submerged, BO:21)
 *   acceleration    = |gravityY| * (density * submergedVolume)                 BO:26-29, the upward ForceMode.Acceleration
 * out4[i] = (waterHeight, heightSubmerged, acceleration, submergedVolume).  The drag terms BO:31-33 need the rigid body's
 * velocities and stay with the caller (apply them when heightSubmerged > 0).  density = WaterBody.density (WB:18). */
int ocean_buoyancy(OceanHandle* h, const OceanBuoyantBox* boxes, int n, int mode /*OceanQueryMode*/, float density,
                   float gravityY, float* out4, int boxesOnDevice, int outOnDevice);

/* ocean_buoyancy without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, same kernels and bit-identical results as ocean_buoyancy, but the intermediate arrays of the chain (query
 * points, heights, and the staging of boxes / answers when they live in host memory) sit in a file-scope
 * `static __device__ __align__(256)` array of the library instead of cudaMalloc'ed scratch: the call makes no cudaMalloc,
 * cudaFree or allocating library call.  Limits and rules:
 *   - n <= OCEAN_NOALLOC_MAX_BOXES; a larger n is rejected with OCEAN_ERR_INVALID_ARGUMENT (split the batch);
 *   - the array is 52 B x OCEAN_NOALLOC_MAX_BOXES = 3.25 MiB and every GPU that loads the library holds it until exit;
 *   - every handle on a device shares the array and it keeps the previous call's contents; each call overwrites every slot
 *     it later reads, so nothing needs clearing, but two calls that OVERLAP ON THE GPU (handles on different streams) race:
 *     serialise them (one stream, or ocean_sync between them);
 *   - with host buffers the call returns when out4 is filled; with device pointers it is asynchronous on the handle's
 *     stream, exactly as ocean_buoyancy. */
#define OCEAN_NOALLOC_MAX_BOXES 65536
int ocean_buoyancy_noalloc(OceanHandle* h, const OceanBuoyantBox* boxes, int n, int mode /*OceanQueryMode*/, float density,
                           float gravityY, float* out4, int boxesOnDevice, int outOnDevice);

/* WaterBody.Update (WB:284-297) + BuoyantObject.FixedUpdate (BO:18-30) with the queries answered on the owning GPU:
 * one frame at `time`, then -- on the handle's copy stream, released as soon as the columns of layer 0 are done -- the n
 * boxes are copied to the device, looked up in that frame's height plane exactly as OCEAN_QUERY_NEAREST_REF does
 * (WB:195-209) and run through BO:19-29; out4 receives n x (waterHeight, heightSubmerged, acceleration, submergedVolume).
 * Instead of the reference's N*N*16 B layer per frame, 24 B per box travel in and 16 B per box travel out.
 * Fully asynchronous: `boxes` and `out4` are host buffers (pinned for real overlap) that must stay untouched until
 * ocean_sync(h) -- or until three later ocean_update / ocean_update_buoyancy calls on h have been issued, the depth of
 * the readback ring.  The frame's layer is NOT shipped: OCEAN_QUERY_READBACK_REF keeps answering from the last
 * ocean_update. */
int ocean_update_buoyancy(OceanHandle* h, float time, const OceanBuoyantBox* boxes, int n, float density, float gravityY,
                          float* out4);

/* ---- foam state: the persistence implicit in the turbulence RenderTexture RF:28,32 --------- */
int ocean_set_turbulence(OceanHandle* h, const float* host /* C*N*N*4 */);
int ocean_reset_foam(OceanHandle* h);
int64_t ocean_frame_index(OceanHandle* h); /* frames stepped since create / reset */

/* ---- queries: WaterBody.GetWaterHeight(Vector3) WB:195-209, called from BO:18 -------------- */

/* n points (x, z) -> n heights.  pointsOnDevice / heightsOnDevice select host or device buffers. */
int ocean_query_heights(OceanHandle* h, const float* worldXZ, int n, int mode /*OceanQueryMode*/,
                        float* heights, int pointsOnDevice, int heightsOnDevice);

/* ---- plumbing ------------------------------------------------------------------------------ */
int ocean_get_stream(OceanHandle* h, void** cudaStream);
int ocean_set_stream(OceanHandle* h, void* cudaStream); /* share one stream between handles (NULL = own) */
/* The second stream of the handle (cudaStream_t as void*): readbacks and ocean_update_buoyancy answers complete on it. */
int ocean_get_copy_stream(OceanHandle* h, void** cudaStream);
/* Kernel launches issued by this library on the calling thread's behalf since load (bench bookkeeping). */
uint64_t ocean_launch_count(void);
/* Names of the most recent launch configuration, for DESIGN.md / bench: writes a short text. */
int ocean_describe(OceanHandle* h, char* buf, size_t bufBytes);

/* ---- ensembles: many independent oceans (BASELINE config 5) -------------------------------- */

/* A pooled runner for `batch` oceans of (N, nCascades) at a time on one device.  Every state
 * shares the cascade set; ocean parameters and the noise seed vary per state -- except OceanParams::lambda,
 * which the frame kernels take as one scalar per wave of `batch` states: ocean_ensemble_run returns
 * OCEAN_ERR_INVALID_PARAMS for a wave whose states do not share one lambda. */
int ocean_ensemble_create(int textureSize, const OceanCascade* cascades, int nCascades, int batch,
                          int device, OceanEnsemble** out);
void ocean_ensemble_destroy(OceanEnsemble* e);
/* For each state s in [0, nStates): noise(seed[s]) -> init spectrum -> one frame at `time` from
 * zero turbulence -> 64-bit checksum of the three output arrays (sum of the fp32 bit patterns).
 * Processes the states in waves of `batch`.  checksums: nStates x 3 uint64 on the host.
 * Asynchronous work is drained before returning. */
int ocean_ensemble_run(OceanEnsemble* e, const OceanParams* states, const uint64_t* seeds, int nStates,
                       float time, uint64_t* checksums);
/* Device pointers of the pooled arrays of the last wave ([batch*C][N][N][4]): the three outputs and
 * OCEAN_INITIAL_SPECTRUM.  Waves alternate between two sets of pooled buffers (two lanes, so that the
 * ALU-bound spectrum kernels of one wave overlap the memory-bound frame of the other); the pointers are
 * those of the lane that ran the last wave.  OCEAN_WAVES_DATA is not kept by ensembles, and with
 * OCEAN_ENSEMBLE_FUSED_INIT=1 in the environment at create time (initial spectrum fused into the row pass: no
 * pooled h0 / waveHot buffers, 40 B per texel-cascade less memory, ~9 % slower on B200) neither is
 * OCEAN_INITIAL_SPECTRUM. */
int ocean_ensemble_get_output(OceanEnsemble* e, int which, void** devPtr, size_t* bytes);
int ocean_ensemble_copy_output(OceanEnsemble* e, int which, int slot, float* host /* C*N*N*4 */);

#ifdef __cplusplus
}
#endif
#endif /* OCEANFFT_H */
