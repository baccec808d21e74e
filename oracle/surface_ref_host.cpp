/*
 * surface_ref_host.cpp -- "Oracle A" for the consumers of the output arrays.  TEST INFRASTRUCTURE ONLY.
 *
 * oracle/Makefile (target `ref`) streams, straight from /root/reference through sed into this translation unit
 * (-include'd as generated text on stdin; no reference source is written into the repo):
 *     SURFACE_DOMAIN_LINES    = Assets/Shaders/Water.shader:322-326   (Domain: displacement summed over cascades)
 *     SURFACE_FRAGMENT_LINES  = Assets/Shaders/Water.shader:337-347   (Fragment: derivatives / turbulence sums, slope, normalOS)
 *     WATERBODY_GETWATERHEIGHT = Assets/Scripts/Water/WaterBody.cs:195-209 (GetWaterHeight, C#)
 *     BUOYANT_FIXEDUPDATE     = Assets/Scripts/Water/BuoyantObject.cs:17-37 (FixedUpdate, C#)
 * and this file supplies what those lines lean on: the HLSL vector types (hlsl_compat.h + the float3 / float4
 * operators below), Unity's Mathf / Vector3 / Rigidbody / Physics / Time surface, and the texture sampler.
 *
 * The SAMPLER is a restatement (it is Unity / driver behaviour, not reference text): Repeat wrap + trilinear as
 * WaterBody.cs:104-128 configures it, texel centres at (i + 0.5) / n, bilinear inside a level, linear between
 * levels -- the same definition oracle/surface_oracle.py::sample_lod documents, operation for operation.  The mip
 * levels themselves are inputs (Unity's GenerateMips filter is not in the reference tree).
 * normalize(float3) is v * (1 / sqrt((x*x + y*y) + z*z)), the rsqrt-multiply form of the HLSL intrinsic.
 *
 * tests/test_surface_oracle.py asserts that oracle/surface_oracle.py equals this library bit for bit: that pins
 * the numpy restatement the CUDA kernels are compared with to the reference's own lines.
 */
#include "hlsl_compat.h"

extern "C" float floorf(float);

namespace hlsl {
inline float3 operator*(const float3& a, float s) { return float3(a.x * s, a.y * s, a.z * s); }
inline float3 operator+(const float3& a, const float3& b) { return float3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline float4 operator+(const float4& a, const float4& b) { return float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline float4 operator-(const float4& a, const float4& b) { return float4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
inline float4 operator*(const float4& a, float s) { return float4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline float4& operator+=(float4& a, const float4& b) { a = a + b; return a; }
inline float3& operator+=(float3& a, const float4& b) { a = float3(a.x + b.x, a.y + b.y, a.z + b.z); return a; }  // float4 -> float3 truncation
inline float saturate(float x) { return x < 0.0f ? 0.0f : (x > 1.0f ? 1.0f : x); }
inline float3 normalize(const float3& v) { return v * (1.0f / ::sqrtf((v.x * v.x + v.y * v.y) + v.z * v.z)); }
}  // namespace hlsl
using namespace hlsl;

// ---- the sampler (restatement; see the header comment) --------------------------------------------------------
struct TextureArray {
    const float4* level0 = nullptr;  // [layer][N][N]
    const float4* mips = nullptr;    // [layer][level 1 | level 2 | ... ] packed, (N*N-1)/3 texels per layer; nullptr: no chain
    int N = 0;
};
static float4 lerp4(const float4& a, const float4& b, float t) { return a + (b - a) * t; }
static float4 sample_bilinear(const float4* img, int n, float u, float v) {
    const float fx = u * (float)n - 0.5f, fy = v * (float)n - 0.5f;
    const float x0f = ::floorf(fx), y0f = ::floorf(fy);
    const float wx = fx - x0f, wy = fy - y0f;
    const int x0 = (int)x0f & (n - 1), y0 = (int)y0f & (n - 1);
    const int x1 = (x0 + 1) & (n - 1), y1 = (y0 + 1) & (n - 1);
    const float4 top = lerp4(img[(long)y0 * n + x0], img[(long)y0 * n + x1], wx);
    const float4 bot = lerp4(img[(long)y1 * n + x0], img[(long)y1 * n + x1], wx);
    return lerp4(top, bot, wy);
}
static const float4* level_ptr(const TextureArray& t, int layer, int level) {
    if (level == 0) return t.level0 + (long)layer * t.N * t.N;
    const long m = t.N >> (level - 1);
    return t.mips + (long)layer * (((long)t.N * t.N - 1) / 3) + ((long)t.N * t.N - m * m) / 3;
}
static float4 sample_lod(const TextureArray& t, float2 uv, int layer, float lod) {
    const float u = uv.x - ::floorf(uv.x), v = uv.y - ::floorf(uv.y);
    int top = 0;
    if (t.mips) while ((t.N >> top) > 1) ++top;
    if (top == 0 || !(lod > 0.0f)) return sample_bilinear(level_ptr(t, layer, 0), t.N, u, v);
    if (lod > (float)top) lod = (float)top;
    const int l0 = (int)lod;
    const float f = lod - (float)l0;
    const float4 s0 = sample_bilinear(level_ptr(t, layer, l0), t.N >> l0, u, v);
    if (f == 0.0f || l0 >= top) return s0;
    return lerp4(s0, sample_bilinear(level_ptr(t, layer, l0 + 1), t.N >> (l0 + 1), u, v), f);
}
struct SamplerState {};
#define SAMPLE_TEXTURE2D_ARRAY_LOD(tex, samp, uv, layer, lod) sample_lod(tex, uv, layer, lod)

// ---- Water.shader uniforms (WS:132-139) and the interpolants the streamed lines touch ---------------------------
static int _NbCascades;
static TextureArray _DisplacementsTextures, _DerivativesTextures, _TurbulenceTextures;
static SamplerState sampler_DisplacementsTextures, sampler_DerivativesTextures, sampler_TurbulenceTextures;
static float _Wavelengths[5];
struct Domain2FragmentData {
    float3 positionWS;
    float2 worldUV;
    float lodLevel;
};

static float3 domain_ref(Domain2FragmentData& output) {
    SURFACE_DOMAIN_LINES
    return displacement;
}
static void fragment_ref(const Domain2FragmentData& input, float3& normalOut, float& turbulenceOut) {
    SURFACE_FRAGMENT_LINES
    normalOut = normalOS;
    turbulenceOut = turbulence;
}

// ---- Unity's C# surface for the two streamed methods ------------------------------------------------------------
struct Vector3 {
    float x = 0, y = 0, z = 0;
    Vector3() = default;
    Vector3(float a, float b, float c) : x(a), y(b), z(c) {}
};
static Vector3 operator-(const Vector3& a) { return Vector3(-a.x, -a.y, -a.z); }
static Vector3 operator*(const Vector3& a, float s) { return Vector3(a.x * s, a.y * s, a.z * s); }
struct Color { float r, g, b, a; };
static struct MathfT {
    float Max(float a, float b) const { return a > b ? a : b; }
    float Abs(float a) const { return ::fabsf(a); }
    float Clamp(float v, float lo, float hi) const { return v < lo ? lo : (v > hi ? hi : v); }
    int Clamp(int v, int lo, int hi) const { return v < lo ? lo : (v > hi ? hi : v); }
    float InverseLerp(float a, float b, float v) const {  // Unity: a != b ? Clamp01((v - a) / (b - a)) : 0
        return a != b ? Clamp((v - a) / (b - a), 0.0f, 1.0f) : 0.0f;
    }
} Mathf;
static struct { Vector3 gravity; } Physics;
static struct { float fixedDeltaTime = 0.02f; } Time;
static struct { int Acceleration = 0, VelocityChange = 1; } ForceMode;
static Color* const null = nullptr;

struct TextureSize { int width = 0, height = 0; };
struct WaterBodyRef {
    int texturesSize = 0;
    TextureSize displacementsTextures;
    const Color* buoyancyData = nullptr;
    float density = 1.0f, drag = 0.0f, angularDrag = 0.0f;
    float lastHeight = 0.0f;
    WATERBODY_GETWATERHEIGHT
};
struct RigidbodyRef {
    Vector3 velocity, angularVelocity;
    Vector3 acceleration;  // sum of the ForceMode.Acceleration forces of this FixedUpdate, gravity excluded
    int calls = 0;
    void AddForce(const Vector3& f, int mode) {
        if (mode == 0 && calls == 0) acceleration = f;  // the first Acceleration force is the buoyant one (BO:29)
        ++calls;
    }
    void AddTorque(const Vector3&, int) {}
};
struct BuoyantObjectRef {
    struct { Vector3 position, localScale; } transform;
    WaterBodyRef waterBody;
    RigidbodyRef rb;
    BUOYANT_FIXEDUPDATE
};

extern "C" {

// pts [n][2] world (x, z) -> disp3 [n][3]                                         Water.shader:322-326
void surface_ref_domain(const float* disp, int N, int C, const float* wavelengths, const float* pts, int n, float lod, float* disp3) {
    _NbCascades = C;
    _DisplacementsTextures = TextureArray{(const float4*)disp, nullptr, N};  // the displacement array has no chain (WB:227)
    for (int i = 0; i < C; ++i) _Wavelengths[i] = wavelengths[i];
    for (int i = 0; i < n; ++i) {
        Domain2FragmentData d;
        d.positionWS = float3(pts[2 * i], 0.0f, pts[2 * i + 1]);
        d.worldUV = float2(pts[2 * i], pts[2 * i + 1]);  // output.positionWS.xz  WS:317
        d.lodLevel = lod;
        const float3 r = domain_ref(d);
        disp3[3 * i] = r.x; disp3[3 * i + 1] = r.y; disp3[3 * i + 2] = r.z;
    }
}

// pts -> normal3 [n][3] (normalOS) and foam [n] (the `turbulence` sum compared with _FoamThreshold)   Water.shader:337-347
void surface_ref_fragment(const float* deriv, const float* derivMips, const float* turb, const float* turbMips, int N, int C,
                          const float* wavelengths, const float* pts, int n, float lod, float* normal3, float* foam) {
    _NbCascades = C;
    _DerivativesTextures = TextureArray{(const float4*)deriv, (const float4*)derivMips, N};
    _TurbulenceTextures = TextureArray{(const float4*)turb, (const float4*)turbMips, N};
    for (int i = 0; i < C; ++i) _Wavelengths[i] = wavelengths[i];
    for (int i = 0; i < n; ++i) {
        Domain2FragmentData d;
        d.worldUV = float2(pts[2 * i], pts[2 * i + 1]);
        d.lodLevel = lod;
        float3 nrm;
        float t;
        fragment_ref(d, nrm, t);
        normal3[3 * i] = nrm.x; normal3[3 * i + 1] = nrm.y; normal3[3 * i + 2] = nrm.z;
        foam[i] = t;
    }
}

// WaterBody.GetWaterHeight WB:195-209 on a read-back layer (layer0 = N*N RGBA texels, or NULL = no readback yet)
void surface_ref_get_water_height(const float* layer0, int N, const float* pts, int n, float* heights) {
    WaterBodyRef wb;
    wb.texturesSize = N;
    wb.displacementsTextures = TextureSize{N, N};
    wb.buoyancyData = (const Color*)layer0;
    for (int i = 0; i < n; ++i) heights[i] = wb.GetWaterHeight(Vector3(pts[2 * i], 0.0f, pts[2 * i + 1]));
}

// BuoyantObject.FixedUpdate BO:17-37 for n boxes (position xyz, localScale xyz): out2 [n][2] = (waterHeight it looked
// up, y of the buoyant ForceMode.Acceleration force it applied; 0 when it applied none)
void surface_ref_buoyancy(const float* layer0, int N, const float* boxes, int n, float density, float gravityY, float* out2) {
    Physics.gravity = Vector3(0.0f, gravityY, 0.0f);
    for (int i = 0; i < n; ++i) {
        BuoyantObjectRef bo;
        bo.waterBody.texturesSize = N;
        bo.waterBody.displacementsTextures = TextureSize{N, N};
        bo.waterBody.buoyancyData = (const Color*)layer0;
        bo.waterBody.density = density;
        const float* b = boxes + 6 * i;
        bo.transform.position = Vector3(b[0], b[1], b[2]);
        bo.transform.localScale = Vector3(b[3], b[4], b[5]);
        bo.FixedUpdate();
        out2[2 * i] = bo.waterBody.GetWaterHeight(bo.transform.position);
        out2[2 * i + 1] = bo.rb.calls > 1 ? bo.rb.acceleration.y : 0.0f;  // > 1: the submerged branch ran before the gravity force
    }
}

}  // extern "C"
