/*
 * hlsl_compat.h -- just enough HLSL for the reference's four water compute shaders to compile
 * as C++ ("Oracle A").  TEST INFRASTRUCTURE ONLY.
 *
 * oracle/Makefile, target `ref` streams each .compute file from /root/reference through sed
 * (strip [numthreads]/semantics, add the `f` suffix HLSL implies on float literals) straight
 * into g++; no reference source is copied into the repo.  This header supplies the vector
 * types with swizzles, texture/buffer objects with operator[], and fp32 intrinsics.
 *
 * No <cmath>/<cstdlib> here on purpose: they would put ::pow(double,double), ::abs(int)...
 * beside the fp32 overloads below and make HLSL calls such as log(2) or pow(x, 5) ambiguous
 * or silently double precision.
 */
#pragma once

extern "C" {
float sqrtf(float); float powf(float, float); float expf(float); float logf(float);
float tanhf(float); float coshf(float); float atan2f(float, float); float cosf(float);
float sinf(float); float fabsf(float);
}

typedef unsigned int uint;

namespace hlsl {

struct float2; struct uint2;

struct uint2 {
    union { struct { uint x, y; }; struct { uint r, g; }; };
    uint2() = default;
    uint2(uint a, uint b) : x(a), y(b) {}
    explicit uint2(const float2& f);
};

struct float2 {
    union { struct { float x, y; }; struct { float r, g; }; };
    float2() = default;
    float2(float a) : x(a), y(a) {}
    float2(float a, float b) : x(a), y(b) {}
};
inline uint2::uint2(const float2& f) : x((uint)f.x), y((uint)f.y) {}

// Swizzle views (.xy .zw .rg .ba): GCC forbids members with constructors inside anonymous
// structs, so the views are constructor-less PODs that convert to the real vector type.
struct float2_view { float x, y; operator float2() const { return float2(x, y); } };
struct uint2_view { uint x, y; operator uint2() const { return uint2(x, y); } };

struct float3 {
    union { struct { float x, y, z; }; struct { float r, g, b; }; struct { float2_view xy; float _z; }; };
    float3() = default;
    float3(float a) : x(a), y(a), z(a) {}
    float3(float a, float b, float c) : x(a), y(b), z(c) {}
};

struct float4 {
    union {
        struct { float x, y, z, w; };
        struct { float r, g, b, a; };
        struct { float2_view xy; float2_view zw; };
        struct { float2_view rg; float2_view ba; };
    };
    float4() = default;
    float4(float v) : x(v), y(v), z(v), w(v) {}
    float4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
    float4(const float2& p, float c, float d) : x(p.x), y(p.y), z(c), w(d) {}
    float4(const float2& p, const float2& q) : x(p.x), y(p.y), z(q.x), w(q.y) {}
};

struct uint3 {
    union { struct { uint x, y, z; }; struct { uint2_view xy; uint _z; }; };
    uint3() = default;
    uint3(uint a, uint b, uint c) : x(a), y(b), z(c) {}
    uint3(const uint2& p, uint c) : x(p.x), y(p.y), z(c) {}
};

// float2 arithmetic (component-wise, scalar broadcast)
inline float2 operator+(const float2& a, const float2& b) { return float2(a.x + b.x, a.y + b.y); }
inline float2 operator-(const float2& a, const float2& b) { return float2(a.x - b.x, a.y - b.y); }
inline float2 operator-(const float2& a) { return float2(-a.x, -a.y); }
inline float2 operator*(const float2& a, float s) { return float2(a.x * s, a.y * s); }
inline float2 operator*(float s, const float2& a) { return float2(s * a.x, s * a.y); }
inline float2 operator*(const float2& a, uint s) { return a * (float)s; }
inline float2 operator*(const float2& a, const float2& b) { return float2(a.x * b.x, a.y * b.y); }
inline float2 operator/(const float2& a, float s) { return float2(a.x / s, a.y / s); }
inline float2 operator/(const float2& a, uint s) { return a / (float)s; }

// fp32 intrinsics.  Integer arguments (log(2), pow(x, 5), min(v, 20)) convert to float as in HLSL.
inline float sqrt(float x) { return ::sqrtf(x); }
inline float pow(float x, float y) { return ::powf(x, y); }
inline float exp(float x) { return ::expf(x); }
inline float log(float x) { return ::logf(x); }
inline float log(int x) { return ::logf((float)x); }
inline float tanh(float x) { return ::tanhf(x); }
inline float cosh(float x) { return ::coshf(x); }
inline float atan2(float y, float x) { return ::atan2f(y, x); }
inline float cos(float x) { return ::cosf(x); }
inline float sin(float x) { return ::sinf(x); }
inline float abs(float x) { return ::fabsf(x); }
inline float min(float a, float b) { return a < b ? a : b; }
inline float length(const float2& v) { return ::sqrtf(v.x * v.x + v.y * v.y); }
inline float2 normalize(const float2& v) { return v / length(v); }

// Resource objects: a view (pointer + extents) over host arrays laid out x-fastest.
template <class T> struct Texture2D {
    T* data = nullptr; uint width = 0, height = 0;
    T& operator[](const uint2& p) const { return data[(unsigned long)p.y * width + p.x]; }
};
template <class T> using RWTexture2D = Texture2D<T>;
template <class T> struct Texture2DArray {
    T* data = nullptr; uint width = 0, height = 0, depth = 0;
    T& operator[](const uint3& p) const { return data[((unsigned long)p.z * height + p.y) * width + p.x]; }
};
template <class T> using RWTexture2DArray = Texture2DArray<T>;
template <class T> struct StructuredBuffer {
    const T* data = nullptr;
    const T& operator[](uint i) const { return data[i]; }
};

}  // namespace hlsl
