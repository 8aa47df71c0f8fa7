// TEST INFRASTRUCTURE ONLY — part of the parity oracle, never linked into the product.
//
// ocl_shim_types.h — OpenCL C vector/scalar types as plain C++ structs, global scope.
// Lane-wise operators follow OpenCL C semantics (every lane, including .w).
#ifndef OCL_SHIM_TYPES_DEFINED
#define OCL_SHIM_TYPES_DEFINED

#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdio>

typedef unsigned int uint;
typedef unsigned long ulong;  // 64-bit on LP64, as OpenCL's ulong
typedef unsigned char uchar;

struct float2
{
    float x, y;
    float2() = default;
    float2(float _x, float _y) : x(_x), y(_y) {}
    float2 &operator+=(float2 b) { x += b.x; y += b.y; return *this; }
    float2 &operator+=(float s) { x += s; y += s; return *this; }
    float2 &operator-=(float s) { x -= s; y -= s; return *this; }
    float2 &operator*=(float s) { x *= s; y *= s; return *this; }
    float2 &operator/=(float s) { x /= s; y /= s; return *this; }
    float2 &operator/=(float2 b) { x /= b.x; y /= b.y; return *this; }
};
static inline float2 operator+(float2 a, float2 b) { return float2{a.x + b.x, a.y + b.y}; }
static inline float2 operator*(float2 a, float s) { return float2{a.x * s, a.y * s}; }

struct int2 { int x, y; int2() = default; int2(int _x, int _y) : x(_x), y(_y) {} };
struct int4 { int x, y, z, w; int4() = default; int4(int a, int b, int c, int d) : x(a), y(b), z(c), w(d) {} };
struct uchar4 { uchar x, y, z, w; };

struct alignas(16) float4
{
    float x, y, z, w;
    float4() = default;
    float4(float _x, float _y, float _z, float _w) : x(_x), y(_y), z(_z), w(_w) {}
    float4 &operator=(float s) { x = y = z = w = s; return *this; }  // OpenCL scalar widening
    float4 &operator+=(float4 b) { x += b.x; y += b.y; z += b.z; w += b.w; return *this; }
    float4 &operator-=(float4 b) { x -= b.x; y -= b.y; z -= b.z; w -= b.w; return *this; }
    float4 &operator*=(float4 b) { x *= b.x; y *= b.y; z *= b.z; w *= b.w; return *this; }
    float4 &operator*=(float s) { x *= s; y *= s; z *= s; w *= s; return *this; }
    float4 &operator/=(float s) { x /= s; y /= s; z /= s; w /= s; return *this; }
};
static inline float4 operator+(float4 a, float4 b) { return float4{a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
static inline float4 operator-(float4 a, float4 b) { return float4{a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w}; }
static inline float4 operator*(float4 a, float4 b) { return float4{a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w}; }
static inline float4 operator/(float4 a, float4 b) { return float4{a.x / b.x, a.y / b.y, a.z / b.z, a.w / b.w}; }
static inline float4 operator*(float4 a, float s) { return float4{a.x * s, a.y * s, a.z * s, a.w * s}; }
static inline float4 operator*(float s, float4 a) { return float4{s * a.x, s * a.y, s * a.z, s * a.w}; }
static inline float4 operator/(float4 a, float s) { return float4{a.x / s, a.y / s, a.z / s, a.w / s}; }
static inline float4 operator-(float4 a) { return float4{-a.x, -a.y, -a.z, -a.w}; }

typedef float4 RGBAColor_t;

namespace ocl_shim
{
extern thread_local size_t gid[2];
}

#define CLK_LOCAL_MEM_FENCE 1

#endif
