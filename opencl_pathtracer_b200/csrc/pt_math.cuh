// pt_math.cuh — the canonical arithmetic of the path (PARITY mode follows it operation for operation).
//
// The reference kernel is OpenCL C, whose dot/length/normalize on float4 act on all four lanes
// (SURVEY.md F8). The association order below is the one the oracle defines
// (oracle/shim/ocl_as_cxx.h): dot4 = ((x*x' + y*y') + z*z') + w*w', normalize = v / sqrt(dot4(v,v)).
// Translation units that must be bit-exact are compiled with -fmad=false; divisions and square roots
// are IEEE (nvcc default -prec-div=true -prec-sqrt=true, no --use_fast_math).
#pragma once

#include <cuda_runtime.h>

namespace ptc
{
__device__ __forceinline__ float4 f4(float x, float y, float z, float w) { return make_float4(x, y, z, w); }
__device__ __forceinline__ float4 operator+(float4 a, float4 b) { return f4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ float4 operator-(float4 a, float4 b) { return f4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
__device__ __forceinline__ float4 operator*(float4 a, float4 b) { return f4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
__device__ __forceinline__ float4 operator*(float4 a, float s) { return f4(a.x * s, a.y * s, a.z * s, a.w * s); }
__device__ __forceinline__ float4 operator*(float s, float4 a) { return f4(s * a.x, s * a.y, s * a.z, s * a.w); }
__device__ __forceinline__ float4 operator/(float4 a, float s) { return f4(a.x / s, a.y / s, a.z / s, a.w / s); }
__device__ __forceinline__ float4 operator-(float4 a) { return f4(-a.x, -a.y, -a.z, -a.w); }

__device__ __forceinline__ float dot4(float4 a, float4 b) { return ((a.x * b.x + a.y * b.y) + a.z * b.z) + a.w * b.w; }
__device__ __forceinline__ float length4(float4 a) { return sqrtf(dot4(a, a)); }
// a / l where l is a length (a square root: >= 0 or NaN). A zero numerator sends the IEEE division to its slow path (FCHK
// fails, a subroutine call), and the w lane of every direction is 0, as are lanes of the tangent frame. +-0 / l is +-0 for
// every l > 0 and NaN otherwise; that case is answered directly and the division itself never sees the zero (a select on a
// plain `a / l` would still evaluate it). Bit-identical results for every finite outcome.
__device__ __forceinline__ float div_by_length(float a, float l)
{
    const bool z = a == 0.f;
    const float t = (z ? 1.f : a) / l;
    return z ? (l > 0.f ? a : __int_as_float(0x7fffffff)) : t;
}
__device__ __forceinline__ float4 normalize4(float4 a)
{
    float l = length4(a);
    return f4(div_by_length(a.x, l), div_by_length(a.y, l), div_by_length(a.z, l), div_by_length(a.w, l));
}
__device__ __forceinline__ float4 cross3(float4 a, float4 b)
{
    return f4(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x, 0.f);
}
// 256-bit read-only global load (sm_100a: LDG.E.256): one L1 request per lane for half a 64-byte record instead of two.
// p must be 32-byte aligned.
__device__ __forceinline__ void ldg256(const void *p, float4 &a, float4 &b)
{
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
                 : "l"(p));
}

// OpenCL max(x, y) = (x < y) ? y : x  — NOT fmaxf (differs on NaN)
__device__ __forceinline__ float clmax(float x, float y) { return x < y ? y : x; }

// reference Kernel/PathTracer_FullKernel_header.cl:246-253: seed = (16807 * seed) & 0x7FFFFFFF in 64-bit,
// u = (float)seed / (float)0x7FFFFFFF; (float)0x7FFFFFFF rounds to 2^31. Only the low 31 bits of the
// product survive the mask, so 32-bit wrapping multiplication gives the same state.
__device__ __forceinline__ float rng_next(int &seed)
{
    seed = (int)((16807u * (unsigned)seed) & 0x7FFFFFFFu);
    return (float)seed * 4.656612873077392578125e-10f;  // = / 2^31 exactly (a power of two; the division's range check would fail)
}

// reference header.cl:255-264 (int arithmetic wraps; only the low 32 bits matter)
__device__ __forceinline__ int rng_init(unsigned x, unsigned y, unsigned iteration, unsigned W, unsigned H)
{
    unsigned s = x + y * W + iteration * W * H;
    s *= 2011u;
    s *= s;
    if (s == 0) s = 1;
    return (int)s;
}

// reference header.cl:237-244
__device__ __forceinline__ void put_in_same_hemisphere(float4 &v, float4 n)
{
    float dp = dot4(v, n);
    if (dp < 0.001f) v = v + n * (0.01f - dp);
}

}  // namespace ptc
