// pt_shade.cuh — per-hit shading arithmetic of the shade stage (compiled with -fmad=false).
//
// Reference functions restated here (Kernel/PathTracer_FullKernel.cl unless noted):
//   texture_lookup            Texture_GetPixelColorValue  header.cl:430-459
//   sky_color                 Sky_GetColorValue / Sky_GetFaceColorValue :438-512
//   fresnel, fresnel_varnish  Material_Fresnel*ReflectionFraction :192-292
//   material_brdf             Material_BRDF :166-190
//   light_power_toward        Light_PowerToward header.cl:403-421
//   concentric_sample_disk    Material_ConcentricSampleDisk :339-416
//   cosine_sample_hemisphere  Material_CosineSampleHemisphere :303-336
#pragma once

#include "pt_device.cuh"
#include "pt_math.cuh"

namespace ptc
{
constexpr float kPi = 3.14159265f;  // PATH_PI
constexpr float kPiInv = 0.31830988618f;

// x / 255.f of convert_float4(uchar4) / 255 (header.cl:455). The IEEE division with a zero numerator is a subroutine call, not the
// inline sequence — and a black texel is common (every sky face without a sky, `loadSky=false`, is the texel (0,0,0,0): ncu v14 showed
// 4 such calls per missed ray, 3.8 % of k_shade's instructions). A zero lane divides 1 instead and answers +0, which is what 0 / 255 is.
__device__ __forceinline__ float over255(float x)
{
    const bool z = x == 0.f;
    const float q = (z ? 1.f : x) / 255.f;
    return z ? 0.f : q;
}

// nearest texel, wrap by truncation, alpha inverted
__device__ __forceinline__ float4 texture_lookup(const SceneDev &sc, uint32_t w, uint32_t h, uint32_t off, float u, float v)
{
    u = u - ((int)u) + (u < 0 ? 1 : 0);
    v = v - ((int)v) + (v < 0 ? 1 : 0);
    const uint32_t x = (uint32_t)(u * (w - 1));
    const uint32_t y = (uint32_t)(v * (h - 1));
    PTC_CHECK(sc.faultSink, kFaultTexel, off + y * w + x, sc.nTexels);
    const uchar4 c = __ldg(&sc.texels[off + y * w + x]);
    float4 bgra = f4(over255((float)c.x), over255((float)c.y), over255((float)c.z), over255((float)c.w));
    bgra.w = 1.f - bgra.w;
    return bgra;
}

// SURVEY.md 8(f) N4 (PTC_FLAG_TEXTURE_BILINEAR, un-oracled): the same wrap and texel grid, the four texels around the
// continuous position blended; at texel centres it returns the nearest-texel value.
__device__ __forceinline__ float4 texture_lookup_bilinear(const SceneDev &sc, uint32_t w, uint32_t h, uint32_t off, float u, float v)
{
    u = u - ((int)u) + (u < 0 ? 1 : 0);
    v = v - ((int)v) + (v < 0 ? 1 : 0);
    const float fx = u * (w - 1), fy = v * (h - 1);
    const uint32_t x0 = (uint32_t)fx, y0 = (uint32_t)fy;
    const uint32_t x1 = x0 + 1 < w ? x0 + 1 : x0, y1 = y0 + 1 < h ? y0 + 1 : y0;
    const float ax = fx - (float)x0, ay = fy - (float)y0;
    PTC_CHECK(sc.faultSink, kFaultTexel, off + y1 * w + x1, sc.nTexels);
    const uchar4 c00 = __ldg(&sc.texels[off + y0 * w + x0]), c10 = __ldg(&sc.texels[off + y0 * w + x1]);
    const uchar4 c01 = __ldg(&sc.texels[off + y1 * w + x0]), c11 = __ldg(&sc.texels[off + y1 * w + x1]);
    const float w00 = (1 - ax) * (1 - ay), w10 = ax * (1 - ay), w01 = (1 - ax) * ay, w11 = ax * ay;
    float4 bgra = f4(over255((w00 * c00.x + w10 * c10.x) + (w01 * c01.x + w11 * c11.x)), over255((w00 * c00.y + w10 * c10.y) + (w01 * c01.y + w11 * c11.y)),
                     over255((w00 * c00.z + w10 * c10.z) + (w01 * c01.z + w11 * c11.z)), over255((w00 * c00.w + w10 * c10.w) + (w01 * c01.w + w11 * c11.w)));
    bgra.w = 1.f - bgra.w;
    return bgra;
}

__device__ __forceinline__ float4 sky_color(const SceneDev &sc, float4 dir)
{
    const float x = sc.sky.cosRot * dir.x - sc.sky.sinRot * dir.y;
    const float y = sc.sky.sinRot * dir.x + sc.sky.cosRot * dir.y;
    const float z = dir.z;
    int face = 0;
    float u = 0, v = 0;
    if (fabsf(z) > fabsf(x) && fabsf(z) > fabsf(y))
    {
        if (z > 0) { face = 5; u = (1 - x / z) / 2; v = (1 + y / z) / 2; }
        else { face = 0; u = (1 + x / z) / 2; v = (1 + y / z) / 2; }
    }
    else if (fabsf(x) > fabsf(y) && fabsf(x) > fabsf(z))
    {
        if (x > 0) { face = 1; u = (1 - y / x) / 2; v = (1 + z / x) / 2; }
        else { face = 3; u = (1 - y / x) / 2; v = (1 - z / x) / 2; }
    }
    else if (fabsf(y) > fabsf(x) && fabsf(y) > fabsf(z))
    {
        if (y > 0) { face = 4; u = (1 + x / y) / 2; v = (1 + z / y) / 2; }
        else { face = 2; u = (1 + x / y) / 2; v = (1 - z / y) / 2; }
    }
    return texture_lookup(sc, sc.sky.w[face], sc.sky.h[face], sc.sky.off[face], u, v);
}

// The three Fresnel fractions share one body; n1/n2 differ. Returns 1 on total reflection.
__device__ __forceinline__ float fresnel(float n1, float n2, float cos1, float &cos2)
{
    const float sin1 = sqrtf(1 - cos1 * cos1);
    const float sin2 = n1 * sin1 / n2;
    cos2 = 0.f;
    if (sin2 >= 1) return 1;
    cos2 = sqrtf(1 - sin2 * sin2);
    const float rPara = (n2 * cos1 - n1 * cos2) / (n2 * cos1 + n1 * cos2);
    const float rPerp = (n1 * cos1 - n2 * cos2) / (n1 * cos1 + n2 * cos2);
    return (rPara * rPara + rPerp * rPerp) / 2.0f;
}

// :256-292 with isInVarnish = false
__device__ __forceinline__ float fresnel_varnish(float4 incident, float4 N)
{
    const float cos1 = fmaxf(0.f, fminf(1.f, -dot4(incident, N)));
    float cos2;
    return fresnel(1.0f, 3.f, cos1, cos2);
}

__device__ __forceinline__ float material_brdf(int type, float4 incident, float4 N, float4 reflected)
{
    if (type == 0) return kPiInv;
    if (type == 2) return 1;
    if (type == 1)
    {
        const float denom = 1 + 0.8f * dot4(incident, reflected);
        return (1 - 0.8f * 0.8f) / (4 * kPi * denom * denom);
    }
    if (type == 3) return (1 - fresnel_varnish(incident, N)) * kPiInv;
    return 1;
}

// AREA: also evaluates a resolved PTC_LIGHT_AREA record (extension; only instantiated for scenes that hold one)
template <bool AREA = false>
__device__ __forceinline__ float light_power_toward(const DevLight &l, float4 p, float4 N)
{
    if (l.type == 0) return l.power * fmaxf(dot4(-l.direction, N), 0.f);
    if (l.type == 1)
    {
        const float4 t = p - l.position;
        return l.power / dot4(t, t) * fmaxf(dot4(normalize4(l.position - p), N), 0.f);
    }
    if (l.type == 2)
    {
        const float4 ld = normalize4(p - l.position);
        const float cosAngle = dot4(ld, l.direction);
        if (cosAngle > l.cosInner) return l.power * fmaxf(-dot4(ld, N), 0.f);
        if (cosAngle < l.cosOuter) return 0.0f;
        return l.power * (cosAngle - l.cosOuter) / (l.cosInner - l.cosOuter) * fmaxf(-dot4(ld, N), 0);
    }
    if (AREA && l.type == 4)
    {
        // extension: one sampled point of an area light (light_resolve below): power already carries the area, direction is the
        // emitter's normal; geometry term cos(light) * cos(surface) / distance^2
        const float4 t = p - l.position;
        const float4 w = normalize4(l.position - p);
        return l.power / dot4(t, t) * fmaxf(-dot4(w, l.direction), 0.f) * fmaxf(dot4(w, N), 0.f);
    }
    return 0.f;
}

// PTC_LIGHT_AREA (extension, FAST only; include/pathtracer_structs.h): turns the parallelogram record into the point of it
// that this path connects to at this bounce. (a, b) are draws 2k+1 and 2k+2 of the path's stream after `seed`, k = index
// of the light among the area lights: the stream is a multiplicative congruence (header.cl:246-253), so "skip 2k draws" is
// one multiplication by 16807^(2k), which ptc_upload_scene stored in the reserved lane. The connect stage and the shade
// stage both call this with the seed the path carries into the bounce and therefore see the same point; k_shade then
// advances the stream past all area-light draws (SceneDev::areaJump).
__device__ __forceinline__ DevLight light_resolve(DevLight l, int seed)
{
    if (l.type != 4) return l;
    int s = (int)(((unsigned)seed * (unsigned)__float_as_int(l.cosOuter)) & 0x7FFFFFFFu);
    const float a = rng_next(s), b = rng_next(s);
    const float4 eu = f4(l.direction.x, l.direction.y, l.direction.z, 0.f), ev = f4(l.position.w, l.direction.w, l.cosInner, 0.f);
    const float4 c = cross3(eu, ev);
    const float area = length4(c);
    l.position = f4(l.position.x + (a * eu.x + b * ev.x), l.position.y + (a * eu.y + b * ev.y), l.position.z + (a * eu.z + b * ev.z), 1.f);
    l.direction = f4(c.x / area, c.y / area, c.z / area, 0.f);
    l.power = l.power * area;
    return l;
}

// PTC_LIGHT_AREA emitters seen directly (extension): the nearest emitter the ray (o, d) crosses from its emitting side before
// `tBest` (the ray parameter of the geometric hit, INFINITY on a miss). Three-lane arithmetic: a w lane the reference's quirks
// may have left in o or d plays no part here. Used for camera rays and after specular bounces only — a diffuse
// vertex has already sampled the emitter itself (next-event estimation), so its bounce ray passes through it.
__device__ __forceinline__ bool emitter_seen(const SceneDev &sc, float4 o, float4 d, float tBest, float4 &radiance)
{
    bool seen = false;
    for (uint32_t li = 0; li < sc.nLights; li++)
    {
        const DevLight l = sc.lights[li];
        if (l.type != 4) continue;
        const float4 eu = f4(l.direction.x, l.direction.y, l.direction.z, 0.f), ev = f4(l.position.w, l.direction.w, l.cosInner, 0.f);
        const float4 n = cross3(eu, ev);
        const float denom = (n.x * d.x + n.y * d.y) + n.z * d.z;
        if (!(denom < 0.f)) continue;  // one-sided
        const float4 c = f4(l.position.x - o.x, l.position.y - o.y, l.position.z - o.z, 0.f);
        const float t = ((n.x * c.x + n.y * c.y) + n.z * c.z) / denom;
        if (!(t > 0.001f) || !(t < tBest)) continue;
        const float4 p = f4(d.x * t - c.x, d.y * t - c.y, d.z * t - c.z, 0.f);
        const float nn = (n.x * n.x + n.y * n.y) + n.z * n.z;
        const float4 pa = cross3(p, ev), pb = cross3(eu, p);
        const float a = ((pa.x * n.x + pa.y * n.y) + pa.z * n.z) / nn, b = ((pb.x * n.x + pb.y * n.y) + pb.z * n.z) / nn;
        if (a < 0.f || a > 1.f || b < 0.f || b > 1.f) continue;
        tBest = t;
        radiance = l.color * l.power;
        seen = true;
    }
    return seen;
}

// cos/sin: the reference calls OpenCL's cos/sin (implementation-defined accuracy); the oracle build uses
// glibc cosf/sinf (< 1 ulp). They are evaluated in double here and rounded once, which yields the
// correctly rounded float and therefore agrees with the oracle in all but a ~1e-7 fraction of calls.
__device__ __forceinline__ void concentric_sample_disk(int &seed, float &dx, float &dy)
{
    const float u1 = rng_next(seed);
    const float u2 = rng_next(seed);
    float r, theta;
    const float sx = 2 * u1 - 1;
    const float sy = 2 * u2 - 1;
    if (fabsf(sx) < 0.0001f) { r = sy; theta = 0; }
    else if (fabsf(sy) < 0.0001f) { r = sx; theta = 2; }
    else if (sx > -sy)
    {
        if (sx > sy) { r = sx; if (sy > 0) theta = sy / sx; else theta = 8.f + sy / sx; }
        else { r = sy; theta = 2.f - sx / sy; }
    }
    else
    {
        if (sx < sy) { r = -sx; theta = 4.f + sy / sx; }
        else { r = -sy; theta = 6.f - sy / sx; }
    }
    theta *= kPi / 4.f;
    r = (float)((double)r * 0.999);  // `r *= 0.999;` with a double literal (:408)
    double sn, cs;
    sincos((double)theta, &sn, &cs);
    dx = r * (float)cs;
    dy = r * (float)sn;
}

__device__ __forceinline__ float4 cosine_sample_hemisphere(int &seed, float4 N)
{
    float x, y, z;
    concentric_sample_disk(seed, x, y);
    z = 1 - x * x - y * y;
    z = (z < 0) ? 0 : sqrtf(z);
    const float4 v = f4(x, y, z, 0);
    if (N.z > 0.9999f) return v;
    if (N.z < -0.9999f) return -v;
    const float4 sn = normalize4(f4(-N.y, N.x, 0, 0));
    const float4 tn = normalize4(cross3(N, sn));
    return normalize4(f4(dot4(f4(sn.x, tn.x, N.x, 0), v), dot4(f4(sn.y, tn.y, N.y, 0), v), dot4(f4(sn.z, tn.z, N.z, 0), v), 0));
}

// Shading normal of a hit (:1267-1274): interpolated vertex normals, turned towards the incoming ray. One function for the
// shade stage and for the light cull of the connect stage, so both see the same bits.
__device__ __forceinline__ float4 shading_normal(const DevShade *sh, float4 N, float4 d, float s, float t)
{
    const bool same = dot4(d, N) > 0;
    const float4 n1 = __ldg(&sh->n1), n2 = __ldg(&sh->n2), n3 = __ldg(&sh->n3);
    float4 Ns = normalize4((n2 * s) + (n3 * t) + (n1 * (1 - s - t)));
    if (same) Ns = -Ns;
    put_in_same_hemisphere(Ns, -d);
    return normalize4(Ns);
}

// Shadow ray of one light from the hit point q (:932-938): direction = normalize(full), bound =
// length(full) (INFINITY for a directional light) — a LENGTH handed to a parameter that the triangle
// test compares with squared distances (F9b), reproduced as is in both modes.
__device__ __forceinline__ void shadow_setup(const DevLight &l, float4 q, float4 &dir, float &bound)
{
    const float4 full = (l.type == 0) ? -l.direction : (l.position - q);
    dir = normalize4(full);
    bound = (l.type == 0) ? INFINITY : length4(full);
    // extension: a sampled area-light point is occluded by anything in front of it — the bound is the squared distance the
    // triangle test compares it with (the reference's lights keep their quirk: occluders beyond sqrt(distance) are not seen)
    if (l.type == 4) bound = dot4(full, full);
}

}  // namespace ptc
