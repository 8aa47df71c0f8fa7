// PathTracer_Types.h — host-side data model, namespace PathTracerNS.
//
// Mirrors the names and memory layout of the reference's host data model
// (reference Controleur/PathTracer_Structs.h:16-189 and Controleur/PathTracer_Utils.h:69-162) so that
// code written against the reference's GlobalVars (importers, PathTracer_Main) compiles against this
// header unchanged, on any platform (no <CL/cl.h>, no __declspec). Every record is the corresponding
// C-ABI record of include/pathtracer_structs.h under its reference name.
#ifndef PATHTRACER_B200_TYPES_H
#define PATHTRACER_B200_TYPES_H

#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>

#include "pathtracer_structs.h"

namespace PathTracerNS
{
typedef unsigned int uint;
typedef unsigned char uchar;

// --- vectors ---------------------------------------------------------------------------------------
// Host vector math follows reference Controleur/PathTracer_Utils.h:129-144: dot() is 3-lane, cross()
// returns w = 1, normalize() divides xyz and w by the 3-lane length and then restores w.
struct Float2 : ptc_float2
{
    Float2() { x = 0; y = 0; }
    Float2(float _x, float _y) { x = _x; y = _y; }
};

struct alignas(16) Float4 : ptc_float4
{
    Float4() { x = y = z = w = 0; }
    Float4(float _x, float _y, float _z, float _w) { x = _x; y = _y; z = _z; w = _w; }
    Float4(const ptc_float4 &v) { x = v.x; y = v.y; z = v.z; w = v.w; }
};
typedef Float4 RGBAColor;

inline Float4 operator+(const Float4 &a, const Float4 &b) { return Float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline Float4 operator-(const Float4 &a, const Float4 &b) { return Float4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
inline Float4 operator-(const Float4 &a) { return Float4(-a.x, -a.y, -a.z, -a.w); }
inline Float4 operator*(const Float4 &a, float s) { return Float4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline Float4 operator/(const Float4 &a, float s) { return Float4(a.x / s, a.y / s, a.z / s, a.w / s); }
inline bool operator==(const Float4 &a, const Float4 &b) { return a.x == b.x && a.y == b.y && a.z == b.z && a.w == b.w; }

inline float dot(const Float4 &a, const Float4 &b) { return (a.x * b.x) + (a.y * b.y) + (a.z * b.z); }
inline Float4 cross(const Float4 &a, const Float4 &b)
{
    return Float4((a.y * b.z) - (a.z * b.y), (a.z * b.x) - (a.x * b.z), (a.x * b.y) - (a.y * b.x), 1);
}
inline float length(const Float4 &v) { return std::sqrt(dot(v, v)); }
inline Float4 normalize(const Float4 &v)
{
    Float4 n = v / length(v);
    n.w = v.w;
    return n;
}
inline Float4 ppmin(const Float4 &a, const Float4 &b)
{
    return Float4(b.x < a.x ? b.x : a.x, b.y < a.y ? b.y : a.y, b.z < a.z ? b.z : a.z, b.w < a.w ? b.w : a.w);
}
inline Float4 ppmax(const Float4 &a, const Float4 &b)
{
    return Float4(a.x < b.x ? b.x : a.x, a.y < b.y ? b.y : a.y, a.z < b.z ? b.z : a.z, a.w < b.w ? b.w : a.w);
}
inline bool Vector_LexLessThan(const Float4 &a, const Float4 &b)
{
    return (a.x < b.x) || (a.x == b.x && a.y < b.y) || (a.x == b.x && a.y == b.y && a.z < b.z);
}

struct Uchar4 : ptc_uchar4
{
    Uchar4() { x = y = z = w = 0; }
    Uchar4(uchar r, uchar g, uchar b, uchar a) { x = r; y = g; z = b; w = a; }
};

// --- records (reference names; layouts asserted below) --------------------------------------------------
typedef ptc_bounding_box BoundingBox;
typedef ptc_light Light;
typedef ptc_material Material;
typedef ptc_node Node;
typedef ptc_texture Texture;
typedef ptc_triangle Triangle;
typedef ptc_sky Sky;

enum LightType { LIGHT_DIRECTIONNAL = PTC_LIGHT_DIRECTIONNAL, LIGHT_POINT = PTC_LIGHT_POINT, LIGHT_SPOT = PTC_LIGHT_SPOT, LIGHT_UNKNOWN = PTC_LIGHT_UNKNOWN };
enum MaterialType { MAT_STANDART = PTC_MAT_STANDART, MAT_WATER = PTC_MAT_WATER, MAT_GLASS = PTC_MAT_GLASS, MAT_VARNHISHED = PTC_MAT_VARNHISHED, MAT_METAL = PTC_MAT_METAL, MAT_UNKNOWN = PTC_MAT_UNKNOWN };
enum NodeStopType { NODE_BAD_SAH = PTC_NODE_BAD_SAH, NODE_LEAF_MAX_SIZE = PTC_NODE_LEAF_MAX_SIZE, NODE_LEAF_MIN_DIAG = PTC_NODE_LEAF_MIN_DIAG };
enum Sampler { JITTERED = 0, RANDOM = 1, UNIFORM = 2 };

static const uint MAX_INTERSETCION_NUMBER = 5000;  // reference Controleur/PathTracer_PreProc.h:17 (spelling kept)
static const uint BVH_MAX_DEPTH = 30;
static const uint MAX_LIGHT_SIZE = 30;

class PathTracerDialog;
class PathTracerImporter;

// reference Controleur/PathTracer_Structs.h:145-189, same member names
struct GlobalVars
{
    PathTracerDialog *window = nullptr;
    PathTracerImporter *importer = nullptr;

    Float4 cameraDirection;
    Float4 cameraRight;
    Float4 cameraUp;
    Float4 cameraPosition;

    Node *bvh = nullptr;
    Triangle *triangulation = nullptr;
    Light *lights = nullptr;
    Material *materiaux = nullptr;
    Texture *textures = nullptr;
    Uchar4 *texturesData = nullptr;

    uint bvhSize = 0;
    uint triangulationSize = 0;
    uint lightsSize = 0;
    uint materiauxSize = 0;
    uint texturesSize = 0;
    uint texturesDataSize = 0;
    uint bvhMaxDepth = 0;
    Sampler sampler = JITTERED;
    bool printLogInfos = false;
    bool superSampling = false;

    Sky sky{};

    uint imageWidth = 0;
    uint imageHeight = 0;
    uint imageSize = 0;
    uint rayMaxDepth = 10;
    RGBAColor *imageColor = nullptr;
    float *imageRayNb = nullptr;
    uint *rayDepths = nullptr;
    uint *rayIntersectedBBx = nullptr;
    uint *rayIntersectedTri = nullptr;
};

static_assert(sizeof(Float4) == 16 && alignof(Float4) == 16, "Float4 layout");
static_assert(sizeof(Uchar4) == 4, "Uchar4 layout");

// reference Controleur/PathTracer_Structs.h:197-205
inline void BoundingBox_Create(BoundingBox &bb, const Float4 &s1, const Float4 &s2, const Float4 &s3)
{
    bb.isEmpty = false;
    Float4 lo = ppmin(ppmin(s1, s2), s3), hi = ppmax(ppmax(s1, s2), s3);
    bb.pMin = lo;
    bb.pMax = hi;
    bb.centroid = (lo + hi) / 2;
}

}  // namespace PathTracerNS

#endif
