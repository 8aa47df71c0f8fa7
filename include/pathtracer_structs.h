/* pathtracer_structs.h — the scene records as the reference lays them out in memory.
 *
 * This is the INPUT VIEW of the PathTracer_CUDA boundary: byte-compatible with the records the
 * reference host code fills and hands to its OpenCL kernel as `void __global*`
 * (reference: Controleur/PathTracer_Structs.h:16-128, mirrored by
 * Kernel/PathTracer_FullKernel_header.cl:89-223; MSVC `__declspec(align(16))`, x64).
 * A reference build can pass its `GlobalVars` arrays straight through these types.
 * The device never sees these layouts: ptc_upload_scene() flattens them into the packed
 * 64-byte node / triangle records described in DESIGN.md.
 *
 * Plain C, no dependencies. Sizes are asserted at the bottom and checked by tests against the
 * reference's own structs compiled with g++ (oracle/_ref).
 */
#ifndef PATHTRACER_STRUCTS_H
#define PATHTRACER_STRUCTS_H

#include <stdint.h>

#if defined(_MSC_VER)
#define PTC_ALIGN(n) __declspec(align(n))
#else
#define PTC_ALIGN(n) __attribute__((aligned(n)))
#endif

#ifdef __cplusplus
extern "C" {
#endif

typedef struct PTC_ALIGN(16) ptc_float4 { float x, y, z, w; } ptc_float4;
typedef struct ptc_float2 { float x, y; } ptc_float2;
typedef struct ptc_uchar4 { uint8_t x, y, z, w; } ptc_uchar4;

/* Controleur/PathTracer_Structs.h:16-22 — 64 bytes */
typedef struct PTC_ALIGN(16) ptc_bounding_box
{
    ptc_float4 pMin;
    ptc_float4 pMax;
    ptc_float4 centroid;
    char isEmpty;
} ptc_bounding_box;

/* Controleur/PathTracer_Structs.h:24-30 */
enum { PTC_LIGHT_DIRECTIONNAL = 0, PTC_LIGHT_POINT = 1, PTC_LIGHT_SPOT = 2, PTC_LIGHT_UNKNOWN = 3,
       /* Extension (the reference has no area or emissive lights): a one-sided parallelogram emitter,
        * FAST mode only, filled by ptc_light_make_area below. */
       PTC_LIGHT_AREA = 4 };

/* Controleur/PathTracer_Structs.h:32-41 — 64 bytes */
typedef struct PTC_ALIGN(16) ptc_light
{
    ptc_float4 position;
    ptc_float4 direction;
    ptc_float4 color;
    float power;
    float cosOfInnerFallOffAngle;
    float cosOfOuterFallOffAngle;
    int32_t type;
} ptc_light;

/* PTC_LIGHT_AREA record: the parallelogram corner + a*eu + b*ev, (a, b) in [0,1)^2, emitting `power` x `color` as
 * radiance towards cross(eu, ev) only. Each lit hit connects to ONE point of it, drawn from the path's own random stream
 * (next-event estimation); the estimator is power * color * BRDF * cos(surface) * cos(light) * area / distance^2. The
 * emitter has no triangles of its own; camera rays and rays leaving a specular event (glass, water, the mirror branch of
 * varnish) that cross it from its emitting side before their geometric hit end there with its radiance (it is visible and
 * is reflected); rays leaving a diffuse vertex pass through it, since that vertex has sampled it already.
 * Its shadow rays are bounded by the squared distance to the sampled point, i.e. every occluder in front of it counts (the
 * reference's own light types hand a LENGTH to a squared-distance test, Kernel/PathTracer_FullKernel.cl:938-944, which both
 * modes reproduce for them).
 * The edge ev does not fit the 64-byte reference record beside eu, so it rides in the three lanes a positional light
 * leaves unused (position.w, direction.w, cosOfInnerFallOffAngle); cosOfOuterFallOffAngle is reserved (the library
 * writes the light's offset into the random stream there at upload). */
static inline void ptc_light_make_area(ptc_light *l, const float corner[3], const float eu[3], const float ev[3],
                                       const float rgb[3], float power)
{
    l->position.x = corner[0]; l->position.y = corner[1]; l->position.z = corner[2]; l->position.w = ev[0];
    l->direction.x = eu[0]; l->direction.y = eu[1]; l->direction.z = eu[2]; l->direction.w = ev[1];
    l->color.x = rgb[0]; l->color.y = rgb[1]; l->color.z = rgb[2]; l->color.w = 1.0f;
    l->power = power;
    l->cosOfInnerFallOffAngle = ev[2];
    l->cosOfOuterFallOffAngle = 0.0f;
    l->type = PTC_LIGHT_AREA;
}

/* Controleur/PathTracer_Structs.h:43-51 */
enum { PTC_MAT_STANDART = 0, PTC_MAT_WATER = 1, PTC_MAT_GLASS = 2, PTC_MAT_VARNHISHED = 3, PTC_MAT_METAL = 4, PTC_MAT_UNKNOWN = 5 };

/* Controleur/PathTracer_Structs.h:54-63 — 48 bytes (the 8-byte host pointer is part of the layout) */
typedef struct PTC_ALIGN(16) ptc_material
{
    ptc_float4 simpleColor;
    const char *textureName; /* host-only; ignored by the device */
    float opacity;
    int32_t textureId;
    int32_t type;
    uint8_t isSimpleColor; /* C++ bool in the reference */
    uint8_t hasAlphaMap;
} ptc_material;

/* Controleur/PathTracer_Structs.h:66-71 */
enum { PTC_NODE_BAD_SAH = 0, PTC_NODE_LEAF_MAX_SIZE = 1, PTC_NODE_LEAF_MIN_DIAG = 2 };

/* Controleur/PathTracer_Structs.h:74-85 — 160 bytes */
typedef struct PTC_ALIGN(16) ptc_node
{
    ptc_bounding_box trianglesAABB;
    ptc_bounding_box centroidsAABB;
    uint32_t cutAxis;
    uint32_t triangleStartIndex;
    uint32_t nbTriangles;
    uint32_t son1Id;
    uint32_t son2Id;
    int32_t comments;
    char isLeaf;
} ptc_node;

/* Controleur/PathTracer_Structs.h:96-101 — 12 bytes */
typedef struct ptc_texture
{
    uint32_t width;
    uint32_t height;
    uint32_t offset; /* first texel in texturesData */
} ptc_texture;

/* Controleur/PathTracer_Structs.h:103-117 — 336 bytes */
typedef struct PTC_ALIGN(16) ptc_triangle
{
    ptc_float4 S1, S2, S3;    /* vertices */
    ptc_float4 N1, N2, N3;    /* vertex normals */
    ptc_float4 T1, T2, T3;    /* tangents   (unused by the kernel) */
    ptc_float4 BT1, BT2, BT3; /* bitangents (unused by the kernel) */
    ptc_float4 N;             /* face normal; .w is 1 with the Maya importer */
    ptc_float2 UVP1, UVP2, UVP3;
    ptc_float2 UVN1, UVN2, UVN3;
    ptc_bounding_box AABB;
    uint32_t materialWithPositiveNormalIndex;
    uint32_t materialWithNegativeNormalIndex;
    uint32_t id;
} ptc_triangle;

/* Controleur/PathTracer_Structs.h:120-128 — 92 bytes */
typedef struct ptc_sky
{
    ptc_texture skyTextures[6];
    float groundScale;
    float exposantFactorX;
    float exposantFactorY;
    float cosRotationAngle;
    float sinRotationAngle;
} ptc_sky;

#ifdef __cplusplus
}
static_assert(sizeof(ptc_bounding_box) == 64, "BoundingBox layout");
static_assert(sizeof(ptc_light) == 64, "Light layout");
static_assert(sizeof(ptc_material) == 48, "Material layout");
static_assert(sizeof(ptc_node) == 160, "Node layout");
static_assert(sizeof(ptc_texture) == 12, "Texture layout");
static_assert(sizeof(ptc_triangle) == 336, "Triangle layout");
static_assert(sizeof(ptc_sky) == 92, "Sky layout");
#endif

#endif /* PATHTRACER_STRUCTS_H */
