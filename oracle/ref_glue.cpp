// TEST INFRASTRUCTURE ONLY — C entry points around the reference's own code compiled for the CPU.
// Built by oracle/build_ref.py into oracle/_ref/libptref.so. Never linked into the product.
//
// Two pieces of the reference are compiled here (from generated copies, see build_ref.py):
//   * the megakernel  Kernel/PathTracer_FullKernel.cl (+ _header.cl), three times, once per sampler
//     (`-D SAMPLE_*` is a compile-time switch in the reference, Controleur/PathTracer_OpenCL.cpp:297-303),
//     each inside its own namespace together with the shim that gives OpenCL C a C++ meaning;
//   * the host BVH builder  Controleur/PathTracer_BVH.cpp  (separate object, linked in).
// The launch loop below restates Controleur/PathTracer_OpenCL.cpp:71-107 (one Kernel_Main call per
// pixel per iteration, iterations strictly in order) with ZERO-initialised output buffers
// (the reference relies on the Intel runtime handing out zero pages).

#include "ocl_shim_types.h"

#include <cstring>
#include <cstdlib>
#include <vector>
#include <omp.h>

// reference host data model + BVH builder (generated copy of Controleur/*.h)
#include "PathTracer_BVH.h"

namespace ocl_shim
{
thread_local size_t gid[2] = {0, 0};
}

// "build options" of the reference (Controleur/PathTracer_OpenCL.cpp:305-308) become run-time values
static int g_image_width = 1, g_image_height = 1, g_lights_size = 0, g_max_reflection = 10;
#define IMAGE_WIDTH g_image_width
#define IMAGE_HEIGHT g_image_height
#define LIGHTS_SIZE g_lights_size
#define MAX_REFLECTION_NUMBER g_max_reflection

// OpenCL qualifiers with no C++ meaning. `volatile` is dropped as well: the reference reads float4
// values through volatile pointers, which C++ struct copy cannot do; no std header follows this line.
#define __global
#define __local
#define __kernel
#define __read_only
#define volatile
#undef NULL

namespace refk_jittered
{
#include "ocl_as_cxx.h"
#define SAMPLE_JITTERED
#include "PathTracer_FullKernel.cl"
#undef SAMPLE_JITTERED
}  // namespace refk_jittered
#undef PATHTRACER_KERNEL
#undef NULL

namespace refk_random
{
#include "ocl_as_cxx.h"
#define SAMPLE_RANDOM
#include "PathTracer_FullKernel.cl"
#undef SAMPLE_RANDOM
}  // namespace refk_random
#undef PATHTRACER_KERNEL
#undef NULL

// SAMPLE_JITTERED again with `-D SUPER_SAMPLING` (Controleur/PathTracer_OpenCL.cpp:311): the adaptive early-out of
// Kernel_Main (:1219-1222, superSamplingStopCriteria :1152-1172) against the supplied X2inv.cl table
namespace refk_jittered_ss
{
#include "ocl_as_cxx.h"
#define SAMPLE_JITTERED
#define SUPER_SAMPLING 1
#include "PathTracer_FullKernel.cl"
#undef SUPER_SAMPLING
#undef SAMPLE_JITTERED
}  // namespace refk_jittered_ss
#undef PATHTRACER_KERNEL
#undef NULL

namespace refk_uniform
{
#include "ocl_as_cxx.h"
#define SAMPLE_UNIFORM
#include "PathTracer_FullKernel.cl"
#undef SAMPLE_UNIFORM
}  // namespace refk_uniform
#undef volatile

extern "C" {

struct ptref_scene
{
    const void *bvh;            // reference Node[]      (160 B each)
    const void *triangulation;  // reference Triangle[]  (336 B each), BVH order
    const void *lights;         // reference Light[]     (64 B)
    const void *materiaux;      // reference Material[]  (48 B)
    const void *textures;       // reference Texture[]   (12 B)
    const void *texturesData;   // uchar4[]
    const void *sky;            // reference Sky         (92 B)
    float camera[16];           // position, direction, right, up (float4 each)
    uint32_t lightsSize;
};

// sizeof of every record on both sides of the reference's host/kernel boundary
int ptref_sizes(uint32_t *out)
{
    out[0] = sizeof(PathTracerNS::BoundingBox);
    out[1] = sizeof(PathTracerNS::Node);
    out[2] = sizeof(PathTracerNS::Triangle);
    out[3] = sizeof(PathTracerNS::Light);
    out[4] = sizeof(PathTracerNS::Material);
    out[5] = sizeof(PathTracerNS::Texture);
    out[6] = sizeof(PathTracerNS::Sky);
    out[7] = sizeof(refk_jittered::BoundingBox);
    out[8] = sizeof(refk_jittered::Node);
    out[9] = sizeof(refk_jittered::Triangle);
    out[10] = sizeof(refk_jittered::Light);
    out[11] = sizeof(refk_jittered::Material);
    out[12] = sizeof(refk_jittered::Texture);
    out[13] = sizeof(refk_jittered::Sky);
    out[14] = sizeof(refk_jittered::Ray3D);
    return 15;
}

// Reference BVH_Create (Controleur/PathTracer_BVH.cpp:12-37). `triangulation` is permuted in place;
// `nodes_out` must hold 2n-1 Nodes; its bytes are zeroed first and only the fields copied, because the
// reference leaves padding and Node.comments of inner nodes uninitialised.
int ptref_bvh_create(void *triangulation, uint32_t n, void *nodes_out, uint32_t *bvhSize, uint32_t *bvhMaxDepth)
{
    using namespace PathTracerNS;
    if (n == 0) return 1;
    GlobalVars gv;
    memset((void *)&gv, 0, sizeof(gv));
    gv.triangulation = (Triangle *)triangulation;
    gv.triangulationSize = n;
    BVH_Create(gv);
    Node *dst = (Node *)nodes_out;
    memset((void *)dst, 0, sizeof(Node) * (size_t)gv.bvhSize);
    for (uint32_t i = 0; i < gv.bvhSize; i++)
    {
        const Node &s = gv.bvh[i];
        Node &d = dst[i];
        d.trianglesAABB.pMin = s.trianglesAABB.pMin;
        d.trianglesAABB.pMax = s.trianglesAABB.pMax;
        d.trianglesAABB.centroid = s.trianglesAABB.centroid;
        d.trianglesAABB.isEmpty = s.trianglesAABB.isEmpty;
        d.centroidsAABB.pMin = s.centroidsAABB.pMin;
        d.centroidsAABB.pMax = s.centroidsAABB.pMax;
        d.centroidsAABB.centroid = s.centroidsAABB.centroid;
        d.centroidsAABB.isEmpty = s.centroidsAABB.isEmpty;
        d.cutAxis = s.cutAxis;
        d.triangleStartIndex = s.triangleStartIndex;
        d.nbTriangles = s.nbTriangles;
        d.son1Id = s.son1Id;
        d.son2Id = s.son2Id;
        d.comments = s.isLeaf ? s.comments : NODE_BAD_SAH;
        d.isLeaf = s.isLeaf;
    }
    *bvhSize = gv.bvhSize;
    *bvhMaxDepth = gv.bvhMaxDepth;
    delete[] gv.bvh;
    return 0;
}

static void set_options(uint32_t W, uint32_t H, uint32_t maxDepth, uint32_t lights)
{
    g_image_width = (int)W;
    g_image_height = (int)H;
    g_max_reflection = (int)maxDepth;
    g_lights_size = (int)lights;
}

#define LAUNCH(NS)                                                                                              \
    NS::Kernel_Main(it, cam[0], cam[1], cam[2], cam[3], (float4 *)imageColor, imageRayNb, (float4 *)imageV, \
                    rayDepths, bbx, tri, s->bvh, s->triangulation, s->lights, s->materiaux, s->textures,      \
                    (const uchar4 *)s->texturesData, s->sky)

// Launch loop: Controleur/PathTracer_OpenCL.cpp:76-107. Output buffers are accumulated into (the caller
// zeroes them before the first call), so consecutive iteration ranges can be rendered by consecutive calls.
// sampler: 0 JITTERED, 1 RANDOM, 2 UNIFORM (Controleur/PathTracer_Structs.h:130-135). RANDOM scatters to
// arbitrary pixels without atomics in the reference, so it is run on one thread here. 3 = JITTERED built with
// `-D SUPER_SAMPLING`. ptref_x2inv exposes the table that build reads (oracle/shim/X2inv.cl).
int ptref_render_rows(const ptref_scene *s, uint32_t W, uint32_t H, uint32_t maxDepth, int sampler, uint32_t iterBegin,
                      uint32_t iterEnd, uint32_t rowBegin, uint32_t rowEnd, float *imageColor, float *imageRayNb, float *imageV,
                      uint32_t *rayDepths, uint32_t *bbx, uint32_t *tri, int nthreads);

int ptref_render(const ptref_scene *s, uint32_t W, uint32_t H, uint32_t maxDepth, int sampler, uint32_t iterBegin,
                 uint32_t iterEnd, float *imageColor, float *imageRayNb, float *imageV, uint32_t *rayDepths,
                 uint32_t *bbx, uint32_t *tri, int nthreads)
{
    return ptref_render_rows(s, W, H, maxDepth, sampler, iterBegin, iterEnd, 0, H, imageColor, imageRayNb, imageV, rayDepths, bbx, tri,
                             nthreads);
}

// Same loop restricted to image rows [rowBegin, rowEnd): a bounded sample of a large frame for the timed CPU
// baseline (work-items of other rows are simply not launched; every launched work-item is unchanged).
int ptref_render_rows(const ptref_scene *s, uint32_t W, uint32_t H, uint32_t maxDepth, int sampler, uint32_t iterBegin,
                      uint32_t iterEnd, uint32_t rowBegin, uint32_t rowEnd, float *imageColor, float *imageRayNb, float *imageV,
                      uint32_t *rayDepths, uint32_t *bbx, uint32_t *tri, int nthreads)
{
    set_options(W, H, maxDepth, s->lightsSize);
    float4 cam[4];
    memcpy(cam, s->camera, sizeof(cam));
    if (nthreads <= 0) nthreads = omp_get_max_threads();
    if (sampler == 1) nthreads = 1;
    for (uint32_t it = iterBegin; it < iterEnd; it++)
    {
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads)
        for (int64_t y = (int64_t)rowBegin; y < (int64_t)(rowEnd < H ? rowEnd : H); y++)
        {
            ocl_shim::gid[1] = (size_t)y;
            for (uint32_t x = 0; x < W; x++)
            {
                ocl_shim::gid[0] = x;
                if (sampler == 0) LAUNCH(refk_jittered);
                else if (sampler == 1) LAUNCH(refk_random);
                else if (sampler == 3) LAUNCH(refk_jittered_ss);  // JITTERED + SUPER_SAMPLING
                else LAUNCH(refk_uniform);
            }
        }
    }
    return 0;
}

// First lines of Kernel_Main (Kernel/PathTracer_FullKernel.cl:1208-1215) + one BVH_IntersectRay call
// (:1252), made with the reference's own functions, to expose what Kernel_Main keeps private:
// the primary ray and the primary hit. Per pixel: hit = BVH-order triangle index or -1; st = (s,t);
// dist2 = |q - origin|^2 recomputed as Triangle_Intersects does (:540-541); dir = ray direction (4 lanes);
// seedAfter = RNG state after the sampler; counts = (numIntersectedBBx, numIntersectedTri).
#define PRIMARY(NS)                                                                                          \
    {                                                                                                        \
        NS::Ray3D r;                                                                                         \
        int seedValue = NS::InitializeRandomSeed(it);                                                        \
        float2 sample = NS::sampler(&r, &seedValue, it);                                                     \
        float4 shot = cam[1] + (cam[2] * sample.x) + (cam[3] * sample.y);                                    \
        NS::Ray3D_Create(&r, &cam[0], &shot, false);                                                         \
        r.sample = sample;                                                                                   \
        bool h = NS::BVH_IntersectRay(&seedValue, (const NS::Node *)s->bvh,                                  \
                                      (const NS::Triangle *)s->triangulation, (const NS::Light *)s->lights,  \
                                      (const NS::Material *)s->materiaux, (const NS::Texture *)s->textures,  \
                                      (const uchar4 *)s->texturesData, (const NS::Sky *)s->sky, &r);         \
        size_t p = (size_t)y * W + x;                                                                        \
        hit[p] = h ? (int32_t)r.intersectedTriangleId : -1;                                                  \
        st[2 * p] = h ? r.s : 0.f;                                                                           \
        st[2 * p + 1] = h ? r.t : 0.f;                                                                       \
        float4 fr = r.intersectionPoint - r.origin;                                                          \
        dist2[p] = h ? NS::dot(fr, fr) : INFINITY;                                                           \
        memcpy(dir + 4 * p, &r.direction, 16);                                                               \
        seedAfter[p] = seedValue;                                                                            \
        counts[2 * p] = r.numIntersectedBBx;                                                                 \
        counts[2 * p + 1] = r.numIntersectedTri;                                                             \
    }

int ptref_primary(const ptref_scene *s, uint32_t W, uint32_t H, int sampler, uint32_t it, int32_t *hit, float *st,
                  float *dist2, float *dir, int32_t *seedAfter, uint32_t *counts)
{
    set_options(W, H, 10, s->lightsSize);
    float4 cam[4];
    memcpy(cam, s->camera, sizeof(cam));
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t y = 0; y < (int64_t)H; y++)
    {
        ocl_shim::gid[1] = (size_t)y;
        for (uint32_t x = 0; x < W; x++)
        {
            ocl_shim::gid[0] = x;
            if (sampler == 0) PRIMARY(refk_jittered)
            else if (sampler == 1) PRIMARY(refk_random)
            else PRIMARY(refk_uniform)
        }
    }
    return 0;
}

// Ray probe: Ray3D_Create (header.cl:286-305, which normalises the direction) + one BVH_IntersectRay call (:620-702) for
// arbitrary rays, so that traversal parity can be checked on rays that start on surfaces (the bounces), not only on
// camera rays. Per ray: hit = BVH-order triangle index or -1; st; dist2 as Triangle_Intersects forms it; dirOut = the
// normalised direction the traversal used; counts = (numIntersectedBBx, numIntersectedTri).
int ptref_trace_rays(const ptref_scene *s, uint32_t n, const float *origin, const float *direction, int32_t *hit, float *st,
                     float *dist2, float *dirOut, uint32_t *counts)
{
    namespace K = refk_jittered;
    set_options(64, 64, 10, s->lightsSize);
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t p = 0; p < (int64_t)n; p++)
    {
        ocl_shim::gid[0] = 0;
        ocl_shim::gid[1] = 0;
        float4 o, d;
        memcpy(&o, origin + 4 * p, 16);
        memcpy(&d, direction + 4 * p, 16);
        K::Ray3D r;
        int seedValue = 1;
        K::Ray3D_Create(&r, &o, &d, false);
        bool h = K::BVH_IntersectRay(&seedValue, (const K::Node *)s->bvh, (const K::Triangle *)s->triangulation,
                                     (const K::Light *)s->lights, (const K::Material *)s->materiaux, (const K::Texture *)s->textures,
                                     (const uchar4 *)s->texturesData, (const K::Sky *)s->sky, &r);
        hit[p] = h ? (int32_t)r.intersectedTriangleId : -1;
        st[2 * p] = h ? r.s : 0.f;
        st[2 * p + 1] = h ? r.t : 0.f;
        float4 fr = r.intersectionPoint - r.origin;
        dist2[p] = h ? K::dot(fr, fr) : INFINITY;
        memcpy(dirOut + 4 * p, &r.direction, 16);
        counts[2 * p] = r.numIntersectedBBx;
        counts[2 * p + 1] = r.numIntersectedTri;
    }
    return 0;
}

// The stand-in chi-square table the SUPER_SAMPLING build indexes (the reference's Kernel/X2inv.cl is empty).
int ptref_x2inv(float *out, int capacity)
{
    const int n = (int)(sizeof(refk_jittered_ss::X2inv) / sizeof(float));
    for (int i = 0; i < n && i < capacity; i++) out[i] = refk_jittered_ss::X2inv[i];
    return n;
}

// RNG known-answer helper: InitializeRandomSeed (header.cl:255-264) then n draws of random() (:246-253).
int ptref_rng(uint32_t x, uint32_t y, uint32_t it, uint32_t W, uint32_t H, int n, int32_t *seed0, int32_t *seeds,
              float *us)
{
    set_options(W, H, 10, 0);
    ocl_shim::gid[0] = x;
    ocl_shim::gid[1] = y;
    int sd = refk_jittered::InitializeRandomSeed(it);
    *seed0 = sd;
    for (int i = 0; i < n; i++)
    {
        us[i] = refk_jittered::random(&sd);
        seeds[i] = sd;
    }
    return 0;
}

// Material_CosineSampleHemisphere (kernel.cl:303-336) for one seed/normal: trig-sensitive KAT.
int ptref_cosine_sample(int32_t seed, const float *N, float *out, int32_t *seedAfter)
{
    float4 n{N[0], N[1], N[2], N[3]};
    int sd = seed;
    float4 v = refk_jittered::Material_CosineSampleHemisphere(&sd, &n);
    memcpy(out, &v, 16);
    *seedAfter = sd;
    return 0;
}

// Debug probe: Kernel_Main's path loop (kernel.cl:1248-1315) for ONE work-item, made with the reference's own functions,
// logging per bounce: origin(4) direction(4) | hit tri, material type | boxes, triangles tested by the closest-hit
// traversal, boxes, triangles by the shadow rays | s, t. Returns the number of bounces logged.
int ptref_path_debug(const ptref_scene *s, uint32_t W, uint32_t H, uint32_t maxDepth, uint32_t x, uint32_t y, uint32_t it,
                     float *rays /*[maxDepth][8]*/, int32_t *ids /*[maxDepth][2]*/, uint32_t *work /*[maxDepth][4]*/, float *st)
{
    namespace K = refk_jittered;
    set_options(W, H, maxDepth, s->lightsSize);
    float4 cam[4];
    memcpy(cam, s->camera, sizeof(cam));
    ocl_shim::gid[0] = x;
    ocl_shim::gid[1] = y;
    K::Ray3D r;
    int seedValue = K::InitializeRandomSeed(it);
    int *seed = &seedValue;
    float2 sample = K::sampler(&r, seed, it);
    float4 shot = cam[1] + (cam[2] * sample.x) + (cam[3] * sample.y);
    K::Ray3D_Create(&r, &cam[0], &shot, false);
    r.sample = sample;
    const K::Node *bvh = (const K::Node *)s->bvh;
    const K::Triangle *tris = (const K::Triangle *)s->triangulation;
    const K::Light *lights = (const K::Light *)s->lights;
    const K::Material *mats = (const K::Material *)s->materiaux;
    const K::Texture *tex = (const K::Texture *)s->textures;
    const uchar4 *texd = (const uchar4 *)s->texturesData;
    const K::Sky *sky = (const K::Sky *)s->sky;
    float4 T{1, 1, 1, 1};
    int n = 0;
    bool active = true;
    while (active && r.reflectionId < maxDepth)
    {
        memcpy(rays + 8 * n, &r.origin, 16);
        memcpy(rays + 8 * n + 4, &r.direction, 16);
        uint32_t b0 = r.numIntersectedBBx, t0 = r.numIntersectedTri;
        bool h = K::BVH_IntersectRay(seed, bvh, tris, lights, mats, tex, texd, sky, &r);
        work[4 * n] = r.numIntersectedBBx - b0;
        work[4 * n + 1] = r.numIntersectedTri - t0;
        work[4 * n + 2] = work[4 * n + 3] = 0;
        ids[2 * n] = h ? (int32_t)r.intersectedTriangleId : -1;
        ids[2 * n + 1] = -1;
        st[2 * n] = h ? r.s : 0;
        st[2 * n + 1] = h ? r.t : 0;
        if (h)
        {
            K::Triangle tri = tris[r.intersectedTriangleId];
            K::Material mat = mats[r.intersectedMaterialId];
            ids[2 * n + 1] = mat.type;
            bool same = (K::dot(r.direction, tri.N) > 0);
            float4 Ng = K::Triangle_GetNormal(&tri, !same);
            float4 Ns = K::Triangle_GetSmoothNormal(&tri, !same, r.s, r.t);
            float4 ro = -r.direction;
            K::Vector_PutInSameHemisphereAs(&Ns, &ro);
            Ns = K::normalize(Ns);
            uint32_t b1 = r.numIntersectedBBx, t1 = r.numIntersectedTri;
            float4 L = K::Scene_ComputeDirectIllumination(seed, bvh, tris, lights, mats, tex, texd, sky, &r, &mat, &Ns);
            work[4 * n + 2] = r.numIntersectedBBx - b1;
            work[4 * n + 3] = r.numIntersectedTri - t1;
            K::Scene_ComputeRadiance(seed, bvh, tris, lights, mats, tex, texd, sky, &r, &tri, &mat, &L, &T, &Ng, &Ns);
            r.reflectionId++;
        }
        else
            active = false;
        n++;
        if (active && K::Vector_Max(&T) <= 0.001f) active = false;
    }
    return n;
}

int ptref_max_threads() { return omp_get_max_threads(); }

}  // extern "C"
