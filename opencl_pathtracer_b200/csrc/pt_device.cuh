// pt_device.cuh — device-side records and launch parameters of the wavefront path tracer.
//
// HBM layout (all arrays 256-byte aligned by cudaMalloc, every record a multiple of 16 bytes so that
// each field group is one LDG.128):
//   DevNode   64 B  one per INNER node of the reference BVH2: both children's boxes + both child
//                   references, so one node visit = one 64-byte fetch (reference: 2 x 160-byte Node
//                   records, Kernel/PathTracer_FullKernel.cl:660-672)
//   DevTri    64 B  one per triangle, only what Triangle_Intersects needs (reference: a 336-byte copy,
//                   kernel.cl:640)
//   DevShade 128 B  one per triangle, what the shade stage needs once per hit (normals, UVs, materials)
//   DevMaterial 32 B, DevLight 64 B (reference layout), DevTexture 16 B
// Path state is SoA, indexed by path slot = localIteration * W*H + pixel.
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

namespace ptc
{
// child reference word: bit 31 = leaf. inner: index into DevNode[]. leaf: bits 30..4 = first triangle,
// bits 3..0 = triangle count (1..15), 0 = "big leaf" whose count is bigLeafCount[first].
constexpr uint32_t kLeafBit = 0x80000000u;
constexpr uint32_t kMaxTriangles = 1u << 27;
constexpr uint32_t kInvalidRef = 0xffffffffu;  // empty scene

struct __align__(16) DevNode
{
    float4 a;  // L.min.x L.min.y L.min.z L.max.x
    float4 b;  // L.max.y L.max.z R.min.x R.min.y
    float4 c;  // R.min.z R.max.x R.max.y R.max.z
    uint4 d;   // refL refR meta 0      meta: bits 1..0 cutAxis, bit 2 L empty, bit 3 R empty
};
static_assert(sizeof(DevNode) == 64, "DevNode");

struct __align__(16) DevTri
{
    float4 n;   // N.xyzw (w lane kept: SURVEY.md F8)
    float4 p;   // S1.xyz, d = dot4(N, S1)
    float4 e1;  // u = S2 - S1 (xyz), denom = 1 / (uv*uv - uu*vv)
    float4 e2;  // v = S3 - S1 (xyz), w = bits: MaterialType of the positive side (3..0) and of the negative side (7..4)
};
static_assert(sizeof(DevTri) == 64, "DevTri");

struct __align__(16) DevShade
{
    float4 n1, n2, n3;       // vertex normals
    float4 uvp12;            // UVP1.xy UVP2.xy
    float4 uvp3_uvn1;        // UVP3.xy UVN1.xy
    float4 uvn23;            // UVN2.xy UVN3.xy
    uint4 ids;               // materialWithPositiveNormalIndex, materialWithNegativeNormalIndex, id, 0
    uint4 pad;
};
static_assert(sizeof(DevShade) == 128, "DevShade");

struct __align__(16) DevMaterial
{
    float4 color;        // simpleColor
    float opacity;
    int textureId;
    int type;
    int isSimpleColor;
};
static_assert(sizeof(DevMaterial) == 32, "DevMaterial");

struct __align__(16) DevLight
{
    float4 position, direction, color;
    float power, cosInner, cosOuter;
    int type;
};
static_assert(sizeof(DevLight) == 64, "DevLight");

struct DevSky
{
    uint32_t w[6], h[6], off[6];
    float cosRot, sinRot;
};

struct SceneDev
{
    const DevNode *nodes;
    const DevTri *tris;
    const DevShade *shade;
    const DevMaterial *materials;
    const DevLight *lights;
    const uint4 *textures;  // width height offset 0
    const uchar4 *texels;
    const uint32_t *bigLeafCount;  // may be null
    // FAST mode with the wide BVH (pt_wide.cuh); null / 0 otherwise
    const struct DevNode8P *nodes8;  // 96-byte padded records (pt_wide.cuh)
    const DevTri *tris8;           // the intersection records again, in wide order
    const uint32_t *tri8Index;     // wide order -> BVH order
    uint32_t nNodes8;
    uint32_t rootRef;
    uint32_t nLights;
    uint32_t nTriangles;
    uint32_t areaJump;  // 16807^(2 * number of PTC_LIGHT_AREA lights) mod 2^32; 1 without area lights (pt_shade.cuh light_resolve)
    DevSky sky;
    // array bounds for the CHECKED build (-DPTC_CHECKED, opencl_pathtracer_b200/_lib_checked): every index the kernels form is
    // compared with these and a violation is recorded in Counters::faults (compute-sanitizer is closed on the GPU pool)
    uint32_t nInner, nMaterials, nTextures, nTexels, stackDepth;
    struct Counters *faultSink;
};

struct Camera
{
    float4 position, direction, right, up;
};

struct PathState
{
    float4 *org;      // ray origin (4 lanes)
    float4 *dir;      // ray direction (4 lanes, normalised)
    float4 *thr;      // transferFunction
    float4 *rad;      // radianceToCompute
    float4 *hit;      // (bits: triangle index or -1, s, t, tpar)   q = org + dir * tpar
    int *seed;        // RNG state
    float2 *sample;   // pixel-plane sample, kept for accumulation (kernel.cl:1333)
    uint32_t *meta;   // bits 7..0 reflectionId, bit 8 isInWater, bit 9 finished
    uint32_t *nbbx;   // numIntersectedBBx (camera ray + its shadow rays)
    uint32_t *ntri;   // numIntersectedTri
    uint32_t *occl;   // bit i = light i occluded at the current hit
    // the queue of the current bounce re-written by k_classify with every tile of 1024 entries partitioned: the paths whose hit reads
    // the direct light first (connect + the lit half of shade), then the others (misses; FAST: hits on glass / water, which never
    // read it); tileLit[tile] = lit entries of the tile
    uint32_t *qpart;
    uint32_t *tileLit;
    uint32_t capacity;  // path slots (bound of every queue entry; CHECKED build)
    uint32_t partitioned;  // connect and shade draw from qpart (LaunchCfg::partition)
};

// queue / cursor block, one set per bounce so nothing has to be reset inside a batch
constexpr int kMaxBounces = 64;
struct Counters
{
    uint32_t qcount[kMaxBounces + 1];     // live paths entering bounce b
    uint32_t curExtend[kMaxBounces + 1];  // work cursors of the persistent kernels
    uint32_t curConnect[kMaxBounces + 1];
    unsigned long long extensionRays, shadowRays, boxTests, triTests, paths;
    unsigned long long extBoxTests, extTriTests;  // share of boxTests / triTests spent by closest-hit rays
    uint32_t faults;        // CHECKED build: bit c set = check c failed somewhere (codes below); always 0 in the normal build
    uint32_t faultInfo[3];  // the first failure: code, offending value, bound
};

// CHECKED build (compute-sanitizer is closed on the GPU pool; SURVEY.md section 5 asks for memcheck evidence): bounds checks on
// every index the kernels form — queue entries, path slots, node / triangle / material / texture / texel indices, traversal stack
// levels, target pixels. A failure sets a bit; nothing traps, so one run reports every class of violation.
enum FaultCode : uint32_t
{
    kFaultQueueItem = 0, kFaultNodeIndex, kFaultTriIndex, kFaultStack, kFaultHitIndex, kFaultQueueAppend, kFaultMaterial, kFaultTexture,
    kFaultTexel, kFaultPixel, kFaultLight
};
#ifdef PTC_CHECKED
__device__ __forceinline__ void ptc_fault(Counters *cn, uint32_t code, uint32_t value, uint32_t bound)
{
    if (atomicOr(&cn->faults, 1u << code) == 0u)
    {
        cn->faultInfo[0] = code; cn->faultInfo[1] = value; cn->faultInfo[2] = bound;
    }
}
#define PTC_CHECK(cn, code, value, bound)                                               \
    do                                                                                  \
    {                                                                                   \
        if (!((uint32_t)(value) < (uint32_t)(bound))) ptc_fault(cn, code, (uint32_t)(value), (uint32_t)(bound)); \
    } while (0)
#else
#define PTC_CHECK(cn, code, value, bound) \
    do                                    \
    {                                     \
    } while (0)
#endif


struct FrameParams
{
    uint32_t width, height;
    uint32_t maxDepth;
    uint32_t sampler;
    uint32_t firstIteration;  // of this batch
    uint32_t nIterations;     // in this batch
    uint32_t flags;
    // pixel shard (ptc_set_pixel_stripes): this context renders the rows y with (y / stripeRows) % stripeWorld == stripeRank;
    // stripeRows = 0: the whole frame. ownedPix = pixels of the shard = path slots per iteration.
    uint32_t stripeRows, stripeRank, stripeWorld;
    uint32_t ownedPix;
    uint32_t countWork;  // per-path work counters are kept (st.nbbx / st.ntri feed the two 5000-bin histograms)
};

// pixel `i` of the shard (row-major over the owned rows) -> pixel of the frame
__device__ __forceinline__ uint32_t shard_pixel(const FrameParams &fp, uint32_t i)
{
    if (fp.stripeRows == 0u) return i;
    const uint32_t oy = i / fp.width, x = i - oy * fp.width;
    const uint32_t stripe = oy / fp.stripeRows, within = oy - stripe * fp.stripeRows;
    return ((stripe * fp.stripeWorld + fp.stripeRank) * fp.stripeRows + within) * fp.width + x;
}

struct ImageDev
{
    float4 *color;
    float *rayNb;
    float4 *variance;  // may be null
    uint32_t *rayDepths;
    uint32_t *rayBbx;
    uint32_t *rayTri;
};

}  // namespace ptc
