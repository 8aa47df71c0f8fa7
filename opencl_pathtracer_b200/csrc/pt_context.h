// pt_context.h — the context behind the opaque `ptc_ctx` of include/pathtracer_cuda.h, shared by pathtracer_cuda.cu (one
// device) and pt_multi.cu (a group of per-device contexts driven by one host thread).
#pragma once

#include <cstdarg>
#include <cstdio>
#include <string>
#include <vector>

#include "pathtracer_cuda.h"
#include "pt_launch.h"
#include "pt_wide.cuh"

namespace ptc
{
struct StageEvent
{
    int stage;  // 0 generate, 1 extend, 2 connect, 3 shade, 4 accumulate
    cudaEvent_t a, b;
};

template <typename T>
struct DevBuf
{
    T *p = nullptr;
    size_t n = 0;
    cudaError_t alloc(size_t count)
    {
        release();
        n = count;
        if (count == 0) count = 1;  // the reference clamps empty buffers to 1 byte (PathTracer_OpenCL.cpp:165-171)
        return cudaMalloc((void **)&p, count * sizeof(T));
    }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr;
        n = 0;
    }
    void adopt(T *ptr, size_t count)  // takes over a cudaMalloc'ed buffer
    {
        release();
        p = ptr;
        n = count;
    }
};

struct MultiGroup;  // pt_multi.cu
}  // namespace ptc

struct ptc_ctx
{
    // a multi-device context (ptc_create_multi) is a handle that owns one plain context per device; everything below the
    // group pointer is unused in the handle itself
    ptc::MultiGroup *group = nullptr;

    ptc_config cfg{};
    int device = 0;
    cudaStream_t stream = nullptr;
    bool ownStream = false;
    bool fast = false;
    bool sceneLoaded = false;
    std::string err;

    uint32_t nPix = 0, batchIters = 1;
    uint32_t stripeRows = 0, stripeRank = 0, stripeWorld = 1, ownedPix = 0;  // pixel shard (ptc_set_pixel_stripes); 0 rows = whole frame
    size_t capacity = 0;  // path slots
    ptc::LaunchCfg lc{};
    int smCount = 0;

    // scene
    ptc::DevBuf<ptc::DevNode> nodes;
    ptc::DevBuf<ptc::DevTri> tris;
    ptc::DevBuf<ptc::DevShade> shade;
    ptc::DevBuf<ptc::DevMaterial> materials;
    ptc::DevBuf<ptc::DevLight> lights;
    ptc::DevBuf<uint4> textures;
    ptc::DevBuf<uchar4> texels;
    ptc::DevBuf<uint32_t> bigLeaf;
    ptc::DevBuf<ptc::DevNode8P> nodes8;  // FAST + wide BVH only; padded records, see pt_wide.cuh
    ptc::DevBuf<ptc::DevTri> tris8;
    ptc::DevBuf<uint32_t> tri8Index;
    bool wide = false;
    bool gpuBvh = false;      // PTC_FLAG_GPU_BVH: the wide BVH is built on the device, the caller's BVH is ignored
    uint32_t wideDepth = 0;
    float msGpuBuild = 0.f;
    ptc::SceneDev sc{};
    ptc::Camera cam{};

    // path state + queues
    ptc::DevBuf<float4> org, dir, thr, rad, hit;
    ptc::DevBuf<int> seed;
    ptc::DevBuf<float2> sample;
    ptc::DevBuf<uint32_t> meta, nbbx, ntri, occl, queue[2], qpart, tileLit;
    ptc::PathState st{};
    ptc::DevBuf<ptc::Counters> counters;

    // accumulators
    ptc::DevBuf<float4> imgColor, imgV;
    ptc::DevBuf<float> imgRayNb;
    ptc::DevBuf<uint32_t> rayDepths, rayBbx, rayTri;
    ptc::ImageDev img{};
    ptc::DevBuf<uint8_t> rgb8;
    ptc::DevBuf<float> x2inv;  // ptc_set_supersampling: the chi-square table of the reference's `-D SUPER_SAMPLING` build; empty = off
    uint32_t x2invSize = 0;

    // host-side counters
    uint64_t launches = 0, launchesExtend = 0, launchesConnect = 0;

    // A launch-bound batch (progressive frames: ~34 kernels of a few microseconds each) is captured once into a CUDA graph and
    // replayed; only the generate and accumulate nodes carry per-batch arguments (iteration numbers, camera).
    struct BatchGraph
    {
        cudaGraph_t graph = nullptr;
        cudaGraphExec_t exec = nullptr;
        cudaGraphNode_t genNode = nullptr, accNode = nullptr;
        cudaKernelNodeParams genParams{}, accParams{};
        uint32_t nIter = 0, ownedPix = 0, stripeRows = 0, stripeRank = 0, stripeWorld = 0;
        uint64_t sceneVersion = 0;
        uint64_t launches = 0, launchesExtend = 0, launchesConnect = 0;  // per replay
    } batchGraph;
    uint64_t sceneVersion = 0;
    bool useGraphs = true;
    double msStage[5] = {0, 0, 0, 0, 0};
    std::vector<ptc::StageEvent> pending;
    std::vector<cudaEvent_t> eventPool;
};

namespace ptc
{
// error message of a failed ptc_create (no context exists yet): per thread
std::string &create_error();

inline int fail(ptc_ctx *c, int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (c) c->err = buf;
    else create_error() = buf;
    return code;
}

#define CK(call)                                                                                              \
    do                                                                                                        \
    {                                                                                                         \
        cudaError_t e__ = (call);                                                                             \
        if (e__ != cudaSuccess) return ptc::fail(ctx, PTC_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e__)); \
    } while (0)

// ---- multi-device groups (pt_multi.cu); every function takes the HANDLE context
int multi_upload_scene(ptc_ctx *h, const ptc_scene_view *scene);
int multi_set_camera(ptc_ctx *h, const float camera[16]);
int multi_reset(ptc_ctx *h);
int multi_render(ptc_ctx *h, uint32_t first, uint32_t n);
int multi_synchronize(ptc_ctx *h);
int multi_reduce(ptc_ctx *h);            // sum the members' accumulators into member 0 (no-op when nothing was rendered since)
int multi_read_counters(ptc_ctx *h, ptc_counters *out);
int multi_read_rgb8(ptc_ctx *h, uint8_t *rgb);  // rgb == nullptr: convert only (ptc_convert_rgb8)
ptc_ctx *multi_primary(ptc_ctx *h);      // member 0: the context that holds the reduced image and answers the probes
void multi_destroy(ptc_ctx *h);
// scene replication (pathtracer_cuda.cu): dst receives copies of src's device scene (peer copies), launch configuration and camera
int replicate_scene(ptc_ctx *dst, const ptc_ctx *src);
}  // namespace ptc
