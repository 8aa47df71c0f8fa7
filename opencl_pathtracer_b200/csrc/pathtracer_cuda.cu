// pathtracer_cuda.cu — PathTracer_CUDA: implementation of the C ABI declared in include/pathtracer_cuda.h.
//
// Replaces the OpenCL context / buffer / launch-loop code of the reference
// (Controleur/PathTracer_OpenCL.cpp:56-198, :316-402). Host side only: the kernels are in pt_kernels.cu.
// There is no CPU path in this library: without a CUDA device ptc_create fails.
#include "pathtracer_cuda.h"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "pt_context.h"
#include "pt_lbvh.h"
#include "pt_wide_build.h"

using namespace ptc;

namespace ptc
{
std::string &create_error()
{
    thread_local std::string e;
    return e;
}
}  // namespace ptc

namespace
{
// A multi-device handle (ptc_create_multi, pt_multi.cu) answers the single-device calls through its member 0 — after the members'
// accumulators have been summed into it where the call reads them — and carries the member's error message over.
template <typename F>
int onPrimary(ptc_ctx *h, bool reduceFirst, F f)
{
    if (reduceFirst)
    {
        const int rc = multi_reduce(h);
        if (rc != PTC_OK) return rc;
    }
    ptc_ctx *p = multi_primary(h);
    const int rc = f(p);
    if (rc != PTC_OK) h->err = p->err;
    return rc;
}

inline float4 to4(const ptc_float4 &v) { return make_float4(v.x, v.y, v.z, v.w); }
inline float hdot4(const ptc_float4 &a, const ptc_float4 &b) { return ((a.x * b.x + a.y * b.y) + a.z * b.z) + a.w * b.w; }

cudaEvent_t getEvent(ptc_ctx *c)
{
    if (!c->eventPool.empty())
    {
        cudaEvent_t e = c->eventPool.back();
        c->eventPool.pop_back();
        return e;
    }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
}

struct StageTimer
{
    ptc_ctx *c;
    StageEvent ev{};
    bool on;
    StageTimer(ptc_ctx *ctx, int stage) : c(ctx), on((ctx->cfg.flags & PTC_FLAG_STAGE_TIMING) != 0)
    {
        if (!on) return;
        ev.stage = stage;
        ev.a = getEvent(c);
        ev.b = getEvent(c);
        cudaEventRecord(ev.a, c->stream);
    }
    ~StageTimer()
    {
        if (!on) return;
        cudaEventRecord(ev.b, c->stream);
        c->pending.push_back(ev);
    }
};

void drainTimers(ptc_ctx *c)
{
    for (auto &e : c->pending)
    {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, e.a, e.b) == cudaSuccess) c->msStage[e.stage] += ms;
        c->eventPool.push_back(e.a);
        c->eventPool.push_back(e.b);
    }
    c->pending.clear();
}

void dropBatchGraph(ptc_ctx *ctx)
{
    auto &g = ctx->batchGraph;
    if (g.exec) cudaGraphExecDestroy(g.exec);
    if (g.graph) cudaGraphDestroy(g.graph);
    g = ptc_ctx::BatchGraph{};
}
}  // namespace

// ================================================================================================ create
extern "C" {

#ifdef PTC_CHECKED
const char *ptc_version(void) { return "PathTracer_CUDA 0.2 (sm_100a) checked"; }
#else
const char *ptc_version(void) { return "PathTracer_CUDA 0.2 (sm_100a)"; }
#endif

int ptc_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

const char *ptc_last_error(const ptc_ctx *ctx) { return ctx ? ctx->err.c_str() : create_error().c_str(); }

void *ptc_host_alloc(size_t bytes)
{
    void *p = nullptr;
    if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) return nullptr;
    return p;
}
void ptc_host_free(void *p)
{
    if (p) cudaFreeHost(p);
}

int ptc_measure_read_bandwidth(ptc_ctx *ctx, size_t bytes, uint32_t repeats, double *gb_per_s)
{
    if (!ctx || !gb_per_s || bytes < 32 || repeats == 0) return PTC_ERR_INVALID;
    if (ctx->group) return onPrimary(ctx, false, [&](ptc_ctx *p) { return ptc_measure_read_bandwidth(p, bytes, repeats, gb_per_s); });
    CK(cudaSetDevice(ctx->device));
    DevBuf<uint4> buf;
    DevBuf<uint32_t> sink;
    bytes &= ~(size_t)31;
    CK(buf.alloc(bytes / 16));
    CK(sink.alloc(1));
    CK(cudaMemsetAsync(buf.p, 0x5a, bytes, ctx->stream));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    const int grid = ctx->smCount * 8;
    launch_read_stream(ctx->stream, grid, buf.p, bytes, 2, sink.p);  // warm-up (fills the L2 when the buffer fits)
    cudaEventRecord(e0, ctx->stream);
    launch_read_stream(ctx->stream, grid, buf.p, bytes, repeats, sink.p);
    cudaEventRecord(e1, ctx->stream);
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    buf.release();
    sink.release();
    ctx->launches += 2;
    if (e != cudaSuccess) return fail(ctx, PTC_ERR_CUDA, "ptc_measure_read_bandwidth: %s", cudaGetErrorString(e));
    *gb_per_s = ms > 0.f ? (double)bytes * repeats / (ms * 1e-3) / 1e9 : 0.0;
    return PTC_OK;
}

void ptc_destroy(ptc_ctx *ctx)
{
    if (!ctx) return;
    if (ctx->group)
    {
        multi_destroy(ctx);
        return;
    }
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    drainTimers(ctx);
    dropBatchGraph(ctx);
    for (auto e : ctx->eventPool) cudaEventDestroy(e);
    ctx->nodes.release(); ctx->tris.release(); ctx->shade.release(); ctx->materials.release(); ctx->lights.release();
    ctx->textures.release(); ctx->texels.release(); ctx->bigLeaf.release();
    ctx->nodes8.release(); ctx->tris8.release(); ctx->tri8Index.release();
    ctx->org.release(); ctx->dir.release(); ctx->thr.release(); ctx->rad.release(); ctx->hit.release();
    ctx->seed.release(); ctx->sample.release(); ctx->meta.release(); ctx->nbbx.release(); ctx->ntri.release();
    ctx->occl.release(); ctx->queue[0].release(); ctx->queue[1].release(); ctx->qpart.release(); ctx->tileLit.release(); ctx->counters.release();
    ctx->imgColor.release(); ctx->imgV.release(); ctx->imgRayNb.release(); ctx->rayDepths.release();
    ctx->rayBbx.release(); ctx->rayTri.release(); ctx->rgb8.release();
    if (ctx->ownStream && ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int ptc_reset_accumulation(ptc_ctx *ctx)
{
    if (!ctx) return PTC_ERR_INVALID;
    if (ctx->group) return multi_reset(ctx);
    CK(cudaSetDevice(ctx->device));
    launch_clear(ctx->stream, ctx->lc, ctx->img, ctx->nPix, ctx->counters.p);
    {
        const cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return fail(ctx, PTC_ERR_CUDA, "ptc_reset_accumulation: %s", cudaGetErrorString(e));
    }
    if (ctx->cfg.flags & PTC_FLAG_STAGE_TIMING)  // pending stage timers must be read before their totals are cleared
    {
        CK(cudaStreamSynchronize(ctx->stream));
        drainTimers(ctx);
    }
    ctx->launches = ctx->launchesExtend = ctx->launchesConnect = 0;
    for (double &m : ctx->msStage) m = 0;
    return PTC_OK;
}

// `-D SUPER_SAMPLING` of the reference (Controleur/PathTracer_OpenCL.cpp:311; Kernel/PathTracer_FullKernel.cl:1152-1172, 1219-1222).
int ptc_set_supersampling(ptc_ctx *ctx, const float *x2inv, uint32_t count)
{
    if (!ctx) return PTC_ERR_INVALID;
    if (ctx->group)
        return count == 0 ? PTC_OK : fail(ctx, PTC_ERR_STATE, "supersampling decides on the accumulators of the iterations before: not available on a multi-device context, whose devices render iteration ranges side by side");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    if (count == 0)
    {
        ctx->x2inv.release();
        ctx->x2invSize = 0;
        return PTC_OK;
    }
    if (!ctx->imgV.p) return fail(ctx, PTC_ERR_STATE, "supersampling reads imageV: the context was created with PTC_FLAG_NO_VARIANCE");
    std::vector<float> table(count);
    if (x2inv) std::memcpy(table.data(), x2inv, sizeof(float) * count);
    else
    {
        // 95 % quantile of the chi-square law with n degrees of freedom (Wilson-Hilferty); entry 0 is never the divisor of a
        // finite variance (a pixel holds >= 6 samples before the criterion is first evaluated)
        table[0] = 1.0f;
        for (uint32_t n = 1; n < count; n++)
        {
            const double a = 2.0 / (9.0 * n), c = 1.0 - a + 1.6448536269514722 * std::sqrt(a);
            table[n] = (float)(n * c * c * c);
        }
    }
    CK(ctx->x2inv.alloc(count));
    CK(cudaMemcpy(ctx->x2inv.p, table.data(), sizeof(float) * count, cudaMemcpyHostToDevice));
    ctx->x2invSize = count;
    return PTC_OK;
}

int ptc_create(const ptc_config *config, ptc_ctx **out_ctx)
{
    ptc_ctx *ctx = nullptr;  // CK() reports into create_error() while ctx is null
    if (!config || !out_ctx) return fail(nullptr, PTC_ERR_INVALID, "ptc_create: null argument");
    *out_ctx = nullptr;
    if (config->struct_size != sizeof(ptc_config)) return fail(nullptr, PTC_ERR_INVALID, "ptc_create: struct_size %u != %zu", config->struct_size, sizeof(ptc_config));
    if (config->image_width == 0 || config->image_height == 0)
        return fail(nullptr, PTC_ERR_INVALID, "Image dimensions must be positive");  // reference PathTracer.cpp:147-153
    if ((uint64_t)config->image_width * config->image_height >= (1ull << 27)) return fail(nullptr, PTC_ERR_INVALID, "image too large (one wavefront pass holds fewer than 2^27 paths)");
    if (config->ray_max_depth > (uint32_t)kMaxBounces) return fail(nullptr, PTC_ERR_LIMIT, "ray_max_depth %u exceeds %d", config->ray_max_depth, kMaxBounces);
    if (config->sampler > 2u || config->mode > 1u) return fail(nullptr, PTC_ERR_INVALID, "unknown sampler or mode");
    int nDev = 0;
    if (cudaGetDeviceCount(&nDev) != cudaSuccess || nDev == 0)
        return fail(nullptr, PTC_ERR_NO_DEVICE, "no CUDA device: PathTracer_CUDA has no CPU path");
    if (config->device < 0 || config->device >= nDev) return fail(nullptr, PTC_ERR_INVALID, "device %d out of range (%d devices)", config->device, nDev);
    CK(cudaSetDevice(config->device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, config->device));
    if (prop.major < 10) return fail(nullptr, PTC_ERR_NO_DEVICE, "device %d is sm_%d%d; this library is built for sm_100a only", config->device, prop.major, prop.minor);

    ptc_ctx *c = new ptc_ctx();
    c->cfg = *config;
    c->device = config->device;
    c->fast = config->mode == PTC_MODE_FAST;
    c->wide = c->fast && !(config->flags & PTC_FLAG_BVH2);
    if (const char *e = std::getenv("PTC_WIDE")) c->wide = c->fast && std::atoi(e) != 0;
    if ((config->flags & PTC_FLAG_RUSSIAN_ROULETTE) && !c->fast)
    {
        delete c;
        return fail(nullptr, PTC_ERR_INVALID, "PTC_FLAG_RUSSIAN_ROULETTE needs PTC_MODE_FAST (PARITY reproduces the reference's random streams)");
    }
    if ((config->flags & PTC_FLAG_TEXTURE_BILINEAR) && !c->fast)
    {
        delete c;
        return fail(nullptr, PTC_ERR_INVALID, "PTC_FLAG_TEXTURE_BILINEAR needs PTC_MODE_FAST (PARITY reproduces the reference's nearest-texel lookup)");
    }
    if (config->flags & PTC_FLAG_NO_GRAPH) c->useGraphs = false;
    c->gpuBvh = (config->flags & PTC_FLAG_GPU_BVH) != 0;
    if (c->gpuBvh && (!c->fast || (config->flags & PTC_FLAG_BVH2) || !c->wide))
    {
        delete c;
        return fail(nullptr, PTC_ERR_INVALID, "PTC_FLAG_GPU_BVH needs PTC_MODE_FAST without PTC_FLAG_BVH2 (PARITY walks the reference BVH2)");
    }
    c->nPix = config->image_width * config->image_height;
    c->ownedPix = c->nPix;
    c->smCount = prop.multiProcessorCount;
    if (config->stream) c->stream = (cudaStream_t)config->stream;
    else
    {
        if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess)
        {
            delete c;
            return fail(nullptr, PTC_ERR_CUDA, "cudaStreamCreate failed");
        }
        c->ownStream = true;
    }
    // iterations kept in flight per wavefront pass: auto = enough for >= 16M paths (1.9 GB of path state). Measured on the
    // bench scene at 1080p: 1 / 2 / 4 / 8 / 16 iterations per pass = 117 / 124 / 128 / 130 / 131 Msamples/s (the tail of every
    // persistent launch is amortised over more rays; profiles/r01_sweeps.txt)
    uint32_t b = config->iterations_per_batch;
    if (b == 0) b = ((1u << 24) + c->nPix - 1) / c->nPix;
    if (b > 64) b = 64;
    while ((uint64_t)b * c->nPix >= (1ull << 27) && b > 1) b--;  // staged items carry the path slot in 27 bits
    c->batchIters = b;
    c->capacity = (size_t)b * c->nPix;
    c->lc.stackDepth = 32;
    c->lc.wide = false;
    c->lc.count = !c->fast || (config->flags & PTC_FLAG_WORK_COUNTERS);
    c->lc.traverseGrid = c->smCount * traverse_ctas_per_sm(c->fast, false, c->lc.stackDepth, false);
    c->lc.connectGrid = c->smCount * traverse_ctas_per_sm(c->fast, false, c->lc.stackDepth, true);
    c->lc.streamGrid = c->smCount * 8;
    c->lc.refillMin = 4;  // idle lanes that trigger a refill from the warp's staged rays (profiles/r01_sweeps.txt)
    if (const char *e = std::getenv("PTC_GRAPHS")) c->useGraphs = std::atoi(e) != 0;
    if (const char *e = std::getenv("PTC_REFILL_MIN")) c->lc.refillMin = (uint32_t)std::max(1, std::min(32, std::atoi(e)));
    {
        // phase votes of k_trace, in eighths of the warp's rays: node steps while >= voteNode/8 wait for a node, triangle steps
        // (after the first) while > voteTri/8 wait for a triangle
        // (sweeps in profiles/r01_sweeps.txt: BVH2 walk 4 / 4; wide walk, whose lanes can park a triangle group, 5 / 3)
        uint32_t vn = c->wide ? 5 : 4, vt = c->wide ? 3 : 4;
        if (const char *e = std::getenv("PTC_VOTE_NODE")) vn = (uint32_t)std::max(0, std::min(8, std::atoi(e)));
        if (const char *e = std::getenv("PTC_VOTE_TRI")) vt = (uint32_t)std::max(0, std::min(8, std::atoi(e)));
        c->lc.refillMin |= (vn << 8) | (vt << 16);
    }
    if (const char *e = std::getenv("PTC_CTAS_PER_SM")) c->lc.traverseGrid = c->lc.connectGrid = c->smCount * std::max(1, std::atoi(e));

    cudaError_t e = cudaSuccess;
    auto A = [&](cudaError_t r) { if (e == cudaSuccess) e = r; };
    const size_t n = c->capacity;
    A(c->org.alloc(n)); A(c->dir.alloc(n)); A(c->thr.alloc(n)); A(c->rad.alloc(n)); A(c->hit.alloc(n));
    A(c->seed.alloc(n)); A(c->sample.alloc(n)); A(c->meta.alloc(n)); A(c->nbbx.alloc(n)); A(c->ntri.alloc(n));
    A(c->occl.alloc(n)); A(c->queue[0].alloc(n)); A(c->queue[1].alloc(n)); A(c->qpart.alloc(n)); A(c->tileLit.alloc(n / 1024 + 1)); A(c->counters.alloc(1));
    A(c->imgColor.alloc(c->nPix)); A(c->imgRayNb.alloc(c->nPix));
    if (!(config->flags & PTC_FLAG_NO_VARIANCE)) A(c->imgV.alloc(c->nPix));
    A(c->rayDepths.alloc(kMaxBounces + 1)); A(c->rayBbx.alloc(PTC_MAX_INTERSECTION_NUMBER)); A(c->rayTri.alloc(PTC_MAX_INTERSECTION_NUMBER));
    if (e != cudaSuccess)
    {
        std::string msg = std::string("device allocation failed: ") + cudaGetErrorString(e);
        ptc_destroy(c);
        return fail(nullptr, PTC_ERR_CUDA, "%s", msg.c_str());
    }
    c->st = PathState{c->org.p, c->dir.p, c->thr.p, c->rad.p, c->hit.p, c->seed.p, c->sample.p, c->meta.p, c->nbbx.p, c->ntri.p, c->occl.p, c->qpart.p, c->tileLit.p, (uint32_t)c->capacity, 0u};
#ifdef PTC_CHECKED
    // self-test of the checked build: with a deliberately wrong bound every queue entry beyond slot 0 must be reported
    if (std::getenv("PTC_CHECKED_SELFTEST")) c->st.capacity = 1u;
#endif
    c->img = ImageDev{c->imgColor.p, c->imgRayNb.p, c->imgV.p, c->rayDepths.p, c->rayBbx.p, c->rayTri.p};
    // the reference never zeroes its accumulators; this library always does
    int rc = ptc_reset_accumulation(c);
    if (rc != PTC_OK)
    {
        create_error() = c->err;
        ptc_destroy(c);
        return rc;
    }
    *out_ctx = c;
    return PTC_OK;
}

// ================================================================================================ scene upload
// Flattening: reference AoS records (include/pathtracer_structs.h) -> packed device records (pt_device.cuh).
//   * one 64-byte DevNode per INNER node: both children's pMin/pMax + both child references + cutAxis
//   * one 64-byte DevTri per triangle: N (4 lanes), S1 + d = dot4(N,S1), u + denom, v   (Triangle_Intersects :528-556)
//   * one 128-byte DevShade per triangle: vertex normals, UVs, material ids
// d and denom are ray independent; they are computed here with the same IEEE single-precision operations
// and association order the kernel of the reference uses (this file is compiled without FMA contraction
// on the host), so the device sees the values the reference would recompute per test.
int ptc_upload_scene(ptc_ctx *ctx, const ptc_scene_view *s)
{
    if (!ctx || !s) return PTC_ERR_INVALID;
    if (ctx->group) return multi_upload_scene(ctx, s);
    if (s->struct_size != sizeof(ptc_scene_view)) return fail(ctx, PTC_ERR_INVALID, "ptc_upload_scene: struct_size mismatch");
    const uint32_t nT = s->triangulation_size, nN = ctx->gpuBvh ? 0u : s->bvh_size, nL = s->lights_size, nM = s->materiaux_size, nX = s->textures_size;
    const uint32_t nTexel = s->textures_data_size;
    if ((nT && !s->triangulation) || (nN && !s->bvh) || (nL && !s->lights) || (nM && !s->materiaux) || (nX && !s->textures) ||
        (nTexel && !s->textures_data) || !s->sky)
        return fail(ctx, PTC_ERR_INVALID, "ptc_upload_scene: null array with non-zero size");
    if (nL >= PTC_MAX_LIGHT_SIZE)  // reference PathTracer.cpp:60-65
        return fail(ctx, PTC_ERR_LIMIT, "The scene has too many lights : it has %u and should be under %u", nL, PTC_MAX_LIGHT_SIZE);
    if (nT >= kMaxTriangles) return fail(ctx, PTC_ERR_LIMIT, "too many triangles (%u, limit %u)", nT, kMaxTriangles);
    const bool gpuBvh = ctx->gpuBvh;
    if (nT && nN == 0 && !gpuBvh) return fail(ctx, PTC_ERR_INVALID, "triangles without a BVH");
    if ((uint64_t)ctx->capacity * (nL ? nL : 1) >= (1ull << 32))
        return fail(ctx, PTC_ERR_LIMIT, "paths in flight (%zu) x lights (%u) exceeds 2^32: lower iterations_per_batch", ctx->capacity, nL);
    CK(cudaSetDevice(ctx->device));

    // ---- BVH: validate, measure depth, number the inner nodes in array (= depth-first) order
    std::vector<uint32_t> innerIndex(nN, 0xffffffffu);
    uint32_t nInner = 0, maxDepth = 0;
    std::vector<uint32_t> bigLeaf;
    if (nT && !gpuBvh)
    {
        std::vector<uint8_t> seen(nN, 0);
        std::vector<std::pair<uint32_t, uint32_t>> stack;  // node, depth
        stack.push_back({0u, 0u});
        while (!stack.empty())
        {
            auto [id, depth] = stack.back();
            stack.pop_back();
            if (id >= nN) return fail(ctx, PTC_ERR_INVALID, "BVH child index %u out of range (%u nodes)", id, nN);
            if (seen[id]) return fail(ctx, PTC_ERR_INVALID, "BVH node %u reachable twice", id);
            seen[id] = 1;
            const ptc_node &n = s->bvh[id];
            if (n.isLeaf)
            {
                if ((uint64_t)n.triangleStartIndex + n.nbTriangles > nT) return fail(ctx, PTC_ERR_INVALID, "BVH leaf %u range out of bounds", id);
                if (depth > maxDepth) maxDepth = depth;
            }
            else
            {
                stack.push_back({n.son2Id, depth + 1});
                stack.push_back({n.son1Id, depth + 1});
            }
        }
        if (maxDepth >= PTC_BVH_MAX_DEPTH)  // reference PathTracer.cpp:54-59
            return fail(ctx, PTC_ERR_LIMIT, "BVH is too deep : it has a size of %u and should be under %u", maxDepth, PTC_BVH_MAX_DEPTH);
        for (uint32_t i = 0; i < nN; i++)
            if (seen[i] && !s->bvh[i].isLeaf) innerIndex[i] = nInner++;
    }
    auto childRef = [&](uint32_t id) -> uint32_t {
        const ptc_node &n = s->bvh[id];
        if (!n.isLeaf) return innerIndex[id];
        uint32_t cnt = n.nbTriangles;
        if (cnt == 0 || cnt > 15)
        {
            if (bigLeaf.empty()) bigLeaf.assign(nT + 1, 0u);
            bigLeaf[n.triangleStartIndex] = cnt;  // a zero-triangle leaf shares its start with another leaf at most when cnt == 0
            cnt = 0;
        }
        return kLeafBit | (n.triangleStartIndex << 4) | cnt;
    };
    std::vector<DevNode> hNodes(nInner);
    uint32_t rootRef = kInvalidRef;
    if (nT && !gpuBvh)
    {
        // zero-triangle leaves: give them an empty range that cannot alias a real big leaf
        for (uint32_t i = 0; i < nN; i++)
        {
            if (innerIndex[i] == 0xffffffffu) continue;
            const ptc_node &n = s->bvh[i];
            const ptc_node &L = s->bvh[n.son1Id], &R = s->bvh[n.son2Id];
            DevNode d;
            d.a = make_float4(L.trianglesAABB.pMin.x, L.trianglesAABB.pMin.y, L.trianglesAABB.pMin.z, L.trianglesAABB.pMax.x);
            d.b = make_float4(L.trianglesAABB.pMax.y, L.trianglesAABB.pMax.z, R.trianglesAABB.pMin.x, R.trianglesAABB.pMin.y);
            d.c = make_float4(R.trianglesAABB.pMin.z, R.trianglesAABB.pMax.x, R.trianglesAABB.pMax.y, R.trianglesAABB.pMax.z);
            uint32_t meta = n.cutAxis & 3u;
            // an empty child box fails BoundingBox_Intersects (:68); a zero-triangle leaf has nothing to test either
            if (L.trianglesAABB.isEmpty || (L.isLeaf && L.nbTriangles == 0)) meta |= 4u;
            if (R.trianglesAABB.isEmpty || (R.isLeaf && R.nbTriangles == 0)) meta |= 8u;
            if (n.cutAxis > 2u) return fail(ctx, PTC_ERR_INVALID, "BVH node %u has cutAxis %u", i, n.cutAxis);
            d.d = make_uint4((meta & 4u) ? (kLeafBit | 1u) : childRef(n.son1Id), (meta & 8u) ? (kLeafBit | 1u) : childRef(n.son2Id), meta, 0u);
            hNodes[innerIndex[i]] = d;
        }
        const ptc_node &root = s->bvh[0];
        if (root.isLeaf && root.nbTriangles == 0) rootRef = kInvalidRef;
        else rootRef = childRef(0);
    }

    // ---- triangles
    std::vector<DevTri> hTris(nT);
    std::vector<DevShade> hShade(nT);
    std::vector<float> hVerts(gpuBvh ? 9 * (size_t)nT : 0);  // device builder input: S1, S2, S3
    int badW = -1, badMat = -1;
#pragma omp parallel for schedule(static) reduction(max : badW, badMat)
    for (long long i = 0; i < (long long)nT; i++)
    {
        const ptc_triangle &t = s->triangulation[i];
        if (!(t.S1.w == t.S2.w && t.S1.w == t.S3.w)) badW = std::max(badW, (int)i);
        if (t.materialWithPositiveNormalIndex >= nM || t.materialWithNegativeNormalIndex >= nM) badMat = std::max(badMat, (int)i);
        const ptc_float4 u = {t.S2.x - t.S1.x, t.S2.y - t.S1.y, t.S2.z - t.S1.z, t.S2.w - t.S1.w};
        const ptc_float4 v = {t.S3.x - t.S1.x, t.S3.y - t.S1.y, t.S3.z - t.S1.z, t.S3.w - t.S1.w};
        const float uv = hdot4(u, v), uu = hdot4(u, u), vv = hdot4(v, v);
        const float denom = 1 / (uv * uv - uu * vv);
        DevTri d;
        d.n = to4(t.N);
        d.p = make_float4(t.S1.x, t.S1.y, t.S1.z, hdot4(t.N, t.S1));
        d.e1 = make_float4(u.x, u.y, u.z, denom);
        uint32_t types = 0x55u;  // MAT_UNKNOWN on both sides if the indices are bad (reported below)
        if (t.materialWithPositiveNormalIndex < nM && t.materialWithNegativeNormalIndex < nM)
            types = ((uint32_t)s->materiaux[t.materialWithPositiveNormalIndex].type & 15u) |
                    (((uint32_t)s->materiaux[t.materialWithNegativeNormalIndex].type & 15u) << 4);
        float typesAsFloat;
        std::memcpy(&typesAsFloat, &types, 4);
        d.e2 = make_float4(v.x, v.y, v.z, typesAsFloat);
        hTris[i] = d;
        DevShade h;
        h.n1 = to4(t.N1); h.n2 = to4(t.N2); h.n3 = to4(t.N3);
        h.uvp12 = make_float4(t.UVP1.x, t.UVP1.y, t.UVP2.x, t.UVP2.y);
        h.uvp3_uvn1 = make_float4(t.UVP3.x, t.UVP3.y, t.UVN1.x, t.UVN1.y);
        h.uvn23 = make_float4(t.UVN2.x, t.UVN2.y, t.UVN3.x, t.UVN3.y);
        h.ids = make_uint4(t.materialWithPositiveNormalIndex, t.materialWithNegativeNormalIndex, t.id, 0u);
        h.pad = make_uint4(0, 0, 0, 0);
        hShade[i] = h;
        if (gpuBvh)
        {
            float *v = &hVerts[9 * (size_t)i];
            v[0] = t.S1.x; v[1] = t.S1.y; v[2] = t.S1.z; v[3] = t.S2.x; v[4] = t.S2.y; v[5] = t.S2.z; v[6] = t.S3.x; v[7] = t.S3.y; v[8] = t.S3.z;
        }
    }
    if (badW >= 0) return fail(ctx, PTC_ERR_UNSUPPORTED, "triangle %d: S1.w, S2.w, S3.w differ (vertices must share one w lane)", badW);
    if (badMat >= 0) return fail(ctx, PTC_ERR_INVALID, "triangle %d: material index out of range", badMat);

    // ---- materials, textures, lights, sky
    std::vector<DevMaterial> hMat(nM);
    for (uint32_t i = 0; i < nM; i++)
    {
        const ptc_material &m = s->materiaux[i];
        hMat[i].color = to4(m.simpleColor);
        hMat[i].opacity = m.opacity;
        hMat[i].textureId = m.textureId;
        hMat[i].type = m.type;
        hMat[i].isSimpleColor = m.isSimpleColor ? 1 : 0;
        if (!m.isSimpleColor && (m.textureId < 0 || (uint32_t)m.textureId >= nX))
            return fail(ctx, PTC_ERR_INVALID, "material %u: texture index %d out of range", i, m.textureId);
    }
    auto texOk = [&](const ptc_texture &t) { return t.width && t.height && (uint64_t)t.offset + (uint64_t)t.width * t.height <= nTexel; };
    std::vector<uint4> hTex(nX);
    for (uint32_t i = 0; i < nX; i++)
    {
        if (!texOk(s->textures[i])) return fail(ctx, PTC_ERR_INVALID, "texture %u exceeds texturesData", i);
        hTex[i] = make_uint4(s->textures[i].width, s->textures[i].height, s->textures[i].offset, 0u);
    }
    DevSky sky{};
    for (int f = 0; f < 6; f++)
    {
        if (!texOk(s->sky->skyTextures[f])) return fail(ctx, PTC_ERR_INVALID, "sky face %d exceeds texturesData", f);
        sky.w[f] = s->sky->skyTextures[f].width;
        sky.h[f] = s->sky->skyTextures[f].height;
        sky.off[f] = s->sky->skyTextures[f].offset;
    }
    sky.cosRot = s->sky->cosRotationAngle;
    sky.sinRot = s->sky->sinRotationAngle;
    static_assert(sizeof(DevLight) == sizeof(ptc_light), "Light is uploaded as is");
    // PTC_LIGHT_AREA (extension, include/pathtracer_structs.h): the reserved lane receives the light's offset into the path's
    // random stream, 16807^(2k) for the k-th area light (light_resolve in pt_shade.cuh)
    std::vector<ptc_light> hLights(s->lights, s->lights + nL);
    uint32_t areaJump = 1u;
    for (uint32_t i = 0; i < nL; i++)
    {
        if (hLights[i].type != PTC_LIGHT_AREA) continue;
        if (!ctx->fast) return fail(ctx, PTC_ERR_UNSUPPORTED, "light %u is an area light: an extension of FAST mode, the reference has none", i);
        const ptc_light &l = hLights[i];
        const float cx = l.direction.y * l.cosOfInnerFallOffAngle - l.direction.z * l.direction.w;
        const float cy = l.direction.z * l.position.w - l.direction.x * l.cosOfInnerFallOffAngle;
        const float cz = l.direction.x * l.direction.w - l.direction.y * l.position.w;
        if (!(cx * cx + cy * cy + cz * cz > 0.f)) return fail(ctx, PTC_ERR_INVALID, "area light %u has no area", i);
        std::memcpy(&hLights[i].cosOfOuterFallOffAngle, &areaJump, sizeof(float));
        areaJump *= 16807u * 16807u;
    }

    // ---- FAST mode: the 8-wide compressed BVH (pt_wide_build.cu) and the intersection records in its order
    WideBvh wideBvh;
    std::vector<DevTri> hTris8;
    if (ctx->wide && !gpuBvh)
    {
        std::string werr;
        if (!build_wide_bvh(s, wideBvh, werr)) return fail(ctx, PTC_ERR_INVALID, "%s", werr.c_str());
        if (wideBvh.depth + 1 > 64u) return fail(ctx, PTC_ERR_LIMIT, "wide BVH of %u levels", wideBvh.depth);
        hTris8.resize(wideBvh.triOrder.size());
#pragma omp parallel for schedule(static)
        for (long long i = 0; i < (long long)hTris8.size(); i++)
        {
            if (wideBvh.triOrder[i] != kNoTriangle) hTris8[i] = hTris[wideBvh.triOrder[i]];
            else std::memset(&hTris8[i], 0, sizeof(DevTri));  // never fetched (no bit in DevNode8::valid)
        }
    }

    // ---- upload (pending renders use the old buffers: wait for them first)
    CK(cudaStreamSynchronize(ctx->stream));
    CK(ctx->nodes8.alloc(wideBvh.nodes.size())); CK(ctx->tris8.alloc(hTris8.size())); CK(ctx->tri8Index.alloc(wideBvh.triOrder.size()));
    CK(ctx->nodes.alloc(nInner)); CK(ctx->tris.alloc(nT)); CK(ctx->shade.alloc(nT)); CK(ctx->materials.alloc(nM));
    CK(ctx->lights.alloc(nL)); CK(ctx->textures.alloc(nX)); CK(ctx->texels.alloc(nTexel)); CK(ctx->bigLeaf.alloc(bigLeaf.size()));
    auto up = [&](void *dst, const void *src, size_t bytes) { return bytes ? cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream) : cudaSuccess; };
    CK(up(ctx->nodes.p, hNodes.data(), sizeof(DevNode) * nInner));
    CK(up(ctx->tris.p, hTris.data(), sizeof(DevTri) * nT));
    CK(up(ctx->shade.p, hShade.data(), sizeof(DevShade) * nT));
    CK(up(ctx->materials.p, hMat.data(), sizeof(DevMaterial) * nM));
    CK(up(ctx->lights.p, hLights.data(), sizeof(DevLight) * nL));
    CK(up(ctx->textures.p, hTex.data(), sizeof(uint4) * nX));
    CK(up(ctx->texels.p, s->textures_data, sizeof(uchar4) * nTexel));
    CK(up(ctx->bigLeaf.p, bigLeaf.data(), sizeof(uint32_t) * bigLeaf.size()));
    if (!wideBvh.nodes.empty())
        CK(cudaMemcpy2DAsync(ctx->nodes8.p, sizeof(DevNode8P), wideBvh.nodes.data(), sizeof(DevNode8), sizeof(DevNode8), wideBvh.nodes.size(),
                             cudaMemcpyHostToDevice, ctx->stream));
    CK(up(ctx->tris8.p, hTris8.data(), sizeof(DevTri) * hTris8.size()));
    CK(up(ctx->tri8Index.p, wideBvh.triOrder.data(), sizeof(uint32_t) * wideBvh.triOrder.size()));
    CK(cudaStreamSynchronize(ctx->stream));  // host staging vectors die at return
    uint32_t nNodes8 = (uint32_t)wideBvh.nodes.size(), wideDepth = wideBvh.depth;
    ctx->msGpuBuild = 0.f;
    if (gpuBvh)
    {
        // ---- PTC_FLAG_GPU_BVH: Morton-order LBVH + collapse on the device (pt_lbvh.cu)
        DevBuf<float> dVerts;
        CK(dVerts.alloc(hVerts.size()));
        CK(up(dVerts.p, hVerts.data(), sizeof(float) * hVerts.size()));
        DeviceWideBvh dw;
        std::string werr;
        const bool ok = build_wide_bvh_device(ctx->stream, dVerts.p, ctx->tris.p, nT, dw, werr);
        dVerts.release();
        if (!ok) return fail(ctx, PTC_ERR_CUDA, "%s", werr.c_str());
        if (dw.depth + 2 > 64u)
        {
            cudaFree(dw.nodes); cudaFree(dw.tris8); cudaFree(dw.tri8Index);
            return fail(ctx, PTC_ERR_LIMIT, "wide BVH of %u levels", dw.depth);
        }
        {
            // the device builder emits 80-byte records: re-pitch them to the padded layout the traversal reads
            cudaError_t e = ctx->nodes8.alloc(dw.nNodes);
            if (e == cudaSuccess && dw.nNodes)
                e = cudaMemcpy2DAsync(ctx->nodes8.p, sizeof(DevNode8P), dw.nodes, sizeof(DevNode8), sizeof(DevNode8), dw.nNodes, cudaMemcpyDeviceToDevice, ctx->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
            cudaFree(dw.nodes);
            if (e != cudaSuccess)
            {
                cudaFree(dw.tris8); cudaFree(dw.tri8Index);
                return fail(ctx, PTC_ERR_CUDA, "wide BVH re-pitch: %s", cudaGetErrorString(e));
            }
        }
        ctx->tris8.adopt(dw.tris8, (size_t)dw.nNodes * kTriSlots);
        ctx->tri8Index.adopt(dw.tri8Index, (size_t)dw.nNodes * kTriSlots);
        nNodes8 = dw.nNodes;
        wideDepth = dw.depth;
        ctx->msGpuBuild = dw.msBuild;
    }
    ctx->wideDepth = wideDepth;

    SceneDev &sc = ctx->sc;
    sc.nodes = ctx->nodes.p; sc.tris = ctx->tris.p; sc.shade = ctx->shade.p; sc.materials = ctx->materials.p;
    sc.lights = ctx->lights.p; sc.textures = ctx->textures.p; sc.texels = ctx->texels.p;
    sc.bigLeafCount = bigLeaf.empty() ? nullptr : ctx->bigLeaf.p;
    sc.nodes8 = ctx->wide ? ctx->nodes8.p : nullptr; sc.tris8 = ctx->wide ? ctx->tris8.p : nullptr;
    sc.tri8Index = ctx->wide ? ctx->tri8Index.p : nullptr; sc.nNodes8 = ctx->wide ? nNodes8 : 0u;
    ctx->lc.wide = ctx->wide;
    sc.rootRef = rootRef; sc.nLights = nL; sc.areaJump = areaJump; sc.nTriangles = nT; sc.sky = sky;
    sc.nInner = nInner; sc.nMaterials = nM; sc.nTextures = nX; sc.nTexels = nTexel; sc.faultSink = ctx->counters.p;
    ctx->cam = Camera{to4(s->camera_position), to4(s->camera_direction), to4(s->camera_right), to4(s->camera_up)};
    // traversal stack: one entry per level below the root is enough (a push happens only when descending)
    if (!std::getenv("PTC_FULL_STACK"))
    {
        ctx->lc.stackDepth = (int)std::min<uint32_t>(32u, std::max<uint32_t>(4u, maxDepth + 1));
        // The wide walk pushes the unvisited siblings of the child it descends into, and the root has none: at most wideDepth - 1 sibling
        // groups are ever stacked, above the kPark slots of parked triangle groups (one spare level; the CHECKED build tests the bound).
        // Round 2 sized this wideDepth + 2: the two unused levels (2 KB per CTA) pushed connect's 7 CTAs per SM past the 164 KB shared-memory
        // carve-out, i.e. halved its L1 (profiles/r02_sweeps.txt T)
        if (ctx->wide) ctx->lc.stackDepth = (int)std::max<uint32_t>(3u, wideDepth + WideLane::kPark + WideLane::kItemLevels);  // + the item word's level
        if (const char *e = std::getenv("PTC_STACK_EXTRA")) ctx->lc.stackDepth = std::max(3, ctx->lc.stackDepth + std::atoi(e));  // experiments
        // a tree too deep for one CTA's shared memory is refused here, with a message, not at the first launch with "invalid argument"
        if (traverse_shared_bytes(ctx->wide, ctx->lc.stackDepth, true) > 227u * 1024u)
            return fail(ctx, PTC_ERR_LIMIT, "BVH of %d levels: the traversal stacks (%zu bytes per CTA) exceed the 227 KB of shared memory", ctx->lc.stackDepth,
                        traverse_shared_bytes(ctx->wide, ctx->lc.stackDepth, true));
        if (!std::getenv("PTC_CTAS_PER_SM"))
        {
            ctx->lc.traverseGrid = ctx->smCount * traverse_ctas_per_sm(ctx->fast, ctx->wide, ctx->lc.stackDepth, false);
            ctx->lc.connectGrid = ctx->smCount * traverse_ctas_per_sm(ctx->fast, ctx->wide, ctx->lc.stackDepth, true);
        }
    }
    sc.stackDepth = (uint32_t)ctx->lc.stackDepth;
    // class-partitioned queues (k_classify): where the scene is small enough to live in L1 the traversal is cheap, shade is the
    // largest stage and instruction bound, and class-uniform warps pay (+28 % samples/s on the Cornell box); on larger scenes shade
    // is a DRAM-bound minority stage and the partition costs sector efficiency (-6 % on the 262 k-triangle atrium). FAST only: in PARITY
    // every hit is lit, only misses would separate, and the pass costs more than it saves (Cornell PARITY 940 -> 900 Msamples/s).
    // PTC_PARTITION=0/1 overrides.
    ctx->lc.partition = ctx->fast && nT <= 4096u;
    if (const char *e = std::getenv("PTC_PARTITION")) ctx->lc.partition = std::atoi(e) != 0;
    // connect alone gains from the pass on every scene: it stages shadow rays by groups of 32 paths, and in the partitioned queue a
    // group holds 32 LIT paths (positions tell the class) instead of whatever the bounce's queue holds — connect -3.5 % on the bench
    // scene, -5 % on the 10 M-triangle scene, for 0.05 ms of k_classify per bounce (profiles/r02_sweeps.txt Y). PTC_PARTITION_CONNECT=0/1.
    ctx->lc.partitionConnect = ctx->fast && ctx->sc.nLights > 0;
    if (const char *e = std::getenv("PTC_PARTITION_CONNECT")) ctx->lc.partitionConnect = std::atoi(e) != 0;
    ctx->st.partitioned = (ctx->lc.partition || ctx->lc.partitionConnect) ? 1u : 0u;
    ctx->sceneLoaded = true;
    ctx->sceneVersion++;  // device pointers and launch configuration may have changed: a captured batch graph is stale
    return PTC_OK;
}

}  // extern "C"

// Multi-device contexts: the scene is flattened and uploaded once (member 0) and copied device to device to the other members
// (cudaMemcpyPeer: NVLink where peer access exists), so the host-side work of ptc_upload_scene — validation, wide-BVH collapse,
// 10 M-triangle flattening — is not repeated per GPU.
namespace ptc
{
int replicate_scene(ptc_ctx *dst, const ptc_ctx *src)
{
    ptc_ctx *ctx = dst;
    CK(cudaSetDevice(dst->device));
    CK(cudaStreamSynchronize(dst->stream));
    cudaError_t e = cudaSuccess;
    auto rep = [&](auto &d, const auto &s_) {
        if (e != cudaSuccess) return;
        e = d.alloc(s_.n);
        if (e == cudaSuccess && s_.n)
            e = cudaMemcpyPeerAsync(d.p, dst->device, s_.p, src->device, s_.n * sizeof(*s_.p), dst->stream);
    };
    rep(dst->nodes, src->nodes); rep(dst->tris, src->tris); rep(dst->shade, src->shade); rep(dst->materials, src->materials);
    rep(dst->lights, src->lights); rep(dst->textures, src->textures); rep(dst->texels, src->texels); rep(dst->bigLeaf, src->bigLeaf);
    rep(dst->nodes8, src->nodes8); rep(dst->tris8, src->tris8); rep(dst->tri8Index, src->tri8Index);
    if (e == cudaSuccess) e = cudaStreamSynchronize(dst->stream);
    if (e != cudaSuccess) return fail(dst, PTC_ERR_CUDA, "scene replication to device %d: %s", dst->device, cudaGetErrorString(e));
    SceneDev sc = src->sc;
    sc.nodes = dst->nodes.p; sc.tris = dst->tris.p; sc.shade = dst->shade.p; sc.materials = dst->materials.p;
    sc.lights = dst->lights.p; sc.textures = dst->textures.p; sc.texels = dst->texels.p;
    sc.bigLeafCount = src->sc.bigLeafCount ? dst->bigLeaf.p : nullptr;
    sc.nodes8 = src->sc.nodes8 ? dst->nodes8.p : nullptr; sc.tris8 = src->sc.tris8 ? dst->tris8.p : nullptr;
    sc.tri8Index = src->sc.tri8Index ? dst->tri8Index.p : nullptr;
    sc.faultSink = dst->counters.p;
    dst->sc = sc;
    dst->cam = src->cam;
    dst->wideDepth = src->wideDepth;
    dst->msGpuBuild = src->msGpuBuild;
    dst->lc.wide = src->lc.wide;
    dst->lc.stackDepth = src->lc.stackDepth;
    dst->lc.partition = src->lc.partition;
    dst->lc.partitionConnect = src->lc.partitionConnect;
    dst->st.partitioned = src->st.partitioned;
    dst->lc.traverseGrid = dst->smCount * traverse_ctas_per_sm(dst->fast, dst->wide, dst->lc.stackDepth, false);  // also opts this device's kernels in to large stacks
    dst->lc.connectGrid = dst->smCount * traverse_ctas_per_sm(dst->fast, dst->wide, dst->lc.stackDepth, true);
    if (const char *env = std::getenv("PTC_CTAS_PER_SM")) dst->lc.traverseGrid = dst->lc.connectGrid = dst->smCount * std::max(1, std::atoi(env));
    dst->sceneLoaded = true;
    dst->sceneVersion++;
    return PTC_OK;
}
}  // namespace ptc

extern "C" {

int ptc_wide_bvh_build(const ptc_scene_view *scene, ptc_wide_bvh *out)
{
    if (!scene || !out || scene->struct_size != sizeof(ptc_scene_view)) return PTC_ERR_INVALID;
    std::memset(out, 0, sizeof(*out));
    WideBvh w;
    std::string err;
    if (!build_wide_bvh(scene, w, err)) return fail(nullptr, PTC_ERR_INVALID, "%s", err.c_str());
    out->n_nodes = (uint32_t)w.nodes.size();
    out->n_triangles = (uint32_t)w.triOrder.size();
    out->depth = w.depth;
    out->nodes = std::malloc(std::max<size_t>(1, sizeof(DevNode8) * w.nodes.size()));
    out->tri_order = (uint32_t *)std::malloc(std::max<size_t>(1, sizeof(uint32_t) * w.triOrder.size()));
    if (!out->nodes || !out->tri_order)
    {
        ptc_wide_bvh_free(out);
        return fail(nullptr, PTC_ERR_INVALID, "ptc_wide_bvh_build: out of memory");
    }
    std::memcpy(out->nodes, w.nodes.data(), sizeof(DevNode8) * w.nodes.size());
    std::memcpy(out->tri_order, w.triOrder.data(), sizeof(uint32_t) * w.triOrder.size());
    return PTC_OK;
}

int ptc_wide_bvh_read(ptc_ctx *ctx, ptc_wide_bvh *out)
{
    if (!ctx || !out) return PTC_ERR_INVALID;
    if (ctx->group) return onPrimary(ctx, false, [&](ptc_ctx *p) { return ptc_wide_bvh_read(p, out); });
    std::memset(out, 0, sizeof(*out));
    if (!ctx->sceneLoaded) return fail(ctx, PTC_ERR_STATE, "ptc_wide_bvh_read before ptc_upload_scene");
    if (!ctx->wide) return fail(ctx, PTC_ERR_STATE, "ptc_wide_bvh_read: this context walks the BVH2");
    CK(cudaSetDevice(ctx->device));
    const size_t nNodes = ctx->nodes8.n, nRec = ctx->tri8Index.n;
    out->n_nodes = (uint32_t)nNodes;
    out->n_triangles = (uint32_t)nRec;
    out->depth = ctx->wideDepth;
    out->reserved = (uint32_t)(ctx->msGpuBuild * 1000.f);
    out->nodes = std::malloc(std::max<size_t>(1, sizeof(DevNode8) * nNodes));
    out->tri_order = (uint32_t *)std::malloc(std::max<size_t>(1, sizeof(uint32_t) * nRec));
    if (!out->nodes || !out->tri_order)
    {
        ptc_wide_bvh_free(out);
        return fail(ctx, PTC_ERR_INVALID, "ptc_wide_bvh_read: out of memory");
    }
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (e == cudaSuccess && nNodes) e = cudaMemcpy2D(out->nodes, sizeof(DevNode8), ctx->nodes8.p, sizeof(DevNode8P), sizeof(DevNode8), nNodes, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && nRec) e = cudaMemcpy(out->tri_order, ctx->tri8Index.p, sizeof(uint32_t) * nRec, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess)
    {
        ptc_wide_bvh_free(out);
        return fail(ctx, PTC_ERR_CUDA, "ptc_wide_bvh_read: %s", cudaGetErrorString(e));
    }
    return PTC_OK;
}

void ptc_wide_bvh_free(ptc_wide_bvh *bvh)
{
    if (!bvh) return;
    std::free(bvh->nodes);
    std::free(bvh->tri_order);
    std::memset(bvh, 0, sizeof(*bvh));
}

int ptc_set_camera(ptc_ctx *ctx, const float camera[16])
{
    if (!ctx || !camera) return PTC_ERR_INVALID;
    if (ctx->group) return multi_set_camera(ctx, camera);
    std::memcpy(&ctx->cam, camera, sizeof(Camera));
    return PTC_OK;
}

// ================================================================================================ render
namespace
{
FrameParams frameParams(const ptc_ctx *ctx, uint32_t first, uint32_t nIter)
{
    FrameParams fp{};
    fp.width = ctx->cfg.image_width;
    fp.height = ctx->cfg.image_height;
    fp.maxDepth = ctx->cfg.ray_max_depth;
    fp.sampler = ctx->cfg.sampler;
    fp.firstIteration = first;
    fp.nIterations = nIter;
    fp.flags = ctx->cfg.flags;
    fp.stripeRows = ctx->stripeRows;
    fp.stripeRank = ctx->stripeRank;
    fp.stripeWorld = ctx->stripeWorld;
    fp.ownedPix = ctx->stripeRows ? ctx->ownedPix : ctx->nPix;
    fp.countWork = ctx->lc.count ? 1u : 0u;
    return fp;
}

// One wavefront pass over `nIter` iterations starting at `first` (state must hold nIter * nPix slots).
// accumulate = false leaves the per-path radiance in st.rad (parity probes). maxBounces < 0 = config depth.
int runBatch(ptc_ctx *ctx, uint32_t first, uint32_t nIter, bool accumulate, int maxBounces)
{
    const FrameParams fp = frameParams(ctx, first, nIter);
    const uint32_t nPaths = nIter * fp.ownedPix;
    if (nPaths == 0) return PTC_OK;  // a shard without rows
    Counters *cn = ctx->counters.p;
    cudaStream_t s = ctx->stream;
    // supersampling: the queue of bounce 0 holds the paths that survive the reference's early-out instead of being the identity
    const bool superSampling = accumulate && ctx->x2invSize != 0 && first + nIter > 6u;
    launch_reset(s, cn, superSampling ? 0u : nPaths);
    {
        StageTimer t(ctx, 0);
        launch_generate(s, ctx->lc, fp, ctx->cam, ctx->st);
        if (superSampling)
        {
            launch_supersample(s, ctx->lc, fp, ctx->st, ctx->img, ctx->x2inv.p, ctx->x2invSize, ctx->queue[0].p, cn);
            ctx->launches++;
        }
    }
    ctx->launches += 2;
    const int depth = maxBounces >= 0 ? maxBounces : (int)fp.maxDepth;
    for (int b = 0; b < depth; b++)
    {
        const uint32_t *qin = b == 0 ? (superSampling ? ctx->queue[0].p : nullptr) : ctx->queue[b & 1].p;
        uint32_t *qout = ctx->queue[(b + 1) & 1].p;
        {
            StageTimer t(ctx, 1);
            launch_extend(s, ctx->lc, ctx->fast, ctx->sc, ctx->st, qin, cn, b);
        }
        if (ctx->lc.partition || (ctx->lc.partitionConnect && ctx->sc.nLights))
        {
            StageTimer t(ctx, 3);  // charged to the shade stage, whose input it prepares
            launch_classify(s, ctx->lc, ctx->st, qin, cn, b);
            ctx->launches++;
        }
        if (ctx->sc.nLights)
        {
            StageTimer t(ctx, 2);
            launch_connect(s, ctx->lc, ctx->fast, ctx->sc, ctx->st, qin, cn, b);
            ctx->launches++;
            ctx->launchesConnect++;
        }
        {
            StageTimer t(ctx, 3);
            launch_shade(s, ctx->lc, ctx->fast, ctx->sc, fp, ctx->st, qin, qout, cn, b);
        }
        ctx->launches += 2;
        ctx->launchesExtend++;
    }
    if (accumulate)
    {
        StageTimer t(ctx, 4);
        launch_accumulate(s, ctx->lc, fp, ctx->st, ctx->img, cn);
        ctx->launches++;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(ctx, PTC_ERR_CUDA, "kernel launch failed: %s", cudaGetErrorString(e));
    return PTC_OK;
}
// Batches of at most this many paths are launch-bound (tens of kernels of a few microseconds): replayed as one graph
constexpr uint64_t kGraphMaxPaths = 1ull << 22;

// runBatch through a CUDA graph: captured on first use of a batch shape, afterwards one cudaGraphLaunch with the generate and
// accumulate nodes re-parameterised. Falls back to plain launches when capture is not possible.
int runBatchGraph(ptc_ctx *ctx, uint32_t first, uint32_t nIter)
{
    auto &g = ctx->batchGraph;
    const FrameParams fp = frameParams(ctx, first, nIter);
    const bool fresh = g.exec && g.nIter == nIter && g.ownedPix == fp.ownedPix && g.stripeRows == fp.stripeRows && g.stripeRank == fp.stripeRank &&
                       g.stripeWorld == fp.stripeWorld && g.sceneVersion == ctx->sceneVersion;
    if (!fresh)
    {
        dropBatchGraph(ctx);
        const uint64_t l0 = ctx->launches, e0 = ctx->launchesExtend, c0 = ctx->launchesConnect;
        if (cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeRelaxed) != cudaSuccess)
        {
            cudaGetLastError();
            return runBatch(ctx, first, nIter, true, -1);
        }
        const int rc = runBatch(ctx, first, nIter, true, -1);
        cudaGraph_t graph = nullptr;
        const cudaError_t ce = cudaStreamEndCapture(ctx->stream, &graph);
        g.launches = ctx->launches - l0; g.launchesExtend = ctx->launchesExtend - e0; g.launchesConnect = ctx->launchesConnect - c0;
        ctx->launches = l0; ctx->launchesExtend = e0; ctx->launchesConnect = c0;  // nothing ran yet
        if (rc != PTC_OK || ce != cudaSuccess || !graph)
        {
            if (graph) cudaGraphDestroy(graph);
            cudaGetLastError();
            ctx->useGraphs = false;
            return rc != PTC_OK ? rc : runBatch(ctx, first, nIter, true, -1);
        }
        g.graph = graph;
        size_t n = 0;
        cudaGraphGetNodes(graph, nullptr, &n);
        std::vector<cudaGraphNode_t> nodes(n);
        cudaGraphGetNodes(graph, nodes.data(), &n);
        for (cudaGraphNode_t node : nodes)
        {
            cudaGraphNodeType t;
            if (cudaGraphNodeGetType(node, &t) != cudaSuccess || t != cudaGraphNodeTypeKernel) continue;
            cudaKernelNodeParams p{};
            if (cudaGraphKernelNodeGetParams(node, &p) != cudaSuccess) continue;
            if (p.func == generate_kernel()) { g.genNode = node; g.genParams = p; }
            else if (p.func == accumulate_kernel(fp.sampler)) { g.accNode = node; g.accParams = p; }
        }
        if (!g.genNode || !g.accNode || cudaGraphInstantiate(&g.exec, graph, 0) != cudaSuccess)
        {
            cudaGetLastError();
            dropBatchGraph(ctx);
            ctx->useGraphs = false;
            return runBatch(ctx, first, nIter, true, -1);
        }
        g.nIter = nIter; g.ownedPix = fp.ownedPix; g.stripeRows = fp.stripeRows; g.stripeRank = fp.stripeRank; g.stripeWorld = fp.stripeWorld;
        g.sceneVersion = ctx->sceneVersion;
    }
    // per-batch arguments (the captured values are those of the capture call)
    FrameParams fpArg = fp;
    Camera cam = ctx->cam;
    PathState st = ctx->st;
    ImageDev img = ctx->img;
    Counters *cn = ctx->counters.p;
    void *genArgs[3] = {&fpArg, &cam, &st};
    void *accArgs[4] = {&fpArg, &st, &img, &cn};
    cudaKernelNodeParams gp = g.genParams, ap = g.accParams;
    gp.kernelParams = genArgs; gp.extra = nullptr;
    ap.kernelParams = accArgs; ap.extra = nullptr;
    cudaError_t e = cudaGraphExecKernelNodeSetParams(g.exec, g.genNode, &gp);
    if (e == cudaSuccess) e = cudaGraphExecKernelNodeSetParams(g.exec, g.accNode, &ap);
    if (e == cudaSuccess) e = cudaGraphLaunch(g.exec, ctx->stream);
    if (e != cudaSuccess) return fail(ctx, PTC_ERR_CUDA, "batch graph: %s", cudaGetErrorString(e));
    ctx->launches += g.launches; ctx->launchesExtend += g.launchesExtend; ctx->launchesConnect += g.launchesConnect;
    return PTC_OK;
}
}  // namespace

int ptc_render(ptc_ctx *ctx, uint32_t first_iteration, uint32_t n_iterations)
{
    if (!ctx) return PTC_ERR_INVALID;
    if (ctx->group) return multi_render(ctx, first_iteration, n_iterations);
    if (!ctx->sceneLoaded) return fail(ctx, PTC_ERR_STATE, "ptc_render before ptc_upload_scene");
    CK(cudaSetDevice(ctx->device));
    uint32_t done = 0;
    while (done < n_iterations)
    {
        uint32_t n = std::min(ctx->batchIters, n_iterations - done);
        if (ctx->x2invSize)
        {
            // supersampling reads the accumulators of the iterations before: PARITY = the reference's sequential loop, one iteration
            // per pass; FAST decides on the statistics at the start of a pass, and no pass straddles iteration 6, the first to decide
            const uint32_t at = first_iteration + done;
            if (!ctx->fast) n = 1;
            else if (at < 6u) n = std::min(n, 6u - at);
        }
        const uint64_t paths = (uint64_t)n * (ctx->stripeRows ? ctx->ownedPix : ctx->nPix);
        const bool graph = ctx->useGraphs && !ctx->x2invSize && !(ctx->cfg.flags & PTC_FLAG_STAGE_TIMING) && paths > 0 && paths <= kGraphMaxPaths;
        int rc = graph ? runBatchGraph(ctx, first_iteration + done, n) : runBatch(ctx, first_iteration + done, n, true, -1);
        if (rc != PTC_OK) return rc;
        done += n;
    }
    return PTC_OK;
}

int ptc_synchronize(ptc_ctx *ctx)
{
    if (!ctx) return PTC_ERR_INVALID;
    if (ctx->group) return multi_synchronize(ctx);
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    drainTimers(ctx);
    return PTC_OK;
}

static int readBack(ptc_ctx *ctx, void *dst, const void *src, size_t bytes)
{
    if (!dst) return PTC_OK;
    CK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    return PTC_OK;
}

int ptc_read_image(ptc_ctx *ctx, float *image_color, float *image_ray_nb)
{
    if (!ctx) return PTC_ERR_INVALID;
    if (ctx->group) return onPrimary(ctx, true, [&](ptc_ctx *p) { return ptc_read_image(p, image_color, image_ray_nb); });
    CK(cudaSetDevice(ctx->device));
    int rc = readBack(ctx, image_color, ctx->imgColor.p, sizeof(float4) * ctx->nPix);
    if (rc == PTC_OK) rc = readBack(ctx, image_ray_nb, ctx->imgRayNb.p, sizeof(float) * ctx->nPix);
    if (rc != PTC_OK) return rc;
    return ptc_synchronize(ctx);
}

int ptc_read_variance(ptc_ctx *ctx, float *image_v)
{
    if (!ctx) return PTC_ERR_INVALID;
    if (ctx->group) return fail(ctx, PTC_ERR_STATE, "imageV is not maintained by a multi-device context (a variance accumulator is not linearly summable)");
    if (!ctx->imgV.p) return fail(ctx, PTC_ERR_STATE, "context created with PTC_FLAG_NO_VARIANCE");
    CK(cudaSetDevice(ctx->device));
    int rc = readBack(ctx, image_v, ctx->imgV.p, sizeof(float4) * ctx->nPix);
    if (rc != PTC_OK) return rc;
    return ptc_synchronize(ctx);
}

int ptc_read_stats(ptc_ctx *ctx, uint32_t *ray_depths, uint32_t *ray_bbx, uint32_t *ray_tri)
{
    if (!ctx) return PTC_ERR_INVALID;
    if (ctx->group) return onPrimary(ctx, true, [&](ptc_ctx *p) { return ptc_read_stats(p, ray_depths, ray_bbx, ray_tri); });
    CK(cudaSetDevice(ctx->device));
    int rc = readBack(ctx, ray_depths, ctx->rayDepths.p, sizeof(uint32_t) * (ctx->cfg.ray_max_depth + 1));
    if (rc == PTC_OK) rc = readBack(ctx, ray_bbx, ctx->rayBbx.p, sizeof(uint32_t) * PTC_MAX_INTERSECTION_NUMBER);
    if (rc == PTC_OK) rc = readBack(ctx, ray_tri, ctx->rayTri.p, sizeof(uint32_t) * PTC_MAX_INTERSECTION_NUMBER);
    if (rc != PTC_OK) return rc;
    return ptc_synchronize(ctx);
}

int ptc_read_rgb8(ptc_ctx *ctx, uint8_t *rgb)
{
    if (!ctx || !rgb) return PTC_ERR_INVALID;
    if (ctx->group) return multi_read_rgb8(ctx, rgb);
    CK(cudaSetDevice(ctx->device));
    int rc = ptc_convert_rgb8(ctx);
    if (rc == PTC_OK) rc = readBack(ctx, rgb, ctx->rgb8.p, (size_t)3 * ctx->nPix);
    if (rc != PTC_OK) return rc;
    return ptc_synchronize(ctx);
}

int ptc_convert_rgb8(ptc_ctx *ctx)
{
    if (!ctx) return PTC_ERR_INVALID;
    if (ctx->group) return multi_read_rgb8(ctx, nullptr);
    CK(cudaSetDevice(ctx->device));
    if (!ctx->rgb8.p) CK(ctx->rgb8.alloc((size_t)3 * ctx->nPix));
    launch_rgb8(ctx->stream, ctx->img, frameParams(ctx, 0, 0), ctx->rgb8.p);
    ctx->launches++;
    return PTC_OK;
}

int ptc_set_pixel_stripes(ptc_ctx *ctx, uint32_t stripe_rows, uint32_t rank, uint32_t world)
{
    if (!ctx) return PTC_ERR_INVALID;
    if (ctx->group) return fail(ctx, PTC_ERR_STATE, "a multi-device context shards by itself (iteration ranges, or pixel stripes when a call renders fewer iterations than it has devices)");
    if (stripe_rows == 0)
    {
        ctx->stripeRows = 0; ctx->stripeRank = 0; ctx->stripeWorld = 1; ctx->ownedPix = ctx->nPix;
        return PTC_OK;
    }
    if (world == 0 || rank >= world) return fail(ctx, PTC_ERR_INVALID, "ptc_set_pixel_stripes: rank %u of %u", rank, world);
    uint32_t rows = 0;
    for (uint32_t y = 0; y < ctx->cfg.image_height; y++) rows += (y / stripe_rows) % world == rank;
    ctx->stripeRows = stripe_rows; ctx->stripeRank = rank; ctx->stripeWorld = world;
    ctx->ownedPix = rows * ctx->cfg.image_width;
    return PTC_OK;
}

int ptc_add_image(ptc_ctx *ctx, const float *image_color, const float *image_ray_nb)
{
    if (!ctx) return PTC_ERR_INVALID;
    if (ctx->group) return onPrimary(ctx, false, [&](ptc_ctx *p) { return ptc_add_image(p, image_color, image_ray_nb); });
    CK(cudaSetDevice(ctx->device));
    DevBuf<float4> c;
    DevBuf<float> n;
    cudaError_t e = cudaSuccess;
    if (image_color)
    {
        e = c.alloc(ctx->nPix);
        if (e == cudaSuccess) e = cudaMemcpyAsync(c.p, image_color, sizeof(float4) * ctx->nPix, cudaMemcpyHostToDevice, ctx->stream);
    }
    if (image_ray_nb && e == cudaSuccess)
    {
        e = n.alloc(ctx->nPix);
        if (e == cudaSuccess) e = cudaMemcpyAsync(n.p, image_ray_nb, sizeof(float) * ctx->nPix, cudaMemcpyHostToDevice, ctx->stream);
    }
    if (e == cudaSuccess)
    {
        launch_add_image(ctx->stream, ctx->img, ctx->nPix, c.p, n.p);
        ctx->launches++;
    }
    const cudaError_t es = cudaStreamSynchronize(ctx->stream);  // also on the error paths: the temporaries may be in use
    if (e == cudaSuccess) e = es;
    c.release();
    n.release();
    if (e != cudaSuccess) return fail(ctx, PTC_ERR_CUDA, "ptc_add_image: %s", cudaGetErrorString(e));
    return PTC_OK;
}

int ptc_read_counters(ptc_ctx *ctx, ptc_counters *out)
{
    if (!ctx || !out) return PTC_ERR_INVALID;
    if (ctx->group) return multi_read_counters(ctx, out);
    int rc = ptc_synchronize(ctx);
    if (rc != PTC_OK) return rc;
    Counters h;
    CK(cudaMemcpy(&h, ctx->counters.p, sizeof(Counters), cudaMemcpyDeviceToHost));
    std::memset(out, 0, sizeof(*out));
    out->paths = h.paths;
    out->extension_rays = h.extensionRays;
    out->shadow_rays = h.shadowRays;
    out->box_tests = h.boxTests;
    out->triangle_tests = h.triTests;
    out->kernel_launches = ctx->launches;
    out->ms_generate = ctx->msStage[0];
    out->ms_extend = ctx->msStage[1];
    out->ms_connect = ctx->msStage[2];
    out->ms_shade = ctx->msStage[3];
    out->ms_accumulate = ctx->msStage[4];
    out->launches_extend = ctx->launchesExtend;
    out->launches_connect = ctx->launchesConnect;
    out->scene_bytes_nodes = ctx->nodes.n * sizeof(DevNode);
    out->scene_bytes_triangles = ctx->tris.n * sizeof(DevTri);
    out->ext_box_tests = h.extBoxTests;
    out->ext_triangle_tests = h.extTriTests;
    out->faults = h.faults;
    out->fault_code = h.faultInfo[0]; out->fault_value = h.faultInfo[1]; out->fault_bound = h.faultInfo[2];
    return PTC_OK;
}

int ptc_get_device_buffers(ptc_ctx *ctx, ptc_device_buffers *out)
{
    if (!ctx || !out) return PTC_ERR_INVALID;
    if (ctx->group) return onPrimary(ctx, true, [&](ptc_ctx *p) { return ptc_get_device_buffers(p, out); });
    CK(cudaSetDevice(ctx->device));
    out->image_color = ctx->imgColor.p;
    out->image_ray_nb = ctx->imgRayNb.p;
    out->image_v = ctx->imgV.p;
    out->ray_depths = ctx->rayDepths.p;
    out->ray_bbx = ctx->rayBbx.p;
    out->ray_tri = ctx->rayTri.p;
    out->stream = ctx->stream;
    if (!ctx->rgb8.p) CK(ctx->rgb8.alloc((size_t)3 * ctx->nPix));
    out->rgb8 = ctx->rgb8.p;
    return PTC_OK;
}

// ================================================================================================ parity probes
int ptc_trace_primary(ptc_ctx *ctx, uint32_t iteration, int32_t *hit, float *st, float *dist2, float *dir, int32_t *seed_after,
                      uint32_t *counts)
{
    if (!ctx) return PTC_ERR_INVALID;
    if (ctx->group) return onPrimary(ctx, false, [&](ptc_ctx *p) { return ptc_trace_primary(p, iteration, hit, st, dist2, dir, seed_after, counts); });
    if (!ctx->sceneLoaded) return fail(ctx, PTC_ERR_STATE, "ptc_trace_primary before ptc_upload_scene");
    if (ctx->stripeRows) return fail(ctx, PTC_ERR_STATE, "ptc_trace_primary: not available on a pixel shard (ptc_set_pixel_stripes)");
    CK(cudaSetDevice(ctx->device));
    const FrameParams fp = frameParams(ctx, iteration, 1);
    const uint32_t n = ctx->nPix;
    launch_reset(ctx->stream, ctx->counters.p, n);
    launch_generate(ctx->stream, ctx->lc, fp, ctx->cam, ctx->st);
    LaunchCfg lcProbe = ctx->lc;
    lcProbe.count = true;  // the probe reports per-ray counts in every mode
    launch_extend(ctx->stream, lcProbe, ctx->fast, ctx->sc, ctx->st, nullptr, ctx->counters.p, 0);
    DevBuf<int32_t> dHit, dSeed;
    DevBuf<float2> dSt;
    DevBuf<float> dD2;
    DevBuf<float4> dDir;
    DevBuf<uint2> dCnt;
    if (hit) CK(dHit.alloc(n));
    if (st) CK(dSt.alloc(n));
    if (dist2) CK(dD2.alloc(n));
    if (dir) CK(dDir.alloc(n));
    if (seed_after) CK(dSeed.alloc(n));
    if (counts) CK(dCnt.alloc(n));
    launch_probe_primary(ctx->stream, fp, ctx->st, dHit.p, dSt.p, dD2.p, dDir.p, dSeed.p, dCnt.p);
    ctx->launches += 4;
    int rc = PTC_OK;
    if (rc == PTC_OK && hit) rc = readBack(ctx, hit, dHit.p, sizeof(int32_t) * n);
    if (rc == PTC_OK && st) rc = readBack(ctx, st, dSt.p, sizeof(float2) * n);
    if (rc == PTC_OK && dist2) rc = readBack(ctx, dist2, dD2.p, sizeof(float) * n);
    if (rc == PTC_OK && dir) rc = readBack(ctx, dir, dDir.p, sizeof(float4) * n);
    if (rc == PTC_OK && seed_after) rc = readBack(ctx, seed_after, dSeed.p, sizeof(int32_t) * n);
    if (rc == PTC_OK && counts) rc = readBack(ctx, counts, dCnt.p, sizeof(uint2) * n);
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    dHit.release(); dSeed.release(); dSt.release(); dD2.release(); dDir.release(); dCnt.release();
    if (rc != PTC_OK) return rc;
    if (e != cudaSuccess) return fail(ctx, PTC_ERR_CUDA, "ptc_trace_primary: %s", cudaGetErrorString(e));
    return PTC_OK;
}

int ptc_trace_rays(ptc_ctx *ctx, uint32_t n, const float *origin, const float *direction, int32_t *hit, float *st, float *dist2,
                   float *dir, uint32_t *counts)
{
    if (!ctx || !origin || !direction) return PTC_ERR_INVALID;
    if (ctx->group) return onPrimary(ctx, false, [&](ptc_ctx *p) { return ptc_trace_rays(p, n, origin, direction, hit, st, dist2, dir, counts); });
    if (!ctx->sceneLoaded) return fail(ctx, PTC_ERR_STATE, "ptc_trace_rays before ptc_upload_scene");
    if (n == 0) return PTC_OK;
    if ((size_t)n > ctx->capacity) return fail(ctx, PTC_ERR_LIMIT, "ptc_trace_rays: %u rays, the context holds %zu path slots", n, ctx->capacity);
    CK(cudaSetDevice(ctx->device));
    DevBuf<float4> dOrg, dDirIn, dDir;
    DevBuf<int32_t> dHit;
    DevBuf<float2> dSt;
    DevBuf<float> dD2;
    DevBuf<uint2> dCnt;
    CK(dOrg.alloc(n)); CK(dDirIn.alloc(n));
    if (hit) CK(dHit.alloc(n));
    if (st) CK(dSt.alloc(n));
    if (dist2) CK(dD2.alloc(n));
    if (dir) CK(dDir.alloc(n));
    if (counts) CK(dCnt.alloc(n));
    CK(cudaMemcpyAsync(dOrg.p, origin, sizeof(float4) * n, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(dDirIn.p, direction, sizeof(float4) * n, cudaMemcpyHostToDevice, ctx->stream));
    FrameParams fp{};
    fp.width = n; fp.height = 1; fp.maxDepth = ctx->cfg.ray_max_depth; fp.nIterations = 1; fp.ownedPix = n;
    launch_reset(ctx->stream, ctx->counters.p, n);
    launch_probe_rays(ctx->stream, n, dOrg.p, dDirIn.p, ctx->st);
    LaunchCfg lcProbe = ctx->lc;
    lcProbe.count = true;
    launch_extend(ctx->stream, lcProbe, ctx->fast, ctx->sc, ctx->st, nullptr, ctx->counters.p, 0);
    launch_probe_primary(ctx->stream, fp, ctx->st, dHit.p, dSt.p, dD2.p, dDir.p, nullptr, dCnt.p);
    ctx->launches += 4;
    int rc = PTC_OK;
    if (rc == PTC_OK && hit) rc = readBack(ctx, hit, dHit.p, sizeof(int32_t) * n);
    if (rc == PTC_OK && st) rc = readBack(ctx, st, dSt.p, sizeof(float2) * n);
    if (rc == PTC_OK && dist2) rc = readBack(ctx, dist2, dD2.p, sizeof(float) * n);
    if (rc == PTC_OK && dir) rc = readBack(ctx, dir, dDir.p, sizeof(float4) * n);
    if (rc == PTC_OK && counts) rc = readBack(ctx, counts, dCnt.p, sizeof(uint2) * n);
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    dOrg.release(); dDirIn.release(); dHit.release(); dSt.release(); dD2.release(); dDir.release(); dCnt.release();
    if (rc != PTC_OK) return rc;
    if (e != cudaSuccess) return fail(ctx, PTC_ERR_CUDA, "ptc_trace_rays: %s", cudaGetErrorString(e));
    return PTC_OK;
}

int ptc_render_single(ptc_ctx *ctx, uint32_t iteration, float *radiance)
{
    if (!ctx || !radiance) return PTC_ERR_INVALID;
    if (ctx->group) return onPrimary(ctx, false, [&](ptc_ctx *p) { return ptc_render_single(p, iteration, radiance); });
    if (!ctx->sceneLoaded) return fail(ctx, PTC_ERR_STATE, "ptc_render_single before ptc_upload_scene");
    if (ctx->stripeRows) return fail(ctx, PTC_ERR_STATE, "ptc_render_single: not available on a pixel shard (ptc_set_pixel_stripes)");
    CK(cudaSetDevice(ctx->device));
    int rc = runBatch(ctx, iteration, 1, false, -1);
    if (rc != PTC_OK) return rc;
    rc = readBack(ctx, radiance, ctx->rad.p, sizeof(float4) * ctx->nPix);
    if (rc != PTC_OK) return rc;
    return ptc_synchronize(ctx);
}

// Parity probe: the state of every path of one iteration after `bounces` bounces (0 = the camera rays), without touching the
// accumulators. Lets a test walk a pixel's path bounce by bounce next to the reference's (tests/test_gpu_parity.py classifies
// the pixels whose 1-spp radiance differs by the first bounce at which the two paths part).
int ptc_path_probe(ptc_ctx *ctx, uint32_t iteration, uint32_t bounces, float *origin, float *direction, float *hit, uint32_t *meta)
{
    if (!ctx) return PTC_ERR_INVALID;
    if (ctx->group) return onPrimary(ctx, false, [&](ptc_ctx *p) { return ptc_path_probe(p, iteration, bounces, origin, direction, hit, meta); });
    if (!ctx->sceneLoaded) return fail(ctx, PTC_ERR_STATE, "ptc_path_probe before ptc_upload_scene");
    if (ctx->stripeRows) return fail(ctx, PTC_ERR_STATE, "ptc_path_probe: not available on a pixel shard (ptc_set_pixel_stripes)");
    if (bounces > ctx->cfg.ray_max_depth) return fail(ctx, PTC_ERR_INVALID, "ptc_path_probe: %u bounces, ray_max_depth is %u", bounces, ctx->cfg.ray_max_depth);
    CK(cudaSetDevice(ctx->device));
    int rc = runBatch(ctx, iteration, 1, false, (int)bounces);
    if (rc == PTC_OK) rc = readBack(ctx, origin, ctx->org.p, sizeof(float4) * ctx->nPix);
    if (rc == PTC_OK) rc = readBack(ctx, direction, ctx->dir.p, sizeof(float4) * ctx->nPix);
    if (rc == PTC_OK) rc = readBack(ctx, hit, ctx->hit.p, sizeof(float4) * ctx->nPix);
    if (rc == PTC_OK) rc = readBack(ctx, meta, ctx->meta.p, sizeof(uint32_t) * ctx->nPix);
    if (rc != PTC_OK) return rc;
    return ptc_synchronize(ctx);
}

}  // extern "C"
