// pt_static.cu — ptc_trace_rays_static: the closest-hit probe of ptc_trace_rays (pathtracer_cuda.cu) for callers that must not
// allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// No counterpart in the reference. ptc_trace_rays allocates seven device buffers per
// call and frees them again; this variant drives the SAME kernels (k_probe_rays -> k_trace<extend> -> k_probe_primary, pt_kernels.cu)
// with every scratch buffer — the staged input rays, the path-state records the extend kernel works on, its queue counters and
// the staged outputs — carved out of ONE file-scope __device__ array of fixed size. The price:
//   * every GPU that loads libpathtracer_cuda.so holds PTC_STATIC_SCRATCH_BYTES (4 MiB) until the process exits;
//   * n is limited to PTC_STATIC_MAX_RAYS per call;
//   * all calls on a device, from whatever context, share the array: calls that overlap on the GPU race. The entry point itself
//     returns only after its stream has drained, so calls made one after the other from one host thread are safe.
//
// In host code `g_staticScratch` is a host-side stand-in for the symbol, not a device address: the address on each device comes
// from cudaGetSymbolAddress, resolved once per device with that device current (scratch_base below).
#include <cstdint>
#include <mutex>

#include "pt_context.h"

using namespace ptc;

namespace
{
constexpr size_t kMaxRays = PTC_STATIC_MAX_RAYS;
constexpr size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

// layout of the scratch array; every block starts on a 256-byte boundary
constexpr size_t kOffCounters = 0;
constexpr size_t kOffInOrg = kOffCounters + align256(sizeof(Counters));   // staged caller rays
constexpr size_t kOffInDir = kOffInOrg + align256(sizeof(float4) * kMaxRays);
constexpr size_t kOffOrg = kOffInDir + align256(sizeof(float4) * kMaxRays);  // path state of the extend kernel
constexpr size_t kOffDir = kOffOrg + align256(sizeof(float4) * kMaxRays);
constexpr size_t kOffHit = kOffDir + align256(sizeof(float4) * kMaxRays);
constexpr size_t kOffSeed = kOffHit + align256(sizeof(float4) * kMaxRays);
constexpr size_t kOffMeta = kOffSeed + align256(sizeof(int) * kMaxRays);
constexpr size_t kOffNbbx = kOffMeta + align256(sizeof(uint32_t) * kMaxRays);
constexpr size_t kOffNtri = kOffNbbx + align256(sizeof(uint32_t) * kMaxRays);
constexpr size_t kOffOccl = kOffNtri + align256(sizeof(uint32_t) * kMaxRays);
constexpr size_t kOffOutHit = kOffOccl + align256(sizeof(uint32_t) * kMaxRays);  // staged outputs
constexpr size_t kOffOutSt = kOffOutHit + align256(sizeof(int32_t) * kMaxRays);
constexpr size_t kOffOutD2 = kOffOutSt + align256(sizeof(float2) * kMaxRays);
constexpr size_t kOffOutCnt = kOffOutD2 + align256(sizeof(float) * kMaxRays);
constexpr size_t kOffEnd = kOffOutCnt + align256(sizeof(uint2) * kMaxRays);

constexpr size_t kStaticScratchBytes = PTC_STATIC_SCRATCH_BYTES;
static_assert(kOffEnd <= kStaticScratchBytes, "PTC_STATIC_MAX_RAYS does not fit the static scratch array");
}  // namespace

static __device__ __align__(256) unsigned char g_staticScratch[kStaticScratchBytes];

// The array keeps the last call's contents. What must start at zero is the counter block at its head (queue lengths, work cursors
// of the persistent extend kernel, work totals, the CHECKED build's fault word); device code of this file names the array directly.
static __global__ void k_static_clear_counters()
{
    uint32_t *w = reinterpret_cast<uint32_t *>(g_staticScratch + kOffCounters);
    for (uint32_t i = threadIdx.x; i < sizeof(Counters) / sizeof(uint32_t); i += blockDim.x) w[i] = 0u;
}
static_assert(sizeof(Counters) % sizeof(uint32_t) == 0, "Counters is cleared word by word");

namespace
{
constexpr int kMaxDevices = 64;
std::mutex g_baseLock;
unsigned char *g_base[kMaxDevices] = {};

// device address of g_staticScratch on `device`, which must be the current device; resolved once per device
cudaError_t scratch_base(int device, unsigned char **out)
{
    std::lock_guard<std::mutex> hold(g_baseLock);
    if (!g_base[device])
    {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_staticScratch);
        if (e != cudaSuccess) return e;
        g_base[device] = static_cast<unsigned char *>(p);
    }
    *out = g_base[device];
    return cudaSuccess;
}

// host or device pointer on either side (unified addressing tells them apart)
cudaError_t copy_async(void *dst, const void *src, size_t bytes, cudaStream_t s)
{
    return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, s);
}
}  // namespace

extern "C" {

int ptc_trace_rays_static(ptc_ctx *ctx, uint32_t n, const float *origin, const float *direction, int32_t *hit, float *st, float *dist2,
                          uint32_t *counts)
{
    if (!ctx || !origin || !direction) return PTC_ERR_INVALID;
    if (ctx->group)  // as the other probes: member 0 answers, its error message is carried over to the handle
    {
        ptc_ctx *p = multi_primary(ctx);
        const int rc = ptc_trace_rays_static(p, n, origin, direction, hit, st, dist2, counts);
        if (rc != PTC_OK) ctx->err = p->err;
        return rc;
    }
    if (!ctx->sceneLoaded) return fail(ctx, PTC_ERR_STATE, "ptc_trace_rays_static before ptc_upload_scene");
    if (n > PTC_STATIC_MAX_RAYS)
        return fail(ctx, PTC_ERR_LIMIT, "ptc_trace_rays_static: %u rays, the static scratch holds PTC_STATIC_MAX_RAYS = %u per call (ptc_trace_rays has no such limit)",
                    n, (unsigned)PTC_STATIC_MAX_RAYS);
    if (ctx->device < 0 || ctx->device >= kMaxDevices) return fail(ctx, PTC_ERR_LIMIT, "ptc_trace_rays_static: device %d, at most %d devices", ctx->device, kMaxDevices);
    if (n == 0) return PTC_OK;
    CK(cudaSetDevice(ctx->device));
    unsigned char *base = nullptr;
    CK(scratch_base(ctx->device, &base));

    Counters *cn = reinterpret_cast<Counters *>(base + kOffCounters);
    float4 *inOrg = reinterpret_cast<float4 *>(base + kOffInOrg), *inDir = reinterpret_cast<float4 *>(base + kOffInDir);
    int32_t *oHit = reinterpret_cast<int32_t *>(base + kOffOutHit);
    float2 *oSt = reinterpret_cast<float2 *>(base + kOffOutSt);
    float *oD2 = reinterpret_cast<float *>(base + kOffOutD2);
    uint2 *oCnt = reinterpret_cast<uint2 *>(base + kOffOutCnt);

    PathState ps{};  // thr / rad / sample / qpart / tileLit stay null: the extend stage does not touch them
    ps.org = reinterpret_cast<float4 *>(base + kOffOrg);
    ps.dir = reinterpret_cast<float4 *>(base + kOffDir);
    ps.hit = reinterpret_cast<float4 *>(base + kOffHit);
    ps.seed = reinterpret_cast<int *>(base + kOffSeed);
    ps.meta = reinterpret_cast<uint32_t *>(base + kOffMeta);
    ps.nbbx = reinterpret_cast<uint32_t *>(base + kOffNbbx);
    ps.ntri = reinterpret_cast<uint32_t *>(base + kOffNtri);
    ps.occl = reinterpret_cast<uint32_t *>(base + kOffOccl);
    ps.capacity = PTC_STATIC_MAX_RAYS;
    ps.partitioned = 0;

    cudaStream_t s = ctx->stream;
    // the counter block starts at zero; every per-ray record of [0, n) is written by k_probe_rays / k_trace before it is read
    k_static_clear_counters<<<1, 128, 0, s>>>();
    CK(copy_async(inOrg, origin, sizeof(float4) * n, s));
    CK(copy_async(inDir, direction, sizeof(float4) * n, s));
    FrameParams fp{};
    fp.width = n; fp.height = 1; fp.maxDepth = ctx->cfg.ray_max_depth; fp.nIterations = 1; fp.ownedPix = n;
    launch_reset(s, cn, n);
    launch_probe_rays(s, n, inOrg, inDir, ps);
    LaunchCfg lc = ctx->lc;
    lc.count = true;      // as ptc_trace_rays: the same kernel instantiation, so the two entry points agree bit for bit
    lc.partition = false;
    launch_extend(s, lc, ctx->fast, ctx->sc, ps, nullptr, cn, 0);
    launch_probe_primary(s, fp, ps, hit ? oHit : nullptr, st ? oSt : nullptr, dist2 ? oD2 : nullptr, nullptr, nullptr, counts ? oCnt : nullptr);
    ctx->launches += 5;
    CK(cudaGetLastError());
    if (hit) CK(copy_async(hit, oHit, sizeof(int32_t) * n, s));
    if (st) CK(copy_async(st, oSt, sizeof(float2) * n, s));
    if (dist2) CK(copy_async(dist2, oD2, sizeof(float) * n, s));
    if (counts) CK(copy_async(counts, oCnt, sizeof(uint2) * n, s));
    CK(cudaStreamSynchronize(s));
    return PTC_OK;
}

}  // extern "C"
