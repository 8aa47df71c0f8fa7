// pt_multi.cu — multi-GPU behind the C ABI: ptc_create_multi and the group half of every ptc_* call.
//
// The reference drives one device (Controleur/PathTracer_OpenCL.cpp:316-402 picks one, :56-140 loops on it). The path shards
// by independent units: the random stream of a path depends on (x, y, iteration, W, H) only
// (Kernel/PathTracer_FullKernel_header.cl:255-264), so a group of devices with the scene replicated renders
//   * disjoint ITERATION RANGES of one ptc_render(first, n) call when n >= devices (the split of sharding.iteration_range), or
//   * interleaved 8-row PIXEL STRIPES of all n iterations when n < devices (progressive frames: one iteration cannot be split),
// each into its own accumulators; the one exchange step is a sum into member 0 when the caller reads: ncclReduce over NVLink
// (one communicator per device, one host thread, ncclGroupStart/End) of imageColor, imageRayNb and the three uint32 histograms,
// after which the other members' accumulators are cleared (member 0 holds the total). A display frame of a stripes-only render
// takes the short way: every member converts its own rows to RGB8 and ONE uint8 reduce (3 bytes per pixel) composes the frame.
//
// NCCL is bound at run time (dlopen of libnccl.so.2: in a process that already holds torch's NCCL that one is found), so the
// library has no link-time dependency on it; without NCCL — or with PTC_MULTI_REDUCE=p2p — the sum is made by peer copies into
// member 0 and an add kernel there.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "pt_context.h"

namespace ptc
{
constexpr uint32_t kStripeRows = 8;  // opencl_pathtracer_b200/sharding.py STRIPE_ROWS

struct NcclApi
{
    void *lib = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*Reduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    bool load()
    {
        if (lib) return true;
        for (const char *name : {"libnccl.so.2", "libnccl.so"})
        {
            lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (lib) break;
        }
        if (!lib) return false;
#define PTC_NCCL_SYM(f) *(void **)(&f) = dlsym(lib, "nccl" #f)
        PTC_NCCL_SYM(CommInitAll); PTC_NCCL_SYM(CommDestroy); PTC_NCCL_SYM(GroupStart); PTC_NCCL_SYM(GroupEnd);
        PTC_NCCL_SYM(Reduce); PTC_NCCL_SYM(GetErrorString);
#undef PTC_NCCL_SYM
        return CommInitAll && CommDestroy && GroupStart && GroupEnd && Reduce && GetErrorString;
    }
};

struct MultiGroup
{
    std::vector<ptc_ctx *> members;  // member 0 holds the reduced image
    NcclApi nccl;
    std::vector<ncclComm_t> comms;   // empty = peer-copy reduce
    bool dirty = false;              // members hold samples that are not summed into member 0 yet
    bool stripesOnly = true;         // since the last reset every render was a pixel-stripes render and no float reduce ran:
                                     // each member's accumulators are non-zero on its own rows only
    bool striped = false;            // the members are currently configured as pixel shards
    // peer-copy reduce: staging on member 0's device, events to order the members' streams against member 0's
    DevBuf<float4> tmpColor;
    DevBuf<float> tmpNb;
    DevBuf<uint32_t> tmpHist;
    DevBuf<uint8_t> tmpRgb;
    std::vector<cudaEvent_t> evMember;
    cudaEvent_t evRoot = nullptr;
};

// u32 / u8 sums of the peer-copy reduce (the float sums reuse launch_add_image)
__global__ void k_add_u32(uint32_t *dst, const uint32_t *src, uint32_t n)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) dst[i] += src[i];
}
__global__ void k_add_u8x4(uint32_t *dst, const uint32_t *src, uint32_t nWords)
{
    // display frames of pixel shards are zero outside the shard: a bytewise sum never carries
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nWords; i += gridDim.x * blockDim.x) dst[i] += src[i];
}

namespace
{
constexpr uint32_t kHist = PTC_MAX_INTERSECTION_NUMBER;

int carry(ptc_ctx *h, ptc_ctx *m, int rc)
{
    if (rc != PTC_OK) h->err = m->err;
    return rc;
}

int ncclFail(ptc_ctx *h, const char *what, ncclResult_t r)
{
    return fail(h, PTC_ERR_CUDA, "%s: NCCL %s", what, h->group->nccl.GetErrorString ? h->group->nccl.GetErrorString(r) : "error");
}

int setStripes(ptc_ctx *h, bool on)
{
    MultiGroup *g = h->group;
    if (g->striped == on) return PTC_OK;
    const uint32_t world = (uint32_t)g->members.size();
    for (uint32_t i = 0; i < world; i++)
    {
        const int rc = ptc_set_pixel_stripes(g->members[i], on ? kStripeRows : 0u, i, world);
        if (rc != PTC_OK) return carry(h, g->members[i], rc);
    }
    g->striped = on;
    return PTC_OK;
}

// member streams -> root stream, and back
cudaError_t rootWaitsForMembers(MultiGroup *g)
{
    cudaError_t e = cudaSuccess;
    for (size_t i = 1; i < g->members.size() && e == cudaSuccess; i++)
    {
        ptc_ctx *m = g->members[i];
        e = cudaSetDevice(m->device);
        if (e == cudaSuccess) e = cudaEventRecord(g->evMember[i], m->stream);
        if (e == cudaSuccess) e = cudaSetDevice(g->members[0]->device);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(g->members[0]->stream, g->evMember[i], 0);
    }
    return e;
}
}  // namespace

ptc_ctx *multi_primary(ptc_ctx *h) { return h->group->members[0]; }

void multi_destroy(ptc_ctx *h)
{
    MultiGroup *g = h->group;
    if (g)
    {
        for (ptc_ctx *m : g->members)
            if (m)
            {
                cudaSetDevice(m->device);
                cudaStreamSynchronize(m->stream);
            }
        for (ncclComm_t c : g->comms)
            if (c) g->nccl.CommDestroy(c);
        if (!g->members.empty() && g->members[0]) cudaSetDevice(g->members[0]->device);
        g->tmpColor.release(); g->tmpNb.release(); g->tmpHist.release(); g->tmpRgb.release();
        for (cudaEvent_t e : g->evMember)
            if (e) cudaEventDestroy(e);
        if (g->evRoot) cudaEventDestroy(g->evRoot);
        for (ptc_ctx *m : g->members)
            if (m) ptc_destroy(m);
        delete g;
    }
    delete h;
}

int multi_upload_scene(ptc_ctx *h, const ptc_scene_view *scene)
{
    MultiGroup *g = h->group;
    int rc = multi_synchronize(h);
    if (rc != PTC_OK) return rc;
    rc = ptc_upload_scene(g->members[0], scene);
    if (rc != PTC_OK) return carry(h, g->members[0], rc);
    for (size_t i = 1; i < g->members.size(); i++)
    {
        rc = replicate_scene(g->members[i], g->members[0]);
        if (rc != PTC_OK) return carry(h, g->members[i], rc);
    }
    h->sceneLoaded = true;
    return PTC_OK;
}

int multi_set_camera(ptc_ctx *h, const float camera[16])
{
    for (ptc_ctx *m : h->group->members)
    {
        const int rc = ptc_set_camera(m, camera);
        if (rc != PTC_OK) return carry(h, m, rc);
    }
    return PTC_OK;
}

int multi_reset(ptc_ctx *h)
{
    MultiGroup *g = h->group;
    for (ptc_ctx *m : g->members)
    {
        const int rc = ptc_reset_accumulation(m);
        if (rc != PTC_OK) return carry(h, m, rc);
    }
    g->dirty = false;
    g->stripesOnly = true;
    return PTC_OK;
}

int multi_render(ptc_ctx *h, uint32_t first, uint32_t n)
{
    MultiGroup *g = h->group;
    if (!h->sceneLoaded) return fail(h, PTC_ERR_STATE, "ptc_render before ptc_upload_scene");
    if (n == 0) return PTC_OK;
    const uint32_t world = (uint32_t)g->members.size();
    if (n < world)
    {
        // fewer iterations than devices: every device renders its rows of all of them
        int rc = setStripes(h, true);
        if (rc != PTC_OK) return rc;
        for (ptc_ctx *m : g->members)
        {
            rc = ptc_render(m, first, n);
            if (rc != PTC_OK) return carry(h, m, rc);
        }
    }
    else
    {
        int rc = setStripes(h, false);
        if (rc != PTC_OK) return rc;
        g->stripesOnly = false;
        const uint32_t base = n / world, extra = n % world;  // sharding.iteration_range: earlier members take the extras
        // batches are enqueued round-robin over the devices, so that every device starts at once and none waits for the host to
        // finish queueing another's whole range
        std::vector<uint32_t> begin(world), left(world);
        for (uint32_t i = 0; i < world; i++)
        {
            begin[i] = first + i * base + std::min(i, extra);
            left[i] = base + (i < extra ? 1u : 0u);
        }
        bool any = true;
        while (any)
        {
            any = false;
            for (uint32_t i = 0; i < world; i++)
            {
                if (!left[i]) continue;
                ptc_ctx *m = g->members[i];
                const uint32_t k = std::min(left[i], m->batchIters);
                const int rc = ptc_render(m, begin[i], k);
                if (rc != PTC_OK) return carry(h, m, rc);
                begin[i] += k;
                left[i] -= k;
                any = any || left[i];
            }
        }
    }
    g->dirty = true;
    return PTC_OK;
}

int multi_synchronize(ptc_ctx *h)
{
    for (ptc_ctx *m : h->group->members)
    {
        const int rc = ptc_synchronize(m);
        if (rc != PTC_OK) return carry(h, m, rc);
    }
    return PTC_OK;
}

// accumulators of every member summed into member 0, the others cleared
int multi_reduce(ptc_ctx *h)
{
    MultiGroup *g = h->group;
    if (!g->dirty) return PTC_OK;
    ptc_ctx *root = g->members[0];
    ptc_ctx *ctx = h;  // CK reports into the handle
    const size_t nPix = root->nPix;
    if (!g->comms.empty())
    {
        ncclResult_t r = g->nccl.GroupStart();
        if (r != ncclSuccess) return ncclFail(h, "ncclGroupStart", r);
        for (size_t i = 0; i < g->members.size() && r == ncclSuccess; i++)
        {
            ptc_ctx *m = g->members[i];
            r = g->nccl.Reduce(m->imgColor.p, root->imgColor.p, 4 * nPix, ncclFloat, ncclSum, 0, g->comms[i], m->stream);
            if (r == ncclSuccess) r = g->nccl.Reduce(m->imgRayNb.p, root->imgRayNb.p, nPix, ncclFloat, ncclSum, 0, g->comms[i], m->stream);
            if (r == ncclSuccess) r = g->nccl.Reduce(m->rayDepths.p, root->rayDepths.p, kMaxBounces + 1, ncclUint32, ncclSum, 0, g->comms[i], m->stream);
            if (r == ncclSuccess) r = g->nccl.Reduce(m->rayBbx.p, root->rayBbx.p, kHist, ncclUint32, ncclSum, 0, g->comms[i], m->stream);
            if (r == ncclSuccess) r = g->nccl.Reduce(m->rayTri.p, root->rayTri.p, kHist, ncclUint32, ncclSum, 0, g->comms[i], m->stream);
        }
        const ncclResult_t re = g->nccl.GroupEnd();
        if (r != ncclSuccess) return ncclFail(h, "ncclReduce", r);
        if (re != ncclSuccess) return ncclFail(h, "ncclGroupEnd", re);
    }
    else
    {
        CK(rootWaitsForMembers(g));
        CK(cudaSetDevice(root->device));
        for (size_t i = 1; i < g->members.size(); i++)
        {
            ptc_ctx *m = g->members[i];
            CK(cudaMemcpyPeerAsync(g->tmpColor.p, root->device, m->imgColor.p, m->device, sizeof(float4) * nPix, root->stream));
            CK(cudaMemcpyPeerAsync(g->tmpNb.p, root->device, m->imgRayNb.p, m->device, sizeof(float) * nPix, root->stream));
            launch_add_image(root->stream, root->img, (uint32_t)nPix, g->tmpColor.p, g->tmpNb.p);
            uint32_t *src[3] = {m->rayDepths.p, m->rayBbx.p, m->rayTri.p}, *dst[3] = {root->rayDepths.p, root->rayBbx.p, root->rayTri.p};
            const uint32_t cnt[3] = {(uint32_t)kMaxBounces + 1u, kHist, kHist};
            for (int k = 0; k < 3; k++)
            {
                CK(cudaMemcpyPeerAsync(g->tmpHist.p, root->device, src[k], m->device, sizeof(uint32_t) * cnt[k], root->stream));
                k_add_u32<<<8, 256, 0, root->stream>>>(dst[k], g->tmpHist.p, cnt[k]);
            }
            root->launches += 4;
        }
        CK(cudaEventRecord(g->evRoot, root->stream));
    }
    // member 0 holds the total: the others start again from zero (in stream order after their part of the sum)
    for (size_t i = 1; i < g->members.size(); i++)
    {
        ptc_ctx *m = g->members[i];
        CK(cudaSetDevice(m->device));
        if (g->comms.empty()) CK(cudaStreamWaitEvent(m->stream, g->evRoot, 0));
        CK(cudaMemsetAsync(m->imgColor.p, 0, sizeof(float4) * nPix, m->stream));
        CK(cudaMemsetAsync(m->imgRayNb.p, 0, sizeof(float) * nPix, m->stream));
        CK(cudaMemsetAsync(m->rayDepths.p, 0, sizeof(uint32_t) * (kMaxBounces + 1), m->stream));
        CK(cudaMemsetAsync(m->rayBbx.p, 0, sizeof(uint32_t) * kHist, m->stream));
        CK(cudaMemsetAsync(m->rayTri.p, 0, sizeof(uint32_t) * kHist, m->stream));
    }
    CK(cudaSetDevice(root->device));
    g->dirty = false;
    g->stripesOnly = false;
    return PTC_OK;
}

// ptc_read_rgb8 / ptc_convert_rgb8 of a group. rgb == nullptr: convert only (the frame stays in member 0's display buffer).
int multi_read_rgb8(ptc_ctx *h, uint8_t *rgb)
{
    MultiGroup *g = h->group;
    ptc_ctx *root = g->members[0];
    ptc_ctx *ctx = h;
    const size_t bytes = (size_t)3 * root->nPix;
    if (g->stripesOnly && g->striped && g->dirty)
    {
        // progressive frame: every member turns ITS rows into RGB8 (zero elsewhere), one uint8 sum composes the frame on member 0;
        // the float accumulators stay where they are (they are still disjoint by rows, so later frames may accumulate further)
        for (ptc_ctx *m : g->members)
        {
            const int rc = ptc_convert_rgb8(m);
            if (rc != PTC_OK) return carry(h, m, rc);
        }
        if (!g->comms.empty())
        {
            ncclResult_t r = g->nccl.GroupStart();
            if (r != ncclSuccess) return ncclFail(h, "ncclGroupStart", r);
            for (size_t i = 0; i < g->members.size() && r == ncclSuccess; i++)
                r = g->nccl.Reduce(g->members[i]->rgb8.p, root->rgb8.p, bytes, ncclUint8, ncclSum, 0, g->comms[i], g->members[i]->stream);
            const ncclResult_t re = g->nccl.GroupEnd();
            if (r != ncclSuccess) return ncclFail(h, "ncclReduce", r);
            if (re != ncclSuccess) return ncclFail(h, "ncclGroupEnd", re);
        }
        else
        {
            CK(rootWaitsForMembers(g));
            CK(cudaSetDevice(root->device));
            if (!g->tmpRgb.p) CK(g->tmpRgb.alloc((bytes + 3) & ~(size_t)3));
            for (size_t i = 1; i < g->members.size(); i++)
            {
                ptc_ctx *m = g->members[i];
                CK(cudaMemcpyPeerAsync(g->tmpRgb.p, root->device, m->rgb8.p, m->device, bytes, root->stream));
                k_add_u8x4<<<root->smCount * 4, 256, 0, root->stream>>>(reinterpret_cast<uint32_t *>(root->rgb8.p), reinterpret_cast<const uint32_t *>(g->tmpRgb.p),
                                                                       (uint32_t)(bytes / 4));
                root->launches++;
            }
            // the members' display buffers must not be rewritten before member 0 has copied them
            CK(cudaEventRecord(g->evRoot, root->stream));
            for (size_t i = 1; i < g->members.size(); i++)
            {
                CK(cudaSetDevice(g->members[i]->device));
                CK(cudaStreamWaitEvent(g->members[i]->stream, g->evRoot, 0));
            }
        }
        CK(cudaSetDevice(root->device));
        if (rgb)
        {
            CK(cudaMemcpyAsync(rgb, root->rgb8.p, bytes, cudaMemcpyDeviceToHost, root->stream));
            return carry(h, root, ptc_synchronize(root));
        }
        return PTC_OK;
    }
    // general case: sum the float accumulators, member 0 converts the whole frame
    int rc = multi_reduce(h);
    if (rc != PTC_OK) return rc;
    const uint32_t rows = root->stripeRows, rank = root->stripeRank, world = root->stripeWorld, owned = root->ownedPix;
    root->stripeRows = 0;  // member 0 may be configured as a shard: the frame it holds now is whole
    rc = rgb ? ptc_read_rgb8(root, rgb) : ptc_convert_rgb8(root);
    root->stripeRows = rows; root->stripeRank = rank; root->stripeWorld = world; root->ownedPix = owned;
    return carry(h, root, rc);
}

int multi_read_counters(ptc_ctx *h, ptc_counters *out)
{
    std::memset(out, 0, sizeof(*out));
    bool first = true;
    for (ptc_ctx *m : h->group->members)
    {
        ptc_counters c;
        const int rc = ptc_read_counters(m, &c);
        if (rc != PTC_OK) return carry(h, m, rc);
        out->paths += c.paths; out->extension_rays += c.extension_rays; out->shadow_rays += c.shadow_rays;
        out->box_tests += c.box_tests; out->triangle_tests += c.triangle_tests; out->kernel_launches += c.kernel_launches;
        out->launches_extend += c.launches_extend; out->launches_connect += c.launches_connect;
        out->ext_box_tests += c.ext_box_tests; out->ext_triangle_tests += c.ext_triangle_tests;
        out->faults |= c.faults;
        if (c.faults && !out->fault_bound) { out->fault_code = c.fault_code; out->fault_value = c.fault_value; out->fault_bound = c.fault_bound; }
        if (first)
        {
            // stage times and scene sizes: member 0's (every member holds the same scene)
            out->ms_generate = c.ms_generate; out->ms_extend = c.ms_extend; out->ms_connect = c.ms_connect; out->ms_shade = c.ms_shade;
            out->ms_accumulate = c.ms_accumulate; out->scene_bytes_nodes = c.scene_bytes_nodes; out->scene_bytes_triangles = c.scene_bytes_triangles;
            first = false;
        }
    }
    return PTC_OK;
}
}  // namespace ptc

using namespace ptc;

extern "C" int ptc_create_multi(const ptc_config *config, const int32_t *devices, uint32_t n_devices, ptc_ctx **out_ctx)
{
    if (!config || !out_ctx || !devices || n_devices == 0) return fail(nullptr, PTC_ERR_INVALID, "ptc_create_multi: null argument or no device");
    *out_ctx = nullptr;
    if (config->struct_size != sizeof(ptc_config)) return fail(nullptr, PTC_ERR_INVALID, "ptc_create_multi: struct_size %u != %zu", config->struct_size, sizeof(ptc_config));
    if (n_devices > 64) return fail(nullptr, PTC_ERR_INVALID, "ptc_create_multi: %u devices", n_devices);
    for (uint32_t i = 0; i < n_devices; i++)
        for (uint32_t j = 0; j < i; j++)
            if (devices[i] == devices[j]) return fail(nullptr, PTC_ERR_INVALID, "ptc_create_multi: device %d listed twice", devices[i]);
    if (config->stream && n_devices > 1) return fail(nullptr, PTC_ERR_INVALID, "ptc_create_multi: a caller's stream belongs to one device; leave ptc_config.stream NULL");
    if (n_devices == 1)
    {
        ptc_config one = *config;
        one.device = devices[0];
        return ptc_create(&one, out_ctx);
    }
    ptc_ctx *h = new ptc_ctx();
    h->cfg = *config;
    h->group = new MultiGroup();
    MultiGroup *g = h->group;
    for (uint32_t i = 0; i < n_devices; i++)
    {
        ptc_config one = *config;
        one.device = devices[i];
        one.flags |= PTC_FLAG_NO_VARIANCE;  // imageV is not linearly summable; only supersampling reads it, which a group refuses
        one.flags &= ~(uint32_t)PTC_FLAG_STAGE_TIMING;
        ptc_ctx *m = nullptr;
        const int rc = ptc_create(&one, &m);
        if (rc != PTC_OK)
        {
            multi_destroy(h);  // create_error() keeps the member's message
            return rc;
        }
        g->members.push_back(m);
    }
    h->device = devices[0];
    h->nPix = g->members[0]->nPix;
    // peer access (NVLink): scene replication and the peer-copy reduce go device to device; failure is not an error (copies are staged)
    for (uint32_t i = 0; i < n_devices; i++)
        for (uint32_t j = 0; j < n_devices; j++)
        {
            if (i == j) continue;
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, devices[i], devices[j]) == cudaSuccess && can)
            {
                cudaSetDevice(devices[i]);
                if (cudaDeviceEnablePeerAccess(devices[j], 0) != cudaSuccess) cudaGetLastError();  // already enabled
            }
        }
    const char *how = std::getenv("PTC_MULTI_REDUCE");
    const bool wantNccl = !(how && std::strcmp(how, "p2p") == 0);
    if (wantNccl && g->nccl.load())
    {
        g->comms.assign(n_devices, nullptr);
        std::vector<int> devs(devices, devices + n_devices);
        const ncclResult_t r = g->nccl.CommInitAll(g->comms.data(), (int)n_devices, devs.data());
        if (r != ncclSuccess)
        {
            std::string msg = std::string("ncclCommInitAll: ") + g->nccl.GetErrorString(r);
            g->comms.clear();
            multi_destroy(h);
            return fail(nullptr, PTC_ERR_CUDA, "%s", msg.c_str());
        }
    }
    else if (how && std::strcmp(how, "nccl") == 0)
    {
        multi_destroy(h);
        return fail(nullptr, PTC_ERR_CUDA, "PTC_MULTI_REDUCE=nccl but libnccl.so.2 could not be loaded: %s", dlerror());
    }
    cudaError_t e = cudaSetDevice(devices[0]);
    if (g->comms.empty())
    {
        const size_t nPix = h->nPix;
        if (e == cudaSuccess) e = g->tmpColor.alloc(nPix);
        if (e == cudaSuccess) e = g->tmpNb.alloc(nPix);
        if (e == cudaSuccess) e = g->tmpHist.alloc(kHist);
        g->evMember.assign(n_devices, nullptr);
        for (uint32_t i = 1; i < n_devices && e == cudaSuccess; i++)
        {
            e = cudaSetDevice(devices[i]);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&g->evMember[i], cudaEventDisableTiming);
        }
        if (e == cudaSuccess) e = cudaSetDevice(devices[0]);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&g->evRoot, cudaEventDisableTiming);
    }
    // the display buffers exist up front: the uint8 reduce of a progressive frame names every member's
    for (ptc_ctx *m : g->members)
    {
        if (e != cudaSuccess) break;
        e = cudaSetDevice(m->device);
        if (e == cudaSuccess) e = m->rgb8.alloc((size_t)3 * m->nPix);
    }
    if (e != cudaSuccess)
    {
        std::string msg = std::string("ptc_create_multi: ") + cudaGetErrorString(e);
        multi_destroy(h);
        return fail(nullptr, PTC_ERR_CUDA, "%s", msg.c_str());
    }
    cudaSetDevice(devices[0]);
    *out_ctx = h;
    return PTC_OK;
}

extern "C" int ptc_context_devices(const ptc_ctx *ctx, int32_t *devices, uint32_t capacity)
{
    if (!ctx) return 0;
    if (!ctx->group)
    {
        if (devices && capacity) devices[0] = ctx->device;
        return 1;
    }
    const uint32_t n = (uint32_t)ctx->group->members.size();
    for (uint32_t i = 0; i < n && i < capacity && devices; i++) devices[i] = ctx->group->members[i]->device;
    return (int)n;
}

extern "C" const char *ptc_multi_reduce_kind(const ptc_ctx *ctx)
{
    if (!ctx || !ctx->group) return "none";
    return ctx->group->comms.empty() ? "p2p" : "nccl";
}
