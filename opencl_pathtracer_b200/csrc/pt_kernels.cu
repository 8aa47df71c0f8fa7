// pt_kernels.cu — the wavefront stages that replace the reference megakernel Kernel_Main
// (Kernel/PathTracer_FullKernel.cl:1180-1350). Compiled for sm_100a with -fmad=false (see pt_math.cuh).
//
//   k_reset       per-batch queue counters
//   k_generate    InitializeRandomSeed header.cl:255-264, sampler :1119-1150, camera ray :1208-1215
//   k_extend      BVH_IntersectRay :620-702          (persistent CTAs, warps pull 32-ray chunks)
//   k_connect     BVH_IntersectShadowRay :705-783 for every light of Scene_ComputeDirectIllumination :930-948
//   k_shade       hit setup :1254-1275, direct light sum :901-954, Scene_ComputeRadiance :791-891,
//                 sky :1281-1288, termination :1248, :1294-1304; ballot-compacts survivors into the next queue
//   k_accumulate  statistics :1319-1331 and per-pixel accumulation :1333-1349
#include <cstdlib>

#include "pt_launch.h"
#include "pt_shade.cuh"
#include "pt_traverse.cuh"
#include "pt_wide.cuh"

namespace ptc
{
#ifndef PTC_TRAVERSE_THREADS
#define PTC_TRAVERSE_THREADS 128  // A/B: scripts/build_variant.sh t64 -DPTC_TRAVERSE_THREADS=64
#endif
constexpr int kTraverseThreads = PTC_TRAVERSE_THREADS;
constexpr uint32_t kFlagRussianRoulette = 1u << 5;  // PTC_FLAG_RUSSIAN_ROULETTE (include/pathtracer_cuda.h)
constexpr uint32_t kFlagTextureBilinear = 1u << 6;  // PTC_FLAG_TEXTURE_BILINEAR
constexpr int kStreamThreads = 256;
#ifndef PTC_SHADE_MIN_BLOCKS
#define PTC_SHADE_MIN_BLOCKS 2  // k_shade: 2 CTAs of 16 warps per SM = 64 registers (A/B: scripts/build_variant.sh NAME -DPTC_SHADE_MIN_BLOCKS=.. -DPTC_SHADE_THREADS=..)
#endif

// the partitioned queue of a bounce (k_classify below): tiles of 1024 entries, the lit ones first
constexpr int kClassifyItems = 4;                                  // consecutive queue entries per thread of k_classify
#ifndef PTC_SHADE_THREADS
#define PTC_SHADE_THREADS 512
#endif
// k_shade is 3,200 instructions long and every warp walks most of it (2,080 per item); with the warps of an SM spread over the whole
// kernel the instruction caches thrash: `no_instruction` was its largest stall (4.2 of 13.9 cycles per issue, ncu round 2). A CTA
// barrier per item round keeps the 16 warps of a CTA in phase: shade 16.4 -> 14.9 ms per step on the bench scene, 5.1 -> 4.7 on the
// Cornell box (profiles/r02_sweeps.txt Q: 256 x 4 / 512 x 2 / 1024 x 1 threads x CTAs, with and without the barrier)
#ifndef PTC_SHADE_SYNC
#define PTC_SHADE_SYNC 1
#endif
constexpr int kShadeThreads = PTC_SHADE_THREADS;

constexpr uint32_t kClassTile = kStreamThreads * kClassifyItems;   // 1024
__device__ __forceinline__ bool entry_is_lit(const PathState &st, uint32_t k) { return (k & (kClassTile - 1u)) < st.tileLit[k / kClassTile]; }

// ------------------------------------------------------------------------------------------------ reset
__global__ void k_reset(Counters *cn, uint32_t nPaths)
{
    const int i = threadIdx.x;
    if (i <= kMaxBounces)
    {
        cn->qcount[i] = i == 0 ? nPaths : 0u;
        cn->curExtend[i] = 0;
        cn->curConnect[i] = 0;
    }
}

// ------------------------------------------------------------------------------------------------ generate
// slot k = localIteration * ownedPix + pixel of the shard; bounce 0 uses the identity queue (queue == nullptr).
__global__ void __launch_bounds__(kStreamThreads) k_generate(FrameParams fp, Camera cam, PathState st)
{
    const uint32_t nPix = fp.ownedPix;
    const uint32_t total = nPix * fp.nIterations;
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < total; k += gridDim.x * blockDim.x)
    {
        const uint32_t li = k / nPix, pix = shard_pixel(fp, k - li * nPix);
        const uint32_t y = pix / fp.width, x = pix - y * fp.width;
        const uint32_t iter = fp.firstIteration + li;
        int seed = rng_init(x, y, iter, fp.width, fp.height);
        float sx, sy;
        if (fp.sampler == 2)
        {
            // SAMPLE_UNIFORM :1124-1135
            const int sampleId = (int)(iter % 9u);
            float ox = (float)(sampleId % 3), oy = (float)(sampleId / 3);
            ox += 0.5f; oy += 0.5f;
            ox /= 3.f; oy /= 3.f;
            sx = (float)x + ox; sy = (float)y + oy;
            sx /= (float)fp.width; sy /= (float)fp.height;
            sx -= 0.5f; sy -= 0.5f;
        }
        else if (fp.sampler == 1)
        {
            // SAMPLE_RANDOM :1138-1141
            sx = rng_next(seed); sy = rng_next(seed);
            sx *= 0.9f; sy *= 0.9f;
            sx += 0.05f; sy += 0.05f;
            sx -= 0.5f; sy -= 0.5f;
        }
        else
        {
            // SAMPLE_JITTERED :1145-1146
            sx = ((float)x + 0.9f * rng_next(seed) + 0.05f) / ((float)fp.width) - 0.5f;
            sy = ((float)y + 0.9f * rng_next(seed) + 0.05f) / ((float)fp.height) - 0.5f;
        }
        const float4 shot = cam.direction + (cam.right * sx) + (cam.up * sy);  // :1213
        st.org[k] = cam.position;
        st.dir[k] = normalize4(shot);
        st.thr[k] = f4(1, 1, 1, 1);
        st.rad[k] = f4(0, 0, 0, 0);
        st.seed[k] = seed;
        st.sample[k] = make_float2(sx, sy);
        st.meta[k] = 0;
        st.nbbx[k] = 0;
        st.ntri[k] = 0;
        st.occl[k] = 0;
    }
}

// ------------------------------------------------------------------------------------------------ traversal
// Persistent kernels: warps draw work from a global cursor until the queue of this bounce is exhausted; per-thread
// traversal stacks live in shared memory, laid out [level][thread].
__device__ __forceinline__ void flush_work(Counters *cn, unsigned long long *rayCounter, uint32_t nRays, uint32_t nb, uint32_t nt)
{
    for (int o = 16; o; o >>= 1)
    {
        nb += __shfl_xor_sync(0xffffffffu, nb, o);
        nt += __shfl_xor_sync(0xffffffffu, nt, o);
        nRays += __shfl_xor_sync(0xffffffffu, nRays, o);
    }
    if ((threadIdx.x & 31u) == 0 && nRays)
    {
        atomicAdd(rayCounter, (unsigned long long)nRays);
        atomicAdd(&cn->boxTests, (unsigned long long)nb);
        atomicAdd(&cn->triTests, (unsigned long long)nt);
        if (rayCounter == &cn->extensionRays)
        {
            atomicAdd(&cn->extBoxTests, (unsigned long long)nb);
            atomicAdd(&cn->extTriTests, (unsigned long long)nt);
        }
    }
}

// One kernel for both traversal stages. ANY = false: extend (closest hit, one item per queued path). ANY = true: connect
// (any hit, one item per (queued path, light) pair).
//
// Warp-level scheduling (what keeps the SIMT lanes busy; none of it changes what a single ray does):
//   * ray staging — when a warp's staging buffer is empty, its 32 lanes together draw 32 items from the global cursor
//     (one atomicAdd), load / construct the rays with all lanes active (the long-latency state loads, the shadow-ray
//     setup and the reciprocal direction happen here, once per 32 rays) and ballot-compact the ones that need a
//     traversal into a per-warp buffer in shared memory;
//   * lane refill — a lane whose ray finished takes the next staged ray at once (a few LDS), as soon as `popMin`
//     lanes are idle, instead of waiting for the slowest ray of a fixed 32-ray chunk;
//   * majority vote — the warp executes node steps while at least half of its rays wait for a node, then triangle
//     tests while those are the majority, so the two code paths are never serialised against each other.
// The unlit / miss flags in st.occl (bits 31, 30) are written by extend and let connect drop items at staging time.
constexpr uint32_t kMetaSpecular = 0x200u;  // st.meta bit 9 (area-light variant only): the ray was sampled by a specular event
constexpr uint32_t kOcclUnlit = 0x80000000u, kOcclMiss = 0x40000000u, kOcclMask = 0x3fffffffu;
constexpr uint32_t kLightShift = 27;  // staged item word: path slot | light << 27 (ptc_create keeps slots below 2^27)

struct StagedRays  // one per warp
{
    float4 org[32];
    float4 dir[32];
    float4 invBest[32];  // reciprocal direction (header.cl:280-282) and the initial distance bound
    uint32_t item[32];
    uint32_t laneItem[32];  // the item each lane is tracing: cold state, read back once at completion instead of riding in a register
    uint32_t nRays;         // rays this warp staged (every staged ray is traced): touched once per 32 items, so it lives here, not in a register
    uint32_t pad[3];
};
constexpr size_t kStageBytes = sizeof(StagedRays) * (kTraverseThreads / 32);

// connect only, one per warp: the hit point and shading normal of the 32 paths whose shadow rays are being staged, so that
// the path state is loaded and the normal computed once per path instead of once per (path, light)
struct PathGroup
{
    float4 q[32];
    float4 ns[32];
    uint32_t slot[32];
};
constexpr size_t kGroupBytes = sizeof(PathGroup) * (kTraverseThreads / 32);

// FAST: 0 = PARITY, 1 = FAST, 2 = FAST for a scene with area lights (extension; connect only — scenes without one run the
// kernels they always ran)
// COUNT: keep the reference's per-ray work counters (Ray3D::numIntersectedBBx / numIntersectedTri, header.cl:95-96) and add them to
// the path's (they feed the two 5000-bin histograms of K.cl:1319-1331). PARITY always does; FAST only with PTC_FLAG_WORK_COUNTERS
// and in the parity probes: its counts follow the warp's votes (a parked triangle group shrinks the bound later), are compared
// with nothing, and cost two registers and two atomics per ray. Without COUNT the lane's counters run on across rays and only
// the global totals (ptc_counters.box_tests / triangle_tests) are kept.
#ifdef PTC_LANE_STATS
// Diagnostic build only (scripts/build_variant.sh stats -DPTC_LANE_STATS): what the 32 lanes of a warp are doing at every node step
// and every triangle step, summed over all warps. [0..7] extend, [8..15] connect:
// 0 node steps, 1 lanes stepping, 2 lanes without a ray, 3 lanes finished (waiting for completion), 4 lanes blocked (hold triangles
// and cannot park them), 5 triangle steps, 6 lanes testing, 7 lanes that wait for a node meanwhile
__device__ unsigned long long g_laneStats[16];
extern "C" __attribute__((visibility("default"))) int ptc_debug_lane_stats(unsigned long long *out, int reset)
{
    if (cudaMemcpyFromSymbol(out, g_laneStats, sizeof(g_laneStats)) != cudaSuccess) return 1;
    if (reset)
    {
        unsigned long long z[16] = {0};
        cudaMemcpyToSymbol(g_laneStats, z, sizeof(z));
    }
    return 0;
}
#endif

template <bool ANY, int FAST, typename LaneT, bool COUNT>
__global__ void __launch_bounds__(kTraverseThreads, (ANY ? LaneT::kMinCtasAny : LaneT::kMinCtas) * 128 / kTraverseThreads) k_trace(SceneDev sc, PathState st, const uint32_t *__restrict__ queue,
                                                           Counters *__restrict__ cn, int bounce, uint32_t popMin, int stackDepth)
{
    typedef typename LaneT::StackT StackT;
    extern __shared__ __align__(16) unsigned char smemRaw[];
    const typename LaneT::StackRef stk = LaneT::stack_ref(reinterpret_cast<StackT *>(smemRaw) + threadIdx.x);
    StagedRays &stage = reinterpret_cast<StagedRays *>(smemRaw + sizeof(StackT) * stackDepth * kTraverseThreads)[threadIdx.x >> 5];
    PathGroup &group = reinterpret_cast<PathGroup *>(smemRaw + sizeof(StackT) * stackDepth * kTraverseThreads + kStageBytes)[threadIdx.x >> 5];  // ANY only
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt = (1u << lane) - 1u;
    const unsigned full = 0xffffffffu;
    const int voteNode = (int)((popMin >> 8) & 0xffu), voteTri = (int)((popMin >> 16) & 0xffu);  // in eighths of the warp's rays
    popMin &= 0xffu;
    const uint32_t count = cn->qcount[bounce];
    const bool partitioned = ANY && st.partitioned;
    if (partitioned) queue = st.qpart;  // connect draws from the partitioned queue (k_classify): an entry's class follows from its position
    const uint32_t nL = ANY ? sc.nLights : 1u;
    const uint32_t total = ANY ? count : count * nL;  // work items drawn from the cursor: paths (connect draws groups of 32 paths)
    uint32_t *cursor = ANY ? &cn->curConnect[bounce] : &cn->curExtend[bounce];
    LaneT L;
    L.reset();
    bool exhausted = ANY && nL == 0u;
    uint32_t qHead = 0, qCount = 0;  // staged rays [qHead, qCount) are waiting (warp-uniform)
    uint32_t gMask = 0, gLight = 0;  // connect: lanes of the current path group that are lit, next light to stage for it
    bool lastGroup = false;
    uint32_t nbSum = 0, ntSum = 0;  // COUNT: per-lane sums of the finished rays' counters; else warp totals (uniform, from the votes)
    if (lane == 0) stage.nRays = 0;
#ifdef PTC_LANE_STATS
    unsigned long long ls[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#endif
    while (true)
    {
        // ---- completion: write the result of a finished ray (the counters are fire-and-forget reductions)
        if (L.finished())
        {
            const uint32_t item = lane_item(stk, &stage.laneItem[lane]);
            const uint32_t p = ANY ? (item & ((1u << kLightShift) - 1u)) : item;
            PTC_CHECK(cn, kFaultQueueItem, p, st.capacity);
            const uint32_t li = ANY ? (item >> kLightShift) : 0u;
            if (!ANY)
            {
                uint32_t hmat;
                const int hitIndex = L.hit_index(sc);
                if (hitIndex >= 0) PTC_CHECK(cn, kFaultHitIndex, hitIndex, sc.nTriangles);
                if (LaneT::kStoresHit)
                {
                    // (s, t, parameter) were stored when the hit was accepted (lane_tri_step)
                    if (hitIndex < 0) st.hit[p] = f4(__int_as_float(-1), 0.f, 0.f, 0.f);
                    else reinterpret_cast<int *>(st.hit + p)[0] = hitIndex;
                    hmat = (L.oct >> LaneT::kMatShift) & 15u;
                }
                else
                {
                    float hs, ht, htp;
                    L.hit_record(sc, hs, ht, htp, hmat);
                    st.hit[p] = f4(__int_as_float(hitIndex), hs, ht, htp);
                }
                if (COUNT)
                {
                    atomicAdd(&st.nbbx[p], L.nb);
                    atomicAdd(&st.ntri[p], L.nt);
                }
                uint32_t flags = 0;
                if (L.hitTri < 0) flags = kOcclMiss;
                else if (FAST && !(hmat == 0u || hmat == 3u)) flags = kOcclUnlit;  // GLASS / WATER / METAL / UNKNOWN never read the direct light (:805-843)
                st.occl[p] = flags;
            }
            else
            {
                if (L.found) atomicOr(&st.occl[p], 1u << li);
                if (COUNT)
                {
                    if (L.nb) atomicAdd(&st.nbbx[p], L.nb);  // :946-947 the camera ray inherits its shadow rays' counters
                    if (L.nt) atomicAdd(&st.ntri[p], L.nt);
                }
            }
            if (COUNT)
            {
                nbSum += L.nb; ntSum += L.nt;
                L.nb = L.nt = 0;
            }
            L.set_idle();
        }
        unsigned idle = __ballot_sync(full, L.idle());
        if (idle)
        {
            // ---- staging: the whole warp prepares the next 32 items
            while (qHead == qCount && !exhausted)
            {
                float4 o = f4(0, 0, 0, 0), d = o;
                float bound = INFINITY;
                uint32_t item = 0;
                bool valid = false;
                if (!ANY)
                {
                    uint32_t base = 0;
                    if (lane == 0) base = atomicAdd(cursor, 32u);
                    base = __shfl_sync(full, base, 0);
                    if (base >= total || total - base <= 32u) exhausted = true;
                    const uint32_t k = base + lane;
                    valid = base < total && k < total;
                    if (valid)
                    {
                        item = queue ? queue[k] : k;
                        PTC_CHECK(cn, kFaultQueueItem, item, st.capacity);
                        o = st.org[item];
                        d = st.dir[item];
                    }
                }
                else
                {
                    // a new group of 32 paths: state loads + hit point + shading normal, once per path
                    if (gLight >= nL || gMask == 0u)
                    {
                        uint32_t base = 0;
                        if (lane == 0) base = atomicAdd(cursor, 32u);
                        base = __shfl_sync(full, base, 0);
                        if (base >= total)
                        {
                            exhausted = true;
                            break;
                        }
                        lastGroup = total - base <= 32u;
                        const uint32_t k = base + lane;
                        bool lit = k < total;
                        uint32_t pp = 0;
                        if (lit)
                        {
                            pp = queue ? queue[k] : k;
                            PTC_CHECK(cn, kFaultQueueItem, pp, st.capacity);
                            lit = partitioned ? entry_is_lit(st, k) : !(st.occl[pp] & (kOcclUnlit | kOcclMiss));
                        }
                        if (lit)
                        {
                            const float4 h = st.hit[pp], po = st.org[pp], pd = st.dir[pp];
                            group.q[lane] = f4(po.x + pd.x * h.w, po.y + pd.y * h.w, po.z + pd.z * h.w, po.w + pd.w * h.w);  // :538
                            if (FAST)
                            {
                                const int tri = __float_as_int(h.x);
                                group.ns[lane] = shading_normal(&sc.shade[tri], __ldg(&sc.tris[tri].n), pd, h.y, h.z);
                            }
                            group.slot[lane] = pp;
                        }
                        gMask = __ballot_sync(full, lit);
                        gLight = 0;
                        __syncwarp();
                        if (gMask == 0u)
                        {
                            if (lastGroup) exhausted = true;
                            continue;
                        }
                    }
                    // the shadow rays of light gLight for the group (the group's rays of one light are adjacent items)
                    const uint32_t l = gLight;
                    valid = (gMask >> lane) & 1u;
                    if (valid)
                    {
                        const float4 q = group.q[lane];
                        PTC_CHECK(cn, kFaultLight, l, sc.nLights);
                        DevLight light = sc.lights[l];
                        if (FAST == 2 && light.type == 4) light = light_resolve(light, st.seed[group.slot[lane]]);
                        // FAST: a light that sends exactly zero power towards the shading normal (:403-421: behind the surface,
                        // outside a spot cone) adds +-0 to the pixel whether it is occluded or not: no ray
                        // (the light is also marked as occluded, so that the shade stage skips its term without evaluating it)
                        if (FAST && light_power_toward<FAST == 2>(light, q, group.ns[lane]) == 0.0f)
                        {
                            valid = false;
                            atomicOr(&st.occl[group.slot[lane]], 1u << l);
                        }
                        shadow_setup(light, q, d, bound);
                        o = q;
                        item = group.slot[lane] | (l << kLightShift);
                    }
                    gLight++;
                    if (gLight >= nL && lastGroup) exhausted = true;
                }
                const unsigned m = __ballot_sync(full, valid);
                if (valid)
                {
                    const uint32_t slot = (uint32_t)__popc(m & lt);
                    stage.org[slot] = o;
                    stage.dir[slot] = d;
                    stage.invBest[slot] = f4(1.0f / d.x, 1.0f / d.y, 1.0f / d.z, bound);  // header.cl:280-282
                    stage.item[slot] = item;
                }
                qHead = 0;
                qCount = (uint32_t)__popc(m);
                if (lane == 0) stage.nRays += qCount;
                __syncwarp();
            }
            // ---- refill: idle lanes take staged rays
            const uint32_t nIdle = (uint32_t)__popc(idle);
            if (qHead < qCount && (nIdle >= popMin || exhausted))
            {
                const uint32_t e = qHead + (uint32_t)__popc(idle & lt);
                if (L.idle() && e < qCount)
                {
                    const float4 o = stage.org[e], d = stage.dir[e], ib = stage.invBest[e];
                    lane_set_item(stk, &stage.laneItem[lane], stage.item[e]);
                    Ray r;
                    r.ox = o.x; r.oy = o.y; r.oz = o.z; r.ow = o.w;
                    r.dx = d.x; r.dy = d.y; r.dz = d.z; r.dw = d.w;
                    r.ix = ib.x; r.iy = ib.y; r.iz = ib.z;
                    // extend: the bound is INFINITY for every ray (a literal here, so that the wide lane's sqrtf(bound) folds away: the IEEE
                    // square root of infinity is a subroutine call, which every extension ray took — 2.1 M calls per launch, ncu v14)
                    lane_begin(L, sc, r, ANY ? ib.w : INFINITY);
                }
                qHead = min(qHead + nIdle, qCount);
                __syncwarp();
                idle = __ballot_sync(full, L.idle());
            }
            if (idle == full)
            {
                if (exhausted && qHead == qCount) break;
                continue;
            }
        }
        // ---- node phase: node steps while at least half of the lanes that hold a ray wait for a node
        const int nHave = 32 - __popc(__ballot_sync(full, L.idle() || L.finished()));
        int thrNode = voteNode * nHave;  // kept in a register (the compiler would re-derive it from the kernel parameter in every iteration)
        asm volatile("" : "+r"(thrNode));
        while (true)
        {
            const bool wantNode = L.wants_node();
            const int nN = __popc(__ballot_sync(full, wantNode));
            if (nN == 0 || 8 * nN < thrNode) break;
            if (!COUNT) nbSum += (uint32_t)(LaneT::kBoxesPerNodeStep * nN);
#ifdef PTC_LANE_STATS
            ls[0]++; ls[1] += nN;
            ls[2] += __popc(__ballot_sync(full, L.idle()));
            ls[3] += __popc(__ballot_sync(full, !L.idle() && L.finished()));
            ls[4] += __popc(__ballot_sync(full, !L.idle() && !L.finished() && !wantNode));
#endif
            if (wantNode) lane_node_step<(FAST != 0)>(L, sc, stk, kTraverseThreads);
        }
        // ---- triangle phase: at least one step if any lane waits for a triangle, then while they are the majority
        // (the lanes that hold a ray are counted again: keeping nHave alive across the node loop cost extend its only spill)
        const int thrTri = voteTri * (32 - __popc(__ballot_sync(full, L.idle() || L.finished())));
        while (true)
        {
            const bool wantTri = L.wants_tri();
            const int nT = __popc(__ballot_sync(full, wantTri));
            if (nT == 0) break;
            if (!COUNT) ntSum += (uint32_t)nT;
#ifdef PTC_LANE_STATS
            ls[5]++; ls[6] += nT;
            ls[7] += __popc(__ballot_sync(full, !wantTri && L.wants_node()));
#endif
            if (wantTri) lane_tri_step<ANY>(L, sc, stk, kTraverseThreads, st.hit);
            if (8 * nT <= thrTri) break;
        }
    }
#ifdef PTC_LANE_STATS
    if (lane == 0)
        for (int i = 0; i < 8; i++) atomicAdd(&g_laneStats[(ANY ? 8 : 0) + i], ls[i]);
#endif
    // warp totals live in lane 0 only (the per-lane sums of COUNT are added up by flush_work's reduction)
    flush_work(cn, ANY ? &cn->shadowRays : &cn->extensionRays, lane == 0 ? stage.nRays : 0u, (COUNT || lane == 0) ? nbSum : 0u, (COUNT || lane == 0) ? ntSum : 0u);
}

// ------------------------------------------------------------------------------------------------ classify
// Between extend and connect: the queue of the bounce is re-written tile by tile (1024 entries) with every tile PARTITIONED —
// first the paths whose hit reads the direct light (connect + the lit copy of shade), then the others (misses; FAST: hits on
// glass / water) — by the flags extend left in st.occl; tileLit[tile] = how many of the tile are lit. Later stages take the class
// of an entry from its position: warps are class-uniform except the one that straddles a tile's boundary.
// Why tile-local: ORDER MATTERS. Path slots in a queue ascend (the compaction of k_shade keeps them so within an L2-sized window),
// and the 112 bytes of state per path stream from HBM in that order. Two global class queues — whether appended to by extend in
// the order rays finish, or split off by a pass like this one with one atomic per tile — concatenate ascending sequences, bounce
// after bounce, into ever sparser sweeps of the slot range: k_shade went from 1.9 ms to 2.9 ms per launch by bounce 9, and its
// DRAM traffic up by 50 % (profiles/r02_sweeps.txt, r02_ncu_shade_*). A partition inside the tile moves no entry by more than 1023
// places, needs no atomic and no scan across tiles, and is deterministic.

__global__ void __launch_bounds__(kStreamThreads) k_classify(PathState st, const uint32_t *__restrict__ queue, Counters *__restrict__ cn, int bounce)
{
    __shared__ uint32_t sWarp[kStreamThreads / 32];
    __shared__ uint32_t sTotal;
    const uint32_t count = cn->qcount[bounce];
    const uint32_t nTiles = (count + kClassTile - 1) / kClassTile;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    for (uint32_t tile = blockIdx.x; tile < nTiles; tile += gridDim.x)
    {
        uint32_t p[kClassifyItems];
        uint32_t litMask = 0, valid = 0;
#pragma unroll
        for (int j = 0; j < kClassifyItems; j++)
        {
            const uint32_t k = tile * kClassTile + threadIdx.x * kClassifyItems + j;
            p[j] = 0;
            if (k < count)
            {
                p[j] = queue ? queue[k] : k;
                valid |= 1u << j;
                if (!(st.occl[p[j]] & (kOcclUnlit | kOcclMiss))) litMask |= 1u << j;
            }
        }
        const uint32_t nLit = (uint32_t)__popc(litMask);
        const uint32_t nValid = (uint32_t)__popc(valid);
        // exclusive prefix of the lit counts over the CTA's threads (entries of a thread are consecutive: the order is kept)
        uint32_t inc = nLit;
        for (int o = 1; o < 32; o <<= 1)
        {
            const uint32_t a = __shfl_up_sync(0xffffffffu, inc, o);
            if ((int)lane >= o) inc += a;
        }
        if (lane == 31u) sWarp[warp] = inc;
        __syncthreads();
        if (warp == 0)
        {
            uint32_t a = lane < kStreamThreads / 32 ? sWarp[lane] : 0u;
            const uint32_t own = a;
            for (int o = 1; o < 32; o <<= 1)
            {
                const uint32_t x = __shfl_up_sync(0xffffffffu, a, o);
                if ((int)lane >= o) a += x;
            }
            if (lane < kStreamThreads / 32) sWarp[lane] = a - own;
            if (lane == kStreamThreads / 32 - 1)
            {
                sTotal = a;
                st.tileLit[tile] = a;
            }
        }
        __syncthreads();
        const uint32_t litBefore = sWarp[warp] + inc - nLit;                       // lit entries of the tile before this thread's
        const uint32_t before = threadIdx.x * kClassifyItems;                       // entries of the tile before this thread's (only the last
        (void)nValid;                                                               // tile is ragged, and its missing entries are at the end)
        uint32_t posLit = tile * kClassTile + litBefore;
        uint32_t posOther = tile * kClassTile + sTotal + (before - litBefore);
#pragma unroll
        for (int j = 0; j < kClassifyItems; j++)
        {
            if (!((valid >> j) & 1u)) continue;
            if ((litMask >> j) & 1u) st.qpart[posLit++] = p[j];
            else st.qpart[posOther++] = p[j];
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------ shade
// One path of the shade stage. CLASS 0: a hit that reads the direct light (every hit in PARITY, MAT_STANDART / MAT_VARNHISHED hits
// in FAST); CLASS 1: the others (misses; FAST: hits on glass / water / unknown materials, whose direct light is never read,
// K.cl:805-843); CLASS 2: any path (the queue is not partitioned, see k_classify). The class is a compile-time constant, so the
// class copies carry only the branches their class can take: the lit copy has no sky and, in FAST, no Fresnel transmission; the
// other copy has no light loop and no cosine sampling. Returns whether the path goes on to the next bounce.
template <int FAST, int CLASS>
__device__ __forceinline__ bool shade_path(const SceneDev &sc, const FrameParams &fp, const PathState &st, const uint32_t p)
{
    const float4 o = st.org[p], d = st.dir[p], h = st.hit[p];
    float4 T = st.thr[p], rad = st.rad[p];
    uint32_t meta = st.meta[p];
    const int tri = __float_as_int(h.x);
    PTC_CHECK(sc.faultSink, kFaultQueueItem, p, st.capacity);
    if (tri >= 0) PTC_CHECK(sc.faultSink, kFaultHitIndex, tri, sc.nTriangles);
    float4 emitted;
    // area lights (extension): a camera ray or a specular chain that reaches an emitter ends there with its radiance
    if (FAST == 2 && ((meta & 0xffu) == 0u || (meta & kMetaSpecular)) && emitter_seen(sc, o, d, tri < 0 ? INFINITY : h.w, emitted))
    {
        st.rad[p] = rad + emitted * T;
        return false;
    }
    if (CLASS != 0 && tri < 0)
    {
        // :1281-1288 sky
        st.rad[p] = rad + sky_color(sc, d) * T;
        return false;
    }
    int seed = st.seed[p];
    const float s = h.y, t = h.z;
    const float4 q = f4(o.x + d.x * h.w, o.y + d.y * h.w, o.z + d.z * h.w, o.w + d.w * h.w);
    const float4 N = __ldg(&sc.tris[tri].n);
    const DevShade *sh = &sc.shade[tri];
    const float nd = dot4(N, d);
    const bool positive = nd < 0;  // :568
    const uint4 ids = __ldg(&sh->ids);
    PTC_CHECK(sc.faultSink, kFaultMaterial, positive ? ids.x : ids.y, sc.nMaterials);
    const DevMaterial mat = sc.materials[positive ? ids.x : ids.y];
    // what the class queue promises about the material (FAST: extend sorted the hit by the MaterialType in its triangle record)
    const bool litType = mat.type == 0 || mat.type == 3;
    // :591-602 surface colour of the final hit
    float4 color;
    if (mat.isSimpleColor) color = mat.color;
    else
    {
        const float4 a = __ldg(&sh->uvp12), b = __ldg(&sh->uvp3_uvn1), c = __ldg(&sh->uvn23);
        const float w0 = 1 - s - t;
        float tu, tv;
        if (positive) { tu = (a.x * w0 + a.z * s) + b.x * t; tv = (a.y * w0 + a.w * s) + b.y * t; }
        else { tu = (b.z * w0 + c.x * s) + c.z * t; tv = (b.w * w0 + c.y * s) + c.w * t; }
        PTC_CHECK(sc.faultSink, kFaultTexture, mat.textureId, sc.nTextures);
        const uint4 tx = __ldg(&sc.textures[mat.textureId]);
        color = (FAST && (fp.flags & kFlagTextureBilinear)) ? texture_lookup_bilinear(sc, tx.x, tx.y, tx.z, tu, tv)
                                                               : texture_lookup(sc, tx.x, tx.y, tx.z, tu, tv);
    }
    // :1267-1274 normals
    const bool same = dot4(d, N) > 0;
    const float4 Ng = same ? -N : N;
    const float4 Ns = shading_normal(sh, N, d, s, t);
    // :901-954 direct illumination from the occlusion bits the connect stage left
    float4 L = f4(0, 0, 0, 0);
    if (CLASS == 0 || (CLASS == 2 && (!FAST || litType)))
    {
        const uint32_t occl = st.occl[p] & kOcclMask;
        for (uint32_t li = 0; li < sc.nLights; li++)
        {
            // an occluded light adds nothing (:944-949); neither does one the connect stage found to send zero power
            // (its term would be +-0): both carry their bit, and the term is not evaluated at all
            if ((occl >> li) & 1u) continue;
            DevLight l = sc.lights[li];
            if (FAST == 2 && l.type == 4) l = light_resolve(l, seed);
            const float4 full = (l.type == 0) ? -l.direction : (l.position - q);
            const float4 ldir = normalize4(full);
            const float brdf = material_brdf(mat.type, -ldir, Ns, d);
            L = L + (light_power_toward<FAST == 2>(l, q, Ns) * brdf) * l.color;
        }
    }
    // area lights (extension): the stream moves past the draws their sample points took, lit hit or not
    if (FAST == 2) seed = (int)(((unsigned)seed * sc.areaJump) & 0x7FFFFFFFu);
    // :791-891
    float4 Nref = d, out = d;
    float4 add = f4(0, 0, 0, 0);
    bool inWater = (meta >> 8) & 1u;
    // the cosine-weighted bounce of MAT_STANDART and of the diffuse branch of MAT_VARNHISHED is drawn in one place,
    // after the per-material part, so that the lanes of both materials run it together (same draws, same order)
    bool diffuseBounce = false;
    bool specular = mat.type == 1 || mat.type == 2;  // only read by the area-light variant
    // which material branches this copy can reach: PARITY's lit class holds every material, FAST's only the two lit ones
    const bool mayBeLit = CLASS != 1, mayBeUnlit = CLASS != 0 || !FAST;
    if (mayBeLit && mat.type == 0)
    {
        T = T * color;
        add = L * T;
        diffuseBounce = true;
        Nref = Ns;
    }
    else if (mayBeUnlit && mat.type == 2)
    {
        float cos2;
        const float cos1 = -dot4(d, Ns);
        const float rF = fresnel(1, 1.55f, cos1, cos2);
        if (rng_next(seed) < rF) { out = d - (Ns * (2 * dot4(d, Ns))); Nref = Ng; }
        else { T = T * (color * (1 - mat.opacity)); Nref = d; }
    }
    else if (mayBeUnlit && mat.type == 1)
    {
        const float n1w = inWater ? 1.333f : 1.f, n2w = inWater ? 1.f : 1.333f;
        float cos2;
        const float cos1 = -dot4(d, Ns);
        const float rF = fresnel(n1w, n2w, cos1, cos2);
        // on total reflection the reference leaves refractedDirection unset and never reads it
        const float4 refr = d * (n1w / n2w) + Ns * (n1w / n2w * cos1 - cos2);
        const float coeff = (n2w * n2w) / (n1w * n1w);
        if (rng_next(seed) < rF) { out = d - (Ns * (2 * dot4(d, Ns))); Nref = Ng; }
        else { inWater = !inWater; out = refr; Nref = -Ng; T = T * coeff; }
    }
    else if (mayBeLit && mat.type == 3)
    {
        add = add + L * color * T;
        const float rF = fresnel_varnish(d, Ns);
        if (rng_next(seed) < rF) { out = d - (Ns * (2 * dot4(d, Ns))); specular = true; }
        else { diffuseBounce = true; T = T * color; }
    }
    if (mayBeLit && diffuseBounce) out = cosine_sample_hemisphere(seed, Ns);
    put_in_same_hemisphere(out, Nref);
    rad = rad + add;
    const uint32_t refl = (meta & 0xffu) + 1u;
    meta = refl | (inWater ? 0x100u : 0u) | ((FAST == 2 && specular) ? kMetaSpecular : 0u);
    st.org[p] = q + 0.001f * out;  // :880
    st.dir[p] = normalize4(out);
    st.thr[p] = T;
    if (CLASS != 1) st.rad[p] = rad;  // the other class adds nothing to the radiance at a hit
    st.seed[p] = seed;
    st.meta[p] = meta;
    // :1294-1304 and the loop condition :1248
    const float maxContribution = clmax(T.x, clmax(T.y, T.z));
    bool alive = !(maxContribution <= 0.001f) && refl < fp.maxDepth;
    // Russian roulette (opt-in, FAST only). The reference carries it disabled and commented out
    // (:12-13, :1306-1314: coefficient maxContribution / (reflectionId - MIN_REFLECTION_NUMBER), MIN = 5), and as written
    // there it would be biased (it keeps a path with probability 1 - c yet divides by c). Here: survive with
    // probability c, divide the throughput by c — unbiased; one extra draw from the path's stream.
    if (FAST && (fp.flags & kFlagRussianRoulette) && alive && refl > 5u)
    {
        const float c = maxContribution / (float)(refl - 5u);
        if (c < 1.f)
        {
            alive = rng_next(seed) < c;
            T = T / c;
            st.thr[p] = T;
            st.seed[p] = seed;
        }
    }
    return alive;
}

// PART: the queue was partitioned by k_classify and an entry's class follows from its position, so a warp runs ONE class copy of
// shade_path (the warp that straddles a tile's class boundary runs both); else the bounce's own queue with the generic copy, lanes
// diverging by material as in the reference's megakernel. Both classes go through the SAME launch in queue order because the
// survivors are appended to the next bounce's queue in the order warps get here: two launches (or two class queues) append class
// after class, and every bounce then doubles the number of ascending runs the queue is made of.
// Which one pays is decided per scene (LaunchCfg::partition): class-uniform warps execute 29 % fewer instructions at 21.9 instead of
// 16 lanes, but the stage streams 112 B of path state per path and the partition costs sector efficiency — lit warps read slot
// runs with holes, the others sparse ones: +25 % DRAM sectors. Cornell box (shade is 45 % of a step, instruction bound): 1,036 ->
// 1,322 Msamples/s. Atrium (shade 16 %, DRAM bound at 4.6 TB/s on bounce 0): 164.5 -> 153.6. (profiles/r02_sweeps.txt)
template <int FAST, bool PART>  // FAST as k_trace
__global__ void __launch_bounds__(kShadeThreads, PTC_SHADE_MIN_BLOCKS) k_shade(SceneDev sc, FrameParams fp, PathState st, const uint32_t *__restrict__ qin,
                                                                            uint32_t *__restrict__ qout, Counters *__restrict__ cn, int bounce)
{
    const uint32_t count = cn->qcount[bounce];
    const uint32_t stride = blockDim.x * gridDim.x;
    const uint32_t rounds = (count + stride - 1) / stride;
    for (uint32_t it = 0; it < rounds; it++)
    {
#if PTC_SHADE_SYNC
        __syncthreads();
#endif
        const uint32_t k = it * stride + blockIdx.x * blockDim.x + threadIdx.x;
        bool alive = false;
        uint32_t p = 0;
        if (k < count)
        {
            if (PART)
            {
                p = st.qpart[k];
                if (entry_is_lit(st, k)) alive = shade_path<FAST, 0>(sc, fp, st, p);
                else alive = shade_path<FAST, 1>(sc, fp, st, p);
            }
            else
            {
                p = qin ? qin[k] : k;
                alive = shade_path<FAST, 2>(sc, fp, st, p);
            }
        }
        // warp-aggregated append of the survivors to the next bounce's queue
        const unsigned m = __ballot_sync(0xffffffffu, alive);
        if (m)
        {
            const unsigned lane = threadIdx.x & 31u;
            const int leader = __ffs(m) - 1;
            uint32_t base = 0;
            if ((int)lane == leader) base = atomicAdd(&cn->qcount[bounce + 1], (uint32_t)__popc(m));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (alive) PTC_CHECK(cn, kFaultQueueAppend, base + __popc(m & ((1u << lane) - 1u)), st.capacity);
            if (alive) qout[base + __popc(m & ((1u << lane) - 1u))] = p;
        }
    }
}

// ------------------------------------------------------------------------------------------------ accumulate
// :1333-1336 — the double-precision literal 0.5 makes this a double computation in the reference
__device__ __forceinline__ uint32_t target_pixel(float2 sample, uint32_t W, uint32_t H)
{
    int px = (int)(((double)sample.x + 0.5) * (double)(int)W);
    int py = (int)(((double)sample.y + 0.5) * (double)(int)H);
    px = px < (int)W - 1 ? px : (int)W - 1;
    py = py < (int)H - 1 ? py : (int)H - 1;
    return (uint32_t)(py * (int)W + px);
}

// ------------------------------------------------------------------------------------------------ supersampling
// `-D SUPER_SAMPLING` (Controleur/PathTracer_OpenCL.cpp:311): Kernel_Main :1219-1222 + superSamplingStopCriteria :1152-1172. From
// iteration MIN_REFLECTION_NUMBER + 1 = 6 on, a path is dropped before it is traced — nothing is accumulated or counted for
// it — when one more draw of its stream exceeds 100 * max_rgb(imageV / n) / X2inv[(uint) n] + 0.05, n = the pixel's sample
// count so far. Runs between k_generate and the first extend: survivors are appended to the queue of bounce 0 (which is
// the identity without supersampling), dropped paths carry kMetaSkipped for k_accumulate. The statistics read are those
// of the start of the wavefront pass (PARITY runs one iteration per pass: the reference's sequential loop exactly).
// x2inv: the caller's table (the reference's Kernel/X2inv.cl is an empty file); the index is clamped to its last entry.
constexpr uint32_t kMetaSkipped = 0x80000000u;
__global__ void __launch_bounds__(kStreamThreads) k_supersample(FrameParams fp, PathState st, ImageDev img, const float *__restrict__ x2inv,
                                                                uint32_t x2invSize, uint32_t *__restrict__ queue0, Counters *cn)
{
    const uint32_t nPix = fp.ownedPix;
    const uint32_t total = nPix * fp.nIterations;
    const unsigned lane = threadIdx.x & 31u;
    for (uint32_t k0 = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u; k0 < total; k0 += gridDim.x * blockDim.x)
    {
        const uint32_t k = k0 + lane;
        bool alive = k < total;
        if (alive && fp.firstIteration + k / nPix > 5u)
        {
            int seed = st.seed[k];
            const uint32_t pix = target_pixel(st.sample[k], fp.width, fp.height);
            const float n = img.rayNb[pix];
            const float4 v = img.variance[pix];
            const float sigma2 = fmaxf(fmaxf(v.x / n, v.y / n), v.z / n);
            uint32_t idx = (uint32_t)n;
            idx = idx < x2invSize ? idx : x2invSize - 1u;
            const bool stop = (double)rng_next(seed) > (double)(100.f * sigma2 / x2inv[idx]) + 0.05;
            st.seed[k] = seed;
            if (stop)
            {
                st.meta[k] = kMetaSkipped;
                alive = false;
            }
        }
        const unsigned m = __ballot_sync(0xffffffffu, alive);
        uint32_t base = 0;
        if (lane == 0 && m) base = atomicAdd(&cn->qcount[0], (uint32_t)__popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (alive) queue0[base + __popc(m & ((1u << lane) - 1u))] = k;
    }
}

// :1339-1349 on one pixel (plain read-modify-write: the caller owns the pixel)
__device__ __forceinline__ void add_sample(const ImageDev &img, uint32_t pix, float4 x, uint32_t iteration)
{
    const float4 sumBefore = img.color[pix];
    const float4 sumAfter = sumBefore + x;
    const float nBefore = img.rayNb[pix];
    const float nAfter = nBefore + 1.f;
    img.rayNb[pix] = nAfter;
    img.color[pix] = sumAfter;
    if (img.variance)
    {
        if (iteration == 0) img.variance[pix] = f4(0, 0, 0, 0);
        else img.variance[pix] = img.variance[pix] + (x - sumBefore / nBefore) * (x - sumAfter / nAfter);
    }
}

__device__ __forceinline__ void add_sample_atomic(const ImageDev &img, uint32_t pix, float4 x)
{
    float *c = reinterpret_cast<float *>(img.color + pix);
    atomicAdd(c + 0, x.x); atomicAdd(c + 1, x.y); atomicAdd(c + 2, x.z); atomicAdd(c + 3, x.w);
    atomicAdd(img.rayNb + pix, 1.f);
}

// :1319-1331; the three histograms are privatised in shared memory (most paths of a frame fall into a few bins: global
// atomics on them serialise) and flushed once per CTA
constexpr uint32_t kHistBins = 5000u;  // MAX_INTERSECTION_NUMBER
__device__ __forceinline__ void count_path(const FrameParams &fp, const PathState &st, uint32_t k, uint32_t *sDepth, uint32_t *sBbx, uint32_t *sTri)
{
    atomicAdd(&sDepth[st.meta[k] & 0xffu], 1u);
    if (!fp.countWork) return;  // FAST without PTC_FLAG_WORK_COUNTERS: no per-path work counters exist
    const uint32_t nb = st.nbbx[k], nt = st.ntri[k];
    if (nb < kHistBins) atomicAdd(&sBbx[nb], 1u);
    if (nt < kHistBins) atomicAdd(&sTri[nt], 1u);
}

// ORDERED (jittered / uniform samplers: a path lands in its own pixel): one thread per pixel adds the
// iterations of the batch in iteration order, so sums are bit-identical to the reference's sequential loop.
// Otherwise (SAMPLE_RANDOM scatters to arbitrary pixels; the reference races there): atomics.
template <bool ORDERED>
__global__ void __launch_bounds__(kStreamThreads) k_accumulate(FrameParams fp, PathState st, ImageDev img, Counters *cn)
{
    __shared__ uint32_t sDepth[kMaxBounces + 1];
    __shared__ uint32_t sBbx[kHistBins], sTri[kHistBins];
    for (int i = threadIdx.x; i <= kMaxBounces; i += blockDim.x) sDepth[i] = 0;
    for (uint32_t i = threadIdx.x; i < kHistBins; i += blockDim.x) sBbx[i] = sTri[i] = 0;
    __syncthreads();
    const uint32_t nPix = fp.ownedPix;
    uint32_t nPaths = 0;
    if (ORDERED)
    {
        for (uint32_t opix = blockIdx.x * blockDim.x + threadIdx.x; opix < nPix; opix += gridDim.x * blockDim.x)
        {
            const uint32_t pix = shard_pixel(fp, opix);
            for (uint32_t li = 0; li < fp.nIterations; li++)
            {
                const uint32_t k = li * nPix + opix;
                if (st.meta[k] & kMetaSkipped) continue;  // dropped by k_supersample: the reference's work-item returned at :1221
                const float4 x = st.rad[k];
                const uint32_t tp = target_pixel(st.sample[k], fp.width, fp.height);
                PTC_CHECK(cn, kFaultPixel, tp, fp.width * fp.height);
                count_path(fp, st, k, sDepth, sBbx, sTri);
                if (tp == pix) add_sample(img, pix, x, fp.firstIteration + li);
                else add_sample_atomic(img, tp, x);
                nPaths++;
            }
        }
    }
    else
    {
        const uint32_t total = nPix * fp.nIterations;
        for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < total; k += gridDim.x * blockDim.x)
        {
            if (st.meta[k] & kMetaSkipped) continue;
            count_path(fp, st, k, sDepth, sBbx, sTri);
            add_sample_atomic(img, target_pixel(st.sample[k], fp.width, fp.height), st.rad[k]);
            nPaths++;
        }
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i <= fp.maxDepth && i <= (uint32_t)kMaxBounces; i += blockDim.x)
        if (sDepth[i]) atomicAdd(&img.rayDepths[i], sDepth[i]);
    for (uint32_t i = threadIdx.x; i < kHistBins && fp.countWork; i += blockDim.x)
    {
        if (sBbx[i]) atomicAdd(&img.rayBbx[i], sBbx[i]);
        if (sTri[i]) atomicAdd(&img.rayTri[i], sTri[i]);
    }
    for (int o = 16; o; o >>= 1) nPaths += __shfl_xor_sync(0xffffffffu, nPaths, o);
    if ((threadIdx.x & 31u) == 0 && nPaths) atomicAdd(&cn->paths, (unsigned long long)nPaths);
}

// ------------------------------------------------------------------------------------------------ probes / epilogues
__global__ void k_probe_primary(FrameParams fp, PathState st, int32_t *hit, float2 *stOut, float *dist2, float4 *dir,
                                int32_t *seedAfter, uint2 *counts)
{
    const uint32_t nPix = fp.width * fp.height;
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < nPix; k += gridDim.x * blockDim.x)
    {
        const float4 h = st.hit[k], o = st.org[k], d = st.dir[k];
        const int tri = __float_as_int(h.x);
        if (hit) hit[k] = tri;
        if (stOut) stOut[k] = tri < 0 ? make_float2(0.f, 0.f) : make_float2(h.y, h.z);
        if (dist2)
        {
            // |q - origin|^2 exactly as Triangle_Intersects forms it (:538-541)
            const float4 q = f4(o.x + d.x * h.w, o.y + d.y * h.w, o.z + d.z * h.w, o.w + d.w * h.w);
            const float4 f = q - o;
            dist2[k] = tri < 0 ? INFINITY : dot4(f, f);
        }
        if (dir) dir[k] = d;
        if (seedAfter) seedAfter[k] = st.seed[k];
        if (counts) counts[k] = make_uint2(st.nbbx[k], st.ntri[k]);
    }
}

// ray probe: caller-supplied rays take the place of k_generate's camera rays (Ray3D_Create header.cl:286-305 normalises)
__global__ void k_probe_rays(uint32_t n, const float4 *origin, const float4 *direction, PathState st)
{
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x)
    {
        st.org[k] = origin[k];
        st.dir[k] = normalize4(direction[k]);
        st.seed[k] = 1;
        st.meta[k] = 0;
        st.nbbx[k] = 0;
        st.ntri[k] = 0;
        st.occl[k] = 0;
    }
}

// reference Alone/PathTracer_bitmap.cpp:259-279 (its test reads .x three times; a negative red marks the pixel)
// With a pixel shard (stripeRows != 0) the rows of other shards are written as 0, so that the frames of all shards sum to the
// whole frame (one uint8 reduce composes it).
__global__ void k_rgb8(ImageDev img, uint32_t nPixels, uint8_t *rgb, uint32_t width, uint32_t stripeRows, uint32_t stripeRank, uint32_t stripeWorld)
{
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < nPixels; p += gridDim.x * blockDim.x)
    {
        const float4 c = img.color[p];
        const float n = img.rayNb[p];
        int r, g, b;
        if (stripeRows && ((p / width) / stripeRows) % stripeWorld != stripeRank) { r = g = b = 0; }
        else if (c.x < 0) { r = 255; g = 0; b = 0; }
        else
        {
            r = (int)fminf(c.x * 255.f / n, 255.f);
            g = (int)fminf(c.y * 255.f / n, 255.f);
            b = (int)fminf(c.z * 255.f / n, 255.f);
        }
        rgb[3 * p + 0] = (uint8_t)r;
        rgb[3 * p + 1] = (uint8_t)g;
        rgb[3 * p + 2] = (uint8_t)b;
    }
}

// ptc_reset_accumulation: every accumulator, histogram and counter in ONE launch (seven memsets before: a progressive frame
// resets once per frame, and a multi-device context once per frame and device)
__global__ void __launch_bounds__(kStreamThreads) k_clear(ImageDev img, uint32_t nPixels, Counters *cn)
{
    const float4 z = f4(0, 0, 0, 0);
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < nPixels; p += gridDim.x * blockDim.x)
    {
        img.color[p] = z;
        img.rayNb[p] = 0.f;
        if (img.variance) img.variance[p] = z;
    }
    if (blockIdx.x == 0)
    {
        for (uint32_t i = threadIdx.x; i < kHistBins; i += blockDim.x) img.rayBbx[i] = img.rayTri[i] = 0u;
        for (uint32_t i = threadIdx.x; i <= (uint32_t)kMaxBounces; i += blockDim.x) img.rayDepths[i] = 0u;
        uint32_t *c = reinterpret_cast<uint32_t *>(cn);
        for (uint32_t i = threadIdx.x; i < sizeof(Counters) / 4; i += blockDim.x) c[i] = 0u;
    }
}

__global__ void k_add_image(ImageDev img, uint32_t nPixels, const float4 *color, const float *rayNb)
{
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < nPixels; p += gridDim.x * blockDim.x)
    {
        if (color) img.color[p] = img.color[p] + color[p];
        if (rayNb) img.rayNb[p] = img.rayNb[p] + rayNb[p];
    }
}

// Roofline denominator: read-only streaming over a buffer with the loads the traversal uses (256-bit, L1 bypassed so that a
// buffer smaller than the L2 measures L2 hit bandwidth and a larger one HBM read bandwidth).
__global__ void __launch_bounds__(kStreamThreads) k_read_stream(const uint4 *__restrict__ buf, size_t nVec32, uint32_t repeats, uint32_t *sink)
{
    uint32_t acc = 0;
    for (uint32_t r = 0; r < repeats; r++)
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nVec32; i += (size_t)gridDim.x * blockDim.x)
        {
            uint32_t a, b, c, d, e, f, g, h;
            asm volatile("ld.global.cg.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                         : "=r"(a), "=r"(b), "=r"(c), "=r"(d), "=r"(e), "=r"(f), "=r"(g), "=r"(h)
                         : "l"(reinterpret_cast<const char *>(buf) + i * 32));
            acc ^= a ^ b ^ c ^ d ^ e ^ f ^ g ^ h;
        }
    if (acc == 0x9e3779b9u) *sink = acc;  // keeps the loads alive
}

// ------------------------------------------------------------------------------------------------ launchers
void launch_read_stream(cudaStream_t s, int grid, const void *buf, size_t bytes, uint32_t repeats, uint32_t *sink)
{
    k_read_stream<<<grid, kStreamThreads, 0, s>>>(reinterpret_cast<const uint4 *>(buf), bytes / 32, repeats, sink);
}

static inline size_t stack_bytes(bool wide, int stackDepth, bool any)
{
    return (size_t)stackDepth * kTraverseThreads * (wide ? sizeof(uint2) : sizeof(uint32_t)) + kStageBytes + (any ? kGroupBytes : 0);
}

template <bool ANY, int FAST, typename LaneT>
static int occupancy_of(size_t bytes)
{
    // deep trees: more than the default 48 KB of dynamic shared memory needs the opt-in (the device builder on degenerate
    // geometry can emit them; without it every launch fails with "invalid argument")
    if (bytes > 48u * 1024u)
    {
        cudaFuncSetAttribute(k_trace<ANY, FAST, LaneT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
        cudaFuncSetAttribute(k_trace<ANY, FAST, LaneT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    }
    int n = 0, m = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_trace<ANY, FAST, LaneT, false>, kTraverseThreads, bytes);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&m, k_trace<ANY, FAST, LaneT, true>, kTraverseThreads, bytes);
    return n < m ? n : m;
}

size_t traverse_shared_bytes(bool wide, int stackDepth, bool any) { return stack_bytes(wide, stackDepth, any); }

int traverse_ctas_per_sm(bool fast, bool wide, int stackDepth, bool any)
{
    if (const char *e = std::getenv("PTC_CARVEOUT"))  // experiment: shared-memory carve-out in percent (what is left is L1)
    {
        const int v = std::atoi(e);
        cudaFuncSetAttribute(k_trace<false, 1, WideLane, false>, cudaFuncAttributePreferredSharedMemoryCarveout, v);
        cudaFuncSetAttribute(k_trace<true, 1, WideLane, false>, cudaFuncAttributePreferredSharedMemoryCarveout, v);
    }
    const size_t bytes = stack_bytes(wide, stackDepth, false), bytesAny = stack_bytes(wide, stackDepth, true);
    int n, m;
    if (wide) { n = occupancy_of<false, 1, WideLane>(bytes); m = occupancy_of<true, 1, WideLane>(bytesAny); }
    else if (fast) { n = occupancy_of<false, 1, TraceLane>(bytes); m = occupancy_of<true, 1, TraceLane>(bytesAny); }
    else { n = occupancy_of<false, 0, TraceLane>(bytes); m = occupancy_of<true, 0, TraceLane>(bytesAny); }
    n = any ? m : n;
    return n < 1 ? 1 : n;
}

void launch_reset(cudaStream_t s, Counters *cn, uint32_t nPaths) { k_reset<<<1, 128, 0, s>>>(cn, nPaths); }

void launch_generate(cudaStream_t s, const LaunchCfg &lc, const FrameParams &fp, const Camera &cam, const PathState &st)
{
    k_generate<<<lc.streamGrid, kStreamThreads, 0, s>>>(fp, cam, st);
}

void launch_supersample(cudaStream_t s, const LaunchCfg &lc, const FrameParams &fp, const PathState &st, const ImageDev &img,
                        const float *x2inv, uint32_t x2invSize, uint32_t *queue0, Counters *cn)
{
    k_supersample<<<lc.streamGrid, kStreamThreads, 0, s>>>(fp, st, img, x2inv, x2invSize, queue0, cn);
}

template <bool ANY>
static void launch_trace(cudaStream_t s, const LaunchCfg &lc, bool fast, const SceneDev &sc, const PathState &st, const uint32_t *queue,
                         Counters *cn, int bounce)
{
    const size_t smem = stack_bytes(lc.wide, lc.stackDepth, ANY);
    const bool area = ANY && sc.areaJump != 1u;  // the scene holds area lights: the connect variant that samples them
#define PTC_TRACE_(MODE, LANE, CNT) k_trace<ANY, MODE, LANE, CNT><<<(ANY ? lc.connectGrid : lc.traverseGrid), kTraverseThreads, smem, s>>>(sc, st, queue, cn, bounce, lc.refillMin, lc.stackDepth)
#define PTC_TRACE(MODE, LANE) do { if (lc.count) PTC_TRACE_(MODE, LANE, true); else PTC_TRACE_(MODE, LANE, false); } while (0)
    if (lc.wide) { if (area) PTC_TRACE(ANY ? 2 : 1, WideLane); else PTC_TRACE(1, WideLane); }
    else if (fast) { if (area) PTC_TRACE(ANY ? 2 : 1, TraceLane); else PTC_TRACE(1, TraceLane); }
    else PTC_TRACE_(0, TraceLane, true);
#undef PTC_TRACE
#undef PTC_TRACE_
}

void launch_extend(cudaStream_t s, const LaunchCfg &lc, bool fast, const SceneDev &sc, const PathState &st, const uint32_t *queue,
                   Counters *cn, int bounce)
{
    launch_trace<false>(s, lc, fast, sc, st, queue, cn, bounce);
}

void launch_connect(cudaStream_t s, const LaunchCfg &lc, bool fast, const SceneDev &sc, const PathState &st, const uint32_t *queue,
                    Counters *cn, int bounce)
{
    launch_trace<true>(s, lc, fast, sc, st, queue, cn, bounce);
}

void launch_classify(cudaStream_t s, const LaunchCfg &lc, const PathState &st, const uint32_t *queue, Counters *cn, int bounce)
{
    k_classify<<<lc.streamGrid, kStreamThreads, 0, s>>>(st, queue, cn, bounce);
}

void launch_shade(cudaStream_t s, const LaunchCfg &lc, bool fast, const SceneDev &sc, const FrameParams &fp, const PathState &st,
                  const uint32_t *qin, uint32_t *qout, Counters *cn, int bounce)
{
#define PTC_SHADE(MODE)                                                                                                    \
    do                                                                                                                     \
    {                                                                                                                      \
        if (lc.partition) k_shade<MODE, true><<<lc.streamGrid * kStreamThreads / kShadeThreads, kShadeThreads, 0, s>>>(sc, fp, st, qin, qout, cn, bounce); \
        else k_shade<MODE, false><<<lc.streamGrid * kStreamThreads / kShadeThreads, kShadeThreads, 0, s>>>(sc, fp, st, qin, qout, cn, bounce); \
    } while (0)
    if (fast && sc.areaJump != 1u) PTC_SHADE(2);
    else if (fast) PTC_SHADE(1);
    else PTC_SHADE(0);
#undef PTC_SHADE
}

void launch_accumulate(cudaStream_t s, const LaunchCfg &lc, const FrameParams &fp, const PathState &st, const ImageDev &img,
                       Counters *cn)
{
    const int grid = lc.streamGrid / 2;  // 40 KB of histograms per CTA: 4 CTAs per SM
    if (fp.sampler == 1) k_accumulate<false><<<grid, kStreamThreads, 0, s>>>(fp, st, img, cn);
    else k_accumulate<true><<<grid, kStreamThreads, 0, s>>>(fp, st, img, cn);
}

// The two kernels whose arguments change from one batch to the next (iteration numbers, camera): a captured batch graph is
// replayed with these two nodes re-parameterised (pathtracer_cuda.cu). Argument order as in the launches above.
const void *generate_kernel() { return reinterpret_cast<const void *>(&k_generate); }
const void *accumulate_kernel(uint32_t sampler)
{
    return sampler == 1 ? reinterpret_cast<const void *>(&k_accumulate<false>) : reinterpret_cast<const void *>(&k_accumulate<true>);
}

void launch_probe_primary(cudaStream_t s, const FrameParams &fp, const PathState &st, int32_t *hit, float2 *stOut, float *dist2,
                          float4 *dir, int32_t *seedAfter, uint2 *counts)
{
    k_probe_primary<<<1024, 256, 0, s>>>(fp, st, hit, stOut, dist2, dir, seedAfter, counts);
}

void launch_probe_rays(cudaStream_t s, uint32_t n, const float4 *origin, const float4 *direction, const PathState &st)
{
    k_probe_rays<<<1024, 256, 0, s>>>(n, origin, direction, st);
}

void launch_rgb8(cudaStream_t s, const ImageDev &img, const FrameParams &fp, uint8_t *rgb)
{
    k_rgb8<<<1024, 256, 0, s>>>(img, fp.width * fp.height, rgb, fp.width, fp.stripeRows, fp.stripeRank, fp.stripeWorld);
}

void launch_clear(cudaStream_t s, const LaunchCfg &lc, const ImageDev &img, uint32_t nPixels, Counters *cn)
{
    k_clear<<<lc.streamGrid, kStreamThreads, 0, s>>>(img, nPixels, cn);
}

void launch_add_image(cudaStream_t s, const ImageDev &img, uint32_t nPixels, const float4 *color, const float *rayNb)
{
    k_add_image<<<1024, 256, 0, s>>>(img, nPixels, color, rayNb);
}

}  // namespace ptc
