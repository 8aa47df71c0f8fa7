// pt_lbvh.cu — the FAST-mode wide BVH built entirely on the device. The reference builds its BVH2 on
// one host thread (BVH_Create / BVH_BuildStructure, Controleur/PathTracer_BVH.cpp:12-356); for the 10 M-triangle
// configuration that is tens of seconds before the first sample. This builder takes the triangles as given (no BVH from
// the caller) and produces the same DevNode8 / 24-records-per-node structure pt_wide_build.cu does on the host:
//
//   1. k_scene_bounds     centroid bounds of the scene (block reduction + ordered-integer atomics)
//   2. k_morton           63-bit Morton code of every centroid; cub::DeviceRadixSort orders (code, triangle) pairs
//   3. k_leaves           one leaf per triangle in Morton order (box from the three vertices)
//      k_internal         the n - 1 inner nodes of the binary radix tree, all in parallel (Karras, "Maximizing Parallelism
//                         in the Construction of BVHs, Octrees, and k-d Trees", HPG 2012)
//   4. k_bottom_up        boxes and the collapse costs C(n, 1..7) of pt_wide_collapse.cuh, leaves to root; the second thread
//                         to reach a node continues upwards (one atomic counter per node)
//   5. k_emit             wide nodes level by level: one thread per wide node gathers its <= 8 children from the recorded
//                         decisions, reserves consecutive indices for the inner ones (one atomicAdd), assigns slots and
//                         quantises; the fixed record range 24 * node + 3 * slot + k needs no allocation
//   6. k_gather_records   the intersection records in wide order
//
// A Morton-order tree is a weaker partition than the reference's SAH sweep (more node visits per ray); it is offered as
// an option (PTC_FLAG_GPU_BVH) for scenes whose set-up time matters, not as the default.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_select.cuh>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "pt_lbvh.h"
#include "pt_wide_collapse.cuh"

namespace ptc
{
namespace
{
constexpr int kThreads = 256;

__device__ __forceinline__ uint32_t ordered(float f)
{
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float unordered(uint32_t u)
{
    const uint32_t v = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#if defined(__CUDA_ARCH__)
    return __uint_as_float(v);
#else
    float f;
    memcpy(&f, &v, 4);
    return f;
#endif
}

// bounds[0..2] = min, bounds[3..5] = max of the centroids, as ordered integers (initialised to 0xffffffff / 0)
__global__ void __launch_bounds__(kThreads) k_scene_bounds(const float *__restrict__ verts, uint32_t n, uint32_t *bounds)
{
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    {
        const float *v = verts + 9 * (size_t)i;
        for (int a = 0; a < 3; a++)
        {
            const float c = (v[a] + v[3 + a] + v[6 + a]) * (1.f / 3.f);
            lo[a] = fminf(lo[a], c);
            hi[a] = fmaxf(hi[a], c);
        }
    }
    for (int a = 0; a < 3; a++)
    {
        for (int o = 16; o; o >>= 1)
        {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
        }
        if ((threadIdx.x & 31) == 0 && lo[a] <= hi[a])
        {
            atomicMin(&bounds[a], ordered(lo[a]));
            atomicMax(&bounds[3 + a], ordered(hi[a]));
        }
    }
}

__device__ __forceinline__ uint64_t spread21(uint32_t x)  // bit i -> bit 3 i
{
    uint64_t v = x & 0x1fffffu;
    v = (v | (v << 32)) & 0x1f00000000ffffull;
    v = (v | (v << 16)) & 0x1f0000ff0000ffull;
    v = (v | (v << 8)) & 0x100f00f00f00f00full;
    v = (v | (v << 4)) & 0x10c30c30c30c30c3ull;
    v = (v | (v << 2)) & 0x1249249249249249ull;
    return v;
}

__global__ void __launch_bounds__(kThreads) k_morton(const float *__restrict__ verts, uint32_t n, const uint32_t *__restrict__ bounds,
                                                     uint64_t *keys, uint32_t *vals)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *v = verts + 9 * (size_t)i;
    uint32_t q[3];
    for (int a = 0; a < 3; a++)
    {
        const float lo = unordered(bounds[a]), hi = unordered(bounds[3 + a]);
        const float c = (v[a] + v[3 + a] + v[6 + a]) * (1.f / 3.f);
        const float ext = hi - lo;
        float t = ext > 0.f ? (c - lo) / ext : 0.f;
        t = fminf(fmaxf(t, 0.f), 1.f);
        q[a] = min((uint32_t)(t * 2097152.f), 2097151u);
    }
    keys[i] = (spread21(q[0]) << 2) | (spread21(q[1]) << 1) | spread21(q[2]);
    vals[i] = i;
}

// leaf j (Morton position j) is binary node n - 1 + j
__global__ void __launch_bounds__(kThreads) k_leaves(const float *__restrict__ verts, const uint32_t *__restrict__ perm, uint32_t n, BinNode *b,
                                                     CollapseTables tab)
{
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const float *v = verts + 9 * (size_t)perm[j];
    BinNode x;
    for (int a = 0; a < 3; a++)
    {
        x.lo[a] = fminf(v[a], fminf(v[3 + a], v[6 + a]));
        x.hi[a] = fmaxf(v[a], fmaxf(v[3 + a], v[6 + a]));
    }
    x.left = x.right = -1;
    x.first = j;
    x.count = 1;
    const size_t me = (size_t)n - 1 + j;
    b[me] = x;
    float c[7];
    uint8_t sp[7], s8, leaf;
    collapse_costs_core(x, nullptr, nullptr, c, sp, s8, leaf);
    for (int i = 0; i < 7; i++) { tab.cost[me * 7 + i] = c[i]; tab.split[me * 7 + i] = sp[i]; }
    tab.split8[me] = s8;
    tab.asLeaf[me] = leaf;
}

// ---- PLOC (Meister, Bittner: "Parallel Locally-Ordered Clustering for Bounding Volume Hierarchy Construction", TVCG 2018):
// the clusters (at first the leaves in Morton order) repeatedly merge with their nearest neighbour — nearest = smallest
// surface area of the joint box — within a window of kPlocRadius positions, whenever the choice is mutual. Far better trees
// than the radix tree for a few more passes over a shrinking array.
constexpr int kPlocRadius = 16;
constexpr uint32_t kNoCluster = 0xffffffffu;

__global__ void __launch_bounds__(kThreads) k_ploc_nearest(const BinNode *__restrict__ b, const uint32_t *__restrict__ cluster, uint32_t nc,
                                                           uint32_t *nearest)
{
    __shared__ float tile[kThreads + 2 * kPlocRadius][6];
    const int base = (int)(blockIdx.x * kThreads) - kPlocRadius;
    for (int t = threadIdx.x; t < kThreads + 2 * kPlocRadius; t += kThreads)
    {
        const int g = base + t;
        if (g >= 0 && g < (int)nc)
        {
            const BinNode &x = b[cluster[g]];
            for (int a = 0; a < 3; a++) { tile[t][a] = x.lo[a]; tile[t][3 + a] = x.hi[a]; }
        }
        else
            for (int a = 0; a < 6; a++) tile[t][a] = a < 3 ? INFINITY : -INFINITY;
    }
    __syncthreads();
    const int me = threadIdx.x + kPlocRadius;
    const int g = base + me;
    if (g >= (int)nc) return;
    float best = INFINITY;
    int bestJ = -1;
    for (int d = -kPlocRadius; d <= kPlocRadius; d++)
    {
        const int j = me + d, gj = base + j;
        if (d == 0 || gj < 0 || gj >= (int)nc) continue;
        const float ex = fmaxf(tile[me][3], tile[j][3]) - fminf(tile[me][0], tile[j][0]);
        const float ey = fmaxf(tile[me][4], tile[j][4]) - fminf(tile[me][1], tile[j][1]);
        const float ez = fmaxf(tile[me][5], tile[j][5]) - fminf(tile[me][2], tile[j][2]);
        const float area = ex * ey + ey * ez + ez * ex;  // symmetric in (me, j): both sides see the same value
        if (area < best || bestJ < 0) { best = area; bestJ = gj; }  // ascending d: ties go to the lower index
    }
    nearest[g] = bestJ < 0 ? kNoCluster : (uint32_t)bestJ;
}

// mutual choices merge: the lower position keeps the new node, the higher one leaves the array
__global__ void __launch_bounds__(kThreads) k_ploc_merge(BinNode *b, CollapseTables tab, const uint32_t *__restrict__ cluster,
                                                         const uint32_t *__restrict__ nearest, uint32_t nc, uint32_t *out, uint32_t *nodeCounter)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nc) return;
    const uint32_t j = nearest[i];
    if (j == kNoCluster || nearest[j] != i)
    {
        out[i] = cluster[i];
        return;
    }
    if (i > j)
    {
        out[i] = kNoCluster;
        return;
    }
    const uint32_t idx = atomicAdd(nodeCounter, 1u);
    const int l = (int)cluster[i], r = (int)cluster[j];
    BinNode x;
    for (int a = 0; a < 3; a++)
    {
        x.lo[a] = fminf(b[l].lo[a], b[r].lo[a]);
        x.hi[a] = fmaxf(b[l].hi[a], b[r].hi[a]);
    }
    x.left = l;
    x.right = r;
    x.first = 0;
    x.count = b[l].count + b[r].count;
    b[idx] = x;
    collapse_costs(b, idx, tab);  // the children are final: they were finished by earlier launches
    out[i] = idx;
}

struct IsCluster
{
    __host__ __device__ bool operator()(const uint32_t &v) const { return v != 0xffffffffu; }
};

// length of the common prefix of the keys at Morton positions i and j, ties broken by the position itself; -1 outside
__device__ __forceinline__ int delta(const uint64_t *__restrict__ keys, int n, int i, int j)
{
    if (j < 0 || j >= n) return -1;
    const uint64_t a = keys[i], c = keys[j];
    if (a == c) return 64 + __clz(i ^ j);
    return __clzll((long long)(a ^ c));
}

__global__ void __launch_bounds__(kThreads) k_internal(const uint64_t *__restrict__ keys, int n, BinNode *b, int *parent)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, j);
    int s = 0, t = l;
    do
    {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const int left = (lo == gamma) ? (n - 1 + gamma) : gamma;
    const int right = (hi == gamma + 1) ? (n - 1 + gamma + 1) : (gamma + 1);
    BinNode &x = b[i];
    x.left = left;
    x.right = right;
    x.first = (uint32_t)lo;
    x.count = (uint32_t)(hi - lo + 1);
    parent[left] = i;
    parent[right] = i;
    if (i == 0) parent[0] = -1;
}

__device__ __forceinline__ void load_costs(const float *p, float c[7])
{
    for (int i = 0; i < 7; i++) c[i] = __ldcg(p + i);
}

// One thread per leaf. Everything a thread reads from another thread's node goes through L2 (__ldcg); a node's data is
// fenced before its parent's counter is bumped.
__global__ void __launch_bounds__(kThreads) k_bottom_up(int n, BinNode *b, const int *__restrict__ parent, int *visits, CollapseTables tab)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    int cur = n - 1 + j;
    {
        const BinNode x = b[cur];
        float c[7];
        uint8_t sp[7], s8, leaf;
        collapse_costs_core(x, nullptr, nullptr, c, sp, s8, leaf);
        for (int i = 0; i < 7; i++) { tab.cost[(size_t)cur * 7 + i] = c[i]; tab.split[(size_t)cur * 7 + i] = sp[i]; }
        tab.split8[cur] = s8;
        tab.asLeaf[cur] = leaf;
    }
    while (true)
    {
        const int p = parent[cur];
        if (p < 0) return;
        __threadfence();
        if (atomicAdd(&visits[p], 1) == 0) return;  // the sibling subtree is not finished: its last thread continues
        __threadfence();
        BinNode x;
        x.left = __ldcg(&b[p].left);
        x.right = __ldcg(&b[p].right);
        x.first = __ldcg(&b[p].first);
        x.count = __ldcg(&b[p].count);
        for (int a = 0; a < 3; a++)
        {
            x.lo[a] = fminf(__ldcg(&b[x.left].lo[a]), __ldcg(&b[x.right].lo[a]));
            x.hi[a] = fmaxf(__ldcg(&b[x.left].hi[a]), __ldcg(&b[x.right].hi[a]));
            b[p].lo[a] = x.lo[a];
            b[p].hi[a] = x.hi[a];
        }
        float cl[7], cr[7], c[7];
        uint8_t sp[7], s8, leaf;
        load_costs(tab.cost + (size_t)x.left * 7, cl);
        load_costs(tab.cost + (size_t)x.right * 7, cr);
        collapse_costs_core(x, cl, cr, c, sp, s8, leaf);
        for (int i = 0; i < 7; i++) { tab.cost[(size_t)p * 7 + i] = c[i]; tab.split[(size_t)p * 7 + i] = sp[i]; }
        tab.split8[p] = s8;
        tab.asLeaf[p] = leaf;
        cur = p;
    }
}

// wide nodes [start, end) of one level. nodeBin[i] = the binary node that wide node i stands for.
__global__ void __launch_bounds__(128) k_emit(const BinNode *__restrict__ b, CollapseTables tab, const uint32_t *__restrict__ perm, int *nodeBin,
                                             uint32_t start, uint32_t end, DevNode8 *nodes, uint32_t *order, uint32_t *nNodes, uint32_t capacity,
                                             uint32_t *error)
{
    const uint32_t i = start + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= end) return;
    int child[8], innerChild[8], nInner = 0;
    const int nc = collapse_gather(b, nodeBin[i], tab, child);
    int want = 0;
    for (int k = 0; k < nc; k++) want += tab.asLeaf[child[k]] ? 0 : 1;
    const uint32_t childBase = want ? atomicAdd(nNodes, (uint32_t)want) : 0u;
    if (want && childBase + (uint32_t)want > capacity)
    {
        atomicOr(error, 1u);
        want = -1;
    }
    DevNode8 nd;
    uint32_t rec[kTriSlots];
    if (!collapse_encode(b, tab, child, nc, childBase, perm, nd, rec, innerChild, nInner)) atomicOr(error, 2u);
    if (want < 0) nd.eimask &= 0x00ffffffu;  // overflow: no inner children (the build is repeated with more room)
    nodes[i] = nd;
    for (uint32_t k = 0; k < kTriSlots; k++) order[(size_t)i * kTriSlots + k] = rec[k];
    if (want > 0)
        for (int r = 0; r < nInner; r++) nodeBin[childBase + r] = innerChild[r];
}

__global__ void __launch_bounds__(kThreads) k_gather_records(const uint32_t *__restrict__ order, const DevTri *__restrict__ tris, size_t nRecords,
                                                             DevTri *tris8)
{
    const size_t r = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nRecords) return;
    const uint32_t o = order[r];
    DevTri t;
    if (o != kNoTriangle) t = tris[o];
    else t.n = t.p = t.e1 = t.e2 = make_float4(0.f, 0.f, 0.f, 0.f);  // never fetched (no bit in DevNode8::valid)
    tris8[r] = t;
}

struct Scratch  // freed on every exit path
{
    std::vector<void *> ptrs;
    ~Scratch() { for (void *p : ptrs) cudaFree(p); }
    template <typename T>
    bool alloc(T *&p, size_t count)
    {
        void *q = nullptr;
        if (cudaMalloc(&q, (count ? count : 1) * sizeof(T)) != cudaSuccess) return false;
        ptrs.push_back(q);
        p = (T *)q;
        return true;
    }
    void release(void *p)  // hand a buffer over to the caller
    {
        for (auto &q : ptrs) if (q == p) q = nullptr;
    }
};
}  // namespace

#define LB_CK(call)                                                                                  \
    do                                                                                               \
    {                                                                                                \
        cudaError_t e_ = (call);                                                                     \
        if (e_ != cudaSuccess)                                                                       \
        {                                                                                            \
            err = std::string("device BVH build: ") + #call + ": " + cudaGetErrorString(e_);         \
            return false;                                                                            \
        }                                                                                            \
    } while (0)

bool build_wide_bvh_device(cudaStream_t stream, const float *verts, const DevTri *tris, uint32_t nT, DeviceWideBvh &out, std::string &err)
{
    out = DeviceWideBvh();
    if (nT == 0) return true;
    if (nT >= (1u << 30)) { err = "device BVH build: too many triangles"; return false; }
    const int n = (int)nT;
    const size_t nBin = 2 * (size_t)nT - 1;
    Scratch S;
    uint32_t *bounds = nullptr, *vals = nullptr, *perm = nullptr, *counters = nullptr;
    uint64_t *keys = nullptr, *keysSorted = nullptr;
    BinNode *b = nullptr;
    int *parent = nullptr, *visits = nullptr, *nodeBin = nullptr;
    CollapseTables tab{};
    void *sortTemp = nullptr;
    size_t sortBytes = 0;
    if (!S.alloc(bounds, 6) || !S.alloc(vals, nT) || !S.alloc(perm, nT) || !S.alloc(keys, nT) || !S.alloc(keysSorted, nT) || !S.alloc(b, nBin) ||
        !S.alloc(parent, nBin) || !S.alloc(visits, nT) || !S.alloc(tab.cost, nBin * 7) || !S.alloc(tab.split, nBin * 7) ||
        !S.alloc(tab.split8, nBin) || !S.alloc(tab.asLeaf, nBin) || !S.alloc(counters, 2))
    {
        err = "device BVH build: out of device memory";
        return false;
    }
    LB_CK(cub::DeviceRadixSort::SortPairs(nullptr, sortBytes, keys, keysSorted, vals, perm, n, 0, 63, stream));
    {
        char *t = nullptr;
        if (!S.alloc(t, sortBytes)) { err = "device BVH build: out of device memory"; return false; }
        sortTemp = t;
    }
    cudaEvent_t e0, e1;
    LB_CK(cudaEventCreate(&e0));
    LB_CK(cudaEventCreate(&e1));
    struct EventGuard { cudaEvent_t a, b; ~EventGuard() { cudaEventDestroy(a); cudaEventDestroy(b); } } guard{e0, e1};
    LB_CK(cudaEventRecord(e0, stream));

    const int grid = (n + kThreads - 1) / kThreads;
    // ---- 1 + 2: Morton order
    const uint32_t initBounds[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
    LB_CK(cudaMemcpyAsync(bounds, initBounds, sizeof(initBounds), cudaMemcpyHostToDevice, stream));
    k_scene_bounds<<<std::min(grid, 148 * 8), kThreads, 0, stream>>>(verts, nT, bounds);
    k_morton<<<grid, kThreads, 0, stream>>>(verts, nT, bounds, keys, vals);
    LB_CK(cub::DeviceRadixSort::SortPairs(sortTemp, sortBytes, keys, keysSorted, vals, perm, n, 0, 63, stream));
    // ---- 3: binary radix tree
    LB_CK(cudaMemsetAsync(parent, 0xff, sizeof(int) * nBin, stream));
    LB_CK(cudaMemsetAsync(visits, 0, sizeof(int) * nT, stream));
    k_leaves<<<grid, kThreads, 0, stream>>>(verts, perm, nT, b, tab);
    int rootBin = 0;  // inner node 0 of the radix tree, or the only leaf (n - 1 + 0 = 0)
    const char *which = std::getenv("PTC_GPU_BVH_BUILDER");
    if (which && std::strcmp(which, "lbvh") == 0)
    {
        if (n > 1) k_internal<<<grid, kThreads, 0, stream>>>(keysSorted, n, b, parent);
        // ---- 4: boxes + collapse costs
        k_bottom_up<<<grid, kThreads, 0, stream>>>(n, b, parent, visits, tab);
        LB_CK(cudaGetLastError());
    }
    else if (n > 1)
    {
        // ---- 3' + 4': PLOC; boxes and collapse costs are finished as each node is created
        uint32_t *cl[2] = {nullptr, nullptr}, *nearest = nullptr, *merged = nullptr, *nSelected = nullptr;
        void *selTemp = nullptr;
        size_t selBytes = 0;
        if (!S.alloc(cl[0], nT) || !S.alloc(cl[1], nT) || !S.alloc(nearest, nT) || !S.alloc(merged, nT) || !S.alloc(nSelected, 1))
        {
            err = "device BVH build: out of device memory";
            return false;
        }
        LB_CK(cub::DeviceSelect::If(nullptr, selBytes, merged, cl[1], nSelected, n, IsCluster(), stream));
        {
            char *t = nullptr;
            if (!S.alloc(t, selBytes)) { err = "device BVH build: out of device memory"; return false; }
            selTemp = t;
        }
        // cluster[j] = leaf j; node indices 0 .. n - 2 are handed out as merges happen (counters[0] doubles as that counter here)
        {
            std::vector<uint32_t> init(nT);
            for (uint32_t j = 0; j < nT; j++) init[j] = nT - 1 + j;
            LB_CK(cudaMemcpyAsync(cl[0], init.data(), sizeof(uint32_t) * nT, cudaMemcpyHostToDevice, stream));
            LB_CK(cudaMemsetAsync(counters, 0, sizeof(uint32_t) * 2, stream));
            LB_CK(cudaStreamSynchronize(stream));
        }
        uint32_t nc = nT;
        int cur = 0;
        for (int pass = 0; nc > 1; pass++)
        {
            if (pass > 4096) { err = "device BVH build: clustering did not converge"; return false; }
            const int g = (int)((nc + kThreads - 1) / kThreads);
            k_ploc_nearest<<<g, kThreads, 0, stream>>>(b, cl[cur], nc, nearest);
            k_ploc_merge<<<g, kThreads, 0, stream>>>(b, tab, cl[cur], nearest, nc, merged, counters);
            LB_CK(cub::DeviceSelect::If(selTemp, selBytes, merged, cl[cur ^ 1], nSelected, (int)nc, IsCluster(), stream));
            uint32_t left = 0;
            LB_CK(cudaMemcpyAsync(&left, nSelected, sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
            LB_CK(cudaStreamSynchronize(stream));
            if (left >= nc) { err = "device BVH build: clustering made no progress"; return false; }
            nc = left;
            cur ^= 1;
        }
        uint32_t root = 0;
        LB_CK(cudaMemcpyAsync(&root, cl[cur], sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
        LB_CK(cudaStreamSynchronize(stream));
        rootBin = (int)root;
    }

    // ---- 5: wide nodes, level by level. The node count is not known in advance: start with room for n / 3 nodes (the
    //         measured ratio is ~ n / 8.4) and repeat with twice as much if it did not fit.
    size_t capacity = std::max<size_t>(64, (size_t)nT / 3 + 16);
    DevNode8 *nodes = nullptr;
    uint32_t *order = nullptr;
    uint32_t nNodes = 0, depth = 0;
    for (int attempt = 0;; attempt++)
    {
        if (attempt == 4) { err = "device BVH build: node capacity exceeded"; return false; }
        if (!S.alloc(nodes, capacity) || !S.alloc(order, capacity * kTriSlots) || !S.alloc(nodeBin, capacity))
        {
            err = "device BVH build: out of device memory";
            return false;
        }
        const uint32_t init[2] = {1u, 0u};  // nodes so far (the root), error bits
        LB_CK(cudaMemcpyAsync(counters, init, sizeof(init), cudaMemcpyHostToDevice, stream));
        LB_CK(cudaMemcpyAsync(nodeBin, &rootBin, sizeof(int), cudaMemcpyHostToDevice, stream));
        uint32_t start = 0, end = 1;
        uint32_t state[2] = {1u, 0u};
        depth = 0;
        while (start < end)
        {
            depth++;
            k_emit<<<(end - start + 127) / 128, 128, 0, stream>>>(b, tab, perm, nodeBin, start, end, nodes, order, counters, (uint32_t)capacity,
                                                                  counters + 1);
            LB_CK(cudaMemcpyAsync(state, counters, sizeof(state), cudaMemcpyDeviceToHost, stream));
            LB_CK(cudaStreamSynchronize(stream));
            if (state[1]) break;
            start = end;
            end = state[0];
            if (depth > 128) { err = "device BVH build: tree deeper than 128 levels"; return false; }
        }
        if (state[1] & 2u) { err = "wide BVH: quantisation overflow"; return false; }
        if (!(state[1] & 1u)) { nNodes = state[0]; break; }
        for (void *p : {(void *)nodes, (void *)order, (void *)nodeBin})
        {
            S.release(p);
            cudaFree(p);
        }
        capacity *= 2;
    }
    // ---- 6: the intersection records in wide order
    const size_t nRecords = (size_t)nNodes * kTriSlots;
    if (nRecords >= (1ull << 31)) { err = "wide BVH: too many nodes"; return false; }
    DevTri *tris8 = nullptr;
    if (!S.alloc(tris8, nRecords)) { err = "device BVH build: out of device memory"; return false; }
    k_gather_records<<<(unsigned)((nRecords + kThreads - 1) / kThreads), kThreads, 0, stream>>>(order, tris, nRecords, tris8);
    LB_CK(cudaGetLastError());
    LB_CK(cudaEventRecord(e1, stream));
    LB_CK(cudaStreamSynchronize(stream));
    LB_CK(cudaEventElapsedTime(&out.msBuild, e0, e1));
    S.release(nodes);
    S.release(order);
    S.release(tris8);
    out.nodes = nodes;
    out.tri8Index = order;
    out.tris8 = tris8;
    out.nNodes = nNodes;
    out.depth = depth;
    return true;
}

}  // namespace ptc
