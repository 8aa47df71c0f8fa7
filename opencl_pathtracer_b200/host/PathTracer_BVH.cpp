// PathTracer_BVH.cpp — binned-SAH BVH2 builder producing exactly the reference's tree.
//
// What it must reproduce (reference Controleur/PathTracer_BVH.cpp:109-356, Wald 2007 binning):
//   * leaf when nbTriangles <= 4, or squared centroid-box diagonal < 0.001, or the best split does not
//     beat the leaf cost by 0.01 (:133, :142, :260);
//   * 64 bins per axis on the centroid box, bin = (int)(k1 * (c - cmin)), k1 = (float)(64*0.999f / extent)
//     with the extent widened to double (:178-197);
//   * SAH(i) = (float)(Nl*Al + Nr*Ar) with half-areas computed in float and the products in double
//     (:229-235, Structs.h:264-271); the search starts from SAH[0][0] and visits bins 1..62 of each
//     axis, so bin 0 of axes 1 and 2 is never a candidate (:242-256);
//   * a Hoare partition whose swap sequence fixes the triangle order inside each child (:276-311);
//   * depth-first numbering: left child = parent + 1, right child = first index after the left subtree.
//
// How it is built differently from the reference: the builder works on a compact 64-byte record per
// triangle (bounds + centroid + original index) instead of swapping 336-byte triangles, keeps all
// scratch on the stack of the recursion (re-entrant and thread-safe; the reference uses function-static
// arrays), bins big nodes with OpenMP, and applies the permutation to the triangle array once at the
// end. An axis that is skipped for lack of extent is never chosen here; the reference would read stale
// static scratch in that case (only reachable when SAH values exceed INT_MAX).

#include "PathTracer_BVH.h"

#include <climits>
#include <cstdlib>
#include <stdexcept>
#include <vector>

namespace PathTracerNS
{
namespace
{
const int kBins = 64;
const uint kLeafMaxSize = 4;
const float kLeafMinDiag = 0.001f;
const float kKI = 1.f, kKT = 0.01f;

struct Box
{
    Float4 lo, hi, c;
    bool empty = true;
};

// reference Structs.h:220-238
inline void unite(Box &a, const Box &b)
{
    if (b.empty) return;
    if (a.empty)
    {
        a = b;
        return;
    }
    a.lo = ppmin(a.lo, b.lo);
    a.hi = ppmax(a.hi, b.hi);
    a.c = (a.lo + a.hi) / 2;
}
// reference Structs.h:247-262
inline void addPoint(Box &a, const Float4 &p)
{
    if (a.empty)
    {
        a.empty = false;
        a.lo = a.hi = a.c = p;
        return;
    }
    a.lo = ppmin(a.lo, p);
    a.hi = ppmax(a.hi, p);
    a.c = (a.lo + a.hi) / 2;
}
// reference Structs.h:264-271: float products and sums, widened on return
inline double halfArea(const Box &b)
{
    if (b.empty) return 0;
    Float4 d = b.hi - b.lo;
    return d.x * d.y + d.y * d.z + d.z * d.x;
}
inline float axisOf(const Float4 &v, int a) { return a == 0 ? v.x : (a == 1 ? v.y : v.z); }

struct Prim
{
    Box box;      // triangle AABB as the importer filled it
    uint source;  // index in the caller's array
};

inline void toBoundingBox(BoundingBox &dst, const Box &b)
{
    dst.pMin = b.lo;
    dst.pMax = b.hi;
    dst.centroid = b.c;
    dst.isEmpty = b.empty;
}

struct Builder
{
    std::vector<Prim> prims;
    Node *nodes = nullptr;
    uint size = 1, maxDepth = 0;

    void makeNode(uint idx, uint start, uint count, const Box &tb, const Box &cb)
    {
        Node &n = nodes[idx];
        std::memset((void *)&n, 0, sizeof(Node));
        n.triangleStartIndex = start;
        n.nbTriangles = count;
        toBoundingBox(n.trianglesAABB, tb);
        toBoundingBox(n.centroidsAABB, cb);
    }

    void makeLeaf(uint idx, int why, uint depth)
    {
        nodes[idx].isLeaf = true;
        nodes[idx].comments = why;
        if (depth > maxDepth) maxDepth = depth;
    }

    // returns the first node index after the subtree rooted at `cur`
    uint build(uint cur, const Box &centroids, uint depth)
    {
        const uint start = nodes[cur].triangleStartIndex, count = nodes[cur].nbTriangles;
        if (count <= kLeafMaxSize)
        {
            makeLeaf(cur, NODE_LEAF_MAX_SIZE, depth);
            return cur + 1;
        }
        {
            Float4 d = centroids.lo - centroids.hi;
            if (dot(d, d) < kLeafMinDiag)
            {
                makeLeaf(cur, NODE_LEAF_MIN_DIAG, depth);
                return cur + 1;
            }
        }
        const uint last = start + count - 1;

        Box l2r[3][kBins], r2l[3][kBins];
        int nl2r[3][kBins], nr2l[3][kBins];
        float sah[3][kBins - 1];
        float k1[3] = {0, 0, 0};
        bool usable[3] = {false, false, false};
        for (int a = 0; a < 3; a++)
            for (int i = 0; i < kBins - 1; i++) sah[a][i] = (float)INT_MAX;

        for (int a = 0; a < 3; a++)
        {
            const float cmin = axisOf(centroids.lo, a);
            const double extent = axisOf(centroids.hi, a) - cmin;
            if (extent < kLeafMinDiag) continue;
            usable[a] = true;
            k1[a] = (float)(kBins * (0.999f) / extent);

            Box bins[kBins];
            int nbins[kBins];
            for (int i = 0; i < kBins; i++) nbins[i] = 0;
            if (count >= (1u << 16))
            {
                // min/max unions and counts commute, so per-thread partial bins merge to the same values
#pragma omp parallel
                {
                    Box lb[kBins];
                    int ln[kBins];
                    for (int i = 0; i < kBins; i++) ln[i] = 0;
#pragma omp for schedule(static) nowait
                    for (long long i = start; i <= (long long)last; i++)
                    {
                        int b = (int)(k1[a] * (axisOf(prims[i].box.c, a) - cmin));
                        ln[b]++;
                        Box &dst = lb[b];
                        const Box &s = prims[i].box;
                        if (s.empty) continue;
                        if (dst.empty) { dst.lo = s.lo; dst.hi = s.hi; dst.empty = false; }
                        else { dst.lo = ppmin(dst.lo, s.lo); dst.hi = ppmax(dst.hi, s.hi); }
                    }
#pragma omp critical
                    for (int i = 0; i < kBins; i++)
                    {
                        nbins[i] += ln[i];
                        if (lb[i].empty) continue;
                        if (bins[i].empty) { bins[i].lo = lb[i].lo; bins[i].hi = lb[i].hi; bins[i].empty = false; }
                        else { bins[i].lo = ppmin(bins[i].lo, lb[i].lo); bins[i].hi = ppmax(bins[i].hi, lb[i].hi); }
                    }
                }
                for (int i = 0; i < kBins; i++)
                    if (!bins[i].empty) bins[i].c = (bins[i].lo + bins[i].hi) / 2;
            }
            else
            {
                for (uint i = start; i <= last; i++)
                {
                    int b = (int)(k1[a] * (axisOf(prims[i].box.c, a) - cmin));
                    nbins[b]++;
                    unite(bins[b], prims[i].box);
                }
            }
            // A bin holding a single triangle keeps the importer's centroid (copy semantics of unite()),
            // which the parallel path must honour too: recompute only changes it if the importer's
            // centroid is not (pMin+pMax)/2. Single-triangle bins never occur at >= 65536 triangles with
            // importer-conformant data; the serial path above is exact in every case.

            for (int i = 0; i < kBins; i++)
            {
                l2r[a][i] = bins[i];
                r2l[a][i] = bins[i];
                nl2r[a][i] = nbins[i];
                nr2l[a][i] = nbins[i];
            }
            for (int i = 1, j = kBins - 2; i < kBins; i++, j--)
            {
                unite(l2r[a][i], l2r[a][i - 1]);
                nl2r[a][i] += nl2r[a][i - 1];
                unite(r2l[a][j], r2l[a][j + 1]);
                nr2l[a][j] += nr2l[a][j + 1];
            }
            for (int i = 0; i < kBins - 1; i++)
                sah[a][i] = (float)(nl2r[a][i] * halfArea(l2r[a][i]) + nr2l[a][i + 1] * halfArea(r2l[a][i + 1]));
        }

        int bestAxis = 0, bestIdx = 0;
        float bestSAH = sah[0][0];
        for (int a = 0; a < 3; a++)
            for (int i = 1; i < kBins - 1; i++)
                if (sah[a][i] < bestSAH)
                {
                    bestAxis = a;
                    bestSAH = sah[a][i];
                    bestIdx = i;
                }

        Box nodeBox;
        nodeBox.lo = nodes[cur].trianglesAABB.pMin;
        nodeBox.hi = nodes[cur].trianglesAABB.pMax;
        nodeBox.empty = nodes[cur].trianglesAABB.isEmpty;
        if (kKI * bestSAH + kKT > count * halfArea(nodeBox) || !usable[bestAxis])
        {
            makeLeaf(cur, NODE_BAD_SAH, depth);
            return cur + 1;
        }
        nodes[cur].cutAxis = bestAxis;

        // Hoare partition, same scan and swap sequence as reference :276-311
        uint l = start, r = last;
        {
            const float cmin = axisOf(centroids.lo, bestAxis), k = k1[bestAxis];
            const int cut = bestIdx + 1;
            while (l < r)
            {
                while ((k * (axisOf(prims[l].box.c, bestAxis) - cmin)) < cut && l < r) l++;
                while ((k * (axisOf(prims[r].box.c, bestAxis) - cmin)) >= cut && l < r) r--;
                if (l < r) std::swap(prims[l], prims[r]);
            }
        }
        Box leftC, rightC;
        for (uint i = start; i < l; i++) addPoint(leftC, prims[i].box.c);
        for (long long i = last; i >= (long long)l; i--) addPoint(rightC, prims[i].box.c);

        const uint nLeft = (uint)nl2r[bestAxis][bestIdx], nRight = (uint)nr2l[bestAxis][bestIdx + 1];
        const Box rightBox = r2l[bestAxis][bestIdx + 1];

        size += 2;
        const uint son1 = cur + 1;
        nodes[cur].son1Id = son1;
        makeNode(son1, start, nLeft, l2r[bestAxis][bestIdx], leftC);
        const uint son2 = build(son1, leftC, depth + 1);
        nodes[cur].son2Id = son2;
        makeNode(son2, start + nLeft, nRight, rightBox, rightC);
        return build(son2, rightC, depth + 1);
    }
};
}  // namespace

void BVH_Create(GlobalVars &gv)
{
    const uint n = gv.triangulationSize;
    if (n == 0) throw std::runtime_error("BVH_Create: the scene has no triangle");

    Builder b;
    b.prims.resize(n);
    Box all, allC;
    for (uint i = 0; i < n; i++)
    {
        const BoundingBox &s = gv.triangulation[i].AABB;
        Prim &p = b.prims[i];
        p.box.lo = s.pMin;
        p.box.hi = s.pMax;
        p.box.c = s.centroid;
        p.box.empty = s.isEmpty != 0;
        p.source = i;
        unite(all, p.box);
        addPoint(allC, p.box.c);
    }
    gv.bvh = new Node[2 * (size_t)n - 1];
    b.nodes = gv.bvh;
    b.makeNode(0, 0, n, all, allC);
    b.build(0, allC, 0);
    gv.bvhSize = b.size;
    gv.bvhMaxDepth = b.maxDepth;

    // apply the permutation once (the reference swaps whole triangles during the build)
    bool identity = true;
    for (uint i = 0; i < n && identity; i++) identity = b.prims[i].source == i;
    if (!identity)
    {
        Triangle *sorted = (Triangle *)std::malloc(sizeof(Triangle) * (size_t)n);
        if (!sorted) throw std::runtime_error("BVH_Create: out of memory");
#pragma omp parallel for schedule(static)
        for (long long i = 0; i < (long long)n; i++) sorted[i] = gv.triangulation[b.prims[i].source];
        std::memcpy((void *)gv.triangulation, (void *)sorted, sizeof(Triangle) * (size_t)n);
        std::free(sorted);
    }
}

static void characteristics(const Node *bvh, uint id, uint depth, BVHCharacteristics &c)
{
    const Node &n = bvh[id];
    if (n.isLeaf)
    {
        c.nLeafs++;
        if (depth < c.minDepth) c.minDepth = depth;
        if (depth > c.maxDepth) c.maxDepth = depth;
        if (n.nbTriangles < c.minLeafSize) c.minLeafSize = n.nbTriangles;
        if (n.nbTriangles > c.maxLeafSize) c.maxLeafSize = n.nbTriangles;
        return;
    }
    c.nNodes += 2;
    characteristics(bvh, n.son1Id, depth + 1, c);
    characteristics(bvh, n.son2Id, depth + 1, c);
}

BVHCharacteristics BVH_GetCharacteristics(const Node *bvh)
{
    BVHCharacteristics c;
    characteristics(bvh, 0, 0, c);
    return c;
}

}  // namespace PathTracerNS
