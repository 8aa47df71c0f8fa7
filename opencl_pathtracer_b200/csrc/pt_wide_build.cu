// pt_wide_build.cu — host side of the FAST-mode wide BVH: collapses the reference BVH2 (the output of BVH_Create,
// Controleur/PathTracer_BVH.cpp:12-356, handed over in the reference's Node layout) into the 8-wide compressed
// nodes of pt_wide.cuh. Nothing of the reference is restated here: the BVH2 is only read (child links, leaf ranges);
// boxes are recomputed from the triangle vertices so that they do not depend on the importer's AABB fields.
//
//   1. binary tree down to single triangles: the reference tree, each of its leaves (<= 4 triangles as a rule) continued by
//      the bipartitions of smallest surface-area cost (array-order halving before: PTC_LEAF_SPLIT=0);
//   2. collapse by dynamic programming on the surface-area cost (Ylitie et al. 2017, section 3.1): C(n, i) = cheapest
//      way to stand for the subtree of n with at most i children of a wide node, i = 1..7; one child is either a leaf
//      (<= 3 triangles, cost area x triangles x kTriCost) or a wide node (area x kNodeCost + the best split of 8 child
//      slots between the two subtrees). Compared with a greedy "open the largest child" collapse it fills the nodes
//      (bench scene: 39.5k nodes with 4.2 children on average before), which is what a ray pays per visit;
//   3. children are assigned to the 8 slots so that, for a ray of octant o, visiting slots by decreasing (slot ^ (7-o))
//      is approximately front to back (greedy assignment on centroid offsets, after Ylitie et al. 2017);
//   4. child boxes are quantised outwards to 8 bits on the node's grid; inner children get consecutive node indices;
//      triangle k of the leaf child in slot s of node i is record 24 * i + 3 * s + k of the wide triangle order (a
//      sparse array: unused records are never fetched, so they cost address space, not cache or bandwidth).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "pt_wide_build.h"
#include "pt_wide_collapse.cuh"

namespace ptc
{
namespace
{
struct Builder
{
    const ptc_scene_view *s;
    std::vector<BinNode> b;
    bool sahLeaves = true;

    void triBox(uint32_t i, BinNode &n) const
    {
        const ptc_triangle &t = s->triangulation[i];
        const float v[3][3] = {{t.S1.x, t.S1.y, t.S1.z}, {t.S2.x, t.S2.y, t.S2.z}, {t.S3.x, t.S3.y, t.S3.z}};
        for (int a = 0; a < 3; a++)
        {
            n.lo[a] = std::min(v[0][a], std::min(v[1][a], v[2][a]));
            n.hi[a] = std::max(v[0][a], std::max(v[1][a], v[2][a]));
        }
    }

    int makeLeaf(uint32_t tri)
    {
        BinNode n;
        triBox(tri, n);
        n.left = n.right = -1;
        n.first = tri;
        n.count = 1;
        b.push_back(n);
        return (int)b.size() - 1;
    }

    static double halfArea(const float lo[3], const float hi[3])
    {
        const double x = (double)hi[0] - lo[0], y = (double)hi[1] - lo[1], z = (double)hi[2] - lo[2];
        return x * y + y * z + z * x;
    }

    // The triangles of one reference leaf (at most 4 unless the reference gave up on a split) as a binary tree down to single
    // triangles. PTC_LEAF_SPLIT=0: halved in array order; default: the partition with the smallest surface-area cost
    // area(L) * |L| + area(R) * |R| — all of them for up to 6 triangles, a sweep along the widest axis of the centroids beyond.
    int makeLeaves(std::vector<uint32_t> &t, uint32_t begin, uint32_t count)
    {
        if (count == 1) return makeLeaf(t[begin]);
        uint32_t nl = count / 2;
        if (sahLeaves)
        {
            std::vector<BinNode> box(count);
            for (uint32_t i = 0; i < count; i++) triBox(t[begin + i], box[i]);
            auto grow = [](float lo[3], float hi[3], const BinNode &x) {
                for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], x.lo[a]); hi[a] = std::max(hi[a], x.hi[a]); }
            };
            double best = 1e300;
            if (count <= 6)
            {
                uint32_t bestMask = 0;
                for (uint32_t m = 1; m < (1u << (count - 1)); m++)  // triangle count-1 always goes right: every bipartition once
                {
                    float lo[2][3] = {{3e38f, 3e38f, 3e38f}, {3e38f, 3e38f, 3e38f}}, hi[2][3] = {{-3e38f, -3e38f, -3e38f}, {-3e38f, -3e38f, -3e38f}};
                    int n[2] = {0, 0};
                    for (uint32_t i = 0; i < count; i++)
                    {
                        const int side = (m >> i) & 1u ? 0 : 1;
                        grow(lo[side], hi[side], box[i]);
                        n[side]++;
                    }
                    const double c = halfArea(lo[0], hi[0]) * n[0] + halfArea(lo[1], hi[1]) * n[1];
                    if (c < best) { best = c; bestMask = m; }
                }
                std::vector<uint32_t> l, r;
                for (uint32_t i = 0; i < count; i++) ((bestMask >> i) & 1u ? l : r).push_back(t[begin + i]);
                nl = (uint32_t)l.size();
                std::copy(l.begin(), l.end(), t.begin() + begin);
                std::copy(r.begin(), r.end(), t.begin() + begin + nl);
            }
            else
            {
                float clo[3] = {3e38f, 3e38f, 3e38f}, chi[3] = {-3e38f, -3e38f, -3e38f};
                for (uint32_t i = 0; i < count; i++)
                    for (int a = 0; a < 3; a++)
                    {
                        const float c = box[i].lo[a] + box[i].hi[a];
                        clo[a] = std::min(clo[a], c); chi[a] = std::max(chi[a], c);
                    }
                int ax = 0;
                for (int a = 1; a < 3; a++) if (chi[a] - clo[a] > chi[ax] - clo[ax]) ax = a;
                std::vector<uint32_t> idx(count);
                for (uint32_t i = 0; i < count; i++) idx[i] = i;
                std::stable_sort(idx.begin(), idx.end(), [&](uint32_t x, uint32_t y) { return box[x].lo[ax] + box[x].hi[ax] < box[y].lo[ax] + box[y].hi[ax]; });
                std::vector<double> right(count + 1, 0.0);
                float lo[3] = {3e38f, 3e38f, 3e38f}, hi[3] = {-3e38f, -3e38f, -3e38f};
                for (uint32_t i = count; i-- > 1;) { grow(lo, hi, box[idx[i]]); right[i] = halfArea(lo, hi) * (count - i); }
                float l2[3] = {3e38f, 3e38f, 3e38f}, h2[3] = {-3e38f, -3e38f, -3e38f};
                for (uint32_t i = 1; i < count; i++)
                {
                    grow(l2, h2, box[idx[i - 1]]);
                    const double c = halfArea(l2, h2) * i + right[i];
                    if (c < best) { best = c; nl = i; }
                }
                std::vector<uint32_t> sorted(count);
                for (uint32_t i = 0; i < count; i++) sorted[i] = t[begin + idx[i]];
                std::copy(sorted.begin(), sorted.end(), t.begin() + begin);
            }
        }
        const int l = makeLeaves(t, begin, nl), r = makeLeaves(t, begin + nl, count - nl);
        return join(l, r);
    }

    int makeLeaves(uint32_t first, uint32_t count)
    {
        std::vector<uint32_t> t(count);
        for (uint32_t i = 0; i < count; i++) t[i] = first + i;
        return makeLeaves(t, 0, count);
    }

    int join(int l, int r)  // children are always created before their parent: b is in post-order
    {
        BinNode n;
        for (int a = 0; a < 3; a++)
        {
            n.lo[a] = std::min(b[l].lo[a], b[r].lo[a]);
            n.hi[a] = std::max(b[l].hi[a], b[r].hi[a]);
        }
        n.left = l;
        n.right = r;
        n.first = 0;
        n.count = b[l].count + b[r].count;
        b.push_back(n);
        return (int)b.size() - 1;
    }

    // reference node -> binary node index, or -1 for a subtree without triangles
    int convert(uint32_t id)
    {
        const ptc_node &n = s->bvh[id];
        if (n.isLeaf) return n.nbTriangles ? makeLeaves(n.triangleStartIndex, n.nbTriangles) : -1;
        const int l = convert(n.son1Id), r = convert(n.son2Id);
        if (l < 0) return r;
        if (r < 0) return l;
        return join(l, r);
    }
};
}  // namespace

bool build_wide_bvh(const ptc_scene_view *s, WideBvh &out, std::string &err)
{
    out.nodes.clear();
    out.triOrder.clear();
    out.depth = 0;
    if (s->triangulation_size == 0 || s->bvh_size == 0) return true;
    Builder B;
    B.s = s;
    if (const char *e = std::getenv("PTC_LEAF_SPLIT")) B.sahLeaves = std::atoi(e) != 0;
    B.b.reserve(2 * (size_t)s->triangulation_size);
    const int root = B.convert(0);
    if (root < 0) return true;
    const std::vector<BinNode> &b = B.b;

    // ---- 2. collapse decisions, children before parents (b is in post-order)
    const size_t nb2 = b.size();
    std::vector<float> cost(nb2 * 7);
    std::vector<uint8_t> split(nb2 * 7), split8(nb2), asLeaf(nb2);
    const CollapseTables tab{cost.data(), split.data(), split8.data(), asLeaf.data()};
    for (size_t n = 0; n < nb2; n++) collapse_costs(b.data(), n, tab);

    // ---- 3 + 4. emit the wide nodes top-down
    out.nodes.reserve(s->triangulation_size / 8 + 16);
    struct Item { int bnode; uint32_t index; uint32_t depth; };
    std::vector<Item> todo;
    out.nodes.emplace_back();
    todo.push_back({root, 0u, 1u});
    while (!todo.empty())
    {
        const Item it = todo.back();
        todo.pop_back();
        out.depth = std::max(out.depth, it.depth);
        int child[8], innerChild[8], nInner = 0;
        const int nc = collapse_gather(b.data(), it.bnode, tab, child);
        DevNode8 n;
        uint32_t order[kTriSlots];
        const uint32_t childBase = (uint32_t)out.nodes.size();
        if (!collapse_encode(b.data(), tab, child, nc, childBase, nullptr, n, order, innerChild, nInner))
        {
            err = "wide BVH: quantisation overflow";
            return false;
        }
        out.nodes[it.index] = n;
        if (out.triOrder.size() < (size_t)(it.index + 1) * kTriSlots) out.triOrder.resize((size_t)(it.index + 1) * kTriSlots, kNoTriangle);
        std::memcpy(&out.triOrder[(size_t)it.index * kTriSlots], order, sizeof(order));
        out.nodes.resize(out.nodes.size() + nInner);  // inner children: consecutive nodes in slot order
        for (int r = 0; r < nInner; r++) todo.push_back({innerChild[r], childBase + (uint32_t)r, it.depth + 1u});
    }
    out.triOrder.resize(out.nodes.size() * kTriSlots, kNoTriangle);
    if (out.triOrder.size() >= (1ull << 31)) { err = "wide BVH: too many nodes"; return false; }
    return true;
}

}  // namespace ptc
