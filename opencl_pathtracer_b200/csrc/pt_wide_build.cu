// pt_wide_build.cu — host side of the FAST-mode wide BVH: collapses the reference BVH2 (the output of BVH_Create,
// Controleur/PathTracer_BVH.cpp:12-356, handed over in the reference's Node layout) into the 8-wide compressed
// nodes of pt_wide.cuh. Nothing of the reference is restated here: the BVH2 is only read (child links, leaf ranges);
// boxes are recomputed from the triangle vertices so that they do not depend on the importer's AABB fields.
//
//   1. binary tree down to single triangles: the reference tree, its leaves halved in array order;
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

    int makeLeaves(uint32_t first, uint32_t count)
    {
        if (count == 1)
        {
            BinNode n;
            triBox(first, n);
            n.left = n.right = -1;
            n.first = first;
            n.count = 1;
            b.push_back(n);
            return (int)b.size() - 1;
        }
        const uint32_t half = count / 2;
        const int l = makeLeaves(first, half), r = makeLeaves(first + half, count - half);
        return join(l, r);
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
