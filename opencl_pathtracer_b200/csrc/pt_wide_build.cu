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

namespace ptc
{
namespace
{
struct Box
{
    float lo[3], hi[3];
    void reset() { for (int a = 0; a < 3; a++) { lo[a] = INFINITY; hi[a] = -INFINITY; } }
    void add(const float *p) { for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], p[a]); hi[a] = std::max(hi[a], p[a]); } }
    void add(const Box &b) { for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], b.lo[a]); hi[a] = std::max(hi[a], b.hi[a]); } }
    double area() const
    {
        const double x = (double)hi[0] - lo[0], y = (double)hi[1] - lo[1], z = (double)hi[2] - lo[2];
        return 2.0 * (x * y + y * z + z * x);
    }
};

struct BNode  // binary tree over the BVH-ordered triangle array, leaves hold one triangle
{
    Box box;
    int left, right;        // -1 for leaves
    uint32_t first, count;  // triangles of the subtree: [first, first + count) when `contiguous`
    bool contiguous;
};

constexpr double kNodeCost = 1.0;  // one wide-node visit (~250 instructions) against
constexpr double kTriCost = 0.6;   // one triangle test (~150)
constexpr uint32_t kLeafTriangles = 3;

struct Builder
{
    const ptc_scene_view *s;
    std::vector<BNode> b;

    Box triBox(uint32_t first, uint32_t count) const
    {
        Box r;
        r.reset();
        for (uint32_t i = first; i < first + count; i++)
        {
            const ptc_triangle &t = s->triangulation[i];
            const float v[3][3] = {{t.S1.x, t.S1.y, t.S1.z}, {t.S2.x, t.S2.y, t.S2.z}, {t.S3.x, t.S3.y, t.S3.z}};
            for (int k = 0; k < 3; k++) r.add(v[k]);
        }
        return r;
    }

    int makeLeaves(uint32_t first, uint32_t count)
    {
        if (count == 1)
        {
            BNode n;
            n.box = triBox(first, 1);
            n.left = n.right = -1;
            n.first = first;
            n.count = 1;
            n.contiguous = true;
            b.push_back(n);
            return (int)b.size() - 1;
        }
        const uint32_t half = count / 2;
        const int l = makeLeaves(first, half), r = makeLeaves(first + half, count - half);
        return join(l, r);
    }

    int join(int l, int r)  // children are always created before their parent: b is in post-order
    {
        BNode n;
        n.box = b[l].box;
        n.box.add(b[r].box);
        n.left = l;
        n.right = r;
        n.first = std::min(b[l].first, b[r].first);
        n.count = b[l].count + b[r].count;
        n.contiguous = b[l].contiguous && b[r].contiguous &&
                       (b[l].first + b[l].count == b[r].first || b[r].first + b[r].count == b[l].first);
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

// smallest e with 255 * 2^e >= extent, as an IEEE biased exponent byte
uint32_t gridExponent(double extent)
{
    int e = -126;
    if (extent > 0)
    {
        e = (int)std::ceil(std::log2(extent / 255.0));
        while (std::ldexp(255.0, e) < extent) e++;
        while (e > -126 && std::ldexp(255.0, e - 1) >= extent) e--;
    }
    e = std::max(-126, std::min(127, e));
    return (uint32_t)(e + 127);
}
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
    const std::vector<BNode> &b = B.b;
    out.nodes.reserve(s->triangulation_size / 8 + 16);

    // ---- 2. collapse: cost[n][i-1] = C(n, i); split[n][i-1] = roots given to the left subtree (0: same as C(n, i-1));
    //         asLeaf[n]: C(n, 1) is a leaf
    const size_t nb2 = b.size();
    std::vector<float> cost(nb2 * 7);
    std::vector<uint8_t> split(nb2 * 7, 0), asLeaf(nb2, 0), split8(nb2, 0);
    for (size_t n = 0; n < nb2; n++)  // post-order: children first
    {
        const BNode &x = b[n];
        const double area = x.box.area();
        float *c = &cost[n * 7];
        if (x.left < 0)
        {
            for (int i = 0; i < 7; i++) c[i] = (float)(area * kTriCost);
            asLeaf[n] = 1;
            continue;
        }
        const float *cl = &cost[(size_t)x.left * 7], *cr = &cost[(size_t)x.right * 7];
        auto distribute = [&](int j, uint8_t &k) {  // j = 2..8 roots over both subtrees
            float best = INFINITY;
            for (int a = 1; a < j; a++)
            {
                if (a > 7 || j - a > 7) continue;
                const float v = cl[a - 1] + cr[j - a - 1];
                if (v < best) { best = v; k = (uint8_t)a; }
            }
            return best;
        };
        const float leafCost = (x.contiguous && x.count <= kLeafTriangles) ? (float)(area * x.count * kTriCost) : INFINITY;
        const float innerCost = distribute(8, split8[n]) + (float)(area * kNodeCost);
        asLeaf[n] = leafCost <= innerCost;
        c[0] = asLeaf[n] ? leafCost : innerCost;
        for (int i = 2; i <= 7; i++)
        {
            uint8_t k = 0;
            const float d = distribute(i, k);
            if (d < c[i - 2]) { c[i - 1] = d; split[n * 7 + i - 1] = k; }
            else { c[i - 1] = c[i - 2]; split[n * 7 + i - 1] = 0; }
        }
    }
    // children of the wide node that stands for binary node n: follow the recorded decisions (iterative)
    auto gather = [&](int n, int child[8]) {
        struct G { int n, i; };
        G stack[16];
        int sp = 0, nc = 0;
        if (b[n].left < 0 || asLeaf[n]) { child[nc++] = n; return nc; }  // the whole scene is one small leaf
        stack[sp++] = {b[n].left, (int)split8[n]};
        stack[sp++] = {b[n].right, 8 - (int)split8[n]};
        while (sp)
        {
            G g = stack[--sp];
            while (g.i > 1 && split[(size_t)g.n * 7 + g.i - 1] == 0) g.i--;
            if (g.i == 1) { child[nc++] = g.n; continue; }
            const int k = split[(size_t)g.n * 7 + g.i - 1];
            stack[sp++] = {b[g.n].left, k};
            stack[sp++] = {b[g.n].right, g.i - k};
        }
        return nc;
    };

    struct Item { int bnode; uint32_t index; uint32_t depth; };
    std::vector<Item> todo;
    out.nodes.emplace_back();
    todo.push_back({root, 0u, 1u});
    while (!todo.empty())
    {
        const Item it = todo.back();
        todo.pop_back();
        out.depth = std::max(out.depth, it.depth);
        int child[8];
        const int nc = gather(it.bnode, child);
        Box nb;
        nb.reset();
        for (int i = 0; i < nc; i++) nb.add(b[child[i]].box);
        // ---- 3. slot assignment: slot s should hold the child that a ray of octant s reaches first
        int slotOf[8], childAt[8];
        for (int i = 0; i < 8; i++) { slotOf[i] = -1; childAt[i] = -1; }
        double cost[8][8];
        for (int i = 0; i < nc; i++)
            for (int sl = 0; sl < 8; sl++)
            {
                double c = 0;
                for (int a = 0; a < 3; a++)
                {
                    const double centre = 0.5 * ((double)b[child[i]].box.lo[a] + b[child[i]].box.hi[a]) - 0.5 * ((double)nb.lo[a] + nb.hi[a]);
                    const bool negative = (sl >> (2 - a)) & 1;  // octant bit 2 = x, 1 = y, 0 = z (pt_wide.cuh)
                    c += negative ? -centre : centre;
                }
                cost[i][sl] = c;
            }
        for (int round = 0; round < nc; round++)
        {
            int bi = -1, bs = -1;
            double bc = INFINITY;
            for (int i = 0; i < nc; i++)
            {
                if (slotOf[i] >= 0) continue;
                for (int sl = 0; sl < 8; sl++)
                    if (childAt[sl] < 0 && cost[i][sl] < bc) { bc = cost[i][sl]; bi = i; bs = sl; }
            }
            slotOf[bi] = bs;
            childAt[bs] = bi;
        }
        // ---- 4. encode
        DevNode8 n;
        std::memset(&n, 0, sizeof(n));
        n.px = nb.lo[0]; n.py = nb.lo[1]; n.pz = nb.lo[2];
        uint32_t ebyte[3];
        double step[3];
        for (int a = 0; a < 3; a++)
        {
            ebyte[a] = gridExponent((double)nb.hi[a] - nb.lo[a]);
            step[a] = std::ldexp(1.0, (int)ebyte[a] - 127);
        }
        uint32_t imask = 0, nInner = 0, valid = 0;
        n.childBase = (uint32_t)out.nodes.size();
        if (out.triOrder.size() < (size_t)(it.index + 1) * kTriSlots) out.triOrder.resize((size_t)(it.index + 1) * kTriSlots, kNoTriangle);
        uint8_t qlo[3][8], qhi[3][8];
        std::memset(qlo, 0, sizeof(qlo));
        std::memset(qhi, 0, sizeof(qhi));
        for (int sl = 0; sl < 8; sl++)
        {
            if (childAt[sl] < 0) continue;
            const BNode &c = b[child[childAt[sl]]];
            for (int a = 0; a < 3; a++)
            {
                const double l = std::floor(((double)c.box.lo[a] - nb.lo[a]) / step[a]);
                const double h = std::ceil(((double)c.box.hi[a] - nb.lo[a]) / step[a]);
                qlo[a][sl] = (uint8_t)std::max(0.0, std::min(255.0, l));
                qhi[a][sl] = (uint8_t)std::max(0.0, std::min(255.0, h));
                if (h > 255.0) { err = "wide BVH: quantisation overflow"; return false; }
            }
            if (!asLeaf[child[childAt[sl]]])
            {
                imask |= 1u << sl;
                nInner++;
            }
            else
            {
                // triangle k of the leaf in slot sl: bit 3 * sl + k of `valid`, record it.index * 24 + 3 * sl + k
                for (uint32_t k = 0; k < c.count; k++)
                {
                    valid |= 1u << (3 * sl + (int)k);
                    out.triOrder[(size_t)it.index * kTriSlots + 3 * sl + k] = c.first + k;
                }
            }
        }
        n.eimask = ebyte[0] | (ebyte[1] << 8) | (ebyte[2] << 16) | (imask << 24);
        n.valid = valid;
        for (int a = 0; a < 3; a++) { std::memcpy(n.qlo[a], qlo[a], 8); std::memcpy(n.qhi[a], qhi[a], 8); }
        out.nodes[it.index] = n;
        // inner children: consecutive nodes in slot order
        out.nodes.resize(out.nodes.size() + nInner);
        uint32_t rank = 0;
        for (int sl = 0; sl < 8; sl++)
            if (childAt[sl] >= 0 && !asLeaf[child[childAt[sl]]]) todo.push_back({child[childAt[sl]], n.childBase + rank++, it.depth + 1u});
    }
    out.triOrder.resize(out.nodes.size() * kTriSlots, kNoTriangle);
    if (out.triOrder.size() >= (1ull << 31)) { err = "wide BVH: too many nodes"; return false; }
    return true;
}

}  // namespace ptc
