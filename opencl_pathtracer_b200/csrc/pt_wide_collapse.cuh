// pt_wide_collapse.cuh — the per-node steps of collapsing a binary BVH into the 8-wide compressed BVH of pt_wide.cuh,
// written once for the host builder (pt_wide_build.cu: the reference BVH2 as input) and for the device builder
// (pt_lbvh.cu: a Morton-order LBVH built on the GPU as input). Nothing here restates the reference: the input tree is
// only read.
//
//   collapse_costs   one node of the dynamic programme on the surface-area cost (Ylitie, Karras, Laine 2017, 3.1):
//                    C(n, i) = cheapest way to stand for the subtree of n with at most i children of a wide node
//   collapse_gather  the (at most 8) children of the wide node that stands for binary node n, from the recorded decisions
//   collapse_encode  slot assignment by ray octant + outward quantisation of the child boxes -> one DevNode8 and its 24
//                    triangle records
#pragma once

#include <cmath>
#include <cstdint>

#include "pt_wide.cuh"

#if defined(__CUDACC__)
#define PTC_HD __host__ __device__
#else
#define PTC_HD
#endif

namespace ptc
{
struct BinNode  // binary tree over an ordered triangle array; leaves hold one triangle
{
    float lo[3], hi[3];
    int left, right;   // -1 for leaves
    uint32_t first;    // leaves: position of the triangle in the ordered array (inner nodes: unused)
    uint32_t count;    // triangles in the subtree
};

constexpr float kNodeCost = 1.0f;  // one wide-node visit (~250 instructions) against
constexpr float kTriCost = 0.6f;   // one triangle test (~150); the result is flat between 0.15 and 1.2 (profiles/r01_sweeps.txt)
constexpr uint32_t kLeafTriangles = 3;

struct CollapseTables  // per binary node
{
    float *cost;      // [n][7]   C(n, i), i = 1..7
    uint8_t *split;   // [n][7]   roots given to the left subtree for C(n, i); 0 = same as C(n, i - 1)
    uint8_t *split8;  // [n]      the same for the 8 slots of the wide node that n becomes
    uint8_t *asLeaf;  // [n]      C(n, 1) is a leaf child
};

PTC_HD inline double box_area(const BinNode &x)
{
    const double a = (double)x.hi[0] - x.lo[0], b = (double)x.hi[1] - x.lo[1], c = (double)x.hi[2] - x.lo[2];
    return 2.0 * (a * b + b * c + c * a);
}

// One node of the dynamic programme on values: x = the node, cl / cr = C(left, 1..7) / C(right, 1..7) (unused for a leaf).
PTC_HD inline void collapse_costs_core(const BinNode &x, const float cl[7], const float cr[7], float c[7], uint8_t split[7], uint8_t &split8,
                                       uint8_t &asLeaf)
{
    const double area = box_area(x);
    if (x.left < 0)
    {
        for (int i = 0; i < 7; i++) { c[i] = (float)(area * kTriCost); split[i] = 0; }
        asLeaf = 1;
        split8 = 0;
        return;
    }
    const float kInf = 3.0e38f;
    // best way to give j = 2..8 roots to the two subtrees
    float dist[9];
    uint8_t distK[9];
    for (int j = 2; j <= 8; j++)
    {
        float best = kInf;
        uint8_t k = 1;
        for (int a = 1; a < j; a++)
        {
            if (a > 7 || j - a > 7) continue;
            const float v = cl[a - 1] + cr[j - a - 1];
            if (v < best) { best = v; k = (uint8_t)a; }
        }
        dist[j] = best;
        distK[j] = k;
    }
    const uint32_t cnt = x.count;
    const float leafCost = cnt <= kLeafTriangles ? (float)(area * cnt * kTriCost) : kInf;
    const float innerCost = dist[8] + (float)(area * kNodeCost);
    split8 = distK[8];
    asLeaf = leafCost <= innerCost;
    c[0] = asLeaf ? leafCost : innerCost;
    split[0] = 0;
    for (int i = 2; i <= 7; i++)
    {
        if (dist[i] < c[i - 2]) { c[i - 1] = dist[i]; split[i - 1] = distK[i]; }
        else { c[i - 1] = c[i - 2]; split[i - 1] = 0; }
    }
}

// the same on the tables; children must have been processed before their parent
PTC_HD inline void collapse_costs(const BinNode *b, size_t n, const CollapseTables &t)
{
    const BinNode &x = b[n];
    const float *cl = x.left < 0 ? nullptr : t.cost + (size_t)x.left * 7, *cr = x.left < 0 ? nullptr : t.cost + (size_t)x.right * 7;
    collapse_costs_core(x, cl, cr, t.cost + n * 7, t.split + n * 7, t.split8[n], t.asLeaf[n]);
}

PTC_HD inline int collapse_gather(const BinNode *b, int n, const CollapseTables &t, int child[8])
{
    int stackN[16], stackI[16];
    int sp = 0, nc = 0;
    if (b[n].left < 0 || t.asLeaf[n]) { child[nc++] = n; return nc; }  // the whole scene is one small leaf
    stackN[sp] = b[n].left; stackI[sp++] = (int)t.split8[n];
    stackN[sp] = b[n].right; stackI[sp++] = 8 - (int)t.split8[n];
    while (sp)
    {
        --sp;
        const int g = stackN[sp];
        int i = stackI[sp];
        while (i > 1 && t.split[(size_t)g * 7 + i - 1] == 0) i--;
        if (i == 1) { child[nc++] = g; continue; }
        const int k = t.split[(size_t)g * 7 + i - 1];
        stackN[sp] = b[g].left; stackI[sp++] = k;
        stackN[sp] = b[g].right; stackI[sp++] = i - k;
    }
    return nc;
}

// smallest e with 255 * 2^e >= extent, as an IEEE biased exponent byte
PTC_HD inline uint32_t grid_exponent(double extent)
{
    int e = -126;
    if (extent > 0)
    {
        frexp(extent / 255.0, &e);  // extent / 255 = m * 2^e, m in [0.5, 1)
        while (ldexp(255.0, e) < extent) e++;
        while (e > -126 && ldexp(255.0, e - 1) >= extent) e--;
    }
    e = e < -126 ? -126 : (e > 127 ? 127 : e);
    return (uint32_t)(e + 127);
}

// Encodes the wide node `index` with the given children. order = the 24 triangle records of this node (kNoTriangle where
// unused); perm maps a position of the ordered triangle array to the triangle index to record (null = identity).
// childBase = index the first inner child will get. Returns the mask of inner children; innerChild[rank] = binary node
// of the rank-th inner child in slot order. false = quantisation overflow (cannot happen for finite boxes).
PTC_HD inline bool collapse_encode(const BinNode *b, const CollapseTables &t, const int child[8], int nc, uint32_t childBase,
                                   const uint32_t *perm, DevNode8 &n, uint32_t order[kTriSlots], int innerChild[8], int &nInner)
{
    float nlo[3] = {INFINITY, INFINITY, INFINITY}, nhi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int i = 0; i < nc; i++)
        for (int a = 0; a < 3; a++)
        {
            nlo[a] = fminf(nlo[a], b[child[i]].lo[a]);
            nhi[a] = fmaxf(nhi[a], b[child[i]].hi[a]);
        }
    // slot assignment: slot s should hold the child that a ray of octant s reaches first (greedy, after Ylitie et al.)
    int slotOf[8], childAt[8];
    for (int i = 0; i < 8; i++) { slotOf[i] = -1; childAt[i] = -1; }
    double cost[8][8];
    for (int i = 0; i < nc; i++)
        for (int sl = 0; sl < 8; sl++)
        {
            double c = 0;
            for (int a = 0; a < 3; a++)
            {
                const double centre = 0.5 * ((double)b[child[i]].lo[a] + b[child[i]].hi[a]) - 0.5 * ((double)nlo[a] + nhi[a]);
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
                if (childAt[sl] < 0 && (bi < 0 || cost[i][sl] < bc)) { bc = cost[i][sl]; bi = i; bs = sl; }
        }
        slotOf[bi] = bs;
        childAt[bs] = bi;
    }
    // encode
    n.px = nlo[0]; n.py = nlo[1]; n.pz = nlo[2];
    uint32_t ebyte[3];
    double step[3];
    for (int a = 0; a < 3; a++)
    {
        ebyte[a] = grid_exponent((double)nhi[a] - nlo[a]);
        step[a] = ldexp(1.0, (int)ebyte[a] - 127);
    }
    uint32_t imask = 0, valid = 0;
    uint32_t qlo[3][2] = {{0, 0}, {0, 0}, {0, 0}}, qhi[3][2] = {{0, 0}, {0, 0}, {0, 0}};
    for (uint32_t k = 0; k < kTriSlots; k++) order[k] = kNoTriangle;
    nInner = 0;
    bool ok = true;
    for (int sl = 0; sl < 8; sl++)
    {
        if (childAt[sl] < 0) continue;
        const int ci = child[childAt[sl]];
        const BinNode &c = b[ci];
        for (int a = 0; a < 3; a++)
        {
            const double l = floor(((double)c.lo[a] - nlo[a]) / step[a]);
            const double h = ceil(((double)c.hi[a] - nlo[a]) / step[a]);
            if (!(h <= 255.0)) ok = false;
            const uint32_t ql = (uint32_t)(l < 0.0 ? 0.0 : (l > 255.0 ? 255.0 : l));
            const uint32_t qh = (uint32_t)(h < 0.0 ? 0.0 : (h > 255.0 ? 255.0 : h));
            qlo[a][sl >> 2] |= ql << (8 * (sl & 3));
            qhi[a][sl >> 2] |= qh << (8 * (sl & 3));
        }
        if (!t.asLeaf[ci])
        {
            imask |= 1u << sl;
            innerChild[nInner++] = ci;
        }
        else
        {
            // triangle k of the leaf in slot sl: bit 3 * sl + k of `valid`, record 3 * sl + k of this node. The (at most 3)
            // triangles are the leaves of the binary subtree, left to right.
            int walk[4], wp = 0;
            uint32_t k = 0;
            walk[wp++] = ci;
            while (wp)
            {
                const BinNode &w = b[walk[--wp]];
                if (w.left < 0)
                {
                    if (k < kLeafTriangles)
                    {
                        valid |= 1u << (3 * sl + (int)k);
                        order[3 * sl + k] = perm ? perm[w.first] : w.first;
                    }
                    k++;
                }
                else if (wp + 2 <= 4)
                {
                    walk[wp++] = w.right;
                    walk[wp++] = w.left;
                }
                else ok = false;
            }
            if (k != c.count || k > kLeafTriangles) ok = false;
        }
    }
    n.eimask = ebyte[0] | (ebyte[1] << 8) | (ebyte[2] << 16) | (imask << 24);
    n.childBase = childBase;
    n.valid = valid;
    n.spare[0] = n.spare[1] = 0;
    for (int a = 0; a < 3; a++)
    {
        n.qlo[a][0] = qlo[a][0]; n.qlo[a][1] = qlo[a][1];
        n.qhi[a][0] = qhi[a][0]; n.qhi[a][1] = qhi[a][1];
    }
    return ok;
}

}  // namespace ptc
