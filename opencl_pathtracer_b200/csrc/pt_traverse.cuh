// pt_traverse.cuh — ray/box, ray/triangle and BVH2 traversal of the extend / connect stages.
//
// Stage -> reference function map (Kernel/PathTracer_FullKernel.cl):
//   box_test   BoundingBox_Intersects :64-139
//   tri_test   Triangle_Intersects    :519-566 (geometric half; the material / colour half :568-586 is
//              evaluated once per final hit by the shade stage)
//   traverse   BVH_IntersectRay :620-702 (ANY = false), BVH_IntersectShadowRay :705-783 (ANY = true)
//
// Arithmetic: this translation unit is compiled with -fmad=false; every expression below keeps the
// reference's association order, so all comparisons see bit-identical values in both modes.
//
// Modes (template parameter FAST):
//   PARITY  the last line of BoundingBox_Intersects compares the LINEAR entry distance with the
//           SQUARED best distance (closest hit) or with the LINEAR light distance that the triangle
//           test then compares with squared distances (shadow) — SURVEY.md F9a/F9b, kept as is.
//   FAST    the same threshold is a squared distance in every use, so a box is skipped when
//           tMin^2 > threshold (with a 1e-5 relative margin that keeps the cull conservative under
//           rounding). Traversal order, triangle arithmetic and tie-break are unchanged, so the hit is
//           the one the reference finds whenever the reference's own culling did not discard the true
//           closest hit (only possible for hit distances < 1 scene unit).
#pragma once

#include "pt_device.cuh"
#include "pt_math.cuh"

namespace ptc
{
constexpr int kStackDepth = 32;  // reference: Node* stack[BVH_MAX_DEPTH = 30]; upload rejects deeper trees

struct Ray
{
    float ox, oy, oz, ow;
    float dx, dy, dz, dw;
    float ix, iy, iz;
};

__device__ __forceinline__ Ray make_ray(float4 o, float4 d)
{
    Ray r;
    r.ox = o.x; r.oy = o.y; r.oz = o.z; r.ow = o.w;
    r.dx = d.x; r.dy = d.y; r.dz = d.z; r.dw = d.w;
    r.ix = 1.0f / d.x; r.iy = 1.0f / d.y; r.iz = 1.0f / d.z;  // header.cl:280-282
    return r;
}

// BoundingBox_Intersects as one boolean expression over the same comparisons on the same values (the
// early returns of the reference only decide the result, they have no side effect).
template <bool FAST>
__device__ __forceinline__ bool box_test(float nx, float ny, float nz, float xx, float xy, float xz, const Ray &r, float best)
{
    const bool px = r.dx > 0, py = r.dy > 0, pz = r.dz > 0;
    float tMin = ((px ? nx : xx) - r.ox) * r.ix;
    float tMax = ((px ? xx : nx) - r.ox) * r.ix;
    bool rej = (tMin < 0 && tMax < 0);
    const float tyMin = ((py ? ny : xy) - r.oy) * r.iy;
    const float tyMax = ((py ? xy : ny) - r.oy) * r.iy;
    rej |= (tyMin < 0 && tyMax < 0);
    rej |= (tMin > tyMax || tyMin > tMax);
    if (tyMin > tMin) tMin = tyMin;
    if (tyMax < tMax) tMax = tyMax;
    const float tzMin = ((pz ? nz : xz) - r.oz) * r.iz;
    const float tzMax = ((pz ? xz : nz) - r.oz) * r.iz;
    rej |= (tzMin < 0 && tzMax < 0);
    rej |= (tMin > tzMax || tzMin > tMax);
    if (tzMin > tMin) tMin = tzMin;
    if (FAST) return !rej && (tMin < 0 || !(tMin * tMin > best * 1.00001f));
    return !rej && (tMin < 0 || !(tMin > best));
}

// uv/uu/vv are recomputed from the stored edges with the reference's association order; denom and d are
// stored (ray independent, computed at upload with the same IEEE operations). The w lanes of the edge
// dot products are exactly zero for conformant data (upload checks S1.w == S2.w == S3.w) and omitted.
__device__ __forceinline__ bool tri_test(const DevTri *__restrict__ tris, uint32_t i, const Ray &r, float &best, float &hs,
                                         float &ht, float &htp)
{
    const float4 n = __ldg(&tris[i].n);
    const float4 p = __ldg(&tris[i].p);
    const float4 e1 = __ldg(&tris[i].e1);
    const float4 e2 = __ldg(&tris[i].e2);
    const float nd = ((n.x * r.dx + n.y * r.dy) + n.z * r.dz) + n.w * r.dw;
    if ((nd > -0.00001f) && (nd < 0.00001f)) return false;
    const float no = ((n.x * r.ox + n.y * r.oy) + n.z * r.oz) + n.w * r.ow;
    const float tp = (p.w - no) / nd;
    const float qx = r.ox + r.dx * tp, qy = r.oy + r.dy * tp, qz = r.oz + r.dz * tp, qw = r.ow + r.dw * tp;
    const float fx = qx - r.ox, fy = qy - r.oy, fz = qz - r.oz, fw = qw - r.ow;
    const float nsd = ((fx * fx + fy * fy) + fz * fz) + fw * fw;
    if (nsd > best) return false;
    if (nsd < 0.00001f) return false;
    const float wx = qx - p.x, wy = qy - p.y, wz = qz - p.z;
    const float uv = (e1.x * e2.x + e1.y * e2.y) + e1.z * e2.z;
    const float wv = (wx * e2.x + wy * e2.y) + wz * e2.z;
    const float wu = (wx * e1.x + wy * e1.y) + wz * e1.z;
    const float uu = (e1.x * e1.x + e1.y * e1.y) + e1.z * e1.z;
    const float vv = (e2.x * e2.x + e2.y * e2.y) + e2.z * e2.z;
    const float s = (uv * wv - vv * wu) * e1.w;
    const float t = (uv * wu - uu * wv) * e1.w;
    if (s < 0 || t < 0 || s + t > 1) return false;
    if ((((fx * r.dx + fy * r.dy) + fz * r.dz) + fw * r.dw) < 0) return false;
    best = nsd;
    hs = s; ht = t; htp = tp;
    return true;
}

// One ray through the BVH2 in the reference's order: children ordered by the sign of
// direction[cutAxis] (:660-669), both boxes tested, near descended, far pushed; leaves scanned in
// array order; a candidate at equal distance replaces the previous one (:543 is a strict test).
// `stk` points at this thread's column of a [level][thread] shared-memory stack (conflict-free).
template <bool ANY, bool FAST>
__device__ __forceinline__ bool traverse(const SceneDev &sc, const Ray &r, float best, int &hitTri, float &hs, float &ht,
                                         float &htp, uint32_t &nb, uint32_t &nt, uint32_t *stk, const int stride)
{
    uint32_t cur = sc.rootRef;
    if (cur == kInvalidRef) return false;
    int sp = 0;
    bool hit = false;
    while (true)
    {
        if (cur & kLeafBit)
        {
            const uint32_t start = (cur & 0x7fffffffu) >> 4;
            uint32_t cnt = cur & 15u;
            if (cnt == 0) cnt = __ldg(&sc.bigLeafCount[start]);
            for (uint32_t i = start; i < start + cnt; i++)
            {
                nt++;
                if (tri_test(sc.tris, i, r, best, hs, ht, htp))
                {
                    if (ANY) return true;
                    hit = true;
                    hitTri = (int)i;
                }
            }
            if (sp == 0) break;
            cur = stk[(--sp) * stride];
        }
        else
        {
            const DevNode *nd = &sc.nodes[cur];
            const float4 a = __ldg(&nd->a), b = __ldg(&nd->b), c = __ldg(&nd->c);
            const uint4 d = __ldg(&nd->d);
            nb += 2;
            const bool bL = !(d.z & 4u) && box_test<FAST>(a.x, a.y, a.z, a.w, b.x, b.y, r, best);
            const bool bR = !(d.z & 8u) && box_test<FAST>(b.z, b.w, c.x, c.y, c.z, c.w, r, best);
            const uint32_t axis = d.z & 3u;
            const float da = axis == 0 ? r.dx : (axis == 1 ? r.dy : r.dz);
            const bool leftFirst = da > 0;
            const uint32_t first = leftFirst ? d.x : d.y, second = leftFirst ? d.y : d.x;
            const bool b1 = leftFirst ? bL : bR, b2 = leftFirst ? bR : bL;
            if (b1)
            {
                if (b2) stk[(sp++) * stride] = second;
                cur = first;
            }
            else if (b2)
                cur = second;
            else
            {
                if (sp == 0) break;
                cur = stk[(--sp) * stride];
            }
        }
    }
    return hit;
}

}  // namespace ptc
