// pt_traverse.cuh — ray/box and ray/triangle tests and the warp-cooperative BVH2 traversal loop of the extend /
// connect stages.
//
// Stage -> reference function map (Kernel/PathTracer_FullKernel.cl):
//   box_test / box_test_ref   BoundingBox_Intersects :64-139
//   tri_test                  Triangle_Intersects    :519-566 (geometric half; the material / colour half :568-586
//                             is evaluated once per final hit by the shade stage)
//   TraceLane::node_step / tri_step: one iteration of BVH_IntersectRay :620-702 (ANY = false) or
//                             BVH_IntersectShadowRay :705-783 (ANY = true)
//
// Arithmetic: this translation unit is compiled with -fmad=false; every expression keeps the reference's
// association order, so all comparisons see bit-identical values in both modes.
//
// Modes (template parameter FAST):
//   PARITY  the last line of BoundingBox_Intersects compares the LINEAR entry distance with the SQUARED best
//           distance (closest hit) or with the LINEAR light distance that the triangle test then compares with
//           squared distances (shadow) — SURVEY.md F9a/F9b, kept as is.
//   FAST    the same threshold is a squared distance in every use, so a box is skipped when tMin^2 > threshold
//           (with a 1e-5 relative margin that keeps the cull conservative under rounding). Traversal order,
//           triangle arithmetic and tie-break are unchanged, so the hit is the one the reference finds whenever
//           the reference's own culling did not discard the true closest hit (only possible for hit distances
//           < 1 scene unit).
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

// A ray whose origin and reciprocal direction are all finite can only produce finite or infinite slab
// distances (never NaN: no 0*inf, no inf-inf), which is what the min/max formulation below relies on.
__device__ __forceinline__ bool ray_is_regular(const Ray &r)
{
    const float big = 3.0e38f;
    return fabsf(r.ox) < big && fabsf(r.oy) < big && fabsf(r.oz) < big && fabsf(r.ix) < big && fabsf(r.iy) < big && fabsf(r.iz) < big;
}

template <bool FAST>
__device__ __forceinline__ bool box_final(float tMin, float best)
{
    if (FAST) return tMin < 0 || !(tMin * tMin > best * 1.00001f);
    return tMin < 0 || !(tMin > best);  // :132-138
}

// BoundingBox_Intersects statement by statement (one boolean expression over the same comparisons on the same
// values: the early returns of the reference only decide the result). Used for rays that are not "regular".
template <bool FAST>
__device__ __noinline__ bool box_test_ref(float nx, float ny, float nz, float xx, float xy, float xz, float ox, float oy, float oz,
                                          float dx, float dy, float dz, float ix, float iy, float iz, float best)
{
    const bool px = dx > 0, py = dy > 0, pz = dz > 0;
    float tMin = ((px ? nx : xx) - ox) * ix;
    float tMax = ((px ? xx : nx) - ox) * ix;
    bool rej = (tMin < 0 && tMax < 0);
    const float tyMin = ((py ? ny : xy) - oy) * iy;
    const float tyMax = ((py ? xy : ny) - oy) * iy;
    rej |= (tyMin < 0 && tyMax < 0);
    rej |= (tMin > tyMax || tyMin > tMax);
    if (tyMin > tMin) tMin = tyMin;
    if (tyMax < tMax) tMax = tyMax;
    const float tzMin = ((pz ? nz : xz) - oz) * iz;
    const float tzMax = ((pz ? xz : nz) - oz) * iz;
    rej |= (tzMin < 0 && tzMax < 0);
    rej |= (tMin > tzMax || tzMin > tMax);
    if (tzMin > tMin) tMin = tzMin;
    return !rej && box_final<FAST>(tMin, best);
}

// The same predicate for regular rays. With t1 = (pMin - o) * inv and t2 = (pMax - o) * inv per axis (the very
// products the reference forms; which one it calls "min" depends on the sign of the direction), near = min(t1,t2)
// and far = max(t1,t2) per axis are the reference's tMin/tMax of that axis (near <= far by monotonicity of
// rounding). Its tests are then: far < 0 on any axis -> reject (:85,:99,:121); pairwise interval overlap (:102,
// :124), which for three intervals is max(near) <= min(far); finally the distance test on tMin = max(near).
// No NaN can occur for a regular ray, so every comparison has the reference's outcome.
template <bool FAST>
__device__ __forceinline__ bool box_test(float nx, float ny, float nz, float xx, float xy, float xz, const Ray &r, float best)
{
    const float ax = (nx - r.ox) * r.ix, bx = (xx - r.ox) * r.ix;
    const float ay = (ny - r.oy) * r.iy, by = (xy - r.oy) * r.iy;
    const float az = (nz - r.oz) * r.iz, bz = (xz - r.oz) * r.iz;
    const float tEnter = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
    const float tExit = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
    return !(tExit < 0) && !(tEnter > tExit) && box_final<FAST>(tEnter, best);
}

// uv/uu/vv are recomputed from the stored edges with the reference's association order; denom and d are stored (ray
// independent, computed at upload with the same IEEE operations). The w lanes of the edge dot products are exactly
// zero for conformant data (upload checks S1.w == S2.w == S3.w) and omitted.
// On acceptance hmat receives the MaterialType of the side that was hit (:568): both sides' types ride in the spare
// lane of the record (e2.w, bits 3..0 positive side, bits 7..4 negative side).
// order / prevHit (wide walk only): the reference keeps the LAST candidate it tests among several at exactly the same distance
// (:543 is a strict test), and inside one leaf it tests in array order. The wide walk tests in another order, so on an
// exact tie it keeps the candidate with the higher BVH-order index: what the reference does whenever the candidates share
// a leaf. order = null in the BVH2 walk, whose test order is the reference's.
__device__ __forceinline__ bool tri_test(const DevTri *__restrict__ tris, uint32_t i, const Ray &r, float &best, float &hs,
                                         float &ht, float &htp, uint32_t &hmat, const uint32_t *__restrict__ order = nullptr,
                                         int prevHit = -1)
{
    float4 n, p, e1, e2;
    ldg256(&tris[i].n, n, p);
    ldg256(&tris[i].e1, e1, e2);
    const float nd = ((n.x * r.dx + n.y * r.dy) + n.z * r.dz) + n.w * r.dw;
    if ((nd > -0.00001f) && (nd < 0.00001f)) return false;
    const float no = ((n.x * r.ox + n.y * r.oy) + n.z * r.oz) + n.w * r.ow;
    // A shadow ray starts ON the triangle it was shot from: p.w - no is then a small multiple of an ulp and very often exactly 0.
    // 0 / nd = +-0 puts q on the origin, the squared distance is 0 and the reference's self-hit guard (:545, `< 0.00001f` below)
    // rejects the candidate — the same answer without the IEEE division, whose zero-numerator case is a subroutine call that 2 of a
    // warp's lanes took in 64 % of connect's triangle steps (ncu v14).
    const float num = p.w - no;
    if (num == 0.f) return false;
    const float tp = num / nd;
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
    // the tie rule comes last: `nsd == best` alone is common — every triangle in the plane of the current hit (a flat tessellated
    // surface) gives the same hit point, 16 % of extend's triangle steps on the bench scene — a true tie (also inside) is rare, and
    // the rule costs two dependent loads
    if (order && nsd == best && prevHit >= 0 && __ldg(&order[i]) < __ldg(&order[prevHit])) return false;
    best = nsd;
    hs = s; ht = t; htp = tp;
    const uint32_t types = __float_as_uint(e2.w);
    hmat = nd < 0 ? (types & 15u) : ((types >> 4) & 15u);
    return true;
}

// Per-lane traversal state. One ray is walked through the BVH2 in the reference's order: children ordered by the sign
// of direction[cutAxis] (:660-669), both boxes tested, near descended, far pushed; leaves scanned in array order, one
// triangle per tri_step; a candidate at equal distance replaces the previous one (:543 is a strict test).
// The warp decides whether its lanes take node steps or triangle steps (see k_trace), so the state machine is
// explicit. `cur` is: an inner-node index; or a leaf reference that doubles as the scan cursor (bits 30..4 = next
// triangle, bits 3..0 = triangles left; after a test cur += 15, i.e. next + 1 and left - 1); or a "big leaf"
// reference (count field 0, more than 15 triangles: cur += 16 per test, the end of the scan is looked up on the first
// test and kept in `te`); or kDone (ray finished, result not yet written); or kIdle (no ray).
constexpr uint32_t kDone = 0xfffffffeu, kIdle = 0xffffffffu;

struct TraceLane
{
    typedef uint32_t *StackRef;
    static constexpr bool kStoresHit = false;  // the hit record rides in the lane (hs, ht, htp, hmat) and is written at completion
    static constexpr uint32_t kMatShift = 0;
    static constexpr int kItemLevels = 0;
    uint32_t oct;  // unused (WideLane's flag word)
    static __device__ __forceinline__ StackRef stack_ref(uint32_t *p) { return p; }
    Ray r;
    float best;
    uint32_t cur;
    uint32_t te;       // big leaves only: one past the last triangle of the leaf being scanned, 0 = not looked up yet
    uint32_t dirPos;   // bit a set <=> direction[a] > 0 (child order, :660)
    int sp;
    int hitTri;
    float hs, ht, htp;
    uint32_t hmat;     // MaterialType of the hit side
    uint32_t nb, nt;
    bool regular;
    bool found;        // ANY: an occluder was accepted

    // the interface k_trace drives (shared with WideLane, pt_wide.cuh)
    typedef uint32_t StackT;
    static constexpr int kBoxesPerNodeStep = 2;
    static constexpr int kMinCtasAny = 8;
    static constexpr int kMinCtas = 8;  // 64 registers (PARITY connect spills 16 B): 53.7 -> 58.9 Msamples/s against an unbounded 79 registers / 6 CTAs
    __device__ __forceinline__ void reset() { cur = kIdle; nb = nt = 0; found = false; hitTri = -1; }
    __device__ __forceinline__ bool idle() const { return cur == kIdle; }
    __device__ __forceinline__ void set_idle() { cur = kIdle; }
    __device__ __forceinline__ bool finished() const { return cur == kDone; }
    __device__ __forceinline__ bool wants_node() const { return !(cur & kLeafBit); }
    __device__ __forceinline__ bool wants_tri() const { return (cur & kLeafBit) && cur < kDone; }
    __device__ __forceinline__ int hit_index(const SceneDev &) const { return hitTri; }
    __device__ __forceinline__ void hit_record(const SceneDev &, float &s, float &t, float &tp, uint32_t &mat) const { s = hs; t = ht; tp = htp; mat = hmat; }
};

__device__ __forceinline__ bool lane_at_inner(const TraceLane &L) { return !(L.cur & kLeafBit); }
__device__ __forceinline__ bool lane_at_leaf(const TraceLane &L) { return (L.cur & kLeafBit) && L.cur < kDone; }

// Next subtree from the stack, or kDone when it is empty. Written so that it compiles to predicated code (one LDS).
__device__ __forceinline__ uint32_t stack_pop(int &sp, const uint32_t *stk, const int stride, bool doPop)
{
    const bool has = doPop && sp > 0;
    uint32_t next = kDone;
    if (has) next = stk[(sp - 1) * stride];
    sp -= has ? 1 : 0;
    return next;
}

__device__ __forceinline__ void lane_begin(TraceLane &L, const SceneDev &sc, const Ray &r, float best)
{
    L.r = r;
    L.best = best;
    L.sp = 0;
    L.hitTri = -1;
    L.hs = L.ht = L.htp = 0.f;
    L.hmat = 0;
    L.te = 0;
    L.found = false;
    L.regular = ray_is_regular(r);
    L.dirPos = (r.dx > 0 ? 1u : 0u) | (r.dy > 0 ? 2u : 0u) | (r.dz > 0 ? 4u : 0u);
    L.cur = sc.rootRef == kInvalidRef ? kDone : sc.rootRef;
}

// One inner-node visit: both child boxes tested (:660-672), then the successor chosen without divergent branches:
// both hit -> near child, far child pushed; one hit -> that child; none -> pop.
template <bool FAST>
__device__ __forceinline__ void lane_node_step(TraceLane &L, const SceneDev &sc, uint32_t *stk, const int stride)
{
    PTC_CHECK(sc.faultSink, kFaultNodeIndex, L.cur, sc.nInner);
    const DevNode *nd = &sc.nodes[L.cur];
    float4 a, b, c, dd;
    ldg256(&nd->a, a, b);
    ldg256(&nd->c, c, dd);
    const uint32_t refL = __float_as_uint(dd.x), refR = __float_as_uint(dd.y), meta = __float_as_uint(dd.z);
    L.nb += 2;
    bool hitL, hitR;
    if (L.regular)
    {
        hitL = box_test<FAST>(a.x, a.y, a.z, a.w, b.x, b.y, L.r, L.best);
        hitR = box_test<FAST>(b.z, b.w, c.x, c.y, c.z, c.w, L.r, L.best);
    }
    else
    {
        const Ray &r = L.r;
        hitL = box_test_ref<FAST>(a.x, a.y, a.z, a.w, b.x, b.y, r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, r.ix, r.iy, r.iz, L.best);
        hitR = box_test_ref<FAST>(b.z, b.w, c.x, c.y, c.z, c.w, r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, r.ix, r.iy, r.iz, L.best);
    }
    hitL = hitL && !(meta & 4u);  // an empty child box never intersects (:68)
    hitR = hitR && !(meta & 8u);
    // near child first (:660-669): left is near when direction[cutAxis] > 0
    const bool leftFirst = (L.dirPos >> (meta & 3u)) & 1u;
    const bool both = hitL && hitR;
    const uint32_t nearRef = leftFirst ? refL : refR, farRef = leftFirst ? refR : refL;
    const uint32_t descend = both ? nearRef : (hitL ? refL : refR);
    if (both) PTC_CHECK(sc.faultSink, kFaultStack, L.sp, sc.stackDepth);
    if (both) stk[L.sp * stride] = farRef;  // (prefetching the postponed child into L1 measured 2 % slower)
    L.sp += both ? 1 : 0;
    const uint32_t popped = stack_pop(L.sp, stk, stride, !(hitL || hitR));
    L.cur = (hitL || hitR) ? descend : popped;
}

// the item a lane is tracing: the BVH2 walk keeps it in the warp's staging block (the wide walk in level 0 of its stack)
__device__ __forceinline__ uint32_t lane_item(const uint32_t *, const uint32_t *slot) { return *slot; }
__device__ __forceinline__ void lane_set_item(const uint32_t *, uint32_t *slot, uint32_t v) { *slot = v; }

template <bool ANY>
__device__ __forceinline__ void lane_tri_step(TraceLane &L, const SceneDev &sc, const uint32_t *stk, const int stride, float4 *)
{
    const uint32_t i = (L.cur >> 4) & 0x7ffffffu;
    PTC_CHECK(sc.faultSink, kFaultTriIndex, i, sc.nTriangles);
    const bool big = (L.cur & 15u) == 0u;
    if (big && L.te == 0u) L.te = i + __ldg(&sc.bigLeafCount[i]);  // rare: leaves of more than 15 triangles
    L.nt++;
    bool stop = false;
    if (tri_test(sc.tris, i, L.r, L.best, L.hs, L.ht, L.htp, L.hmat))
    {
        L.hitTri = (int)i;
        if (ANY)
        {
            L.found = true;
            stop = true;
        }
    }
    const bool last = big ? (i + 1u == L.te) : ((L.cur & 15u) == 1u);
    L.cur += big ? 16u : 15u;
    if (last) L.te = 0u;
    const uint32_t popped = stack_pop(L.sp, stk, stride, last && !stop);
    if (last) L.cur = popped;
    if (stop) L.cur = kDone;
}

}  // namespace ptc
