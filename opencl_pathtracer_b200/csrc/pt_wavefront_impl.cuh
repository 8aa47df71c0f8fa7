// pt_wavefront_impl.cuh — extend / connect / shade stages. Included twice:
//   pt_stages_parity.cu  (PT_FAST 0, compiled -fmad=false): reference traversal order, reference
//                         culling quirks, IEEE arithmetic -> bit-exact against the oracle;
//   pt_stages_fast.cu    (PT_FAST 1, FMA allowed): same estimator, distance-ordered traversal, correct
//                         culling, shadow rays with zero contribution skipped.
//
// Stage -> reference function map (all in Kernel/PathTracer_FullKernel.cl unless noted):
//   k_extend   BVH_IntersectRay :620-702, BoundingBox_Intersects :64-139, Triangle_Intersects :519-566
//   k_connect  BVH_IntersectShadowRay :705-783 for every light of Scene_ComputeDirectIllumination :930-948
//   k_shade    hit setup :1254-1275, direct light sum :901-954, Scene_ComputeRadiance :791-891,
//              Material_* :166-416, Sky_* :438-512, Triangle_GetColorValueAt :591-602,
//              Texture_GetPixelColorValue header.cl:430-459, termination :1294-1304
#include "pt_device.cuh"
#include "pt_math.cuh"

#ifndef PT_FAST
#error "define PT_FAST to 0 or 1"
#endif

namespace ptc
{
namespace PT_NS
{
constexpr int kStackDepth = 32;      // reference: Node* stack[BVH_MAX_DEPTH = 30]; upload rejects deeper trees
constexpr int kTraverseThreads = 128;
constexpr float kPi = 3.14159265f;   // PATH_PI
constexpr float kPiInv = 0.31830988618f;

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

#if !PT_FAST
// BoundingBox_Intersects (:64-139) as one boolean expression over the same comparisons on the same
// values. `best` is the SQUARED hit distance in closest-hit traversal and the LINEAR light distance in
// shadow traversal, compared against the linear entry distance — the reference's quirks F9a / F9b.
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
    return !rej && (tMin < 0 || !(tMin > best));
}

// Triangle_Intersects (:519-566). uv/uu/vv are recomputed from the stored edges with the reference's
// association order; denom and d are stored (ray independent, computed once at upload with the same
// operations). The +w lanes of the edge dot products are exactly zero for importer-conformant data
// (upload checks S1.w == S2.w == S3.w) and are omitted.
__device__ __forceinline__ bool tri_test(const DevTri *__restrict__ tris, uint32_t i, const Ray &r, float &best, float &hs,
                                         float &ht, float &htp)
{
    const float4 n = __ldg(&tris[i].n);
    const float nd = ((n.x * r.dx + n.y * r.dy) + n.z * r.dz) + n.w * r.dw;
    if ((nd > -0.00001f) && (nd < 0.00001f)) return false;
    const float4 p = __ldg(&tris[i].p);
    const float no = ((n.x * r.ox + n.y * r.oy) + n.z * r.oz) + n.w * r.ow;
    const float tp = (p.w - no) / nd;
    const float qx = r.ox + r.dx * tp, qy = r.oy + r.dy * tp, qz = r.oz + r.dz * tp, qw = r.ow + r.dw * tp;
    const float fx = qx - r.ox, fy = qy - r.oy, fz = qz - r.oz, fw = qw - r.ow;
    const float nsd = ((fx * fx + fy * fy) + fz * fz) + fw * fw;
    if (nsd > best) return false;
    if (nsd < 0.00001f) return false;
    const float4 e1 = __ldg(&tris[i].e1);
    const float4 e2 = __ldg(&tris[i].e2);
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
// ANY = BVH_IntersectShadowRay: return at the first accepted triangle.
template <bool ANY>
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
            const bool bL = !(d.z & 4u) && box_test(a.x, a.y, a.z, a.w, b.x, b.y, r, best);
            const bool bR = !(d.z & 8u) && box_test(b.z, b.w, c.x, c.y, c.z, c.w, r, best);
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

#else  // ------------------------------------------------------------------------------ FAST

// Slab test returning the entry distance; correct culling against the LINEAR best distance.
__device__ __forceinline__ bool box_test_fast(float nx, float ny, float nz, float xx, float xy, float xz, const Ray &r,
                                              float tbest, float &tnear)
{
    const float ax = (nx - r.ox) * r.ix, bx = (xx - r.ox) * r.ix;
    const float ay = (ny - r.oy) * r.iy, by = (xy - r.oy) * r.iy;
    const float az = (nz - r.oz) * r.iz, bz = (xz - r.oz) * r.iz;
    const float t0 = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.f));
    const float t1 = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), tbest));
    tnear = t0;
    return t0 <= t1;
}

// Same plane / barycentric formulation as the reference (so s, t and the hit point agree to rounding),
// with the rejection tests expressed on the ray parameter: |q - o|^2 = tp^2 for a unit direction.
__device__ __forceinline__ bool tri_test_fast(const DevTri *__restrict__ tris, uint32_t i, const Ray &r, float &tbest, float &hs,
                                              float &ht, float &htp)
{
    const float4 n = __ldg(&tris[i].n);
    const float nd = n.x * r.dx + n.y * r.dy + n.z * r.dz;
    if (fabsf(nd) < 0.00001f) return false;
    const float4 p = __ldg(&tris[i].p);
    const float no = n.x * r.ox + n.y * r.oy + n.z * r.oz + n.w * r.ow;
    const float tp = (p.w - no) / nd;
    if (!(tp <= tbest) || tp * tp < 0.00001f || tp < 0) return false;
    const float4 e1 = __ldg(&tris[i].e1);
    const float4 e2 = __ldg(&tris[i].e2);
    const float wx = r.ox + r.dx * tp - p.x, wy = r.oy + r.dy * tp - p.y, wz = r.oz + r.dz * tp - p.z;
    const float uv = e1.x * e2.x + e1.y * e2.y + e1.z * e2.z;
    const float wv = wx * e2.x + wy * e2.y + wz * e2.z;
    const float wu = wx * e1.x + wy * e1.y + wz * e1.z;
    const float uu = e1.x * e1.x + e1.y * e1.y + e1.z * e1.z;
    const float vv = e2.x * e2.x + e2.y * e2.y + e2.z * e2.z;
    const float s = (uv * wv - vv * wu) * e1.w;
    const float t = (uv * wu - uu * wv) * e1.w;
    if (s < 0 || t < 0 || s + t > 1) return false;
    tbest = tp;
    hs = s; ht = t; htp = tp;
    return true;
}

template <bool ANY>
__device__ __forceinline__ bool traverse(const SceneDev &sc, const Ray &r, float tbest, int &hitTri, float &hs, float &ht,
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
                if (tri_test_fast(sc.tris, i, r, tbest, hs, ht, htp))
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
            float tL, tR;
            const bool bL = !(d.z & 4u) && box_test_fast(a.x, a.y, a.z, a.w, b.x, b.y, r, tbest, tL);
            const bool bR = !(d.z & 8u) && box_test_fast(b.z, b.w, c.x, c.y, c.z, c.w, r, tbest, tR);
            if (bL && bR)
            {
                const bool leftFirst = ANY ? true : (tL <= tR);
                stk[(sp++) * stride] = leftFirst ? d.y : d.x;
                cur = leftFirst ? d.x : d.y;
            }
            else if (bL)
                cur = d.x;
            else if (bR)
                cur = d.y;
            else
            {
                if (sp == 0) break;
                cur = stk[(--sp) * stride];
            }
        }
    }
    return hit;
}
#endif

// ------------------------------------------------------------------------------------------------
// Persistent traversal kernels. Each warp pulls chunks of 32 queue entries from a global cursor
// until the queue of this bounce is exhausted; per-thread traversal stacks live in shared memory,
// laid out [level][thread] so that a warp touching one level is conflict-free.

__device__ __forceinline__ bool fetch_chunk(uint32_t *cursor, uint32_t count, uint32_t &base)
{
    const unsigned lane = threadIdx.x & 31u;
    uint32_t b = 0;
    if (lane == 0) b = atomicAdd(cursor, 32u);
    base = __shfl_sync(0xffffffffu, b, 0);
    return base < count;
}

__global__ void __launch_bounds__(kTraverseThreads) k_extend(SceneDev sc, PathState st, const uint32_t *__restrict__ queue,
                                                            Counters *__restrict__ cn, int bounce)
{
    extern __shared__ uint32_t smemStack[];
    uint32_t *stk = smemStack + threadIdx.x;
    const uint32_t count = cn->qcount[bounce];
    uint32_t nbSum = 0, ntSum = 0, nRays = 0;
    uint32_t base;
    while (fetch_chunk(&cn->curExtend[bounce], count, base))
    {
        const uint32_t k = base + (threadIdx.x & 31u);
        if (k < count)
        {
            const uint32_t p = queue[k];
            const Ray r = make_ray(st.org[p], st.dir[p]);
            int tri = -1;
            float hs = 0.f, ht = 0.f, htp = 0.f;
            uint32_t nb = 0, nt = 0;
            traverse<false>(sc, r, INFINITY, tri, hs, ht, htp, nb, nt, stk, kTraverseThreads);
            st.hit[p] = f4(__int_as_float(tri), hs, ht, htp);
            st.nbbx[p] += nb;
            st.ntri[p] += nt;
            nbSum += nb; ntSum += nt; nRays++;
        }
    }
    // work counters: one atomic per warp
    for (int o = 16; o; o >>= 1)
    {
        nbSum += __shfl_xor_sync(0xffffffffu, nbSum, o);
        ntSum += __shfl_xor_sync(0xffffffffu, ntSum, o);
        nRays += __shfl_xor_sync(0xffffffffu, nRays, o);
    }
    if ((threadIdx.x & 31u) == 0 && nRays)
    {
        atomicAdd(&cn->extensionRays, (unsigned long long)nRays);
        atomicAdd(&cn->boxTests, (unsigned long long)nbSum);
        atomicAdd(&cn->triTests, (unsigned long long)ntSum);
    }
}

// Shadow ray of light `li` from the hit point q (:932-938): direction = normalize(full), bound =
// length(full) (INFINITY for a directional light) — a LENGTH handed to a parameter compared with
// squared distances (F9b), reproduced as is.
__device__ __forceinline__ void shadow_setup(const DevLight &l, float4 q, float4 &dir, float &bound)
{
    const float4 full = (l.type == 0) ? -l.direction : (l.position - q);
    dir = normalize4(full);
    bound = (l.type == 0) ? INFINITY : length4(full);
}

__global__ void __launch_bounds__(kTraverseThreads) k_connect(SceneDev sc, PathState st, const uint32_t *__restrict__ queue,
                                                             Counters *__restrict__ cn, int bounce)
{
    extern __shared__ uint32_t smemStack[];
    uint32_t *stk = smemStack + threadIdx.x;
    const uint32_t count = cn->qcount[bounce];
    uint32_t nbSum = 0, ntSum = 0, nRays = 0;
    uint32_t base;
    while (fetch_chunk(&cn->curConnect[bounce], count, base))
    {
        const uint32_t k = base + (threadIdx.x & 31u);
        if (k >= count) continue;
        const uint32_t p = queue[k];
        const float4 h = st.hit[p];
        if (__float_as_int(h.x) < 0) continue;  // miss: no direct lighting
        const float4 o = st.org[p], d = st.dir[p];
        const float4 q = f4(o.x + d.x * h.w, o.y + d.y * h.w, o.z + d.z * h.w, o.w + d.w * h.w);  // :538
        uint32_t occl = 0, nb = 0, nt = 0;
#if PT_FAST
        // lights that cannot contribute are not traced: GLASS and WATER ignore direct light (:805-843)
        const DevTri *tp = &sc.tris[__float_as_int(h.x)];
        const float4 n = __ldg(&tp->n);
        const float ndir = n.x * d.x + n.y * d.y + n.z * d.z;
        const uint4 ids = __ldg(&sc.shade[__float_as_int(h.x)].ids);
        const int mtype = sc.materials[ndir < 0 ? ids.x : ids.y].type;
        const bool lit = (mtype == 0 || mtype == 3);
#endif
        for (uint32_t li = 0; li < sc.nLights; li++)
        {
#if PT_FAST
            if (!lit) break;
#endif
            float4 ldir;
            float bound;
            shadow_setup(sc.lights[li], q, ldir, bound);
            const Ray r = make_ray(q, ldir);
            int tri;
            float hs, ht, htp;
#if PT_FAST
            // the occlusion predicate of the reference accepts a blocker iff |q' - q|^2 <= bound, i.e. t <= sqrt(bound)
            const float tmax = sqrtf(bound);
            if (traverse<true>(sc, r, tmax, tri, hs, ht, htp, nb, nt, stk, kTraverseThreads)) occl |= 1u << li;
#else
            if (traverse<true>(sc, r, bound, tri, hs, ht, htp, nb, nt, stk, kTraverseThreads)) occl |= 1u << li;
#endif
            nRays++;
        }
        st.occl[p] = occl;
        st.nbbx[p] += nb;  // :946-947 the camera ray inherits its shadow rays' counters
        st.ntri[p] += nt;
        nbSum += nb; ntSum += nt;
    }
    for (int o = 16; o; o >>= 1)
    {
        nbSum += __shfl_xor_sync(0xffffffffu, nbSum, o);
        ntSum += __shfl_xor_sync(0xffffffffu, ntSum, o);
        nRays += __shfl_xor_sync(0xffffffffu, nRays, o);
    }
    if ((threadIdx.x & 31u) == 0 && nRays)
    {
        atomicAdd(&cn->shadowRays, (unsigned long long)nRays);
        atomicAdd(&cn->boxTests, (unsigned long long)nbSum);
        atomicAdd(&cn->triTests, (unsigned long long)ntSum);
    }
}

// ------------------------------------------------------------------------------------------------
// Shade stage

// header.cl:430-459 (nearest texel, wrap by truncation, alpha inverted)
__device__ __forceinline__ float4 texture_lookup(const SceneDev &sc, uint32_t w, uint32_t h, uint32_t off, float u, float v)
{
    u = u - ((int)u) + (u < 0 ? 1 : 0);
    v = v - ((int)v) + (v < 0 ? 1 : 0);
    const uint32_t x = (uint32_t)(u * (w - 1));
    const uint32_t y = (uint32_t)(v * (h - 1));
    const uchar4 c = __ldg(&sc.texels[off + y * w + x]);
    float4 bgra = f4((float)c.x, (float)c.y, (float)c.z, (float)c.w) / 255.f;
    bgra.w = 1.f - bgra.w;
    return bgra;
}

// :438-512
__device__ __forceinline__ float4 sky_color(const SceneDev &sc, float4 dir)
{
    const float x = sc.sky.cosRot * dir.x - sc.sky.sinRot * dir.y;
    const float y = sc.sky.sinRot * dir.x + sc.sky.cosRot * dir.y;
    const float z = dir.z;
    int face = 0;
    float u = 0, v = 0;
    if (fabsf(z) > fabsf(x) && fabsf(z) > fabsf(y))
    {
        if (z > 0) { face = 5; u = (1 - x / z) / 2; v = (1 + y / z) / 2; }
        else { face = 0; u = (1 + x / z) / 2; v = (1 + y / z) / 2; }
    }
    else if (fabsf(x) > fabsf(y) && fabsf(x) > fabsf(z))
    {
        if (x > 0) { face = 1; u = (1 - y / x) / 2; v = (1 + z / x) / 2; }
        else { face = 3; u = (1 - y / x) / 2; v = (1 - z / x) / 2; }
    }
    else if (fabsf(y) > fabsf(x) && fabsf(y) > fabsf(z))
    {
        if (y > 0) { face = 4; u = (1 + x / y) / 2; v = (1 + z / y) / 2; }
        else { face = 2; u = (1 + x / y) / 2; v = (1 - z / y) / 2; }
    }
    return texture_lookup(sc, sc.sky.w[face], sc.sky.h[face], sc.sky.off[face], u, v);
}

// :192-292 — the three Fresnel fractions share one body; n1/n2 differ. Returns 1 on total reflection.
__device__ __forceinline__ float fresnel(float n1, float n2, float cos1, float &cos2)
{
    const float sin1 = sqrtf(1 - cos1 * cos1);
    const float sin2 = n1 * sin1 / n2;
    cos2 = 0.f;
    if (sin2 >= 1) return 1;
    cos2 = sqrtf(1 - sin2 * sin2);
    const float rPara = (n2 * cos1 - n1 * cos2) / (n2 * cos1 + n1 * cos2);
    const float rPerp = (n1 * cos1 - n2 * cos2) / (n1 * cos1 + n2 * cos2);
    return (rPara * rPara + rPerp * rPerp) / 2.0f;
}

// :256-292 with isInVarnish = false
__device__ __forceinline__ float fresnel_varnish(float4 incident, float4 N)
{
    const float cos1 = fmaxf(0.f, fminf(1.f, -dot4(incident, N)));
    float cos2;
    return fresnel(1.0f, 3.f, cos1, cos2);
}

// :166-190
__device__ __forceinline__ float material_brdf(int type, float4 incident, float4 N, float4 reflected)
{
    if (type == 0) return kPiInv;
    if (type == 2) return 1;
    if (type == 1)
    {
        const float denom = 1 + 0.8f * dot4(incident, reflected);
        return (1 - 0.8f * 0.8f) / (4 * kPi * denom * denom);
    }
    if (type == 3) return (1 - fresnel_varnish(incident, N)) * kPiInv;
    return 1;
}

// header.cl:403-421
__device__ __forceinline__ float light_power_toward(const DevLight &l, float4 p, float4 N)
{
    if (l.type == 0) return l.power * fmaxf(dot4(-l.direction, N), 0.f);
    if (l.type == 1)
    {
        const float4 t = p - l.position;
        return l.power / dot4(t, t) * fmaxf(dot4(normalize4(l.position - p), N), 0.f);
    }
    if (l.type == 2)
    {
        const float4 ld = normalize4(p - l.position);
        const float cosAngle = dot4(ld, l.direction);
        if (cosAngle > l.cosInner) return l.power * fmaxf(-dot4(ld, N), 0.f);
        if (cosAngle < l.cosOuter) return 0.0f;
        return l.power * (cosAngle - l.cosOuter) / (l.cosInner - l.cosOuter) * fmaxf(-dot4(ld, N), 0);
    }
    return 0.f;
}

// :339-416. cos/sin: the reference calls OpenCL's cos/sin (implementation-defined accuracy). PARITY
// evaluates them in double and rounds once, which is what the oracle restatement does, so both sides
// produce the correctly rounded float in all but ~1e-8 of the calls.
__device__ __forceinline__ void concentric_sample_disk(int &seed, float &dx, float &dy)
{
    const float u1 = rng_next(seed);
    const float u2 = rng_next(seed);
    float r, theta;
    const float sx = 2 * u1 - 1;
    const float sy = 2 * u2 - 1;
    if (fabsf(sx) < 0.0001f) { r = sy; theta = 0; }
    else if (fabsf(sy) < 0.0001f) { r = sx; theta = 2; }
    else if (sx > -sy)
    {
        if (sx > sy) { r = sx; if (sy > 0) theta = sy / sx; else theta = 8.f + sy / sx; }
        else { r = sy; theta = 2.f - sx / sy; }
    }
    else
    {
        if (sx < sy) { r = -sx; theta = 4.f + sy / sx; }
        else { r = -sy; theta = 6.f - sy / sx; }
    }
    theta *= kPi / 4.f;
#if PT_FAST
    r *= 0.999f;
    float sn, cs;
    __sincosf(theta, &sn, &cs);
    dx = r * cs;
    dy = r * sn;
#else
    r = (float)((double)r * 0.999);  // `r *= 0.999;` with a double literal (:408)
    double sn, cs;
    sincos((double)theta, &sn, &cs);
    dx = r * (float)cs;
    dy = r * (float)sn;
#endif
}

// :303-336
__device__ __forceinline__ float4 cosine_sample_hemisphere(int &seed, float4 N)
{
    float x, y, z;
    concentric_sample_disk(seed, x, y);
    z = 1 - x * x - y * y;
    z = (z < 0) ? 0 : sqrtf(z);
    const float4 v = f4(x, y, z, 0);
    if (N.z > 0.9999f) return v;
    if (N.z < -0.9999f) return -v;
    const float4 sn = normalize4(f4(-N.y, N.x, 0, 0));
    const float4 tn = normalize4(cross3(N, sn));
    return normalize4(f4(dot4(f4(sn.x, tn.x, N.x, 0), v), dot4(f4(sn.y, tn.y, N.y, 0), v), dot4(f4(sn.z, tn.z, N.z, 0), v), 0));
}

constexpr int kShadeThreads = 128;

__global__ void __launch_bounds__(kShadeThreads) k_shade(SceneDev sc, FrameParams fp, PathState st, const uint32_t *__restrict__ qin,
                                                        uint32_t *__restrict__ qout, Counters *__restrict__ cn, int bounce)
{
    const uint32_t count = cn->qcount[bounce];
    const uint32_t rounds = (count + blockDim.x * gridDim.x - 1) / (blockDim.x * gridDim.x);
    for (uint32_t it = 0; it < rounds; it++)
    {
        const uint32_t k = (it * gridDim.x + blockIdx.x) * blockDim.x + threadIdx.x;
        bool alive = false;
        uint32_t p = 0;
        if (k < count)
        {
            p = qin[k];
            const float4 o = st.org[p], d = st.dir[p], h = st.hit[p];
            float4 T = st.thr[p], rad = st.rad[p];
            uint32_t meta = st.meta[p];
            const int tri = __float_as_int(h.x);
            if (tri < 0)
            {
                // :1281-1288 sky
                rad = rad + sky_color(sc, d) * T;
                st.rad[p] = rad;
            }
            else
            {
                int seed = st.seed[p];
                const float s = h.y, t = h.z;
                const float4 q = f4(o.x + d.x * h.w, o.y + d.y * h.w, o.z + d.z * h.w, o.w + d.w * h.w);
                const float4 N = __ldg(&sc.tris[tri].n);
                const DevShade *sh = &sc.shade[tri];
                const float nd = dot4(N, d);
                const bool positive = nd < 0;  // :568
                const uint4 ids = __ldg(&sh->ids);
                const DevMaterial mat = sc.materials[positive ? ids.x : ids.y];
                // :591-602 surface colour
                float4 color;
                if (mat.isSimpleColor) color = mat.color;
                else
                {
                    const float4 a = __ldg(&sh->uvp12), b = __ldg(&sh->uvp3_uvn1), c = __ldg(&sh->uvn23);
                    const float w0 = 1 - s - t;
                    float tu, tv;
                    if (positive) { tu = (a.x * w0 + a.z * s) + b.x * t; tv = (a.y * w0 + a.w * s) + b.y * t; }
                    else { tu = (b.z * w0 + c.x * s) + c.z * t; tv = (b.w * w0 + c.y * s) + c.w * t; }
                    const uint4 tx = __ldg(&sc.textures[mat.textureId]);
                    color = texture_lookup(sc, tx.x, tx.y, tx.z, tu, tv);
                }
                // :1267-1274 normals
                const bool same = dot4(d, N) > 0;
                const float4 Ng = same ? -N : N;
                const float4 n1 = __ldg(&sh->n1), n2 = __ldg(&sh->n2), n3 = __ldg(&sh->n3);
                float4 Ns = normalize4((n2 * s) + (n3 * t) + (n1 * (1 - s - t)));
                if (same) Ns = -Ns;
                put_in_same_hemisphere(Ns, -d);
                Ns = normalize4(Ns);
                // :901-954 direct illumination from the occlusion bits the connect stage left
                float4 L = f4(0, 0, 0, 0);
                const uint32_t occl = st.occl[p];
#if PT_FAST
                if (mat.type == 0 || mat.type == 3)
#endif
                for (uint32_t li = 0; li < sc.nLights; li++)
                {
                    const DevLight l = sc.lights[li];
                    const float4 full = (l.type == 0) ? -l.direction : (l.position - q);
                    const float4 ldir = normalize4(full);
                    const float brdf = material_brdf(mat.type, -ldir, Ns, d);
                    if (!((occl >> li) & 1u)) L = L + (light_power_toward(l, q, Ns) * brdf) * l.color;
                }
                // :791-891
                float4 Nref = d, out = d;
                float4 add = f4(0, 0, 0, 0);
                bool inWater = (meta >> 8) & 1u;
                if (mat.type == 0)
                {
                    T = T * color;
                    add = L * T;
                    out = cosine_sample_hemisphere(seed, Ns);
                    Nref = Ns;
                }
                else if (mat.type == 2)
                {
                    float cos2;
                    const float cos1 = -dot4(d, Ns);
                    const float rF = fresnel(1, 1.55f, cos1, cos2);
                    if (rng_next(seed) < rF) { out = d - (Ns * (2 * dot4(d, Ns))); Nref = Ng; }
                    else { T = T * (color * (1 - mat.opacity)); Nref = d; }
                }
                else if (mat.type == 1)
                {
                    const float n1w = inWater ? 1.333f : 1.f, n2w = inWater ? 1.f : 1.333f;
                    float cos2;
                    const float cos1 = -dot4(d, Ns);
                    const float rF = fresnel(n1w, n2w, cos1, cos2);
                    // when totally reflected the reference leaves refractedDirection unset and never reads it
                    const float4 refr = d * (n1w / n2w) + Ns * (n1w / n2w * cos1 - cos2);
                    const float coeff = (n2w * n2w) / (n1w * n1w);
                    if (rng_next(seed) < rF) { out = d - (Ns * (2 * dot4(d, Ns))); Nref = Ng; }
                    else { inWater = !inWater; out = refr; Nref = -Ng; T = T * coeff; }
                }
                else if (mat.type == 3)
                {
                    add = add + L * color * T;
                    const float rF = fresnel_varnish(d, Ns);
                    if (rng_next(seed) < rF) out = d - (Ns * (2 * dot4(d, Ns)));
                    else { out = cosine_sample_hemisphere(seed, Ns); T = T * color; }
                }
                put_in_same_hemisphere(out, Nref);
                const float4 nd4 = normalize4(out);
                const float4 no = q + 0.001f * out;
                rad = rad + add;
                const uint32_t refl = (meta & 0xffu) + 1u;
                meta = refl | (inWater ? 0x100u : 0u);
                st.org[p] = no;
                st.dir[p] = nd4;
                st.thr[p] = T;
                st.rad[p] = rad;
                st.seed[p] = seed;
                st.meta[p] = meta;
                // :1294-1304 and the loop condition :1248
                alive = !(clmax(T.x, clmax(T.y, T.z)) <= 0.001f) && refl < fp.maxDepth;
            }
        }
        // warp-aggregated append of the survivors to the next bounce's queue
        const unsigned m = __ballot_sync(0xffffffffu, alive);
        if (m)
        {
            const unsigned lane = threadIdx.x & 31u;
            uint32_t base = 0;
            if (lane == (unsigned)(__ffs(m) - 1)) base = atomicAdd(&cn->qcount[bounce + 1], (uint32_t)__popc(m));
            base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
            if (alive) qout[base + __popc(m & ((1u << lane) - 1u))] = p;
        }
    }
}

// ------------------------------------------------------------------------------------------------ launchers
void launch_extend(cudaStream_t s, int grid, const SceneDev &sc, const PathState &st, const uint32_t *queue, Counters *cn, int bounce)
{
    k_extend<<<grid, kTraverseThreads, kStackDepth * kTraverseThreads * sizeof(uint32_t), s>>>(sc, st, queue, cn, bounce);
}
void launch_connect(cudaStream_t s, int grid, const SceneDev &sc, const PathState &st, const uint32_t *queue, Counters *cn, int bounce)
{
    k_connect<<<grid, kTraverseThreads, kStackDepth * kTraverseThreads * sizeof(uint32_t), s>>>(sc, st, queue, cn, bounce);
}
void launch_shade(cudaStream_t s, int grid, const SceneDev &sc, const FrameParams &fp, const PathState &st, const uint32_t *qin,
                  uint32_t *qout, Counters *cn, int bounce)
{
    k_shade<<<grid, kShadeThreads, 0, s>>>(sc, fp, st, qin, qout, cn, bounce);
}
int traverse_ctas_per_sm()
{
    int n = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_extend, kTraverseThreads, kStackDepth * kTraverseThreads * sizeof(uint32_t));
    int m = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&m, k_connect, kTraverseThreads, kStackDepth * kTraverseThreads * sizeof(uint32_t));
    n = n < m ? n : m;
    return n < 1 ? 1 : n;
}

}  // namespace PT_NS
}  // namespace ptc
