// pt_wide.cuh — FAST mode only: traversal of the 8-wide compressed BVH that ptc_upload_scene collapses the reference
// BVH2 into (pt_wide_build.cu). The reference has no such structure (Kernel/PathTracer_FullKernel.cl:620-783 walks the
// BVH2 of Controleur/PathTracer_BVH.cpp); this is the "option of widening" of the north star. What stays the reference's:
// the triangle test (tri_test: Triangle_Intersects :519-566, same operations, same order, same acceptance rule), so a
// hit carries the same (s, t, distance) bits as in PARITY mode. What changes is WHICH triangles get tested and in which
// order: boxes are quantised outwards (never smaller than the exact box) and children are visited front to back by
// ray octant, so the closest hit is the one PARITY finds except between candidates at exactly the same distance.
//
// Node layout (80 bytes, after Ylitie, Karras, Laine: "Efficient Incoherent Ray Traversal on GPUs Through Compressed
// Wide BVHs", HPG 2017): child boxes are 8-bit offsets on a per-node grid p + q * 2^e; inner children (imask) are stored
// consecutively from childBase. Unlike the paper there is no per-child meta byte: a leaf child holds at most 3 triangles
// and triangle k of slot s of node i is record 24 * i + 3 * s + k of a sparse triangle array (bit 3 * s + k of `valid`),
// so the 24-bit triangle hit mask is a bit spread of the 8 slot results and needs no per-child shifts.
#pragma once

#include <cuda_fp16.h>

#include "pt_traverse.cuh"

namespace ptc
{
constexpr uint32_t kTriSlots = 24;               // triangle records reserved per wide node
constexpr uint32_t kNoTriangle = 0xffffffffu;    // unused record of the wide triangle order

struct __align__(16) DevNode8
{
    float px, py, pz;       // grid origin = lower corner of the node
    uint32_t eimask;        // bytes 0..2: IEEE biased exponents of the grid step per axis; byte 3: mask of inner children
    uint32_t childBase;     // index of the first inner child
    uint32_t valid;         // bit 3 * slot + k: the leaf child in `slot` has a triangle k
    uint32_t spare[2];
    uint32_t qlo[3][2];     // 8 x 8 bit per axis
    uint32_t qhi[3][2];
};
static_assert(sizeof(DevNode8) == 80, "DevNode8");

// The node as the traversal kernels see it: padded to 96 bytes so that every record is 32-byte aligned and is fetched by
// three 256-bit loads (LDG.E.256) instead of five 128-bit ones — per warp of incoherent rays 96 instead of 160 L1 wavefronts
// per node visit. The builders, the inspection ABI (ptc_wide_bvh) and the tests keep the 80-byte record; ptc_upload_scene
// copies with a pitch.
struct __align__(32) DevNode8P
{
    DevNode8 n;
    uint32_t pad[4];
};
static_assert(sizeof(DevNode8P) == 96, "DevNode8P");

struct WideLane
{
    Ray r;             // ix, iy, iz clamped to finite values
    float best;        // squared distance bound, the reference's `squaredDistance`
    float tmax;        // the same bound as a ray parameter, padded
    uint32_t gx, gy;   // node group: first child index | bits 31..24 children still to visit (front to back), bits 7..0 imask
    uint32_t tx, ty;   // triangle group: first record (24 * node) | 24 bits of triangles still to test
    uint32_t oct;      // flag word — bits 2..0: 7 - ray octant; 9..8: parked triangle groups; kStrict, kSingle; 15..12: MaterialType of the hit side
    int sp;            // stack entries, -1 = no ray
    int hitTri;        // wide record index of the accepted hit, -1 = none; (s, t, parameter) go to the path's hit record when the hit is
                       // accepted (kStoresHit) instead of riding in three registers through every step
    uint32_t nb, nt;
    bool found;

    typedef uint2 StackT;
    // The lane's stack as a 32-bit shared-window address, made opaque to the compiler once: left to itself it re-derives the
    // address (S2UR CgaCtaId, ULEA, S2R tid, IMAD: 7 instructions) in every iteration of the node loop rather than keep a register.
    struct StackRef
    {
        uint32_t a;
        __device__ __forceinline__ uint2 ld(int level, int stride) const
        {
            uint2 v;
            asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a + (uint32_t)(level * stride) * 8u));
            return v;
        }
        __device__ __forceinline__ void st(int level, int stride, uint2 v) const
        {
            asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(a + (uint32_t)(level * stride) * 8u), "r"(v.x), "r"(v.y));
        }
        // level 0 holds the item (path slot | light) the lane is tracing: cold state that is read when a hit is accepted and at
        // completion — one LDS off the register that addresses the stack, where &stage.laneItem[lane] took six instructions to form
        __device__ __forceinline__ uint32_t item() const
        {
            uint32_t v;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
            return v;
        }
        __device__ __forceinline__ void set_item(uint32_t v) const { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v)); }
    };
    static constexpr int kItemLevels = 1;  // stack levels below the parked groups
    static __device__ __forceinline__ StackRef stack_ref(StackT *p)
    {
        StackRef r;
        r.a = (uint32_t)__cvta_generic_to_shared(p);
        asm volatile("" : "+r"(r.a));
        return r;
    }
    static constexpr int kBoxesPerNodeStep = 8;
#ifndef PTC_WIDE_MIN_CTAS
#define PTC_WIDE_MIN_CTAS 8
#endif
    static constexpr int kMinCtas = PTC_WIDE_MIN_CTAS;  // 64 registers: 8 CTAs of 4 warps per SM
#ifndef PTC_WIDE_MIN_CTAS_ANY
#define PTC_WIDE_MIN_CTAS_ANY 7
#endif
    // connect: 72 registers, 7 CTAs per SM — its staging holds a light record and a shading normal besides the lane state; at 64
    // registers it spilled 16-24 B in those cold paths. Spill-free it measures +0.3 % (165.0 vs 164.5 Msamples/s, profiles/r02_sweeps.txt)
    static constexpr int kMinCtasAny = PTC_WIDE_MIN_CTAS_ANY;
    __device__ __forceinline__ void reset() { sp = -1; gy = ty = 0; nb = nt = 0; found = false; hitTri = -1; }
    __device__ __forceinline__ bool idle() const { return sp < 0; }
    __device__ __forceinline__ void set_idle() { sp = -1; }
    // Parked triangle groups: a lane may take a node step while it still holds untested triangles — it parks the group in one of
    // the kPark lowest levels of its shared-memory stack (bits 9..8 of oct count them) and tests it in a later triangle phase.
#ifndef PTC_PARK_SLOTS
#define PTC_PARK_SLOTS 1
#endif
    static constexpr int kPark = PTC_PARK_SLOTS;
    static_assert(kPark >= 1 && kPark <= 3, "two bits count the parked groups");
    static constexpr uint32_t kParkedOne = 0x100u, kParkedMask = 0x300u;
    // The reference's 4-lane arithmetic lets a refracted direction keep a w lane != 0 and every later origin of that path a
    // w lane != 1. For such a ray the "distance" a triangle is accepted with is no longer the parameter at which
    // the ray crosses the triangle's box, so which candidates survive the distance cull depends on the order of the tests. These
    // rays never park triangles: node step, all of its triangles, next node step — an order that is a function of the ray alone
    // (with parking it followed the warp's votes: about 40 of 2 M pixels of the bench scene changed from run to run).
    static constexpr uint32_t kStrict = 0x400u;
    // A ray (nearly) parallel to a coordinate plane — one reciprocal direction component 2^14 times another — takes the single-precision
    // slab test: the packed half-precision one scales a node's times by its coarsest axis and has no resolution left for the others
    // then (it stays conservative, but such a ray would enter whole slabs of the tree: a camera looking along an axis shoots a column
    // of them). About 2 in 10^4 of random directions, besides the structural ones.
    static constexpr uint32_t kSingle = 0x800u;
    // extend: (s, t, parameter) of an accepted hit go straight to the path's hit record (one 16-byte store per acceptance, the last one
    // stands) and the MaterialType of the side hit rides in bits 15..12 of oct; completion adds the triangle index. Before, the accepted
    // test was repeated at completion — 87 instructions with 7 of 32 lanes active, 4.6 % of extend's warp instructions (ncu, round 2).
    static constexpr bool kStoresHit = true;
    static constexpr uint32_t kMatShift = 12;
    __device__ __forceinline__ bool finished() const { return sp == 0 && ty == 0u && gy <= 0x00ffffffu && !(oct & kParkedMask); }
    // A lane can take a node step while it still holds untested triangles: it parks that group (once) and tests it in the
    // next triangle phase. So a lane often qualifies for both phases and follows the warp's vote instead of idling.
    __device__ __forceinline__ bool wants_node() const { return sp >= 0 && gy > 0x00ffffffu && (ty == 0u || (!(oct & kStrict) && (oct & kParkedMask) < kPark * kParkedOne)); }
    __device__ __forceinline__ bool wants_tri() const { return sp >= 0 && (ty != 0u || (oct & kParkedMask)); }
    // BVH-order index of the hit (what the shade stage and the reference's ids use), -1 = miss
    __device__ __forceinline__ int hit_index(const SceneDev &sc) const { return hitTri < 0 ? -1 : (int)__ldg(&sc.tri8Index[hitTri]); }
    // (s, t, ray parameter, MaterialType of the side hit) of the final hit by repeating the accepted test — the same operations on the
    // same operands, hence the same bits (what completion did before kStoresHit; kept for callers that hold only the lane)
    __device__ __forceinline__ void hit_record(const SceneDev &sc, float &hs, float &ht, float &htp, uint32_t &hmat) const
    {
        hs = ht = htp = 0.f;
        hmat = 0u;
        if (hitTri < 0) return;
        float inf = INFINITY;
        tri_test(sc.tris8, (uint32_t)hitTri, r, inf, hs, ht, htp, hmat);
    }
};

__device__ __forceinline__ float finite_inverse(float inv, float d)
{
    return fabsf(inv) < 1.0e30f ? inv : copysignf(1.0e30f, d);  // d = +-0 or denormal: a huge slope instead of inf (no 0 * inf)
}

// The triangle test accepts by the SQUARED DISTANCE |q - o|^2 of the hit point q = o + d * t as the reference computes it
// (Triangle_Intersects :536-543), the boxes are culled by the ray PARAMETER. The two disagree by the rounding of q: about one
// ulp of the coordinates, whatever t is, so for a hit very close to a far-from-zero origin the squared distances of two
// candidates need not be ordered like their parameters. The cull bound therefore carries an absolute pad of 2^-20 of the
// origin's coordinates besides the relative one: a candidate that the distance test could still accept is never culled,
// whichever of the two was tested first — the hit does not depend on the order in which a warp's votes make a lane test
// its triangles (without the pad about one ray in 10^6 did, run to run).
__device__ __forceinline__ float cull_pad(const Ray &r) { return (fabsf(r.ox) + fabsf(r.oy) + fabsf(r.oz)) * 9.5367431640625e-7f; }
__device__ __forceinline__ float bound_as_parameter(float t, const Ray &r)
{
    return __fmaf_rn(t, 1.00001f, cull_pad(r));  // directions are unit length
}

__device__ __forceinline__ void lane_begin(WideLane &L, const SceneDev &sc, const Ray &r, float best)
{
    L.r = r;
    L.r.ix = finite_inverse(r.ix, r.dx);
    L.r.iy = finite_inverse(r.iy, r.dy);
    L.r.iz = finite_inverse(r.iz, r.dz);
    L.best = best;
    // an unbounded ray (every extension ray, shadow rays of directional lights) skips the IEEE square root: sqrtf(INFINITY) is a
    // subroutine call, not an instruction
    L.tmax = best < 3.0e38f ? bound_as_parameter(sqrtf(best), r) : INFINITY;
    const uint32_t oct = (r.dx < 0 ? 4u : 0u) | (r.dy < 0 ? 2u : 0u) | (r.dz < 0 ? 1u : 0u);
    L.oct = (7u - oct) | ((r.dw == 0.f && r.ow == 1.f) ? 0u : WideLane::kStrict);
    const float ax = fabsf(L.r.ix), ay = fabsf(L.r.iy), az = fabsf(L.r.iz);
    if (!(fmaxf(fmaxf(ax, ay), az) < 16384.f * fminf(fminf(ax, ay), az))) L.oct |= WideLane::kSingle;
    L.gx = 0u;
    L.gy = sc.nNodes8 ? 0x80000000u : 0u;  // "slot 7 ^ octinv of an imask-less group at index 0" = the root
    L.tx = L.ty = 0u;
    L.sp = 0;
    L.hitTri = -1;
    L.found = false;
}

// no inner child left in the current node group: continue with the most recently postponed siblings. Stack level 0 of a
// lane holds its item word, the next kPark levels are the slots of its parked triangle groups, sibling groups live in the sp levels above them.
__device__ __forceinline__ void wide_normalise(WideLane &L, const WideLane::StackRef stk, const int stride)
{
    const bool pop = L.gy <= 0x00ffffffu && L.sp > 0;
    if (pop)
    {
        const uint2 e = stk.ld(L.sp + WideLane::kItemLevels + WideLane::kPark - 1, stride);
        L.gx = e.x;
        L.gy = e.y;
    }
    L.sp -= pop ? 1 : 0;
}

// Quantised planes without the conversion pipe: a byte q placed under the half-precision exponent 0x64 is the half 1024 + q
// (exact), two bytes per PRMT; half -> float is exact as well and runs beside the FMAs. The 1024 is folded into the constant
// term of t(q) = q * a + b. 48 I2F.U8 per node visit (quarter rate, the busiest pipe of round 1's captures: XU 66-72 %) became
// 24 PRMT + 48 HADD2.F32: +1.4 % samples/s on the 262 k-triangle atrium, +1.8 % on the 10 M-triangle terrain, +11 % on the
// 1 M-triangle atrium (profiles/r02_sweeps.txt).
__device__ __forceinline__ void q8x4(uint32_t w, float &f0, float &f1, float &f2, float &f3)
{
    const uint32_t h01 = __byte_perm(w, 0x64646464u, 0x4140), h23 = __byte_perm(w, 0x64646464u, 0x4342);
    const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&h01)), b = __half22float2(*reinterpret_cast<const __half2 *>(&h23));
    f0 = a.x; f1 = a.y; f2 = b.x; f3 = b.y;
}

#ifndef PTC_WIDE_HALF2
#define PTC_WIDE_HALF2 1
#endif

// The 8 slab tests of a node visit in packed half precision: two children per instruction. A byte q over a zero byte IS the
// (denormal) half q * 2^-24, so one PRMT turns two quantised planes into an HFMA2 operand; per pair of children 6 HFMA2 give
// the six plane times, 2 three-input + 2 two-input packed min / max the interval, one HSET2 the verdict: 85 instructions for
// the 8 boxes where the single-precision version (below, PTC_WIDE_HALF2=0) issues 168 (48 HADD2.F32 + 48 FFMA + 32 FMNMX + ...).
//
// Half precision has 11 significant bits and a small exponent range, so times are first made node-relative and node-sized:
// tau = (t - t0) * sigma, t0 = the time the ray enters the node's own box (0 if it starts inside), sigma the power of two that
// brings the coarsest axis' time per grid step to [2^-10, 2^-9). A child the ray really crosses lies inside the node, so its
// times are |tau| <= 255 steps of the FINEST axis: half's relative precision is then 1/8 of a grid step at worst.
//
// CONSERVATIVE — never misses a box the exact test enters — without an error analysis of the HFMA2 results themselves: the
// fused multiply-add rounds the exact q * A + B once, and rounding is monotonic, so x <= y implies rn(x) <= rn(y). It is
// therefore enough that the OPERANDS err outwards: the entry constants are lowered and the exit constants raised by
//   |B| * 2^-11 (their own rounding to half) + 2^-24 (a denormal result's absolute rounding)
//   + 0.126 step (q * (A - rn_half(A)), q <= 255) + 2^-21 |b| (the single-precision rounding of (p - o) / d, as the float version)
// i.e. boxes grow by at most a quarter of a grid step per side on top of the <= 1 step of the 8-bit quantisation. Overflow to
// +-inf errs outwards as well (a crossed child's times are node-sized, see above); a NaN (0 * inf cannot occur: A is finite by
// the choice of sigma) only arises when t0 itself is infinite, i.e. the ray never enters the node.
__device__ __forceinline__ __half2 h2bits(uint32_t u) { return *reinterpret_cast<const __half2 *>(&u); }

__device__ __forceinline__ uint32_t slab_test_half2(const WideLane &L, const uint4 &n0, const uint4 &n2, const uint4 &n3, const uint4 &n4)
{
    // single precision: time per grid step a = 2^e / d and time of plane 0 b = (p - o) / d, per axis
    const float adx = __uint_as_float((n0.w & 0xffu) << 23) * L.r.ix;
    const float ady = __uint_as_float(((n0.w >> 8) & 0xffu) << 23) * L.r.iy;
    const float adz = __uint_as_float(((n0.w >> 16) & 0xffu) << 23) * L.r.iz;
    const float bx = (__uint_as_float(n0.x) - L.r.ox) * L.r.ix;
    const float by = (__uint_as_float(n0.y) - L.r.oy) * L.r.iy;
    const float bz = (__uint_as_float(n0.z) - L.r.oz) * L.r.iz;
    // entry into the node's own box [plane 0, plane 255]^3, not before the origin
    const float t0 = fmaxf(fmaxf(fminf(bx, __fmaf_rn(255.f, adx, bx)), fminf(by, __fmaf_rn(255.f, ady, by))),
                           fmaxf(fminf(bz, __fmaf_rn(255.f, adz, bz)), 0.f));
    // sigma from the exponent of the coarsest axis: A = a * 2^(141 - E) in [2^14, 2^15) multiplies q * 2^-24
    const float amax = fmaxf(fmaxf(fabsf(adx), fabsf(ady)), fabsf(adz));
    const uint32_t E = min(max(__float_as_uint(amax) >> 23, 14u), 240u);
    const float sA = __uint_as_float((268u - E) << 23), sT = __uint_as_float((244u - E) << 23);
    const float Ax = adx * sA, Ay = ady * sA, Az = adz * sA;
    const float Bx = (bx - t0) * sT, By = (by - t0) * sT, Bz = (bz - t0) * sT;
    // the pad (see above); |b| <= |B| / sigma + t0, so the 2^-21 |b| term splits into a share of |B|'s and one of the node's
    const float kHalfUlp = 4.96e-4f, kStep = 7.52e-9f /* 0.126 * 2^-24 */, kDenorm = 5.9604645e-8f /* 2^-24 */;
    const float pn = __fmaf_rn(t0 * sT, 4.76837158e-7f, kDenorm);
    const float px = __fmaf_rn(fabsf(Bx), kHalfUlp, __fmaf_rn(fabsf(Ax), kStep, pn));
    const float py = __fmaf_rn(fabsf(By), kHalfUlp, __fmaf_rn(fabsf(Ay), kStep, pn));
    const float pz = __fmaf_rn(fabsf(Bz), kHalfUlp, __fmaf_rn(fabsf(Az), kStep, pn));
    // the ray's own interval [0, tmax] in node time, erring outwards (t0 >= 0, so tau(0) <= 0)
    const float tb = __fmaf_rn(t0 * sT, -1.001f, -kDenorm);
    const float te0 = (L.tmax - t0) * sT, te = __fmaf_rn(fabsf(te0), 1.0e-3f, te0) + kDenorm;
    // eleven constants in six registers: HFMA2 / HMNMX2 broadcast either half of an operand for free (.H0_H0 / .H1_H1)
    const __half2 hBx = __floats2half2_rn(Bx - px, Bx + px), hBy = __floats2half2_rn(By - py, By + py), hBz = __floats2half2_rn(Bz - pz, Bz + pz);
    const __half2 hAxy = __floats2half2_rn(Ax, Ay), hAzT = __floats2half2_rn(Az, tb), hTe = __floats2half2_rn(te, te);
    const __half2 hAx = __low2half2(hAxy), hAy = __high2half2(hAxy), hAz = __low2half2(hAzT), hT0 = __high2half2(hAzT), hTM = hTe;
    const __half2 hBlx = __low2half2(hBx), hBhx = __high2half2(hBx), hBly = __low2half2(hBy), hBhy = __high2half2(hBy);
    const __half2 hBlz = __low2half2(hBz), hBhz = __high2half2(hBz);
    // entry planes are the low ones along axes the ray travels up, the high ones otherwise
    const bool nx = L.r.ix < 0, ny = L.r.iy < 0, nz = L.r.iz < 0;
    // words: n2 = qlo.x[0] qlo.x[1] qlo.y[0] qlo.y[1]; n3 = qlo.z[0] qlo.z[1] qhi.x[0] qhi.x[1]; n4 = qhi.y[0] qhi.y[1] qhi.z[0] qhi.z[1]
    uint32_t acc = 0;  // child c = 2k: bit c; child c = 2k + 1: bit 16 + c
#pragma unroll
    for (int h = 0; h < 2; h++)
    {
        const uint32_t lox = h ? n2.y : n2.x, loy = h ? n2.w : n2.z, loz = h ? n3.y : n3.x;
        const uint32_t hix = h ? n3.w : n3.z, hiy = h ? n4.y : n4.x, hiz = h ? n4.w : n4.z;
        const uint32_t ex = nx ? hix : lox, fx = nx ? lox : hix;
        const uint32_t ey = ny ? hiy : loy, fy = ny ? loy : hiy;
        const uint32_t ez = nz ? hiz : loz, fz = nz ? loz : hiz;
#pragma unroll
        for (int k = 0; k < 2; k++)
        {
            const uint32_t sel = k ? 0x4342u : 0x4140u;  // bytes (2k, 2k + 1) -> the low bytes of the two halves, zero above
            const __half2 tex = __hfma2(h2bits(__byte_perm(ex, 0u, sel)), hAx, hBlx), tfx = __hfma2(h2bits(__byte_perm(fx, 0u, sel)), hAx, hBhx);
            const __half2 tey = __hfma2(h2bits(__byte_perm(ey, 0u, sel)), hAy, hBly), tfy = __hfma2(h2bits(__byte_perm(fy, 0u, sel)), hAy, hBhy);
            const __half2 tez = __hfma2(h2bits(__byte_perm(ez, 0u, sel)), hAz, hBlz), tfz = __hfma2(h2bits(__byte_perm(fz, 0u, sel)), hAz, hBhz);
            const __half2 tEnter = __hmax2(__hmax2(tex, tey), __hmax2(tez, hT0));
            const __half2 tExit = __hmin2(__hmin2(tfx, tfy), __hmin2(tfz, hTM));
            const int c = 4 * h + 2 * k;
            acc |= __hle2_mask(tEnter, tExit) & ((1u << c) | (1u << (c + 17)));
        }
    }
    return (acc | (acc >> 16)) & 0xffu;
}

// The 8 slab tests in single precision: what every ray took before the packed version above, and what rays flagged kSingle take.
__device__ __forceinline__ uint32_t slab_test_single(const WideLane &L, const uint4 &n0, const uint4 &n2, const uint4 &n3, const uint4 &n4)
{
    // the ray in the node's grid: t(q) = q * (2^e / d) + (p - o) / d, the constant term padded outwards by 2^-21
    const float adx = __uint_as_float((n0.w & 0xffu) << 23) * L.r.ix;
    const float ady = __uint_as_float(((n0.w >> 8) & 0xffu) << 23) * L.r.iy;
    const float adz = __uint_as_float(((n0.w >> 16) & 0xffu) << 23) * L.r.iz;
    const float bx = (__uint_as_float(n0.x) - L.r.ox) * L.r.ix;
    const float by = (__uint_as_float(n0.y) - L.r.oy) * L.r.iy;
    const float bz = (__uint_as_float(n0.z) - L.r.oz) * L.r.iz;
    const float kPad = 4.76837158e-7f;
    // planes arrive as 1024 + q: the constant term takes the -1024 * a and is padded after that
    const float cx = __fmaf_rn(-1024.f, adx, bx), cy = __fmaf_rn(-1024.f, ady, by), cz = __fmaf_rn(-1024.f, adz, bz);
    const float bxl = __fmaf_rn(-fabsf(cx), kPad, cx), bxh = __fmaf_rn(fabsf(cx), kPad, cx);
    const float byl = __fmaf_rn(-fabsf(cy), kPad, cy), byh = __fmaf_rn(fabsf(cy), kPad, cy);
    const float bzl = __fmaf_rn(-fabsf(cz), kPad, cz), bzh = __fmaf_rn(fabsf(cz), kPad, cz);
    // entry planes are the low ones along axes the ray travels up, the high ones otherwise
    const bool nx = L.r.ix < 0, ny = L.r.iy < 0, nz = L.r.iz < 0;
    // words: n2 = qlo.x[0] qlo.x[1] qlo.y[0] qlo.y[1]; n3 = qlo.z[0] qlo.z[1] qhi.x[0] qhi.x[1]; n4 = qhi.y[0] qhi.y[1] qhi.z[0] qhi.z[1]
    uint32_t slotHit = 0;  // bit s: the box in slot s is hit (empty slots may "hit": imask and valid drop them)
#pragma unroll
    for (int h = 0; h < 2; h++)
    {
        const uint32_t lox = h ? n2.y : n2.x, loy = h ? n2.w : n2.z, loz = h ? n3.y : n3.x;
        const uint32_t hix = h ? n3.w : n3.z, hiy = h ? n4.y : n4.x, hiz = h ? n4.w : n4.z;
        const uint32_t ex = nx ? hix : lox, fx = nx ? lox : hix;
        const uint32_t ey = ny ? hiy : loy, fy = ny ? loy : hiy;
        const uint32_t ez = nz ? hiz : loz, fz = nz ? loz : hiz;
        float qex[4], qfx[4], qey[4], qfy[4], qez[4], qfz[4];
        q8x4(ex, qex[0], qex[1], qex[2], qex[3]); q8x4(fx, qfx[0], qfx[1], qfx[2], qfx[3]);
        q8x4(ey, qey[0], qey[1], qey[2], qey[3]); q8x4(fy, qfy[0], qfy[1], qfy[2], qfy[3]);
        q8x4(ez, qez[0], qez[1], qez[2], qez[3]); q8x4(fz, qfz[0], qfz[1], qfz[2], qfz[3]);
#pragma unroll
        for (int j = 0; j < 4; j++)
        {
            const float tex = __fmaf_rn(qex[j], adx, bxl), tfx = __fmaf_rn(qfx[j], adx, bxh);
            const float tey = __fmaf_rn(qey[j], ady, byl), tfy = __fmaf_rn(qfy[j], ady, byh);
            const float tez = __fmaf_rn(qez[j], adz, bzl), tfz = __fmaf_rn(qfz[j], adz, bzh);
            const float tEnter = fmaxf(fmaxf(tex, tey), fmaxf(tez, 0.f));
            const float tExit = fminf(fminf(tfx, tfy), fminf(tfz, L.tmax));
            if (tEnter <= tExit) slotHit |= 1u << (4 * h + j);
        }
    }
    return slotHit;
}

__device__ __forceinline__ uint32_t lane_item(const WideLane::StackRef stk, const uint32_t *) { return stk.item(); }
__device__ __forceinline__ void lane_set_item(const WideLane::StackRef stk, uint32_t *, uint32_t v) { stk.set_item(v); }

// One node visit: take the nearest unvisited child of the current group, postpone its siblings, test its 8 children.
template <bool FAST>
__device__ __forceinline__ void lane_node_step(WideLane &L, const SceneDev &sc, const WideLane::StackRef stk, const int stride)
{
    const uint32_t hits = L.gy;
    const int bit = 31 - __clz((int)hits);
    const uint32_t slot = (uint32_t)(bit - 24) ^ (L.oct & 7u);
    const uint32_t rest = hits & ~(1u << bit);
    const uint32_t rel = (uint32_t)__popc(hits & 0xffu & ((1u << slot) - 1u));
    const uint32_t node = L.gx + rel;
    PTC_CHECK(sc.faultSink, kFaultNodeIndex, node, sc.nNodes8);
    const char *np = reinterpret_cast<const char *>(sc.nodes8 + node);
    uint4 n0, n1, n2, n3, n4, n5;
    ldg256(np, reinterpret_cast<float4 &>(n0), reinterpret_cast<float4 &>(n1));
    ldg256(np + 32, reinterpret_cast<float4 &>(n2), reinterpret_cast<float4 &>(n3));
    ldg256(np + 64, reinterpret_cast<float4 &>(n4), reinterpret_cast<float4 &>(n5));  // n5 = padding
    if (L.ty != 0u)
    {
        stk.st(WideLane::kItemLevels + (WideLane::kPark == 1 ? 0 : (int)((L.oct >> 8) & 3u)), stride, make_uint2(L.tx, L.ty));  // park the untested triangles (wants_node: a slot is free)
        L.oct += WideLane::kParkedOne;
    }
    if (rest > 0x00ffffffu)
    {
        PTC_CHECK(sc.faultSink, kFaultStack, L.sp + WideLane::kItemLevels + WideLane::kPark, sc.stackDepth);
        stk.st(L.sp + WideLane::kItemLevels + WideLane::kPark, stride, make_uint2(L.gx, rest));
    }
    L.sp += rest > 0x00ffffffu ? 1 : 0;
    L.nb += 8;

#if PTC_WIDE_HALF2
    const uint32_t slotHit = (L.oct & WideLane::kSingle) ? slab_test_single(L, n0, n2, n3, n4) : slab_test_half2(L, n0, n2, n3, n4);
#else
    const uint32_t slotHit = slab_test_single(L, n0, n2, n3, n4);
#endif
    // inner children: the bit of slot s goes to position 24 + (s ^ octinv), so that "highest bit first" is front to back
    const uint32_t imask = n0.w >> 24;
    uint32_t ih = slotHit & imask;
    const uint32_t oi = L.oct;
    if (oi & 1u) ih = ((ih & 0x55u) << 1) | ((ih >> 1) & 0x55u);
    if (oi & 2u) ih = ((ih & 0x33u) << 2) | ((ih >> 2) & 0x33u);
    if (oi & 4u) ih = ((ih & 0x0fu) << 4) | (ih >> 4);
    // triangles: bit s -> bits 3s..3s+2 (the multiplications only add disjoint bit sets), then keep the ones that exist
    uint32_t th = slotHit;
    th = (th * 0x101u) & 0x0000f00fu;
    th = (th * 0x11u) & 0x000c30c3u;
    th = (th * 0x5u) & 0x00249249u;
    th = (th * 7u) & n1.y;
    L.tx = node * kTriSlots;
    L.ty = th;
    L.gx = n1.x;
    L.gy = (ih << 24) | imask;
    wide_normalise(L, stk, stride);
}

// One triangle test. Nothing after the test branches: the accepted hit is kept as a record index of the wide order (hit_index
// translates it once per ray) and the new bound follows from the hit's own ray parameter (directions are unit length over
// the four lanes the reference normalises, so parameter and distance coincide).
template <bool ANY>
__device__ __forceinline__ void lane_tri_step(WideLane &L, const SceneDev &sc, const WideLane::StackRef stk, const int stride, float4 *__restrict__ hitOut)
{
    if (L.ty == 0u)
    {
        L.oct -= WideLane::kParkedOne;
        const uint2 e = stk.ld(WideLane::kItemLevels + (WideLane::kPark == 1 ? 0 : (int)((L.oct >> 8) & 3u)), stride);  // the group parked last
        L.tx = e.x;
        L.ty = e.y;
    }
    const int bit = __ffs((int)L.ty) - 1;
    L.ty &= L.ty - 1u;
    const uint32_t i = L.tx + (uint32_t)bit;
    PTC_CHECK(sc.faultSink, kFaultTriIndex, i, sc.nNodes8 * kTriSlots);
    L.nt++;
    float hs, ht, htp = 0.f;
    uint32_t hmat;
    const bool acc = tri_test(sc.tris8, i, L.r, L.best, hs, ht, htp, hmat, ANY ? nullptr : sc.tri8Index, L.hitTri);
    L.hitTri = acc ? (int)i : L.hitTri;
    L.tmax = acc ? bound_as_parameter(htp, L.r) : L.tmax;
    if (!ANY && acc)
    {
        hitOut[stk.item()] = make_float4(0.f, hs, ht, htp);  // .x = the triangle index, written at completion
        L.oct = (L.oct & ~(15u << WideLane::kMatShift)) | (hmat << WideLane::kMatShift);
    }
    if (ANY)
    {
        L.found = L.found || acc;
        L.gy = acc ? 0u : L.gy;
        L.ty = acc ? 0u : L.ty;
        L.sp = acc ? 0 : L.sp;
        L.oct = acc ? (L.oct & 7u) : L.oct;
    }
    wide_normalise(L, stk, stride);
}

}  // namespace ptc
