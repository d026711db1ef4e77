// traverse.cuh -- closest-hit traversal of the 4-wide BVH with the reference's leaf arithmetic.
//
// Stands where hit_scene -> hit_bvh -> {AABB::hit, hit_sphere, hit_plane} stand in the reference
// (include/scene.h:23-35, include/bvh.h:19-65, include/aabb.h:42-65, include/sphere.h:24-53,
// include/plane.h:57-96).  Differences by design:
//   * nodes are 128-byte lines fetched as 7 x 128-bit loads through the L1 (__ldg); staging the top levels in shared memory
//     is implemented (smem_top.cuh) and measured slower, so the production launch stages nothing;
//   * children are visited near-to-far (the reference's order is undefined) with a per-thread stack of packed
//     (entry distance, ref) pairs in LOCAL memory (L1-cached; a shared-memory front part is compiled out, RTB_SMEM_STACK = 0);
//   * the slab test uses a per-ray reciprocal instead of 3 divides per node (include/aabb.h:44);
//   * leaf tests only produce (t, primitive); normals / uv are derived once, after traversal;
//   * acceptance follows the reference exactly: tmin <= t <= closest (Interval::contains, include/interval.h:16).
//     Exact-t ties resolve to the primitive the reference's linear scan (include/scene.h:37-51) would keep: the
//     later one in (spheres, then planes) array order.
#pragma once
#include "device_math.cuh"
#include "leaf_size.h"

namespace rtb {

struct DevScene {
    const float4* __restrict__ nodes;        // WideNode[ ] as 8 x float4
    const float4* __restrict__ spheres;      // re-ordered (cx,cy,cz,r)
    const int2* __restrict__ sphere_meta;    // (material_idx, caller's index)
    const float4* __restrict__ planes;       // re-ordered, 5 x float4: (n,D) (w,type) (u,mat) (v,caller's index) (base,0)
    const float4* __restrict__ planes_fast;  // same order, 3 x float4: (n,D) (A,a0|type) (B,b0) -- fast-mode leaf test (lbvh.cuh)
    const float4* __restrict__ mats;         // 3 x float4: (albedo,type) (emit,fuzz) (absorption,ir)
    const float4* __restrict__ tex;          // RGBA texels, row-major (RT_TEX_CPU lookups)
    cudaTextureObject_t tex_obj;             // RT_TEX_GPU lookups
    int tex_w, tex_h, tex_mat;
    int num_nodes;
    int top_nodes;                           // wide nodes [0, top_nodes) are staged in shared memory
    int num_hoisted;                         // primitives kept out of the tree (wide_bvh.h): tested once per ray in trav_begin
    int hoisted[4];                          // their leaf references
    // uniform grid over a dense field of small spheres (scene_pack.h GridAccel); grid_on = 0: everything is in the tree
    int grid_on;
    const uint2* __restrict__ grid_cells;    // (first ref, count) per cell, x fastest
    const int* __restrict__ grid_refs;       // leaf references of the spheres overlapping a cell
    float grid_lo[3], grid_hi[3], grid_cw[3], grid_icw[3];
    int grid_n[3];
    float tree_lo[3], tree_hi[3];            // bounds of what the tree holds: a ray that misses them skips the tree
};

#ifndef RTB_SMEM_STACK
#define RTB_SMEM_STACK 0
#endif
// The per-thread traversal stack lives in LOCAL memory (thread-interleaved, L1-cached, one 64-bit LDL/STL per entry with
// an immediate-offset address).  A shared-memory front part (RTB_SMEM_STACK > 0 entries, strided by the block size) was
// the first design; measured on B200 it is 10 % SLOWER than all-local (two LDS/STS + address math + a shared/local branch
// per push and pop, and it takes L1 capacity away from the nodes), so the default is 0.  The builder guarantees
// max_stack <= 64.
#ifndef RTB_FAST_RCP
#define RTB_FAST_RCP 1
#endif
// RTB_FAST_PLANE=1: fast mode tests planes on the two-dot-product record (plane_fast_t) and hands only the hits near the
// polygon's outline to the reference-order test; 0: always the reference-order test (plane_t).
#ifndef RTB_FAST_PLANE
#define RTB_FAST_PLANE 1
#endif
#ifndef RTB_FAST_SLAB
#define RTB_FAST_SLAB 1
#endif
#ifndef RTB_FAST_SQRT
#define RTB_FAST_SQRT 1
#endif
#ifndef RTB_SORT_PARTIAL
#define RTB_SORT_PARTIAL 1
#endif
constexpr int kSmemStack = RTB_SMEM_STACK;
constexpr int kLocalStack = 64 - kSmemStack;
constexpr int kTopStride = 9;    // float4 per staged node: 8 + 1 pad (144 B) so 128-bit LDS spread over banks
constexpr int32_t kEmptyChildDev = 0x7fffffff;

struct HitInfo {
    float t;
    int type;   // -1 miss, 0 sphere, 1 plane
    int slot;   // position in the re-ordered array
    int orig;   // caller's index
};

struct WorkCount {
    unsigned int segments, nodes, spheres, planes, hits, samples;
};

// ---- leaf tests ---------------------------------------------------------------------------------------------
// sqrt of the discriminant: IEEE in exact mode (the reference's sqrtf, include/sphere.h:33); in fast mode the 2-instruction
// approximation (2 ulp) instead of the ~10-instruction IEEE sequence -- measured 201.6 vs 203.1 ms on the headline frame
template <bool X>
__device__ __forceinline__ float leaf_sqrt(float D) {
#if RTB_FAST_SQRT && defined(__CUDA_ARCH__)
    if (!X) {
        float s;
        asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(s) : "f"(D));
        return s;
    }
#endif
    return __fsqrt_rn(D);
}

// include/sphere.h:24-47.  a = |d|^2 and (fast mode) 1/a are per-ray constants.
//
// BOTH policies form the discriminant exactly as the reference does -- oc, hb, c and D = hb*hb - a*c as separately rounded
// fp32 operations in its order.  D is the difference of two numbers of size |oc|^2 |d|^2 that agree to all but the last
// few bits (r << |oc|): its value, and for a grazing ray its SIGN, is decided by how the products round.  An FMA-contracted
// evaluation is no less accurate, but it is a different answer: measured on the 10k-sphere 1080p frame it moved t by up
// to 1.8e-4 relative and changed the hit primitive of 2e-4 of the primary rays against the reference (north_star's gates
// are 1e-5 and "ids equal").  What the fast policy keeps is everything after D: the 2-ulp SFU square root and the
// reciprocal multiply instead of the IEEE square root and the double-precision division (relative effect on t: ~3e-7).
template <bool X>
__device__ __forceinline__ bool sphere_t(V3 o, V3 d, float a, float inv_a, float4 s, float tmin, float tmax, float& t_out) {
#ifndef RTB_FAST_SPHERE_D
#define RTB_FAST_SPHERE_D 0
#endif
    constexpr bool XD = X || !RTB_FAST_SPHERE_D;  // arithmetic policy of the discriminant
    V3 oc = sub3<XD>(o, mk3(s.x, s.y, s.z));
    float hb = dot3<XD>(oc, d);
    float cc = fsub<XD>(len2_3<XD>(oc), fmul<XD>(s.w, s.w));
    float D = fsub<XD>(fmul<XD>(hb, hb), fmul<XD>(a, cc));
    if (D < 0.0f) return false;
    float sq = leaf_sqrt<X>(D);
    float t;
    if (X) {
        t = (float)(((double)(-hb) - (double)sq) / (double)a);
        if (!(tmin <= t && t <= tmax)) {
            t = (float)(((double)(-hb) + (double)sq) / (double)a);
            if (!(tmin <= t && t <= tmax)) return false;
        }
    } else {
        t = (-hb - sq) * inv_a;
        if (!(tmin <= t && t <= tmax)) {
            t = (-hb + sq) * inv_a;
            if (!(tmin <= t && t <= tmax)) return false;
        }
    }
    t_out = t;
    return true;
}

// include/plane.h:57-96 incl. the interior tests :30-55.  p -> the plane's 5 float4.
//
// As for the sphere, BOTH policies evaluate t's operands, the hit point and the interior coordinates alpha, beta in the
// reference's operation order without contraction: a ray through the shared edge of two faces of a solid has alpha or beta
// within rounding of 0 (the BASELINE camera looks along such an edge: a whole column of pixel centres has |beta| ~ 3e-8),
// and whether it is accepted, by which face, or falls through the crack is decided by the last bit.  The fast policy
// replaces only the reference's double-precision operations by fp32 ones that round to the same value: the division
// (IEEE fp32 division == the double quotient rounded to float, except on exact double-rounding ties) and the triangle's
// alpha + beta > 1 (fp32 sum, with the exact rounding error consulted when the sum rounds to 1).
template <bool X>
__device__ __forceinline__ bool plane_t(V3 o, V3 d, const float4* __restrict__ p, float tmin, float tmax, float& t_out,
                                        float& alpha_out, float& beta_out) {
    float4 nD = __ldg(p);
    V3 n = mk3(nD.x, nD.y, nD.z);
    float denom = dot3<true>(n, d);
    // the reference compares float |denom| with the DOUBLE literal 1e-8 (include/plane.h:59): true exactly for
    // |denom| <= fp32(1e-8) = 9.99999994e-9
    if (fabsf(denom) <= 1e-8f) return false;
    float num = __fsub_rn(nD.w, dot3<true>(n, o));
    float t = X ? (float)((double)num / (double)denom) : __fdiv_rn(num, denom);
    if (!(tmin <= t && t <= tmax)) return false;
    float4 wT = __ldg(p + 1), u4 = __ldg(p + 2), v4 = __ldg(p + 3), b4 = __ldg(p + 4);
    V3 ip = add3<true>(o, scale3<true>(t, d));
    V3 ph = sub3<true>(ip, mk3(b4.x, b4.y, b4.z));
    V3 w = mk3(wT.x, wT.y, wT.z);
    float alpha = dot3<true>(w, cross3<true>(ph, mk3(v4.x, v4.y, v4.z)));
    float beta = dot3<true>(w, cross3<true>(mk3(u4.x, u4.y, u4.z), ph));
    int type = __float_as_int(wT.w);
    if (type == 0) {  // QUAD
        if (!(0.0f <= alpha && alpha <= 1.0f) || !(0.0f <= beta && beta <= 1.0f)) return false;
    } else if (type == 1) {  // ELLIPSE: powf(a-.5,2)+powf(b-.5,2) > .25 rejects; (double)alpha - 0.5 is exact, so its float is the fp32 difference
        float da = __fsub_rn(alpha, 0.5f), db = __fsub_rn(beta, 0.5f);
        if (__fadd_rn(__fmul_rn(da, da), __fmul_rn(db, db)) > 0.25f) return false;
    } else if (type == 2) {  // TRIANGLE: the reference forms alpha+beta in double (exact) and compares with 1.0
        if (alpha < 0.0f || beta < 0.0f) return false;
        if (X) {
            if ((double)alpha + (double)beta > 1.0) return false;
        } else {
            const float sum = __fadd_rn(alpha, beta);
            if (sum > 1.0f) return false;
            if (sum == 1.0f) {  // the exact sum is 1 + err: TwoSum's error term (Knuth), exact in fp32
                const float bv = __fsub_rn(sum, alpha);
                const float err = __fadd_rn(__fsub_rn(alpha, __fsub_rn(sum, bv)), __fsub_rn(beta, bv));
                if (err > 0.0f) return false;
            }
        }
    }
    t_out = t;
    alpha_out = alpha;
    beta_out = beta;
    return true;
}

// Fast-mode plane test: the 3 x float4 record of lbvh.cuh (plane_fast_record) first -- t from (n, D), the interior
// coordinates as alpha = A.o + t A.d - a0, beta likewise: 48 bytes and ~30 instructions instead of 80 bytes and ~65 -- and the
// reference-order test (plane_t) only for the hits it cannot decide: those within kEdgeEps of the polygon's outline, where the
// two evaluations of alpha / beta (they agree to a few 1e-6) could fall on different sides.  Everywhere else the decision is
// the reference's, and t differs from the reference's by fp32 rounding (~3e-7 relative: no cancellation in t's operands).
constexpr float kEdgeEps = 1e-4f, kTieEps = 1e-6f;
#ifndef RTB_NOINLINE_FALLBACK
#define RTB_NOINLINE_FALLBACK 0
#endif
// the reference-order test as the fast test's fallback: rare (hits within kEdgeEps of an outline, near-parallel rays), and
// inlined it is ~90 instructions at each of plane_fast_t's two exits times every place a leaf is tested
#if RTB_NOINLINE_FALLBACK && defined(__CUDA_ARCH__)
__device__ __noinline__
#else
__device__ __forceinline__
#endif
bool plane_fallback_t(V3 o, V3 d, const float4* __restrict__ p, float tmin, float tmax, float& t_out) {
    float al, be;
    return plane_t<false>(o, d, p, tmin, tmax, t_out, al, be);
}
__device__ __forceinline__ bool plane_fast_t(V3 o, V3 d, const float4* __restrict__ q, const float4* __restrict__ p, float tmin,
                                             float tmax, float& t_out) {
    const float4 nD = __ldg(q);
    const float denom = fmaf(nD.z, d.z, fmaf(nD.y, d.y, nD.x * d.x));
    if (fabsf(denom) < 2e-8f) return plane_fallback_t(o, d, p, tmin, tmax, t_out);  // (almost) parallel: the reference's own |denom| < 1e-8 rule decides
    const float num = nD.w - fmaf(nD.z, o.z, fmaf(nD.y, o.y, nD.x * o.x));
    const float t = __fdividef(num, denom);
    // the same margin on t's interval: a hit within 1e-6 (relative) of tmin or of the closest hit so far is the reference's call
    if (!(tmin * (1.0f - kTieEps) <= t && t <= tmax * (1.0f + kTieEps))) return false;
    const bool t_clear = tmin * (1.0f + kTieEps) <= t && t <= tmax * (1.0f - kTieEps);
    const float4 A = __ldg(q + 1), B = __ldg(q + 2);
    const float ao = fmaf(A.z, o.z, fmaf(A.y, o.y, A.x * o.x)), ad = fmaf(A.z, d.z, fmaf(A.y, d.y, A.x * d.x));
    const float bo = fmaf(B.z, o.z, fmaf(B.y, o.y, B.x * o.x)), bd = fmaf(B.z, d.z, fmaf(B.y, d.y, B.x * d.x));
    const float alpha = fmaf(t, ad, ao) - A.w, beta = fmaf(t, bd, bo) - B.w;
    const int type = __float_as_int(A.w) & 3;
    // m > 0: inside by more than the margin; m < 0: outside by more than the margin; else undecided
    float m;
    if (type == 0) {
        m = fminf(fminf(alpha, 1.0f - alpha), fminf(beta, 1.0f - beta));
    } else if (type == 1) {
        const float da = alpha - 0.5f, db = beta - 0.5f;
        m = 0.25f - fmaf(da, da, db * db);
    } else {
        m = fminf(fminf(alpha, beta), 1.0f - (alpha + beta));
    }
    if (m > kEdgeEps && t_clear) {
        t_out = t;
        return true;
    }
    if (m < -kEdgeEps) return false;
    return plane_fallback_t(o, d, p, tmin, tmax, t_out);
}

// ---- traversal ----------------------------------------------------------------------------------------------
// The traversal is written as a per-ray state (RayTrav) plus three steps -- trav_inner (one wide node), trav_leaf (one
// leaf) and trav_pop -- so that the same code serves two schedulers: trace_closest below (one ray to completion: the
// primary-ray entry point, the wavefront extend kernel, the CPU emulation of the tests) and the megakernel, which
// interleaves the steps of 32 rays so that a warp executes ONE kind of step at a time (kernels.cu).
//
// RayTrav::ref says what the ray does next: >= 0 and < kRefIdle: visit that inner node; < 0: test that leaf;
// kRefDone: traversal finished (closest hit in htype/hslot/horig/closest).  kRefNeed / kRefIdle are the megakernel's
// "lane wants a new path" / "no work left" states.
constexpr int kRefDone = 0x7fffffff;
constexpr int kRefNeed = 0x7ffffffe;
constexpr int kRefIdle = 0x7ffffffd;
constexpr int kRefGrid = 0x7ffffffc;   // the ray is walking the sphere grid (trav_grid_step); the tree comes after it
__device__ __forceinline__ bool ref_is_leaf(int r) { return r < 0; }
__device__ __forceinline__ bool ref_is_inner(int r) { return (unsigned int)r < (unsigned int)kRefGrid; }
__device__ __forceinline__ bool ref_is_grid(int r) { return r == kRefGrid; }
// the ray still has traversal steps to do
__device__ __forceinline__ bool ref_is_active(int r) { return (unsigned int)r - (unsigned int)kRefIdle > 2u; }

struct RayTrav {
    V3 o, d;
    float idx, idy, idz, oox, ooy, ooz;  // reciprocal direction, origin * reciprocal direction
    float a, inv_a;                      // |d|^2 and (fast mode) its reciprocal: per-ray constants of the sphere test
    float tmin, closest, eps;
    int nx, ny, nz;                      // which float4 of the node holds the NEAR slab plane per axis (lo for d >= 0, hi for d < 0)
    int ref, sp;
    int htype, hslot, horig;             // closest hit so far: -1 miss / 0 sphere / 1 plane, re-ordered slot, caller's index
    unsigned int best_key;
    int gcell, gcur, gend;               // grid walk: packed cell (x | y << 10 | z << 20), next / end position in its ref list
};

// per-thread stack: kSmemStack entries in shared memory (strided by the block size), the rest in local memory as
// packed (t, ref) pairs
struct TravStack {
    float* st;
    int* sr;
    float2* lo;
    int tid, nt;
    __device__ __forceinline__ void push(int sp, float t, int r) const {
        if (kSmemStack > 0 && sp < kSmemStack) { st[sp * nt + tid] = t; sr[sp * nt + tid] = r; }
        else lo[sp - kSmemStack] = make_float2(t, __int_as_float(r));
    }
    __device__ __forceinline__ void load(int sp, float& t, int& r) const {
        if (kSmemStack > 0 && sp < kSmemStack) { t = st[sp * nt + tid]; r = sr[sp * nt + tid]; }
        else { float2 e = lo[sp - kSmemStack]; t = e.x; r = __float_as_int(e.y); }
    }
};

// the per-ray constants of a segment (everything RayTrav derives from o, d): a pure function of (o, d), so a scheduler may
// drop them and call this again for a traversal it resumes (render_mega's time-sliced TRACE phase does)
template <bool X>
__device__ __forceinline__ void trav_setup(RayTrav& r, V3 o, V3 d, float tmin) {
    const float eps = 1e-20f;
    float dx = fabsf(d.x) < eps ? copysignf(eps, d.x) : d.x;
    float dy = fabsf(d.y) < eps ? copysignf(eps, d.y) : d.y;
    float dz = fabsf(d.z) < eps ? copysignf(eps, d.z) : d.z;
    r.o = o;
    r.d = d;
#if RTB_FAST_RCP && defined(__CUDA_ARCH__)
    if (!X) {  // fast mode: MUFU.RCP (1 ulp) instead of the IEEE division sequence; these only steer the culling
        r.idx = __fdividef(1.0f, dx); r.idy = __fdividef(1.0f, dy); r.idz = __fdividef(1.0f, dz);
    } else
#endif
    { r.idx = 1.0f / dx; r.idy = 1.0f / dy; r.idz = 1.0f / dz; }
    r.oox = o.x * r.idx; r.ooy = o.y * r.idy; r.ooz = o.z * r.idz;
    r.a = len2_3<true>(d);  // exactly the reference's a = |d|^2 (include/sphere.h:27): it multiplies into the discriminant
#if RTB_FAST_RCP && defined(__CUDA_ARCH__)
    r.inv_a = X ? 0.0f : __fdividef(1.0f, r.a);
#else
    r.inv_a = X ? 0.0f : 1.0f / r.a;
#endif
    r.tmin = tmin;
    // conservative slab test: the far side is widened by a few ulps of the slab products (boxes are padded too)
    // Each axis contributes the rounding error of its own slab products, CAPPED: for an axis the ray is (almost) parallel
    // to, id = 1e20 and o*id = 1e21 -- its products decide nothing unless the origin sits within 1e-6 of a box face, but
    // uncapped they would widen the test so much that such a ray visits every box its origin plane cuts (measured: one
    // such ray in ~4e6 (sincospi gives exact zeros) cost 15 ms, the whole tail of a frame).
    r.eps = 1e-7f * (fminf(fabsf(r.oox), 1e3f) + fminf(fabsf(r.ooy), 1e3f) + fminf(fabsf(r.ooz), 1e3f));
    // taken from the sign of the reciprocal (the same sign as dx), which stays in a register through the traversal: when the
    // register allocator prefers recomputing these in the node loop over keeping them, that is a compare + select on a live
    // value instead of the whole |d| < eps guard (measured: 6 % of the kernel's instructions were that rematerialisation)
    r.nx = __float_as_int(r.idx) < 0 ? 3 : 0;
    r.ny = __float_as_int(r.idy) < 0 ? 4 : 1;
    r.nz = __float_as_int(r.idz) < 0 ? 5 : 2;
}

// start of a traversal: nothing hit, at the root, empty stack
__device__ __forceinline__ void trav_reset(RayTrav& r, float tmax) {
    r.closest = tmax;
    r.ref = 0;  // root
    r.sp = 0;
    r.htype = -1;
    r.hslot = -1;
    r.horig = -1;
    r.best_key = 0;
    r.gcell = 0;
    r.gcur = 0;
    r.gend = 0;
}

template <bool X>
__device__ __forceinline__ void trav_init(RayTrav& r, V3 o, V3 d, float tmin, float tmax) {
    trav_setup<X>(r, o, d, tmin);
    trav_reset(r, tmax);
}

// pops entries until one starts no further than the closest hit; kRefDone when the stack runs out
__device__ __forceinline__ void trav_pop(RayTrav& r, const TravStack& stk) {
    r.ref = kRefDone;
    while (r.sp > 0) {
        --r.sp;
        float t;
        int e;
        stk.load(r.sp, t, e);
        if (t <= r.closest) { r.ref = e; break; }
    }
}

// one inner node: 7 x 128-bit loads, 4 slab tests, near-first ordering.  Returns true when no child was hit (the
// caller pops); otherwise r.ref is the nearest child and the others are on the stack, far to near.
// The near / far plane of each axis is chosen by the LOAD ADDRESS (per-ray constants nx, ny, nz), not by min/max of
// both products; empty slots hold lo = +inf, hi = -inf and fail the test on their own.
template <bool X, bool COUNT, bool TOP = true>
__device__ __forceinline__ bool trav_inner(const DevScene& sc, const float4* __restrict__ s_top, const TravStack& stk, RayTrav& r,
                                           WorkCount& wc) {
    const int ref = r.ref;
    float4 nx4, ny4, nz4, fx4, fy4, fz4, c;
    if (TOP && ref < sc.top_nodes) {
        const float4* p = s_top + ref * kTopStride;
        nx4 = p[r.nx]; ny4 = p[r.ny]; nz4 = p[r.nz]; fx4 = p[3 - r.nx]; fy4 = p[5 - r.ny]; fz4 = p[7 - r.nz];
        c = p[6];
    } else {
        const float4* p = sc.nodes + (size_t)ref * 8;
        nx4 = __ldg(p + r.nx); ny4 = __ldg(p + r.ny); nz4 = __ldg(p + r.nz);
        fx4 = __ldg(p + (3 - r.nx)); fy4 = __ldg(p + (5 - r.ny)); fz4 = __ldg(p + (7 - r.nz));
        c = __ldg(p + 6);
    }
    if (COUNT) wc.nodes++;
    float tn[4];
    int cr[4] = {__float_as_int(c.x), __float_as_int(c.y), __float_as_int(c.z), __float_as_int(c.w)};
    const float nxs[4] = {nx4.x, nx4.y, nx4.z, nx4.w}, nys[4] = {ny4.x, ny4.y, ny4.z, ny4.w}, nzs[4] = {nz4.x, nz4.y, nz4.z, nz4.w};
    const float fxs[4] = {fx4.x, fx4.y, fx4.z, fx4.w}, fys[4] = {fy4.x, fy4.y, fy4.z, fy4.w}, fzs[4] = {fz4.x, fz4.y, fz4.z, fz4.w};
    const float inf = __int_as_float(0x7f800000);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float x0 = fmaf(nxs[i], r.idx, -r.oox), x1 = fmaf(fxs[i], r.idx, -r.oox);
        float y0 = fmaf(nys[i], r.idy, -r.ooy), y1 = fmaf(fys[i], r.idy, -r.ooy);
        float z0 = fmaf(nzs[i], r.idz, -r.ooz), z1 = fmaf(fzs[i], r.idz, -r.ooz);
        float tnear = fmaxf(fmaxf(x0, y0), fmaxf(z0, r.tmin));
        float tfar = fminf(fminf(x1, y1), fminf(z1, r.closest));
#if RTB_FAST_SLAB
        // fast mode: no widening of the far side.  The rounding of the slab products is ~6e-8 * (|origin| + |plane|) in world
        // units, the leaf boxes are padded by >= 1e-4-ish (make_prim_boxes), and inner boxes are unions of them -- so only a
        // ray that grazes a box edge AND the primitive inside within that rounding can be lost; exact mode keeps the guard.
        bool h = X ? tnear <= fmaf(tfar, 1.0000005f, r.eps) : tnear <= tfar;
#else
        bool h = tnear <= fmaf(tfar, 1.0000005f, r.eps);
#endif
        tn[i] = h ? tnear : inf;
    }
    // sorting network on 4 (t, ref) pairs: (0,1)(2,3)(0,2)(1,3)(1,2)
#define RTB_CSWAP(i, j)                                            \
    {                                                              \
        bool s_ = tn[j] < tn[i];                                   \
        float ta_ = s_ ? tn[j] : tn[i], tb_ = s_ ? tn[i] : tn[j];  \
        int ra_ = s_ ? cr[j] : cr[i], rb_ = s_ ? cr[i] : cr[j];    \
        tn[i] = ta_; tn[j] = tb_; cr[i] = ra_; cr[j] = rb_;        \
    }
#if RTB_SORT_PARTIAL
    // three comparators: entry 0 is the nearest child, entry 2 the nearer of the two pair losers' winners; 1 and 3 unordered
    RTB_CSWAP(0, 1) RTB_CSWAP(2, 3) RTB_CSWAP(0, 2)
#undef RTB_CSWAP
    {
        const int order[3] = {3, 1, 2};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const int i = order[k];
            if (tn[i] < inf) {
                stk.push(r.sp, tn[i], cr[i]);
                ++r.sp;
            }
        }
    }
#else
    RTB_CSWAP(0, 1) RTB_CSWAP(2, 3) RTB_CSWAP(0, 2) RTB_CSWAP(1, 3) RTB_CSWAP(1, 2)
#undef RTB_CSWAP
    // push far-to-near: entries 3,2,1 ; continue with entry 0
#pragma unroll
    for (int i = 3; i >= 1; --i) {
        if (tn[i] < inf) {
            stk.push(r.sp, tn[i], cr[i]);
            ++r.sp;
        }
    }
#endif
    if (tn[0] < inf) {
        r.ref = cr[0];
        return false;
    }
    return true;
}

// one leaf: up to 4 same-type primitives; the caller pops afterwards
template <bool X, bool COUNT>
__device__ __forceinline__ void trav_leaf(const DevScene& sc, RayTrav& r, WorkCount& wc, int leaf_ref) {
    unsigned int bits = (unsigned int)leaf_ref;
    const int type = (bits >> 30) & 1, cnt = kMaxLeafPrims == 1 ? 1 : (int)((bits >> 28) & 3) + 1, first = bits & 0x0fffffff;
    for (int k = 0; k < cnt; ++k) {
        float t, al, be;
        bool h;
        if (type == 0) {
            if (COUNT) wc.spheres++;
            h = sphere_t<X>(r.o, r.d, r.a, r.inv_a, __ldg(sc.spheres + first + k), r.tmin, r.closest, t);
        } else {
            if (COUNT) wc.planes++;
#if RTB_FAST_PLANE
            if (!X) h = plane_fast_t(r.o, r.d, sc.planes_fast + 3 * (size_t)(first + k), sc.planes + 5 * (size_t)(first + k), r.tmin, r.closest, t);
            else
#endif
            h = plane_t<X>(r.o, r.d, sc.planes + 5 * (size_t)(first + k), r.tmin, r.closest, t, al, be);
        }
        if (h) {
            if (X) {
                // exact mode reproduces the reference's tie rule (t == closest: the later primitive in array order wins)
                int orig = type == 0 ? __ldg(sc.sphere_meta + first + k).y : __float_as_int(__ldg(sc.planes + 5 * (size_t)(first + k) + 3).w);
                unsigned int key = ((unsigned int)type << 30) | (unsigned int)orig;
                if (t < r.closest || r.htype < 0 || key > r.best_key) {
                    r.closest = t;
                    r.best_key = key;
                    r.htype = type;
                    r.hslot = first + k;
                    r.horig = orig;
                }
            } else {
                // production: no caller's index and no key are carried (two registers and a load less).  An exact tie
                // t == closest has measure zero for generic rays, but the shared edge of two faces of a solid produces it
                // (both faces accept the edge point at the same t); only then are the two callers' indices fetched, and the
                // reference's rule applied: the later primitive in (spheres, then planes) array order wins.
                bool take = true;
                if (t == r.closest && r.htype >= 0) {
                    const int orig = type == 0 ? __ldg(sc.sphere_meta + first + k).y : __float_as_int(__ldg(sc.planes + 5 * (size_t)(first + k) + 3).w);
                    const int best = r.htype == 0 ? __ldg(sc.sphere_meta + r.hslot).y : __float_as_int(__ldg(sc.planes + 5 * (size_t)r.hslot + 3).w);
                    take = (((unsigned int)type << 30) | (unsigned int)orig) > (((unsigned int)r.htype << 30) | (unsigned int)best);
                }
                if (take) {
                    r.closest = t;
                    r.htype = type;
                    r.hslot = first + k;
                }
            }
        }
    }
}

template <bool X, bool COUNT>
__device__ __forceinline__ void trav_leaf(const DevScene& sc, RayTrav& r, WorkCount& wc) {
    trav_leaf<X, COUNT>(sc, r, wc, r.ref);
}

// ---- sphere grid ---------------------------------------------------------------------------------------------
// A ray's walk through the grid is a function of its cell alone: the parameter at which it leaves cell c along axis a is
// t_a(c) = (lo_a + (c_a + [d_a > 0]) cw_a) id_a - o_a id_a, recomputed from the integer cell at every step (two FFMA per
// axis on constants and the per-ray reciprocals the tree walk keeps anyway) instead of accumulated -- so a walk that is
// interrupted (render_mega's time-sliced TRACE) and resumed takes exactly the steps of an uninterrupted one, and the
// state is three integers.  The walk ends behind the first cell boundary that lies beyond the closest hit so far, or
// where it leaves the grid; then the tree (planes, outlying spheres) is entered if the ray reaches its bounds at all.

// slab test of [tmin, closest] against a box; nx/ny/nz-style selection is done with min/max here (boxes are not in node layout)
template <bool X>
__device__ __forceinline__ bool ray_box(const RayTrav& r, float lx, float ly, float lz, float hx, float hy, float hz) {
    const float x0 = fmaf(lx, r.idx, -r.oox), x1 = fmaf(hx, r.idx, -r.oox);
    const float y0 = fmaf(ly, r.idy, -r.ooy), y1 = fmaf(hy, r.idy, -r.ooy);
    const float z0 = fmaf(lz, r.idz, -r.ooz), z1 = fmaf(hz, r.idz, -r.ooz);
    const float tnear = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), r.tmin));
    const float tfar = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), r.closest));
    return tnear <= fmaf(tfar, 1.0000005f, r.eps);  // the exact-mode guard of trav_inner, for both policies: these tests are rare
}

// what follows the grid walk (or the hoisted tests, without a grid): the tree's root, or nothing
template <bool X>
__device__ __forceinline__ int trav_tree_entry(const DevScene& sc, const RayTrav& r) {
    if (!sc.grid_on) return 0;
    return ray_box<X>(r, sc.tree_lo[0], sc.tree_lo[1], sc.tree_lo[2], sc.tree_hi[0], sc.tree_hi[1], sc.tree_hi[2]) ? 0 : kRefDone;
}

// enters the grid at the first cell the ray's [tmin, closest] interval reaches, or goes on to the tree
template <bool X>
__device__ __forceinline__ void trav_grid_begin(const DevScene& sc, RayTrav& r) {
    const float x0 = fmaf(sc.grid_lo[0], r.idx, -r.oox), x1 = fmaf(sc.grid_hi[0], r.idx, -r.oox);
    const float y0 = fmaf(sc.grid_lo[1], r.idy, -r.ooy), y1 = fmaf(sc.grid_hi[1], r.idy, -r.ooy);
    const float z0 = fmaf(sc.grid_lo[2], r.idz, -r.ooz), z1 = fmaf(sc.grid_hi[2], r.idz, -r.ooz);
    const float t0 = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), r.tmin));
    const float t1 = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), r.closest));
    if (!(t0 <= fmaf(t1, 1.0000005f, r.eps))) {
        r.ref = trav_tree_entry<X>(sc, r);
        return;
    }
    // entry cell: the cell of o + t0 d, clamped (a point on a face of the grid belongs to the cell inside)
    const float px = fmaf(t0, r.d.x, r.o.x), py = fmaf(t0, r.d.y, r.o.y), pz = fmaf(t0, r.d.z, r.o.z);
    int cx = (int)floorf((px - sc.grid_lo[0]) * sc.grid_icw[0]);
    int cy = (int)floorf((py - sc.grid_lo[1]) * sc.grid_icw[1]);
    int cz = (int)floorf((pz - sc.grid_lo[2]) * sc.grid_icw[2]);
    cx = cx < 0 ? 0 : (cx >= sc.grid_n[0] ? sc.grid_n[0] - 1 : cx);
    cy = cy < 0 ? 0 : (cy >= sc.grid_n[1] ? sc.grid_n[1] - 1 : cy);
    cz = cz < 0 ? 0 : (cz >= sc.grid_n[2] ? sc.grid_n[2] - 1 : cz);
    r.gcell = cx | (cy << 10) | (cz << 20);
    r.gcur = 0;
    r.gend = -1;  // "list of this cell not fetched yet"
    r.ref = kRefGrid;
}

// one step of the walk: (move to the next cell and fetch its list)? then (one sphere of the list)?
template <bool X, bool COUNT>
__device__ __forceinline__ void trav_grid_step(const DevScene& sc, RayTrav& r, WorkCount& wc) {
    if (r.gcur >= r.gend) {
        int cx = r.gcell & 1023, cy = (r.gcell >> 10) & 1023, cz = r.gcell >> 20;
        if (r.gend >= 0) {
            // leave the cell through the nearest of its three far faces
            const float ex = fmaf((float)(cx + (__float_as_int(r.idx) < 0 ? 0 : 1)), sc.grid_cw[0], sc.grid_lo[0]);
            const float ey = fmaf((float)(cy + (__float_as_int(r.idy) < 0 ? 0 : 1)), sc.grid_cw[1], sc.grid_lo[1]);
            const float ez = fmaf((float)(cz + (__float_as_int(r.idz) < 0 ? 0 : 1)), sc.grid_cw[2], sc.grid_lo[2]);
            const float tx = fmaf(ex, r.idx, -r.oox), ty = fmaf(ey, r.idy, -r.ooy), tz = fmaf(ez, r.idz, -r.ooz);
            const float tn = fminf(tx, fminf(ty, tz));
            // a hit inside the cells walked so far ends the walk.  The guard keeps a hit that lies ON the face (within the
            // rounding of tn) from ending it one cell early: a sphere overlapping the next cell could be nearer by less than that
            bool out = r.closest < tn - fmaf(fabsf(tn), 4e-7f, r.eps);
            if (!out) {
                if (tx <= ty && tx <= tz) { cx += __float_as_int(r.idx) < 0 ? -1 : 1; out = (unsigned int)cx >= (unsigned int)sc.grid_n[0]; }
                else if (ty <= tz) { cy += __float_as_int(r.idy) < 0 ? -1 : 1; out = (unsigned int)cy >= (unsigned int)sc.grid_n[1]; }
                else { cz += __float_as_int(r.idz) < 0 ? -1 : 1; out = (unsigned int)cz >= (unsigned int)sc.grid_n[2]; }
            }
            if (out) {
                r.ref = trav_tree_entry<X>(sc, r);
                return;
            }
            r.gcell = cx | (cy << 10) | (cz << 20);
        }
        if (COUNT) wc.nodes++;  // a cell counts as a node visit
        const uint2 h = __ldg(sc.grid_cells + ((size_t)cz * sc.grid_n[1] + cy) * sc.grid_n[0] + cx);
        r.gcur = (int)h.x;
        r.gend = (int)(h.x + h.y);
    }
    if (r.gcur < r.gend) {
        const int ref = __ldg(sc.grid_refs + r.gcur);
        ++r.gcur;
        // the sphere's padded box first (make_prim_boxes' padding): what a tree would have culled is culled here too, so the
        // fp32 discriminant's phantom hits OUTSIDE the box -- which the reference's own tree never reaches -- stay unreported
        const float4 s = __ldg(sc.spheres + (ref & 0x0fffffff));
        const float ar = fabsf(s.w);
        const float m = fmaxf(fabsf(s.x), fmaxf(fabsf(s.y), fabsf(s.z)));
        const float R = ar + fmaf(1e-3f, ar, 1e-6f * (m + ar));
        if (ray_box<X>(r, s.x - R, s.y - R, s.z - R, s.x + R, s.y + R, s.z + R)) trav_leaf<X, COUNT>(sc, r, wc, ref);
    }
}

// Start of a segment: per-ray constants, then the hoisted primitives (the few that span most of the scene, kept out of the
// tree by the builder).  Callers reach this point with the warp converged (the SHADE phase of the megakernel), so the floor
// quad -- which most rays hit -- is tested at full lane occupancy instead of as a 2-lanes-active leaf step, and `closest`
// already holds its hit when the traversal starts.  Exact mode: the update rule is the leaf step's, so the closest hit and
// its tie-break do not depend on where a primitive is tested.
template <bool X, bool COUNT>
__device__ __forceinline__ void trav_begin(const DevScene& sc, RayTrav& r, V3 o, V3 d, float tmin, float tmax, WorkCount& wc) {
    trav_init<X>(r, o, d, tmin, tmax);
    for (int k = 0; k < sc.num_hoisted; ++k) trav_leaf<X, COUNT>(sc, r, wc, sc.hoisted[k]);
    if (sc.grid_on) trav_grid_begin<X>(sc, r);
}

// leaf step followed by the pop, with the top stack entry requested BEFORE the primitive test so that the two loads
// (primitive, stack entry) are in flight together
template <bool X, bool COUNT>
__device__ __forceinline__ void trav_leaf_pop(const DevScene& sc, RayTrav& r, const TravStack& stk, WorkCount& wc) {
    float t0 = 0.0f;
    int e0 = kRefDone;
    const bool have = r.sp > 0;
    if (have) stk.load(r.sp - 1, t0, e0);
    trav_leaf<X, COUNT>(sc, r, wc, r.ref);
    r.ref = kRefDone;
    if (have) {
        --r.sp;
        if (t0 <= r.closest) r.ref = e0;
        else trav_pop(r, stk);
    }
}

__device__ __forceinline__ HitInfo trav_hit(const RayTrav& r) {
    HitInfo h;
    h.t = r.closest;
    h.type = r.htype;
    h.slot = r.hslot;
    h.orig = r.horig;
    return h;
}

// one ray to completion
template <bool X, bool COUNT>
__device__ __forceinline__ void trace_closest(const DevScene& sc, const float4* __restrict__ s_top, float* __restrict__ s_st,
                                              int* __restrict__ s_sr, V3 o, V3 d, float tmin, float tmax, HitInfo& hit,
                                              WorkCount& wc) {
    float2 lstack[kLocalStack];
    TravStack stk;
    stk.st = s_st; stk.sr = s_sr; stk.lo = lstack;
    stk.tid = threadIdx.x; stk.nt = blockDim.x;
    RayTrav r;
    trav_begin<X, COUNT>(sc, r, o, d, tmin, tmax, wc);
    while (r.ref != kRefDone) {
        if (ref_is_grid(r.ref)) {
            trav_grid_step<X, COUNT>(sc, r, wc);
            continue;
        }
        bool pop = true;
        if (ref_is_leaf(r.ref)) trav_leaf<X, COUNT>(sc, r, wc);
        else pop = trav_inner<X, COUNT>(sc, s_top, stk, r, wc);
        if (pop) trav_pop(r, stk);
    }
    hit = trav_hit(r);
    if (!X && r.htype >= 0)  // fast mode does not carry the caller's index through the traversal: one lookup at the end
        hit.orig = r.htype == 0 ? __ldg(sc.sphere_meta + r.hslot).y : __float_as_int(__ldg(sc.planes + 5 * (size_t)r.hslot + 3).w);
}

}  // namespace rtb
