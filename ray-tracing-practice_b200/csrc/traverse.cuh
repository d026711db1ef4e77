// traverse.cuh -- closest-hit traversal of the 4-wide BVH with the reference's leaf arithmetic.
//
// Stands where hit_scene -> hit_bvh -> {AABB::hit, hit_sphere, hit_plane} stand in the reference
// (include/scene.h:23-35, include/bvh.h:19-65, include/aabb.h:42-65, include/sphere.h:24-53,
// include/plane.h:57-96).  Differences by design:
//   * nodes are 128-byte lines fetched as 8 x 128-bit loads (__ldg), the top levels from shared memory;
//   * children are visited near-to-far (the reference's order is undefined) with a per-thread short
//     stack in shared memory (spill to local memory only beyond kSmemStack entries);
//   * the slab test uses a per-ray reciprocal instead of 3 divides per node (include/aabb.h:44);
//   * leaf tests only produce (t, primitive); normals / uv are derived once, after traversal;
//   * acceptance follows the reference exactly: tmin <= t <= closest (Interval::contains, include/interval.h:16).
//     Exact-t ties resolve to the primitive the reference's linear scan (include/scene.h:37-51) would keep: the
//     later one in (spheres, then planes) array order.
#pragma once
#include "device_math.cuh"

namespace rtb {

struct DevScene {
    const float4* __restrict__ nodes;        // WideNode[ ] as 8 x float4
    const float4* __restrict__ spheres;      // re-ordered (cx,cy,cz,r)
    const int2* __restrict__ sphere_meta;    // (material_idx, caller's index)
    const float4* __restrict__ planes;       // re-ordered, 5 x float4: (n,D) (w,type) (u,mat) (v,caller's index) (base,0)
    const float4* __restrict__ mats;         // 3 x float4: (albedo,type) (emit,fuzz) (absorption,ir)
    const float4* __restrict__ tex;          // RGBA texels, row-major (RT_TEX_CPU lookups)
    cudaTextureObject_t tex_obj;             // RT_TEX_GPU lookups
    int tex_w, tex_h, tex_mat;
    int num_nodes;
    int top_nodes;                           // wide nodes [0, top_nodes) are staged in shared memory
};

constexpr int kSmemStack = 12;   // per-thread stack entries held in shared memory
constexpr int kLocalStack = 52;  // overflow entries (local memory); builder guarantees max_stack <= 64
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
// include/sphere.h:24-47.  a = |d|^2 and (fast mode) 1/a are per-ray constants.
template <bool X>
__device__ __forceinline__ bool sphere_t(V3 o, V3 d, float a, float inv_a, float4 s, float tmin, float tmax, float& t_out) {
    V3 oc = sub3<X>(o, mk3(s.x, s.y, s.z));
    float hb = dot3<X>(oc, d);
    float cc = fsub<X>(len2_3<X>(oc), fmul<X>(s.w, s.w));
    float D = fsub<X>(fmul<X>(hb, hb), fmul<X>(a, cc));
    if (D < 0.0f) return false;
    float sq = __fsqrt_rn(D);
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
template <bool X>
__device__ __forceinline__ bool plane_t(V3 o, V3 d, const float4* __restrict__ p, float tmin, float tmax, float& t_out,
                                        float& alpha_out, float& beta_out) {
    float4 nD = __ldg(p);
    V3 n = mk3(nD.x, nD.y, nD.z);
    float denom = dot3<X>(n, d);
    if (fabsf(denom) < 1e-8f) {
        // reference compares float |denom| with the DOUBLE literal 1e-8: equal to the fp32 compare unless
        // |denom| lies between 1e-8 and fp32(1e-8)
        if (!X || (double)fabsf(denom) < 1e-8) return false;
    }
    float num = fsub<X>(nD.w, dot3<X>(n, o));
    float t = X ? (float)((double)num / (double)denom) : __fdividef(num, denom);
    if (!(tmin <= t && t <= tmax)) return false;
    float4 wT = __ldg(p + 1), u4 = __ldg(p + 2), v4 = __ldg(p + 3), b4 = __ldg(p + 4);
    V3 ip = add3<X>(o, scale3<X>(t, d));
    V3 ph = sub3<X>(ip, mk3(b4.x, b4.y, b4.z));
    V3 w = mk3(wT.x, wT.y, wT.z);
    float alpha = dot3<X>(w, cross3<X>(ph, mk3(v4.x, v4.y, v4.z)));
    float beta = dot3<X>(w, cross3<X>(mk3(u4.x, u4.y, u4.z), ph));
    int type = __float_as_int(wT.w);
    if (type == 0) {  // QUAD
        if (!(0.0f <= alpha && alpha <= 1.0f) || !(0.0f <= beta && beta <= 1.0f)) return false;
    } else if (type == 1) {  // ELLIPSE: powf(a-.5,2)+powf(b-.5,2) > .25 rejects
        float da = X ? (float)((double)alpha - 0.5) : alpha - 0.5f;
        float db = X ? (float)((double)beta - 0.5) : beta - 0.5f;
        if (fadd<X>(fmul<X>(da, da), fmul<X>(db, db)) > 0.25f) return false;
    } else if (type == 2) {  // TRIANGLE: the reference forms alpha+beta in double
        if (alpha < 0.0f || beta < 0.0f) return false;
        if (X ? ((double)alpha + (double)beta > 1.0) : (alpha + beta > 1.0f)) return false;
    }
    t_out = t;
    alpha_out = alpha;
    beta_out = beta;
    return true;
}

// ---- traversal ----------------------------------------------------------------------------------------------
struct Stack {
    float* st;  // shared: [kSmemStack][blockDim] distances
    int* sr;    // shared: [kSmemStack][blockDim] refs
    int nthreads;
    float lt[kLocalStack];
    int lr[kLocalStack];
};

template <bool X, bool COUNT>
__device__ __forceinline__ void trace_closest(const DevScene& sc, const float4* __restrict__ s_top, float* __restrict__ s_st,
                                              int* __restrict__ s_sr, V3 o, V3 d, float tmin, float tmax, HitInfo& hit,
                                              WorkCount& wc) {
    // per-ray constants
    const float eps = 1e-20f;
    float dx = fabsf(d.x) < eps ? copysignf(eps, d.x) : d.x;
    float dy = fabsf(d.y) < eps ? copysignf(eps, d.y) : d.y;
    float dz = fabsf(d.z) < eps ? copysignf(eps, d.z) : d.z;
    const float idx = 1.0f / dx, idy = 1.0f / dy, idz = 1.0f / dz;
    const float oox = o.x * idx, ooy = o.y * idy, ooz = o.z * idz;
    const float a = len2_3<X>(d);
    const float inv_a = X ? 0.0f : 1.0f / a;

    float closest = tmax;
    hit.type = -1;
    hit.slot = -1;
    hit.orig = -1;
    hit.t = tmax;
    unsigned int best_key = 0;

    float lst[kLocalStack];
    int lsr[kLocalStack];
    const int tid = threadIdx.x, nt = blockDim.x;
    int sp = 0;
    int ref = 0;  // root

    for (;;) {
        bool have = true;
        if (ref >= 0) {
            // ---- inner node: 8 x 128-bit loads, 4 slab tests, near-first ordering
            float4 q0, q1, q2, q3, q4, q5;
            int4 ch;
            if (ref < sc.top_nodes) {
                const float4* p = s_top + ref * kTopStride;
                q0 = p[0]; q1 = p[1]; q2 = p[2]; q3 = p[3]; q4 = p[4]; q5 = p[5];
                float4 c = p[6];
                ch = make_int4(__float_as_int(c.x), __float_as_int(c.y), __float_as_int(c.z), __float_as_int(c.w));
            } else {
                const float4* p = sc.nodes + (size_t)ref * 8;
                q0 = __ldg(p); q1 = __ldg(p + 1); q2 = __ldg(p + 2); q3 = __ldg(p + 3); q4 = __ldg(p + 4); q5 = __ldg(p + 5);
                float4 c = __ldg(p + 6);
                ch = make_int4(__float_as_int(c.x), __float_as_int(c.y), __float_as_int(c.z), __float_as_int(c.w));
            }
            if (COUNT) wc.nodes++;
            float tn[4];
            int cr[4] = {ch.x, ch.y, ch.z, ch.w};
            const float lox[4] = {q0.x, q0.y, q0.z, q0.w}, loy[4] = {q1.x, q1.y, q1.z, q1.w}, loz[4] = {q2.x, q2.y, q2.z, q2.w};
            const float hix[4] = {q3.x, q3.y, q3.z, q3.w}, hiy[4] = {q4.x, q4.y, q4.z, q4.w}, hiz[4] = {q5.x, q5.y, q5.z, q5.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float x0 = fmaf(lox[i], idx, -oox), x1 = fmaf(hix[i], idx, -oox);
                float y0 = fmaf(loy[i], idy, -ooy), y1 = fmaf(hiy[i], idy, -ooy);
                float z0 = fmaf(loz[i], idz, -ooz), z1 = fmaf(hiz[i], idz, -ooz);
                float tnear = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), tmin));
                float tfar = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), closest));
                // conservative: widen the far side by a few ulps of the slab products (boxes are padded too)
                bool h = (cr[i] != kEmptyChildDev) && (tnear <= tfar * 1.0000005f + 1e-7f * (fabsf(oox) + fabsf(ooy) + fabsf(ooz)));
                tn[i] = h ? tnear : __int_as_float(0x7f800000);
            }
            // sorting network on 4 (t, ref) pairs: (0,1)(2,3)(0,2)(1,3)(1,2)
#define RTB_CSWAP(i, j)                                   \
    {                                                     \
        bool s_ = tn[j] < tn[i];                          \
        float ta_ = s_ ? tn[j] : tn[i], tb_ = s_ ? tn[i] : tn[j]; \
        int ra_ = s_ ? cr[j] : cr[i], rb_ = s_ ? cr[i] : cr[j];   \
        tn[i] = ta_; tn[j] = tb_; cr[i] = ra_; cr[j] = rb_;       \
    }
            RTB_CSWAP(0, 1) RTB_CSWAP(2, 3) RTB_CSWAP(0, 2) RTB_CSWAP(1, 3) RTB_CSWAP(1, 2)
#undef RTB_CSWAP
            const float inf = __int_as_float(0x7f800000);
            // push far-to-near: entries 3,2,1 ; continue with entry 0
#pragma unroll
            for (int i = 3; i >= 1; --i) {
                if (tn[i] < inf) {
                    if (sp < kSmemStack) { s_st[sp * nt + tid] = tn[i]; s_sr[sp * nt + tid] = cr[i]; }
                    else { lst[sp - kSmemStack] = tn[i]; lsr[sp - kSmemStack] = cr[i]; }
                    ++sp;
                }
            }
            if (tn[0] < inf) { ref = cr[0]; continue; }
            have = false;
        } else {
            // ---- leaf: up to 4 same-type primitives
            unsigned int bits = (unsigned int)ref;
            int type = (bits >> 30) & 1, cnt = ((bits >> 28) & 3) + 1, first = bits & 0x0fffffff;
            for (int k = 0; k < cnt; ++k) {
                float t, al, be;
                bool h;
                if (type == 0) {
                    if (COUNT) wc.spheres++;
                    h = sphere_t<X>(o, d, a, inv_a, __ldg(sc.spheres + first + k), tmin, closest, t);
                } else {
                    if (COUNT) wc.planes++;
                    h = plane_t<X>(o, d, sc.planes + 5 * (size_t)(first + k), tmin, closest, t, al, be);
                }
                if (h) {
                    int orig = type == 0 ? __ldg(sc.sphere_meta + first + k).y : __float_as_int(__ldg(sc.planes + 5 * (size_t)(first + k) + 3).w);
                    unsigned int key = ((unsigned int)type << 30) | (unsigned int)orig;
                    if (t < closest || hit.type < 0 || key > best_key) {
                        closest = t;
                        best_key = key;
                        hit.t = t;
                        hit.type = type;
                        hit.slot = first + k;
                        hit.orig = orig;
                    }
                }
            }
            have = false;
        }
        // ---- pop (skip entries that start beyond the current closest hit)
        while (!have && sp > 0) {
            --sp;
            float t;
            int r;
            if (sp < kSmemStack) { t = s_st[sp * nt + tid]; r = s_sr[sp * nt + tid]; }
            else { t = lst[sp - kSmemStack]; r = lsr[sp - kSmemStack]; }
            if (t <= closest) { ref = r; have = true; }
        }
        if (!have) break;
    }
}

}  // namespace rtb
