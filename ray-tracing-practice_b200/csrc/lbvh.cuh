// lbvh.cuh -- per-element steps of the DEVICE BVH build.
//
// Stands where the reference's host builder stands (include/bvh_builder.h:52-120: recursive median split with
// std::nth_element, ~1.8 s for 1M primitives on one core) and next to this repository's own host builder
// (wide_bvh.cpp: binned SAH, better trees, tens of ms).  This one runs on the GPU in a few hundred microseconds to a
// few milliseconds, for callers that rebuild per frame:
//     1. conservative primitive boxes (the same arithmetic as make_prim_boxes, wide_bvh.cpp) + centroid bounds
//     2. 63-bit Morton key of the box centre, top bit = "large primitive" class so that the floor quad and the
//        polyhedra faces do not inflate the boxes of the small-sphere subtrees they happen to lie in
//     3. radix sort of (key, primitive)
//     4. binary radix tree over the sorted keys, one thread per inner node (Karras 2012), ties broken by position
//     5. bottom-up box fit (second arrival at a node computes it)
//     6. breadth-first collapse to the 128-byte 4-wide nodes the render kernels read; which binary nodes become wide
//        nodes is decided by the same dynamic programme as in the host builder (tables filled in step 5); one primitive
//        per leaf
//     7. per-node depth / worst-case stack need, swept level by level from the leaves up
// Every step is a function of one element, written once here and called from the kernels in bvh_gpu.cu and -- for the
// CPU test tier -- from tests/emul, which runs them in plain loops and checks the tree with validate_wide_bvh.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "ref_layouts.h"

namespace rtb {
namespace lbvh {

#ifdef __CUDA_ARCH__
#define LB_ADD(a, b) __fadd_rn(a, b)
#define LB_SUB(a, b) __fsub_rn(a, b)
#define LB_MUL(a, b) __fmul_rn(a, b)
#define LB_DIV(a, b) __fdiv_rn(a, b)
#else
// host builds (tests/emul, and the host halves of bvh_gpu.cu) are compiled without FMA contraction
#define LB_ADD(a, b) ((a) + (b))
#define LB_SUB(a, b) ((a) - (b))
#define LB_MUL(a, b) ((a) * (b))
#define LB_DIV(a, b) ((a) / (b))
#endif
#define LB_HD __host__ __device__ inline

struct Box {
    float lo[3], hi[3];
};

LB_HD float lb_min(float a, float b) { return a < b ? a : b; }
LB_HD float lb_max(float a, float b) { return a > b ? a : b; }
LB_HD float lb_abs(float a) { return a < 0.f ? -a : a; }

// 2*(dx*dy + dy*dz + dz*dx), 0 for inverted boxes: the ordering key of the collapse (Box3::area in wide_bvh.cpp)
LB_HD float box_area(const Box& b) {
    float dx = LB_SUB(b.hi[0], b.lo[0]), dy = LB_SUB(b.hi[1], b.lo[1]), dz = LB_SUB(b.hi[2], b.lo[2]);
    if (dx < 0 || dy < 0 || dz < 0) return 0.f;
    return LB_MUL(2.f, LB_ADD(LB_ADD(LB_MUL(dx, dy), LB_MUL(dy, dz)), LB_MUL(dz, dx)));
}
LB_HD Box box_union(const Box& a, const Box& b) {
    Box r;
    for (int k = 0; k < 3; ++k) { r.lo[k] = lb_min(a.lo[k], b.lo[k]); r.hi[k] = lb_max(a.hi[k], b.hi[k]); }
    return r;
}

// ---- step 1: conservative boxes.  Bit for bit the boxes of make_prim_boxes (wide_bvh.cpp): supersets of everything the
// reference's leaf tests (include/sphere.h:24, include/plane.h:57) can accept.  i < S: sphere i, else plane i-S.
LB_HD Box prim_box(const SphereIn* sp, int S, const PlaneIn* pl, int i) {
    Box b;
    if (i < S) {
        const SphereIn& s = sp[i];
        float r = lb_abs(s.r);
        float m = lb_max(lb_abs(s.c[0]), lb_max(lb_abs(s.c[1]), lb_abs(s.c[2])));
        float pad = LB_ADD(LB_MUL(1e-3f, r), LB_MUL(1e-6f, LB_ADD(m, r)));
        for (int a = 0; a < 3; ++a) {
            b.lo[a] = LB_SUB(LB_SUB(s.c[a], r), pad);
            b.hi[a] = LB_ADD(LB_ADD(s.c[a], r), pad);
        }
    } else {
        const PlaneIn& p = pl[i - S];
        float m = 0.f;
        for (int a = 0; a < 3; ++a) {
            float c0 = p.base[a], c1 = LB_ADD(p.base[a], p.u[a]), c2 = LB_ADD(p.base[a], p.v[a]);
            float c3 = LB_ADD(LB_ADD(p.base[a], p.u[a]), p.v[a]);
            float lo = lb_min(c0, lb_min(c1, c2)), hi = lb_max(c0, lb_max(c1, c2));
            if (p.type != 2) { lo = lb_min(lo, c3); hi = lb_max(hi, c3); }
            b.lo[a] = lo;
            b.hi[a] = hi;
            m = lb_max(m, lb_max(lb_abs(lo), lb_abs(hi)));
        }
        float pad = LB_ADD(1e-4f, LB_MUL(4e-6f, m));
        for (int a = 0; a < 3; ++a) { b.lo[a] = LB_SUB(b.lo[a], pad); b.hi[a] = LB_ADD(b.hi[a], pad); }
    }
    return b;
}

// ---- step 1b: hoisted primitives (wide_bvh.h: the few whose box has >= kHoistAreaFraction of the scene box's area are
// kept out of the tree and tested once per ray before the traversal).  Same rule as select_hoisted in wide_bvh.cpp.
constexpr int kMaxHoist = 4;              // == kMaxHoisted
constexpr float kHoistFraction = 0.25f;   // == kHoistAreaFraction
constexpr int kHoistCandCap = 1024;       // candidates the device records; more than that (a scene made of scene-sized
                                          // primitives) hoists nothing
LB_HD bool hoist_candidate(const Box& b, float scene_area) { return scene_area > 0.f && box_area(b) >= LB_MUL(kHoistFraction, scene_area); }

// ---- fast-mode plane record (traverse.cuh, plane_fast_t): 12 floats = 3 x float4 per plane,
//     (n, D)  (A, a0)  (B, b0)   with  A = v x w,  a0 = A . base,  B = w x u,  b0 = B . base,
// so that the reference's interior coordinates (include/plane.h:73-76: alpha = w . ((p - base) x v), beta = w . (u x (p - base)))
// become alpha = A . p - a0, beta = B . p - b0 -- two dot products instead of two cross products and two dot products.
// The plane type (0 quad, 1 ellipse, 2 triangle) rides in the two lowest mantissa bits of a0 (moves the alpha = 0 edge
// by < 3 ulp of a0).  Only the production (fast-math) traversal reads these; exact mode keeps the reference's arithmetic.
LB_HD void plane_fast_record(const PlaneIn& p, float out[12]) {
    // separately rounded products (LB_*): the host packer and the device packer produce the same bits
    float A[3], B[3];
    for (int i = 0; i < 3; ++i) {
        const int j = (i + 1) % 3, k = (i + 2) % 3;
        A[i] = LB_SUB(LB_MUL(p.v[j], p.w[k]), LB_MUL(p.v[k], p.w[j]));
        B[i] = LB_SUB(LB_MUL(p.w[j], p.u[k]), LB_MUL(p.w[k], p.u[j]));
    }
    float a0 = LB_ADD(LB_ADD(LB_MUL(A[0], p.base[0]), LB_MUL(A[1], p.base[1])), LB_MUL(A[2], p.base[2]));
    const float b0 = LB_ADD(LB_ADD(LB_MUL(B[0], p.base[0]), LB_MUL(B[1], p.base[1])), LB_MUL(B[2], p.base[2]));
    union { float f; uint32_t u; } c;
    c.f = a0;
    c.u = (c.u & ~3u) | ((uint32_t)p.type & 3u);
    a0 = c.f;
    out[0] = p.n[0]; out[1] = p.n[1]; out[2] = p.n[2]; out[3] = p.D;
    out[4] = A[0]; out[5] = A[1]; out[6] = A[2]; out[7] = a0;
    out[8] = B[0]; out[9] = B[1]; out[10] = B[2]; out[11] = b0;
}

// floats as order-preserving unsigned keys, so scene bounds reduce with integer atomicMin/atomicMax
LB_HD uint32_t float_to_ordered(float f) {
    union { float f; uint32_t u; } c;
    c.f = f;
    return (c.u & 0x80000000u) ? ~c.u : (c.u | 0x80000000u);
}
LB_HD float ordered_to_float(uint32_t u) {
    union { float f; uint32_t u; } c;
    c.u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
    return c.f;
}

// ---- step 2: Morton key
LB_HD uint64_t spread21(uint32_t v) {  // 21 bits -> every third bit of 63
    uint64_t x = v & 0x1fffffu;
    x = (x | (x << 32)) & 0x001f00000000ffffull;
    x = (x | (x << 16)) & 0x001f0000ff0000ffull;
    x = (x | (x << 8)) & 0x100f00f00f00f00full;
    x = (x | (x << 4)) & 0x10c30c30c30c30c3ull;
    x = (x | (x << 2)) & 0x1249249249249249ull;
    return x;
}
// A primitive wider than this fraction of the scene goes to the "large" class (its own subtree under the root).  Measured on
// the 100k-sphere BASELINE scene (wide-node visits per ray, host run of these same steps): no class 15.7; 1/128..1/12 (floor,
// polyhedra faces, light spheres) 12.5; 1/8..1/2 (the floor quad only) 11.3 -- every extra class member is a box all rays test.
constexpr float kLargeFraction = 1.f / 8.f;

// cmin / cmax: bounds of the box centres.  key bit 63 = large class, bits 62..0 = Morton code of the centre.
LB_HD uint64_t morton_key(const Box& b, const float cmin[3], const float cmax[3]) {
    float ext = lb_max(LB_SUB(cmax[0], cmin[0]), lb_max(LB_SUB(cmax[1], cmin[1]), LB_SUB(cmax[2], cmin[2])));
    uint32_t q[3];
    float wide = 0.f;
    for (int a = 0; a < 3; ++a) {
        // each axis is scaled by its own extent (measured on the BASELINE scenes: 12.5 wide-node visits per ray against 13.3
        // with one cubical scale -- the sphere field is a thin slab, and slab-shaped cells follow it)
        float e = LB_SUB(cmax[a], cmin[a]);
        float c = LB_MUL(0.5f, LB_ADD(b.lo[a], b.hi[a]));
        float x = e > 0.f ? LB_MUL(LB_DIV(LB_SUB(c, cmin[a]), e), 2097152.f) : 0.f;
        x = lb_min(lb_max(x, 0.f), 2097151.f);
        q[a] = (uint32_t)x;
        wide = lb_max(wide, LB_SUB(b.hi[a], b.lo[a]));
    }
    uint64_t key = (spread21(q[0]) << 2) | (spread21(q[1]) << 1) | spread21(q[2]);
    if (ext > 0.f && wide > LB_MUL(kLargeFraction, ext)) key |= 1ull << 63;
    return key;
}

// ---- step 4: binary radix tree (Karras, "Maximizing parallelism in the construction of BVHs, octrees and k-d trees",
// HPG 2012).  n sorted keys, inner nodes 0..n-2 (root 0), leaves referenced as (n-1)+position.
LB_HD int lb_clz64(uint64_t x) {
#ifdef __CUDA_ARCH__
    return __clzll((long long)x);
#else
    return x ? __builtin_clzll(x) : 64;
#endif
}
LB_HD int lb_clz32(uint32_t x) {
#ifdef __CUDA_ARCH__
    return __clz((int)x);
#else
    return x ? __builtin_clz(x) : 32;
#endif
}
LB_HD int key_delta(const uint64_t* keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + lb_clz32((uint32_t)i ^ (uint32_t)j);  // equal keys: the position breaks the tie
    return lb_clz64(a ^ b);
}
// children of inner node i: node ids in [0, 2n-1): inner c -> c, leaf at sorted position p -> (n-1)+p
LB_HD void radix_node(const uint64_t* keys, int n, int i, int& left, int& right) {
    int d = key_delta(keys, n, i, i + 1) - key_delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
    int dmin = key_delta(keys, n, i, i - d);
    int lmax = 2;
    while (key_delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (key_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = key_delta(keys, n, i, j);
    int s = 0;
    int t = l;
    do {
        t = (t + 1) / 2;
        if (key_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + (d < 0 ? -1 : 0);
    int lo = i < j ? i : j, hi = i < j ? j : i;
    left = (lo == gamma) ? (n - 1) + gamma : gamma;
    right = (hi == gamma + 1) ? (n - 1) + gamma + 1 : gamma + 1;
}

// ---- step 6: collapse.  `bin` is the binary inner node that roots this wide node; kids[] receives up to four binary
// node ids (-1 unused).  Same rule as the host builder: start from the two children, keep opening the inner child with
// the largest area (first one on ties) until four slots are used.  Returns the number of INNER kids.
LB_HD int collapse_pick(const int2* children, const Box* boxes, int n, int bin, int kids[4]) {
    int nk = 0;
    kids[0] = kids[1] = kids[2] = kids[3] = -1;
    if (n == 1) {  // single-primitive scene: the root wide node holds the one leaf
        kids[0] = 0;
        return 0;
    }
    kids[nk++] = children[bin].x;
    kids[nk++] = children[bin].y;
    while (nk < 4) {
        int pick = -1;
        float pa = -1.f;
        for (int k = 0; k < nk; ++k) {
            if (kids[k] >= n - 1) continue;  // leaf
            float a = box_area(boxes[kids[k]]);
            if (a > pa) { pa = a; pick = k; }
        }
        if (pick < 0) break;
        int open = kids[pick];
        kids[pick] = children[open].x;
        kids[nk++] = children[open].y;
    }
    int inner = 0;
    for (int k = 0; k < nk; ++k) inner += kids[k] < n - 1;
    return inner;
}

// ---- step 5b/6b: WHICH binary nodes become wide nodes, chosen by the same dynamic programme as the host builder
// (wide_bvh.cpp; after Ylitie, Karras, Laine, HPG 2017, section 4.1).  With one primitive per leaf the leaf term of the SAH
// is the same for every collapse, so the cost to minimise is the summed area of the wide nodes:
//     c[0] = area(n) + min_k (cl[k-1] + cr[3-k]),  k = 1..3 roots from the left subtree, 4-k from the right
//     c[i] = min(c[i-1], min_k (cl[k-1] + cr[i-k]))     subtree n represented by at most i+1 roots (k = 0: "use fewer")
// The table of a node is filled by the thread that fits its box (bottom-up), 16 bytes per inner node; leaves cost 0.
struct alignas(16) CollapseDp {
    float c[3];
    uint32_t k;  // choice for i roots in bits 8*i .. 8*i+7
};
LB_HD CollapseDp dp_leaf() {
    CollapseDp d;
    d.c[0] = d.c[1] = d.c[2] = 0.f;
    d.k = 0;
    return d;
}
LB_HD CollapseDp dp_inner(const CollapseDp& l, const CollapseDp& r, float area) {
    CollapseDp d;
    float best = LB_ADD(l.c[0], r.c[2]);
    uint32_t bk = 1;
    for (int k = 2; k <= 3; ++k) {
        float c = LB_ADD(l.c[k - 1], r.c[3 - k]);
        if (c < best) { best = c; bk = (uint32_t)k; }
    }
    d.c[0] = LB_ADD(area, best);
    d.k = bk;
    for (int i = 1; i < 3; ++i) {
        float bi = d.c[i - 1];
        uint32_t ki = 0;
        for (int k = 1; k <= i; ++k) {
            float c = LB_ADD(l.c[k - 1], r.c[i - k]);
            if (c < bi) { bi = c; ki = (uint32_t)k; }
        }
        d.c[i] = bi;
        d.k |= ki << (8 * i);
    }
    return d;
}
// ---- step 5, the work at one inner node once both of its subtrees are final: an optional TREE ROTATION (swap one child with
// a grandchild from the other side when that shrinks the box of the node in between; Kensler, "Tree rotations for improving
// bounding volume hierarchies", RT 2008 -- the radix tree splits at Morton bit boundaries, not where the area says), then
// the node's box and its collapse table.  Everything below `cur` belongs to the calling thread at this point (the other
// arrival at `cur` has left), but was written by other threads of the same launch: on the device it is read with
// ld.global.cg (L2), never from this SM's L1.  Measured on the 100k-sphere scene: 10.35 -> 10.08 wide-node visits per ray.
#ifdef __CUDA_ARCH__
#define LB_LDCG(p) __ldcg(p)
#else
#define LB_LDCG(p) (*(p))
#endif
#ifndef RTB_LBVH_ROTATE
#define RTB_LBVH_ROTATE 1
#endif
LB_HD Box load_box_cg(const Box* b) {
    const float* f = reinterpret_cast<const float*>(b);
    Box r;
    for (int k = 0; k < 3; ++k) { r.lo[k] = LB_LDCG(f + k); r.hi[k] = LB_LDCG(f + 3 + k); }
    return r;
}
LB_HD CollapseDp load_dp_cg(const CollapseDp* dp, int node, int n) {
    if (node >= n - 1) return dp_leaf();
    const float4 v = LB_LDCG(reinterpret_cast<const float4*>(dp + node));
    CollapseDp d;
    d.c[0] = v.x; d.c[1] = v.y; d.c[2] = v.z;
    union { float f; uint32_t u; } c;
    c.f = v.w;
    d.k = c.u;
    return d;
}
LB_HD void store_dp(CollapseDp* dp, int node, const CollapseDp& d) {
    union { float f; uint32_t u; } c;
    c.u = d.k;
    *reinterpret_cast<float4*>(dp + node) = make_float4(d.c[0], d.c[1], d.c[2], c.f);
}
LB_HD void fit_node(int2* children, int* parent, Box* boxes, CollapseDp* dp, int n, int cur) {
    int2 ch = LB_LDCG(children + cur);
    Box bl = load_box_cg(boxes + ch.x), br = load_box_cg(boxes + ch.y);
#if RTB_LBVH_ROTATE
    {
        float best = 0.f;
        int which = -1;
        int2 cl = make_int2(-1, -1), cr = make_int2(-1, -1);
        if (ch.y < n - 1) {  // right child is inner: candidates L <-> RL, L <-> RR
            cr = LB_LDCG(children + ch.y);
            const float ar = box_area(br);
            const float g0 = LB_SUB(ar, box_area(box_union(bl, load_box_cg(boxes + cr.y))));
            const float g1 = LB_SUB(ar, box_area(box_union(load_box_cg(boxes + cr.x), bl)));
            if (g0 > best) { best = g0; which = 0; }
            if (g1 > best) { best = g1; which = 1; }
        }
        if (ch.x < n - 1) {  // left child is inner: candidates R <-> LL, R <-> LR
            cl = LB_LDCG(children + ch.x);
            const float al = box_area(bl);
            const float g2 = LB_SUB(al, box_area(box_union(br, load_box_cg(boxes + cl.y))));
            const float g3 = LB_SUB(al, box_area(box_union(load_box_cg(boxes + cl.x), br)));
            if (g2 > best) { best = g2; which = 2; }
            if (g3 > best) { best = g3; which = 3; }
        }
        if (which >= 0) {
            const int node = which < 2 ? ch.y : ch.x, other = which < 2 ? ch.x : ch.y;
            int2 cn = which < 2 ? cr : cl;
            const int moved = (which & 1) ? cn.y : cn.x;  // the grandchild that comes up to `cur`
            if (which & 1) cn.y = other; else cn.x = other;
            children[node] = cn;
            parent[other] = node;
            parent[moved] = cur;
            if (which < 2) ch.x = moved; else ch.y = moved;
            children[cur] = ch;
            const Box nb = box_union(load_box_cg(boxes + cn.x), load_box_cg(boxes + cn.y));
            boxes[node] = nb;
            store_dp(dp, node, dp_inner(load_dp_cg(dp, cn.x, n), load_dp_cg(dp, cn.y, n), box_area(nb)));
            // the two children of `cur` as they are now (the union below is the same box as before the rotation)
            if (which < 2) { bl = load_box_cg(boxes + moved); br = nb; } else { bl = nb; br = load_box_cg(boxes + moved); }
        }
    }
#endif
    const Box u = box_union(bl, br);
    boxes[cur] = u;
    store_dp(dp, cur, dp_inner(load_dp_cg(dp, ch.x, n), load_dp_cg(dp, ch.y, n), box_area(u)));
}

// kids of the wide node rooted at binary inner node `bin` according to the tables; returns the number of INNER kids
LB_HD int collapse_pick_dp(const int2* children, const CollapseDp* dp, int n, int bin, int kids[4]) {
    int nk = 0;
    kids[0] = kids[1] = kids[2] = kids[3] = -1;
    if (n == 1) {
        kids[0] = 0;
        return 0;
    }
    int st_node[4], st_roots[4], sp = 0;
    const int kl = (int)(dp[bin].k & 0xffu);
    st_node[sp] = children[bin].y; st_roots[sp++] = 4 - kl;
    st_node[sp] = children[bin].x; st_roots[sp++] = kl;
    while (sp > 0) {
        --sp;
        const int node = st_node[sp], roots = st_roots[sp];
        if (node >= n - 1 || roots <= 1) { kids[nk++] = node; continue; }
        const int k = (int)((dp[node].k >> (8 * (roots - 1))) & 0xffu);
        if (k == 0) { st_node[sp] = node; st_roots[sp++] = roots - 1; continue; }
        st_node[sp] = children[node].y; st_roots[sp++] = roots - k;
        st_node[sp] = children[node].x; st_roots[sp++] = k;
    }
    int inner = 0;
    for (int k = 0; k < nk; ++k) inner += kids[k] < n - 1;
    return inner;
}

constexpr int32_t kEmpty = 0x7fffffff;  // == kEmptyChild (wide_bvh.h)
LB_HD int32_t make_leaf_ref(int type, int first) { return (int32_t)(0x80000000u | ((uint32_t)type << 30) | (uint32_t)first); }

// writes the 32 floats of wide node `out` (layout of WideNode, wide_bvh.h).  next_base + off + rank = index of an inner
// kid's wide node (next level, breadth-first); next_items[off + rank] = its binary node.  leaf_refs[p] = leaf reference
// of the primitive at sorted position p.
LB_HD void collapse_write(const Box* boxes, const int32_t* leaf_refs, int n, const int kids[4], int next_base, int off,
                          float* out, int* next_items) {
    const float inf = __builtin_huge_valf();
    int32_t* child = reinterpret_cast<int32_t*>(out + 24);
    int rank = 0;
    for (int i = 0; i < 4; ++i) {
        int k = kids[i];
        if (k < 0) {
            for (int a = 0; a < 3; ++a) { out[a * 4 + i] = inf; out[12 + a * 4 + i] = -inf; }
            child[i] = kEmpty;
        } else {
            const Box& b = boxes[k];
            for (int a = 0; a < 3; ++a) { out[a * 4 + i] = b.lo[a]; out[12 + a * 4 + i] = b.hi[a]; }
            if (k >= n - 1) {
                child[i] = leaf_refs[k - (n - 1)];
            } else {
                child[i] = next_base + off + rank;
                next_items[off + rank] = k;
                ++rank;
            }
        }
        child[4 + i] = 0;  // pad
    }
}

// ---- step 7: depth and worst-case traversal stack need of wide node `w`, from its children's (already final) values
LB_HD void depth_need(const float* node, const int* dep, const int* need, int& d_out, int& n_out) {
    const int32_t* child = reinterpret_cast<const int32_t*>(node + 24);
    int k = 0, dmax = 0, nmax = 0;
    for (int c = 0; c < 4; ++c) {
        if (child[c] == kEmpty) continue;
        ++k;
        if (child[c] >= 0) {
            dmax = dep[child[c]] > dmax ? dep[child[c]] : dmax;
            nmax = need[child[c]] > nmax ? need[child[c]] : nmax;
        }
    }
    d_out = 1 + dmax;
    n_out = (k > 0 ? k - 1 : 0) + nmax;
}

}  // namespace lbvh
}  // namespace rtb
