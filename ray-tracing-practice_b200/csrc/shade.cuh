// shade.cuh -- hit record reconstruction, floor texture lookup and the material scatter functions.
//
// Stands where the tail of hit_sphere / hit_plane (include/sphere.h:44-52, include/plane.h:88-95), ray_color's
// material handling (src/camera.cu:231-249) and material_scatter (include/materials.h:70-140) stand.  The
// ESTIMATOR is the reference's, quirks included: the "Lambertian" lobe is a uniform
// hemisphere direction with no cosine / pdf factor, METAL mixes 80 % fuzzy mirror with 20 % of that lobe,
// DIELECTRIC applies Beer-Lambert only when leaving the medium and then Russian roulette on max(attenuation).
// What changes is the execution: one material fetch of 3 x 128 bits instead of a 64-byte AoS copy, no sphere uv
// unless the material is textured (Q13), direct instead of rejection sampling on the Philox stream.
#pragma once
#include "traverse.cuh"

namespace rtb {

enum { kLambertian = 0, kMetal = 1, kDielectric = 2, kDiffuseLight = 3 };

// RT_TEX_CPU: the reference's tex2D_cpu (include/materials.h:20-51) -- wrap by -floor, py=(1-v)*H, no half texel.
template <bool X>
__device__ __forceinline__ V3 tex_cpu(const DevScene& sc, float u, float v) {
    int w = sc.tex_w, h = sc.tex_h;
    u = fsub<X>(u, floorf(u));
    v = fsub<X>(v, floorf(v));
    float px = fmul<X>(u, (float)w);
    float py = fmul<X>(fsub<X>(1.0f, v), (float)h);
    int x0 = (int)px, y0 = (int)py;
    int x1 = (x0 + 1) % w, y1 = (y0 + 1) % h;
    float dx = fsub<X>(px, (float)x0), dy = fsub<X>(py, (float)y0);
    x0 %= w;  // the reference indexes out of bounds here (UB); the oracle and this kernel wrap
    y0 %= h;
    float4 a = __ldg(sc.tex + (size_t)y0 * w + x0), b = __ldg(sc.tex + (size_t)y0 * w + x1);
    float4 c = __ldg(sc.tex + (size_t)y1 * w + x0), e = __ldg(sc.tex + (size_t)y1 * w + x1);
    float ix = fsub<X>(1.0f, dx), iy = fsub<X>(1.0f, dy);
    V3 top = add3<X>(scale3<X>(ix, mk3(a.x, a.y, a.z)), scale3<X>(dx, mk3(b.x, b.y, b.z)));
    V3 bot = add3<X>(scale3<X>(ix, mk3(c.x, c.y, c.z)), scale3<X>(dx, mk3(e.x, e.y, e.z)));
    return add3<X>(scale3<X>(iy, top), scale3<X>(dy, bot));
}

// include/sphere.h:16-22 (only evaluated for textured materials)
template <bool X>
__device__ __forceinline__ void sphere_uv(V3 p, float& u, float& v) {
    if (X) {
        float theta = (float)acos((double)p.y);
        float phi = (float)((double)(float)atan2((double)-p.z, (double)p.x) + 3.14159265358979323846);
        u = (float)((double)phi / (2 * 3.14159265358979323846));
        v = (float)((double)theta / 3.14159265358979323846);
    } else {
        u = (atan2f(-p.z, p.x) + 3.14159265f) * 0.159154943f;
        v = acosf(p.y) * 0.318309886f;
    }
}

// include/materials.h:64-68
template <bool X>
__device__ __forceinline__ float schlick(float cosine, float ref_idx) {
    float r0 = X ? __fdiv_rn(fsub<X>(1.0f, ref_idx), fadd<X>(1.0f, ref_idx)) : __fdividef(1.0f - ref_idx, 1.0f + ref_idx);
    r0 = fmul<X>(r0, r0);
    return fadd<X>(r0, fmul<X>(fsub<X>(1.0f, r0), pow5_f<X>(fsub<X>(1.0f, cosine))));
}

// One path vertex: consumes the closest hit of segment (o,d), adds emission to L, samples the next segment.
// Returns false when the path ends here (miss is handled by the caller).
template <int MODE, bool X>
__device__ __forceinline__ bool shade_vertex(const DevScene& sc, int tex_mode, const HitInfo& hit, V3& o, V3& d, V3& beta, V3& L,
                                             Rng<MODE>& rng) {
    V3 p, n;
    bool front;
    int mat;
    float u = 0.f, v = 0.f;
    if (hit.type == 0) {
        float4 s = __ldg(sc.spheres + hit.slot);
        mat = __ldg(sc.sphere_meta + hit.slot).x;
        p = add3<X>(o, scale3<X>(hit.t, d));
        V3 outward = div3<X>(sub3<X>(p, mk3(s.x, s.y, s.z)), s.w);
        front = dot3<X>(d, outward) < 0.0f;
        n = front ? outward : neg3(outward);
        if (mat == sc.tex_mat) sphere_uv<X>(outward, u, v);
    } else {
        const float4* q = sc.planes + 5 * (size_t)hit.slot;
        float4 nD = __ldg(q), u4 = __ldg(q + 2);
        mat = __float_as_int(u4.w);
        V3 pn = mk3(nD.x, nD.y, nD.z);
        p = add3<X>(o, scale3<X>(hit.t, d));
        front = dot3<X>(d, pn) < 0.0f;
        n = front ? pn : neg3(pn);
        if (mat == sc.tex_mat) {
            float4 wT = __ldg(q + 1), v4 = __ldg(q + 3), b4 = __ldg(q + 4);
            V3 ph = sub3<X>(p, mk3(b4.x, b4.y, b4.z));
            V3 w = mk3(wT.x, wT.y, wT.z);
            u = dot3<X>(w, cross3<X>(ph, mk3(v4.x, v4.y, v4.z)));
            v = dot3<X>(w, cross3<X>(mk3(u4.x, u4.y, u4.z), ph));
        }
    }
    const float4* mp = sc.mats + 3 * (size_t)mat;
    float4 m0 = __ldg(mp), m1 = __ldg(mp + 1);
    V3 albedo = mk3(m0.x, m0.y, m0.z);
    int mtype = __float_as_int(m0.w);
    if (mat == sc.tex_mat) {
        V3 tx;
        if (tex_mode == 0) {
            tx = tex_cpu<X>(sc, u, v);
        } else {
            float4 t4 = tex2D<float4>(sc.tex_obj, u, v);
            tx = mk3(t4.x, t4.y, t4.z);
        }
        albedo = mul3<X>(albedo, tx);
    }
    L = add3<X>(L, mul3<X>(beta, mk3(m1.x, m1.y, m1.z)));  // final += beta * emit (src/camera.cu:238)

    V3 att = albedo;
    bool alive = true;
    if (mtype == kLambertian || mtype == kMetal) {
        bool diffuse = true;
        if (mtype == kMetal) diffuse = !(rng.next() < 0.8f);
        if (diffuse) {
            V3 dir = sample_hemisphere<MODE, X>(n, rng);
            if (fabsf(dir.x) < 1e-8f && fabsf(dir.y) < 1e-8f && fabsf(dir.z) < 1e-8f) dir = n;
            d = dir;
        } else {
            V3 refl = reflect3<X>(unit3<X>(d), n);
            V3 b = sample_ball<MODE, X>(rng);
            d = add3<X>(refl, scale3<X>(m1.w, b));
            alive = dot3<X>(d, n) > 0.0f;
        }
        o = p;
    } else if (mtype == kDielectric) {
        float4 m2 = __ldg(mp + 2);
        float ir = m2.w;
        float ratio = front ? recip<X>(ir) : ir;
        V3 ud = unit3<X>(d);
        float cos_t = fminf(dot3<X>(neg3(ud), n), 1.0f);
        float sin2 = fsub<X>(1.0f, fmul<X>(cos_t, cos_t));
        float sin_t = X ? __fsqrt_rn(sin2) : fast_sqrt(sin2);
        bool tir = fmul<X>(ratio, sin_t) > 1.0f;
        V3 dir;
        if (tir || schlick<X>(cos_t, ratio) > rng.next()) dir = reflect3<X>(ud, n);
        else dir = refract3<X>(ud, n, ratio);
        att = mk3(1.0f, 1.0f, 1.0f);
        if (!front) {
            float dist = len3<X>(sub3<X>(p, o));
            att = mk3(exp_f<X>(fmul<X>(-m2.x, dist)), exp_f<X>(fmul<X>(-m2.y, dist)), exp_f<X>(fmul<X>(-m2.z, dist)));
            if (X) att = mul3<X>(mk3(1.0f, 1.0f, 1.0f), att);
        }
        float pr = fmaxf(att.x, fmaxf(att.y, att.z));
        if (rng.next() > pr) return false;
        att = scale3<X>(recip<X>(pr), att);
        float sgn = dot3<X>(dir, n) > 0.0f ? 1.0f : -1.0f;
        o = add3<X>(p, scale3<X>(sgn, scale3<X>(1e-4f, n)));
        d = dir;
    } else {
        return false;  // DIFFUSE_LIGHT: emission only
    }
    beta = mul3<X>(beta, att);
    return alive;
}

}  // namespace rtb
