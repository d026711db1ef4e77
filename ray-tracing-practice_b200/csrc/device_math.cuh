// device_math.cuh -- vector math with two arithmetic policies, and the two random streams.
//
// Policy X=true ("exact"): every operation is a separately rounded IEEE op in the order the reference's HOST build
// performs it (include/vec3.h; x86-64 without FMA contraction), including its float->double->float excursions
// (vec3 operator/ multiplies by a DOUBLE reciprocal, include/vec3.h:97).  __fmul_rn/__fadd_rn are never fused by
// nvcc, so results are bit-identical to oracle/oracle.cpp.  Used by rt_trace_primary and by the parity render mode.
// Policy X=false ("fast"): plain fp32 with FMA contraction and reciprocal multiplies -- the production path.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rtb {

struct V3 {
    float x, y, z;
};
__device__ __forceinline__ V3 mk3(float x, float y, float z) { V3 v; v.x = x; v.y = y; v.z = z; return v; }
__device__ __forceinline__ V3 neg3(V3 a) { return mk3(-a.x, -a.y, -a.z); }

template <bool X> __device__ __forceinline__ float fmul(float a, float b) { return X ? __fmul_rn(a, b) : a * b; }
template <bool X> __device__ __forceinline__ float fadd(float a, float b) { return X ? __fadd_rn(a, b) : a + b; }
template <bool X> __device__ __forceinline__ float fsub(float a, float b) { return X ? __fsub_rn(a, b) : a - b; }

template <bool X> __device__ __forceinline__ V3 add3(V3 a, V3 b) { return mk3(fadd<X>(a.x, b.x), fadd<X>(a.y, b.y), fadd<X>(a.z, b.z)); }
template <bool X> __device__ __forceinline__ V3 sub3(V3 a, V3 b) { return mk3(fsub<X>(a.x, b.x), fsub<X>(a.y, b.y), fsub<X>(a.z, b.z)); }
template <bool X> __device__ __forceinline__ V3 mul3(V3 a, V3 b) { return mk3(fmul<X>(a.x, b.x), fmul<X>(a.y, b.y), fmul<X>(a.z, b.z)); }
template <bool X> __device__ __forceinline__ V3 scale3(float t, V3 v) { return mk3(fmul<X>(t, v.x), fmul<X>(t, v.y), fmul<X>(t, v.z)); }
// include/vec3.h:99 -- (a0*b0 + a1*b1) + a2*b2
template <bool X> __device__ __forceinline__ float dot3(V3 a, V3 b) {
    if (X) return __fadd_rn(__fadd_rn(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y)), __fmul_rn(a.z, b.z));
    return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x));
}
template <bool X> __device__ __forceinline__ float len2_3(V3 a) { return dot3<X>(a, a); }
// include/vec3.h:101-103
template <bool X> __device__ __forceinline__ V3 cross3(V3 a, V3 b) {
    return mk3(fsub<X>(fmul<X>(a.y, b.z), fmul<X>(a.z, b.y)), fsub<X>(fmul<X>(a.z, b.x), fmul<X>(a.x, b.z)),
               fsub<X>(fmul<X>(a.x, b.y), fmul<X>(a.y, b.x)));
}
// Fast-mode (production) special functions: the SFU approximations (1-2 ulp, 2 instructions) instead of the IEEE-rounded
// sequences with their slow-path calls.  Exact mode never uses them.  RTB_FAST_SFU=0 restores the IEEE versions.
#ifndef RTB_FAST_SFU
#define RTB_FAST_SFU 1
#endif
__device__ __forceinline__ float fast_sqrt(float x) {
#if RTB_FAST_SFU && defined(__CUDA_ARCH__)
    float s;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(s) : "f"(x));
    return s;
#else
    return sqrtf(x);
#endif
}
__device__ __forceinline__ float fast_rcp(float x) {
#if RTB_FAST_SFU && defined(__CUDA_ARCH__)
    return __fdividef(1.0f, x);
#else
    return __frcp_rn(x);
#endif
}
__device__ __forceinline__ float fast_cbrt01(float u) {  // u in [0, 1]
#if RTB_FAST_SFU && defined(__CUDA_ARCH__)
    return exp2f(__log2f(u) * (1.0f / 3.0f));  // MUFU.LG2, FMUL, MUFU.EX2; u = 0 -> 0
#else
    return cbrtf(u);
#endif
}
__device__ __forceinline__ void fast_sincos2pi(float u, float* sn, float* cs) {  // sin / cos of 2*pi*u, u in [0, 1)
#ifndef RTB_FAST_SINCOS
#define RTB_FAST_SINCOS 1
#endif
#if RTB_FAST_SFU && RTB_FAST_SINCOS && defined(__CUDA_ARCH__)
    __sincosf(6.28318530718f * u, sn, cs);  // MUFU.SIN / MUFU.COS: abs. error ~5e-7 on [0, 2 pi], plenty for a sampling direction
#else
    sincospif(2.0f * u, sn, cs);
#endif
}
// reciprocal as vec3::operator/ forms it (include/vec3.h:97): float(1.0 / double(t))
template <bool X> __device__ __forceinline__ float recip(float t) {
    if (X) return (float)(1.0 / (double)t);
    return fast_rcp(t);
}
template <bool X> __device__ __forceinline__ V3 div3(V3 v, float t) { return scale3<X>(recip<X>(t), v); }
template <bool X> __device__ __forceinline__ float len3(V3 a) { return X ? __fsqrt_rn(len2_3<X>(a)) : fast_sqrt(len2_3<X>(a)); }
template <bool X> __device__ __forceinline__ V3 unit3(V3 v) {
    if (X) return div3<true>(v, len3<true>(v));
    return scale3<false>(rsqrtf(len2_3<false>(v)), v);
}
// include/vec3.h:63 -- v - (2*dot(v,n))*n
template <bool X> __device__ __forceinline__ V3 reflect3(V3 v, V3 n) {
    return sub3<X>(v, scale3<X>(fmul<X>(2.0f, dot3<X>(v, n)), n));
}
// include/vec3.h:65-70
template <bool X> __device__ __forceinline__ V3 refract3(V3 v, V3 n, float eta) {
    float c = fminf(dot3<X>(neg3(v), n), 1.0f);
    V3 perp = scale3<X>(eta, add3<X>(v, scale3<X>(c, n)));
    float k = -__fsqrt_rn(fabsf(fsub<X>(1.0f, len2_3<X>(perp))));  // 1.0-x in double then narrowed == fp32 rn subtraction
    return add3<X>(perp, scale3<X>(k, n));
}
// libm calls of the reference host build (glibc: correctly rounded in practice) restated through double
template <bool X> __device__ __forceinline__ float exp_f(float x) {
    if (X) return (float)exp((double)x);
    return __expf(x);
}
template <bool X> __device__ __forceinline__ float pow5_f(float x) {
    if (X) { double d = x, d2 = d * d; return (float)(d2 * d2 * d); }
    float x2 = x * x;
    return x2 * x2 * x;
}

// ---------------------------------------------------------------- random streams
// (1) the reference stream (include/random_utils.h:7-23): seed = wang_hash(seed); value = float(seed) / 2^32
__device__ __forceinline__ uint32_t wang_hash_u32(uint32_t s) {
    s = (s ^ 61u) ^ (s >> 16);
    s *= 9u;
    s ^= s >> 4;
    s *= 0x27d4eb2du;
    s ^= s >> 15;
    return s;
}
// (2) Philox4x32-10 (Salmon et al., SC'11), counter = (pixel, sample, bounce, block), key = 64-bit seed
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
        uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += W0;
        key.y += W1;
    }
    return ctr;
}

enum { kRngPhilox = 0, kRngReference = 1 };

template <int MODE> struct Rng;

template <> struct Rng<kRngReference> {
    uint32_t s;
    // per-sample seeding exactly as render_kernel / render_cpu: src/camera.cu:25-28 (note x*W + y)
    __device__ __forceinline__ void begin_sample(uint32_t pixel_base, uint32_t sample, uint32_t, uint2) { s = wang_hash_u32(pixel_base + sample); }
    __device__ __forceinline__ void attach(uint32_t, uint32_t, uint2) {}
    __device__ __forceinline__ void begin_bounce(uint32_t) {}
    // the stream's whole state is one word: what a path parks between segments
    __device__ __forceinline__ uint32_t save() const { return s; }
    __device__ __forceinline__ void restore(uint32_t w) { s = w; }
    __device__ __forceinline__ float next() {
        s = wang_hash_u32(s);
        return __fdiv_rn((float)s, 4294967296.0f);
    }
};

template <> struct Rng<kRngPhilox> {
    // ONE Philox block = the four uniforms of one path vertex: the camera jitter of a sample (bounce 0) or one scatter
    // event (bounce b >= 1 needs at most 4: lobe choice + 3 for the fuzz ball).  The block is generated eagerly at
    // begin_sample / begin_bounce -- a single call site that every lane of a shading pass reaches together -- and
    // next() only hands out its words, so no Philox code is replicated into the divergent material branches.
    uint4 c;   // counter (pixel, sample, bounce, 0)
    uint4 v;   // the block
    uint2 key;
    int used;
    // binds the stream to (pixel, sample) without generating anything (what a path does when it resumes)
    __device__ __forceinline__ void attach(uint32_t sample, uint32_t pixel, uint2 k) {
        key = k;
        c = make_uint4(pixel, sample, 0u, 0u);
        used = 4;
    }
    __device__ __forceinline__ void begin_sample(uint32_t, uint32_t sample, uint32_t pixel, uint2 k) {
        attach(sample, pixel, k);
        v = philox4x32_10(c, key);
        used = 0;
    }
    // counter-based: nothing to park, the counter is rebuilt from (pixel, sample, bounce)
    __device__ __forceinline__ uint32_t save() const { return 0u; }
    __device__ __forceinline__ void restore(uint32_t) {}
    __device__ __forceinline__ void begin_bounce(uint32_t bounce) {
        c.z = bounce;
        c.w = 0u;
        v = philox4x32_10(c, key);
        used = 0;
    }
    __device__ __forceinline__ float next() {
        uint32_t r = used == 0 ? v.x : used == 1 ? v.y : used == 2 ? v.z : v.w;  // `used` folds to a constant per path
        used++;
        return (float)(r >> 8) * (1.0f / 16777216.0f);  // [0,1)
    }
};

// uniform point in the unit ball.  Reference stream: the rejection loop of include/random_utils.h:25-32 with
// the draw order of the reference HOST build (first draw -> z, see oracle/oracle.cpp ball()).
template <int MODE, bool X> __device__ __forceinline__ V3 sample_ball(Rng<MODE>& rng) {
    if (MODE == kRngReference) {
        for (;;) {
            float z = fadd<X>(-1.0f, fmul<X>(2.0f, rng.next()));
            float y = fadd<X>(-1.0f, fmul<X>(2.0f, rng.next()));
            float x = fadd<X>(-1.0f, fmul<X>(2.0f, rng.next()));
            V3 c = mk3(x, y, z);
            if (len2_3<X>(c) < 1.0f) return c;
        }
    } else {
        float u1 = rng.next(), u2 = rng.next(), u3 = rng.next();
        float z = 1.0f - 2.0f * u1;
        float r = fast_sqrt(fmaxf(0.0f, 1.0f - z * z));
        float sn, cs;
        fast_sincos2pi(u2, &sn, &cs);
        float rad = fast_cbrt01(u3);
        return mk3(rad * r * cs, rad * r * sn, rad * z);
    }
}
// uniform direction = unit_vector(random_in_unit_sphere) (include/random_utils.h:34)
template <int MODE, bool X> __device__ __forceinline__ V3 sample_unit(Rng<MODE>& rng) {
    if (MODE == kRngReference) {
        return unit3<X>(sample_ball<MODE, X>(rng));
    } else {
        float u1 = rng.next(), u2 = rng.next();
        float z = 1.0f - 2.0f * u1;
        float r = fast_sqrt(fmaxf(0.0f, 1.0f - z * z));
        float sn, cs;
        fast_sincos2pi(u2, &sn, &cs);
        return mk3(r * cs, r * sn, z);
    }
}
// include/random_utils.h:36-42
template <int MODE, bool X> __device__ __forceinline__ V3 sample_hemisphere(V3 n, Rng<MODE>& rng) {
    V3 d = sample_unit<MODE, X>(rng);
    return dot3<X>(d, n) > 0.0f ? d : neg3(d);
}

}  // namespace rtb
