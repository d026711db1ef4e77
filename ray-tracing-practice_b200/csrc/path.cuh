// path.cuh -- one sample's path state and the two steps every kernel variant shares: start a sample
// (CameraData::get_ray, include/camera.cuh:97-110) and advance it by ONE segment (one iteration of ray_color's
// loop, src/camera.cu:223-249).  The megakernel, the wavefront kernels and the CPU emulation used by the tests
// (tests/emul) all call exactly these functions.
#pragma once
#include "shade.cuh"

namespace rtb {

// image of the reference's CameraData (include/camera.cuh:86-96), 76 bytes
struct CameraPod {
    float origin[3], p00[3], du[3], dv[3], bg[3];
    int W, H, spp, depth;
};
static_assert(sizeof(CameraPod) == 76, "CameraData layout");

struct PathState {
    V3 o, d, beta, L;
    int depth;
};

__device__ __forceinline__ V3 ld3(const float* p) { return mk3(p[0], p[1], p[2]); }

// seeds the stream for sample `sample` of pixel (x,y) and shoots its camera ray; direction is NOT normalised
template <int MODE, bool X>
__device__ __forceinline__ void begin_path(const CameraPod& c, int x, int y, int sample, uint32_t pixel_base, uint2 key, Rng<MODE>& rng,
                                           PathState& ps) {
    rng.begin_sample(pixel_base, (uint32_t)sample, (uint32_t)(y * c.W + x), key);
    V3 du = ld3(c.du), dv = ld3(c.dv);
    V3 pc = add3<X>(add3<X>(ld3(c.p00), scale3<X>((float)x, du)), scale3<X>((float)y, dv));
    float ox = fsub<X>(rng.next(), 0.5f);
    float oy = fsub<X>(rng.next(), 0.5f);
    V3 pt = add3<X>(add3<X>(pc, scale3<X>(ox, du)), scale3<X>(oy, dv));
    ps.o = ld3(c.origin);
    ps.d = sub3<X>(pt, ps.o);
    ps.beta = mk3(1, 1, 1);
    ps.L = mk3(0, 0, 0);
    ps.depth = 0;
}

// the reference's per-pixel seed base: wang_hash(i*W + j) with i = x (src/camera.cu:25)
__device__ __forceinline__ uint32_t pixel_seed_base(int x, int y, int W) { return wang_hash_u32((uint32_t)x * (uint32_t)W + (uint32_t)y); }

// second half of a segment, given its closest hit: miss -> background, else emission + scatter.
// Returns true when the sample is finished (ps.L is its radiance).
template <int MODE, bool X, bool COUNT>
__device__ __forceinline__ bool finish_segment(const DevScene& sc, const CameraPod& cam, int tex_mode, const HitInfo& hit, PathState& ps,
                                               Rng<MODE>& rng, WorkCount& wc) {
    wc.segments++;
    bool alive;
    if (hit.type < 0) {
        ps.L = add3<X>(ps.L, mul3<X>(ps.beta, ld3(cam.bg)));
        alive = false;
    } else {
        if (COUNT) wc.hits++;
        if (MODE == kRngPhilox) rng.begin_bounce((uint32_t)ps.depth + 1u);
        alive = shade_vertex<MODE, X>(sc, tex_mode, hit, ps.o, ps.d, ps.beta, ps.L, rng);
    }
    ++ps.depth;
    return !alive || ps.depth >= cam.depth;
}

// one whole segment: closest hit (Interval(0.001f, 1e30f), src/camera.cu:226), then finish_segment
template <int MODE, bool X, bool COUNT>
__device__ __forceinline__ bool advance_path(const DevScene& sc, const CameraPod& cam, int tex_mode, const float4* s_top, float* s_st,
                                             int* s_sr, PathState& ps, Rng<MODE>& rng, WorkCount& wc) {
    HitInfo hit;
    trace_closest<X, COUNT>(sc, s_top, s_st, s_sr, ps.o, ps.d, 0.001f, 1e30f, hit, wc);
    return finish_segment<MODE, X, COUNT>(sc, cam, tex_mode, hit, ps, rng, wc);
}

}  // namespace rtb
