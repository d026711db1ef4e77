#include "scene_pack.h"

#include "lbvh.cuh"

#include <chrono>
#include <cstring>

namespace rtb {
namespace {
inline float i2f(int v) { float f; std::memcpy(&f, &v, 4); return f; }
}  // namespace

PackedScene pack_scene(const SphereIn* sp, int S, const PlaneIn* pl, int P, const MatIn* mt, int M) {
    PackedScene ps;
    auto t0 = std::chrono::steady_clock::now();
    std::vector<PrimBox> prims = make_prim_boxes(sp, S, pl, P);
    ps.bvh = build_wide_bvh(prims, kMaxLeafPrims);
    ps.build_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    const WideBvh& bvh = ps.bvh;
    ps.spheres.resize(bvh.sphere_order.size() * 4);
    ps.sphere_meta.resize(bvh.sphere_order.size() * 2);
    for (size_t k = 0; k < bvh.sphere_order.size(); ++k) {
        const SphereIn& s = sp[bvh.sphere_order[k]];
        float* o = &ps.spheres[4 * k];
        o[0] = s.c[0]; o[1] = s.c[1]; o[2] = s.c[2]; o[3] = s.r;
        ps.sphere_meta[2 * k] = s.mat;
        ps.sphere_meta[2 * k + 1] = bvh.sphere_order[k];
    }
    ps.planes.resize(bvh.plane_order.size() * 20);
    ps.planes_fast.resize(bvh.plane_order.size() * 12);
    for (size_t k = 0; k < bvh.plane_order.size(); ++k) {
        const PlaneIn& p = pl[bvh.plane_order[k]];
        float* o = &ps.planes[20 * k];
        o[0] = p.n[0]; o[1] = p.n[1]; o[2] = p.n[2]; o[3] = p.D;
        o[4] = p.w[0]; o[5] = p.w[1]; o[6] = p.w[2]; o[7] = i2f(p.type);
        o[8] = p.u[0]; o[9] = p.u[1]; o[10] = p.u[2]; o[11] = i2f(p.mat);
        o[12] = p.v[0]; o[13] = p.v[1]; o[14] = p.v[2]; o[15] = i2f(bvh.plane_order[k]);
        o[16] = p.base[0]; o[17] = p.base[1]; o[18] = p.base[2]; o[19] = 0.f;
        lbvh::plane_fast_record(p, &ps.planes_fast[12 * k]);
    }
    ps.mats.resize((size_t)M * 12);
    for (int i = 0; i < M; ++i) {
        float* o = &ps.mats[12 * (size_t)i];
        o[0] = mt[i].albedo[0]; o[1] = mt[i].albedo[1]; o[2] = mt[i].albedo[2]; o[3] = i2f(mt[i].type);
        o[4] = mt[i].emit[0]; o[5] = mt[i].emit[1]; o[6] = mt[i].emit[2]; o[7] = mt[i].fuzz;
        o[8] = mt[i].absorb[0]; o[9] = mt[i].absorb[1]; o[10] = mt[i].absorb[2]; o[11] = mt[i].ir;
    }
    return ps;
}

}  // namespace rtb
