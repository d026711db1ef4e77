// emul.cpp -- TEST INFRASTRUCTURE ONLY: host build of the product's device headers (see emul_shim.h) behind a
// small C surface, so `pytest -m "not gpu"` can compare the kernels' traversal + shading logic with the oracle.
#include "emul_shim.h"
// clang-format off
#include "../../ray-tracing-practice_b200/csrc/path.cuh"
#include "../../ray-tracing-practice_b200/csrc/scene_pack.h"
// clang-format on
#include <atomic>
#include <thread>
#include <vector>

using namespace rtb;

struct EmuScene {
    PackedScene pk;
    std::vector<float> tex;
    DevScene dev{};
};

extern "C" {

int emu_scene_create(const void* sp, int S, const void* pl, int P, const void* mt, int M, const float* tex, int tw, int th,
                     int tex_mat, void** out) {
    EmuScene* s = new EmuScene();
    s->pk = pack_scene((const SphereIn*)sp, S, (const PlaneIn*)pl, P, (const MatIn*)mt, M);
    s->dev.nodes = (const float4*)s->pk.bvh.nodes.data();
    s->dev.spheres = (const float4*)s->pk.spheres.data();
    s->dev.sphere_meta = (const int2*)s->pk.sphere_meta.data();
    s->dev.planes = (const float4*)s->pk.planes.data();
    s->dev.mats = (const float4*)s->pk.mats.data();
    s->dev.tex_mat = -1;
    if (tex) {
        s->tex.assign(tex, tex + (size_t)tw * th * 4);
        s->dev.tex = (const float4*)s->tex.data();
        s->dev.tex_w = tw; s->dev.tex_h = th; s->dev.tex_mat = tex_mat;
    }
    s->dev.num_nodes = (int)s->pk.bvh.nodes.size();
    s->dev.top_nodes = 0;
    *out = s;
    return 0;
}
void emu_scene_destroy(void* h) { delete (EmuScene*)h; }

// structural soundness of the wide BVH + stats: out = nodes, depth, max_stack, leaves
int emu_bvh_check(const void* sp, int S, const void* pl, int P, int* out4, float* sah) {
    std::vector<PrimBox> prims = make_prim_boxes((const SphereIn*)sp, S, (const PlaneIn*)pl, P);
    WideBvh b = build_wide_bvh(prims, 4);
    out4[0] = (int)b.nodes.size(); out4[1] = b.depth; out4[2] = b.max_stack; out4[3] = b.leaves;
    if (sah) *sah = b.sah_cost;
    return validate_wide_bvh(b, prims);
}

static int hw_threads() { int n = (int)std::thread::hardware_concurrency(); return n < 1 ? 1 : n; }

int emu_trace(void* h, const float* rays6, int n, float tmin, float tmax, int exact, int* type, int* index, float* t) {
    EmuScene* s = (EmuScene*)h;
    int nth = n < 4096 ? 1 : hw_threads();
    auto work = [&](int t0) {
        std::vector<float> st(kSmemStack);
        std::vector<int> sr(kSmemStack);
        WorkCount wc{};
        for (int i = t0; i < n; i += nth) {
            const float* r = rays6 + 6 * (size_t)i;
            HitInfo hit;
            if (exact) trace_closest<true, false>(s->dev, nullptr, st.data(), sr.data(), mk3(r[0], r[1], r[2]), mk3(r[3], r[4], r[5]), tmin, tmax, hit, wc);
            else trace_closest<false, false>(s->dev, nullptr, st.data(), sr.data(), mk3(r[0], r[1], r[2]), mk3(r[3], r[4], r[5]), tmin, tmax, hit, wc);
            type[i] = hit.type; index[i] = hit.orig; t[i] = hit.type < 0 ? 0.f : hit.t;
        }
    };
    std::vector<std::thread> th;
    for (int k = 1; k < nth; ++k) th.emplace_back(work, k);
    work(0);
    for (auto& x : th) x.join();
    return 0;
}

}  // extern "C"

template <int MODE, bool X>
static void render_rows(EmuScene* s, const CameraPod& cam, int tex_mode, uint2 key, float* fb, std::atomic<int>& row,
                        unsigned long long* counters6) {
    std::vector<float> st(kSmemStack);
    std::vector<int> sr(kSmemStack);
    WorkCount wc{};
    unsigned long long tot[6] = {0, 0, 0, 0, 0, 0};
    for (;;) {
        int y = row.fetch_add(1);
        if (y >= cam.H) break;
        for (int x = 0; x < cam.W; ++x) {
            V3 acc = mk3(0, 0, 0);
            uint32_t base = pixel_seed_base(x, y, cam.W);
            for (int k = 0; k < cam.spp; ++k) {
                Rng<MODE> rng;
                PathState ps;
                begin_path<MODE, X>(cam, x, y, k, base, key, rng, ps);
                wc.samples++;
                while (!advance_path<MODE, X, true>(s->dev, cam, tex_mode, nullptr, st.data(), sr.data(), ps, rng, wc)) {}
                acc = add3<X>(acc, ps.L);
            }
            float* o = fb + 3 * ((size_t)y * cam.W + x);
            o[0] = acc.x; o[1] = acc.y; o[2] = acc.z;
        }
        tot[0] += wc.segments; tot[1] += wc.nodes; tot[2] += wc.spheres; tot[3] += wc.planes; tot[4] += wc.hits; tot[5] += wc.samples;
        wc = WorkCount{};
    }
    static std::atomic_flag lock = ATOMIC_FLAG_INIT;
    while (lock.test_and_set()) {}
    for (int i = 0; i < 6; ++i) counters6[i] += tot[i];
    lock.clear();
}

extern "C" {

int emu_render(void* h, const void* cam76, int rng_mode, int exact, int tex_mode, unsigned long long seed, float* fb,
               unsigned long long* counters6) {
    EmuScene* s = (EmuScene*)h;
    CameraPod cam;
    memcpy(&cam, cam76, 76);
    uint2 key = make_uint2((unsigned)(seed & 0xffffffffu), (unsigned)(seed >> 32));
    for (int i = 0; i < 6; ++i) counters6[i] = 0;
    std::atomic<int> row{0};
    int nth = hw_threads();
    auto work = [&]() {
        if (rng_mode == 1 && exact) render_rows<kRngReference, true>(s, cam, tex_mode, key, fb, row, counters6);
        else if (rng_mode == 1) render_rows<kRngReference, false>(s, cam, tex_mode, key, fb, row, counters6);
        else if (exact) render_rows<kRngPhilox, true>(s, cam, tex_mode, key, fb, row, counters6);
        else render_rows<kRngPhilox, false>(s, cam, tex_mode, key, fb, row, counters6);
    };
    std::vector<std::thread> th;
    for (int k = 1; k < nth; ++k) th.emplace_back(work);
    work();
    for (auto& x : th) x.join();
    return 0;
}

}  // extern "C"
