// emul.cpp -- TEST INFRASTRUCTURE ONLY: host build of the product's device headers (see emul_shim.h) behind a
// small C surface, so `pytest -m "not gpu"` can compare the kernels' traversal + shading logic with the oracle.
#include "emul_shim.h"
// clang-format off
#include "../../ray-tracing-practice_b200/csrc/path.cuh"
#include "../../ray-tracing-practice_b200/csrc/scene_pack.h"
#include "../../ray-tracing-practice_b200/csrc/lbvh.cuh"
// clang-format on
#include <algorithm>
#include <numeric>
#include <atomic>
#include <thread>
#include <vector>

using namespace rtb;

struct EmuScene {
    PackedScene pk;
    std::vector<float> tex;
    DevScene dev{};
};

// The device BVH build (csrc/bvh_gpu.cu) restated as plain loops over the SAME per-element functions (csrc/lbvh.cuh):
// one loop per kernel, a stable sort where the device uses cub's radix sort.  Replaces pk's tree and primitive streams.
static void lbvh_build_host(const SphereIn* sp, int S, const PlaneIn* pl, int P, PackedScene& pk) {
    using lbvh::Box;
    int n = S + P;
    WideBvh& out = pk.bvh;
    out = WideBvh{};
    pk.spheres.assign((size_t)S * 4, 0.f);
    pk.sphere_meta.assign((size_t)S * 2, 0);
    pk.planes.assign((size_t)P * 20, 0.f);
    pk.planes_fast.assign((size_t)P * 12, 0.f);
    if (n == 0) {
        WideNode root;
        for (int a = 0; a < 3; ++a)
            for (int i = 0; i < 4; ++i) { root.lo[a][i] = __builtin_huge_valf(); root.hi[a][i] = -__builtin_huge_valf(); }
        for (int i = 0; i < 4; ++i) { root.child[i] = kEmptyChild; root.pad[i] = 0; }
        out.nodes.push_back(root);
        out.depth = 1;
        return;
    }
    std::vector<Box> prim_boxes(n);
    float cmin[3] = {__builtin_huge_valf(), __builtin_huge_valf(), __builtin_huge_valf()}, cmax[3] = {-cmin[0], -cmin[0], -cmin[0]};
    for (int i = 0; i < n; ++i) {  // k_boxes
        prim_boxes[i] = lbvh::prim_box(sp, S, pl, i);
        for (int a = 0; a < 3; ++a) {
            float c = 0.5f * (prim_boxes[i].lo[a] + prim_boxes[i].hi[a]);
            // through the same ordered-uint encoding the device reduces with
            float cc = lbvh::ordered_to_float(lbvh::float_to_ordered(c));
            cmin[a] = std::min(cmin[a], cc);
            cmax[a] = std::max(cmax[a], cc);
        }
    }
    // k_hoist + the host's pick (bvh_gpu.cu)
    int hoist[4] = {-1, -1, -1, -1}, nh = 0;
    if (!(getenv("RTB_HOIST") && atoi(getenv("RTB_HOIST")) == 0)) {
        Box sb = prim_boxes[0];
        for (int i = 1; i < n; ++i) sb = lbvh::box_union(sb, prim_boxes[i]);
        std::vector<std::pair<float, int>> byarea;
        for (int i = 0; i < n; ++i)
            if (lbvh::hoist_candidate(prim_boxes[i], lbvh::box_area(sb))) byarea.emplace_back(-lbvh::box_area(prim_boxes[i]), i);
        if ((int)byarea.size() <= lbvh::kHoistCandCap) {
            std::sort(byarea.begin(), byarea.end());
            for (size_t k = 0; k < byarea.size() && nh < lbvh::kMaxHoist; ++k) hoist[nh++] = byarea[k].second;
        }
    }
    const int n_all = n;
    std::vector<uint64_t> keys(n);
    std::vector<unsigned> vals(n);
    for (int i = 0; i < n; ++i) {  // k_keys
        const bool h = i == hoist[0] || i == hoist[1] || i == hoist[2] || i == hoist[3];
        keys[i] = h ? ~0ull : lbvh::morton_key(prim_boxes[i], cmin, cmax);
        vals[i] = (unsigned)i;
    }
    std::stable_sort(vals.begin(), vals.end(), [&](unsigned a, unsigned b) { return keys[a] < keys[b]; });
    std::vector<uint64_t> skeys(n);
    for (int p = 0; p < n; ++p) skeys[p] = keys[vals[p]];
    std::vector<int> sphere_pos(n);
    for (int p = 0, acc = 0; p < n; ++p) { sphere_pos[p] = acc; acc += vals[p] < (unsigned)S; }  // k_flags + scan
    std::vector<Box> boxes(2 * (size_t)n);
    std::vector<int> leaf_refs(n);
    auto i2f = [](int v) { float f; memcpy(&f, &v, 4); return f; };
    const int nt = n - nh;  // primitives in the tree; sorted positions >= nt hold the hoisted ones
    for (int p = 0; p < n; ++p) {  // k_pack
        int id = (int)vals[p];
        if (p < nt) boxes[(nt - 1) + p] = prim_boxes[id];
        if (id < S) {
            int pos = sphere_pos[p];
            float* o = &pk.spheres[4 * (size_t)pos];
            o[0] = sp[id].c[0]; o[1] = sp[id].c[1]; o[2] = sp[id].c[2]; o[3] = sp[id].r;
            pk.sphere_meta[2 * (size_t)pos] = sp[id].mat;
            pk.sphere_meta[2 * (size_t)pos + 1] = id;
            leaf_refs[p] = lbvh::make_leaf_ref(0, pos);
        } else {
            int pos = p - sphere_pos[p];
            const PlaneIn& q = pl[id - S];
            float* o = &pk.planes[20 * (size_t)pos];
            o[0] = q.n[0]; o[1] = q.n[1]; o[2] = q.n[2]; o[3] = q.D;
            o[4] = q.w[0]; o[5] = q.w[1]; o[6] = q.w[2]; o[7] = i2f(q.type);
            o[8] = q.u[0]; o[9] = q.u[1]; o[10] = q.u[2]; o[11] = i2f(q.mat);
            o[12] = q.v[0]; o[13] = q.v[1]; o[14] = q.v[2]; o[15] = i2f(id - S);
            o[16] = q.base[0]; o[17] = q.base[1]; o[18] = q.base[2]; o[19] = 0.f;
            lbvh::plane_fast_record(q, &pk.planes_fast[12 * (size_t)pos]);
            leaf_refs[p] = lbvh::make_leaf_ref(1, pos);
        }
    }
    for (int k = 0; k < nh; ++k) out.hoisted_refs.push_back(leaf_refs[nt + k]);
    auto finish_orders = [&]() {
        out.leaves = n_all;
        out.sphere_order.resize(S);
        out.plane_order.resize(P);
        for (int k = 0; k < S; ++k) out.sphere_order[k] = pk.sphere_meta[2 * (size_t)k + 1];
        for (int k = 0; k < P; ++k) { int v; memcpy(&v, &pk.planes[20 * (size_t)k + 15], 4); out.plane_order[k] = v; }
    };
    if (nt == 0) {  // everything hoisted
        WideNode root;
        for (int a = 0; a < 3; ++a)
            for (int i = 0; i < 4; ++i) { root.lo[a][i] = __builtin_huge_valf(); root.hi[a][i] = -__builtin_huge_valf(); }
        for (int i = 0; i < 4; ++i) { root.child[i] = kEmptyChild; root.pad[i] = 0; }
        out.nodes.push_back(root);
        out.depth = 1;
        finish_orders();
        return;
    }
    n = nt;  // from here on: the tree
    std::vector<int2> children(n);
    std::vector<int> parent(2 * (size_t)n, -1);
    for (int i = 0; i < n - 1; ++i) {  // k_radix
        int l, r;
        lbvh::radix_node(skeys.data(), n, i, l, r);
        children[i] = make_int2(l, r);
        parent[l] = i;
        parent[r] = i;
    }
    std::vector<unsigned> arrived(n, 0);
    std::vector<lbvh::CollapseDp> dp(n);
    for (int p = 0; p < n && n > 1; ++p) {  // k_fit
        int cur = parent[(n - 1) + p];
        while (true) {
            if (arrived[cur]++ == 0) break;
            lbvh::fit_node(children.data(), parent.data(), boxes.data(), dp.data(), n, cur);
            if (cur == 0) break;
            cur = parent[cur];
        }
    }
    static const bool greedy = getenv("RTB_LBVH_COLLAPSE") && atoi(getenv("RTB_LBVH_COLLAPSE")) == 0;
    std::vector<int> cur_items{0}, next_items;
    std::vector<int> level_base, level_count;
    int base = 0;
    while (!cur_items.empty()) {  // k_collapse_count + scan + k_collapse_write
        int count = (int)cur_items.size();
        level_base.push_back(base);
        level_count.push_back(count);
        std::vector<int> kids(4 * (size_t)count), cnt(count), off(count);
        for (int k = 0; k < count; ++k)
            cnt[k] = greedy ? lbvh::collapse_pick(children.data(), boxes.data(), n, cur_items[k], &kids[4 * (size_t)k])
                            : lbvh::collapse_pick_dp(children.data(), dp.data(), n, cur_items[k], &kids[4 * (size_t)k]);
        int total = 0;
        for (int k = 0; k < count; ++k) { off[k] = total; total += cnt[k]; }
        out.nodes.resize((size_t)base + count);
        next_items.assign(total, -1);
        for (int k = 0; k < count; ++k)
            lbvh::collapse_write(boxes.data(), leaf_refs.data(), n, &kids[4 * (size_t)k], base + count, off[k],
                                 reinterpret_cast<float*>(&out.nodes[(size_t)base + k]), next_items.data());
        base += count;
        cur_items.swap(next_items);
    }
    std::vector<int> dep(out.nodes.size(), 0), need(out.nodes.size(), 0);
    for (int l = (int)level_base.size() - 1; l >= 0; --l)  // k_depth
        for (int k = 0; k < level_count[l]; ++k) {
            int w = level_base[l] + k;
            lbvh::depth_need(reinterpret_cast<const float*>(&out.nodes[w]), dep.data(), need.data(), dep[w], need[w]);
        }
    out.depth = dep[0];
    out.max_stack = need[0];
    finish_orders();
}

extern "C" {

// builder: 0 host binned-SAH (wide_bvh.cpp), 1 the device builder's steps run on the host (lbvh_build_host)
int emu_scene_create2(const void* sp, int S, const void* pl, int P, const void* mt, int M, const float* tex, int tw, int th,
                      int tex_mat, int builder, void** out);
int emu_scene_create(const void* sp, int S, const void* pl, int P, const void* mt, int M, const float* tex, int tw, int th,
                     int tex_mat, void** out) {
    return emu_scene_create2(sp, S, pl, P, mt, M, tex, tw, th, tex_mat, 0, out);
}

// structural soundness + stats of the device builder's tree: out = nodes, depth, max_stack, leaves; also checks that
// lbvh::prim_box reproduces make_prim_boxes bit for bit (return 100 + axis if not)
int emu_lbvh_check(const void* sp, int S, const void* pl, int P, int* out4) {
    std::vector<PrimBox> prims = make_prim_boxes((const SphereIn*)sp, S, (const PlaneIn*)pl, P);
    for (int i = 0; i < S + P; ++i) {
        lbvh::Box b = lbvh::prim_box((const SphereIn*)sp, S, (const PlaneIn*)pl, i);
        for (int a = 0; a < 3; ++a)
            if (memcmp(&b.lo[a], &prims[i].lo[a], 4) || memcmp(&b.hi[a], &prims[i].hi[a], 4)) return 100 + a;
    }
    PackedScene pk;
    lbvh_build_host((const SphereIn*)sp, S, (const PlaneIn*)pl, P, pk);
    out4[0] = (int)pk.bvh.nodes.size(); out4[1] = pk.bvh.depth; out4[2] = pk.bvh.max_stack; out4[3] = pk.bvh.leaves;
    return validate_wide_bvh(pk.bvh, prims);
}

int emu_scene_create2(const void* sp, int S, const void* pl, int P, const void* mt, int M, const float* tex, int tw, int th,
                      int tex_mat, int builder, void** out) {
    EmuScene* s = new EmuScene();
    // builder 0: host SAH tree; 1: the device builder's steps on the host; 2: host SAH tree + the sphere grid where it applies
    s->pk = pack_scene((const SphereIn*)sp, S, (const PlaneIn*)pl, P, (const MatIn*)mt, M, builder == 2 ? 1 : 0);
    if (builder == 1) lbvh_build_host((const SphereIn*)sp, S, (const PlaneIn*)pl, P, s->pk);
    if (s->pk.grid.enabled) {
        const GridAccel& g = s->pk.grid;
        s->dev.grid_on = 1;
        s->dev.grid_cells = (const uint2*)g.cells.data();
        s->dev.grid_refs = g.refs.data();
        for (int a = 0; a < 3; ++a) {
            s->dev.grid_lo[a] = g.lo[a]; s->dev.grid_hi[a] = g.hi[a]; s->dev.grid_cw[a] = g.cw[a];
            s->dev.grid_icw[a] = 1.0f / g.cw[a]; s->dev.grid_n[a] = g.n[a];
            s->dev.tree_lo[a] = g.tree_lo[a]; s->dev.tree_hi[a] = g.tree_hi[a];
        }
    }
    s->dev.nodes = (const float4*)s->pk.bvh.nodes.data();
    s->dev.spheres = (const float4*)s->pk.spheres.data();
    s->dev.sphere_meta = (const int2*)s->pk.sphere_meta.data();
    s->dev.planes = (const float4*)s->pk.planes.data();
    s->dev.planes_fast = (const float4*)s->pk.planes_fast.data();
    s->dev.mats = (const float4*)s->pk.mats.data();
    s->dev.tex_mat = -1;
    if (tex) {
        s->tex.assign(tex, tex + (size_t)tw * th * 4);
        s->dev.tex = (const float4*)s->tex.data();
        s->dev.tex_w = tw; s->dev.tex_h = th; s->dev.tex_mat = tex_mat;
    }
    s->dev.num_nodes = (int)s->pk.bvh.nodes.size();
    s->dev.top_nodes = 0;
    s->dev.num_hoisted = (int)s->pk.bvh.hoisted_refs.size();
    for (int k = 0; k < s->dev.num_hoisted; ++k) s->dev.hoisted[k] = s->pk.bvh.hoisted_refs[k];
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
