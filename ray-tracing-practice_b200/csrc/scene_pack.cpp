#include "scene_pack.h"

#include "lbvh.cuh"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>

namespace rtb {
namespace {
inline float i2f(int v) { float f; std::memcpy(&f, &v, 4); return f; }
}  // namespace

// Which spheres go into the grid: a field of many similar, small spheres.  Candidates have a radius of at most 4 x the
// median; the field's extent is the 2 .. 98 percentile range of their centres per axis, widened by a tenth of its width
// (a uniform field is inside it entirely, a stray light high above a thin slab is not); a grid is used when it would hold
// at least 4096 spheres and at least half of the scene's.  in_grid[i] says whether sphere i goes in.
static bool select_grid_spheres(const SphereIn* sp, int S, std::vector<char>& in_grid) {
    in_grid.assign((size_t)S, 0);
    if (S < 4096) return false;
    std::vector<float> rad((size_t)S);
    for (int i = 0; i < S; ++i) rad[i] = std::fabs(sp[i].r);
    std::vector<float> tmp = rad;
    std::nth_element(tmp.begin(), tmp.begin() + S / 2, tmp.end());
    const float rmax = 4.0f * tmp[S / 2];
    if (!(rmax > 0.0f)) return false;
    std::vector<int> cand;
    for (int i = 0; i < S; ++i)
        if (rad[i] <= rmax && rad[i] > 0.0f) cand.push_back(i);
    if (cand.size() < 4096) return false;
    float lo[3], hi[3];
    for (int a = 0; a < 3; ++a) {
        tmp.resize(cand.size());
        for (size_t k = 0; k < cand.size(); ++k) tmp[k] = sp[cand[k]].c[a];
        const size_t k0 = cand.size() / 50, k1 = cand.size() - 1 - cand.size() / 50;
        std::nth_element(tmp.begin(), tmp.begin() + k0, tmp.end());
        const float p0 = tmp[k0];
        std::nth_element(tmp.begin(), tmp.begin() + k1, tmp.end());
        const float p1 = tmp[k1];
        const float w = 0.1f * (p1 - p0) + 2.0f * rmax;
        lo[a] = p0 - w;
        hi[a] = p1 + w;
    }
    size_t n = 0;
    for (int i : cand) {
        bool in = true;
        for (int a = 0; a < 3; ++a) in = in && sp[i].c[a] >= lo[a] && sp[i].c[a] <= hi[a];
        in_grid[i] = in;
        n += in;
    }
    if (n < 4096 || 2 * n < (size_t)S) {
        in_grid.assign((size_t)S, 0);
        return false;
    }
    return true;
}

// cells of ~1.5 mean diameters (measured on the 100k-sphere slab: 4.1 cells + 6.8 sphere candidates per secondary ray;
// cells of 2 / 3 diameters: 3.4 + 8.0 / 2.4 + 12.2), at most 1023 per axis (packed cell coordinates) and 2^25 in all
static void build_grid(const std::vector<PrimBox>& boxes, const std::vector<int>& ids, int first_slot, GridAccel& g) {
    double mean_d = 0.0;
    for (int a = 0; a < 3; ++a) { g.lo[a] = 3.0e38f; g.hi[a] = -3.0e38f; }
    for (int i : ids) {
        const PrimBox& b = boxes[i];
        mean_d += 0.5 * ((double)b.hi[0] - b.lo[0] + (double)b.hi[1] - b.lo[1]);
        for (int a = 0; a < 3; ++a) { g.lo[a] = std::fmin(g.lo[a], b.lo[a]); g.hi[a] = std::fmax(g.hi[a], b.hi[a]); }
    }
    mean_d /= (double)ids.size();
    for (int a = 0; a < 3; ++a) {  // a margin, so that no padded box touches the grid's faces
        const float m = 1e-4f * std::fmax(1.0f, std::fmax(std::fabs(g.lo[a]), std::fabs(g.hi[a])));
        g.lo[a] -= m;
        g.hi[a] += m;
    }
    double cs = 1.5 * mean_d;
    for (;;) {
        double total = 1.0;
        for (int a = 0; a < 3; ++a) {
            int n = (int)std::ceil(((double)g.hi[a] - g.lo[a]) / cs);
            g.n[a] = n < 1 ? 1 : (n > 1023 ? 1023 : n);
            total *= g.n[a];
        }
        if (total <= (double)(1 << 25)) break;
        cs *= 1.26;
    }
    for (int a = 0; a < 3; ++a) g.cw[a] = (g.hi[a] - g.lo[a]) / (float)g.n[a];
    auto cell_of = [&](float v, int a) {
        int k = (int)std::floor((v - g.lo[a]) / g.cw[a]);
        return k < 0 ? 0 : (k >= g.n[a] ? g.n[a] - 1 : k);
    };
    const size_t nc = (size_t)g.n[0] * g.n[1] * g.n[2];
    std::vector<unsigned int> count(nc + 1, 0);
    // a sphere is listed in every cell its padded box overlaps, plus a one-ulp-of-the-cell guard for the rounding of cell_of
    auto range = [&](const PrimBox& b, int a, int& c0, int& c1) {
        const float e = 1e-3f * g.cw[a];
        c0 = cell_of(b.lo[a] - e, a);
        c1 = cell_of(b.hi[a] + e, a);
    };
    for (int pass = 0; pass < 2; ++pass) {
        if (pass == 1) {
            unsigned int run = 0;
            for (size_t c = 0; c <= nc; ++c) { unsigned int k = count[c]; count[c] = run; run += k; }
            g.refs.assign(run, 0);
            g.cells.assign(2 * nc, 0);
            for (size_t c = 0; c < nc; ++c) { g.cells[2 * c] = count[c]; g.cells[2 * c + 1] = count[c + 1] - count[c]; }
        }
        for (size_t k = 0; k < ids.size(); ++k) {
            const PrimBox& b = boxes[ids[k]];
            int x0, x1, y0, y1, z0, z1;
            range(b, 0, x0, x1); range(b, 1, y0, y1); range(b, 2, z0, z1);
            for (int z = z0; z <= z1; ++z)
                for (int y = y0; y <= y1; ++y)
                    for (int x = x0; x <= x1; ++x) {
                        const size_t c = ((size_t)z * g.n[1] + y) * g.n[0] + x;
                        if (pass == 0) count[c]++;
                        else g.refs[count[c]++] = leaf_ref(0, first_slot + (int)k, 1);
                    }
        }
    }
    g.sphere_ids = ids;
    g.enabled = true;
}

PackedScene pack_scene(const SphereIn* sp, int S, const PlaneIn* pl, int P, const MatIn* mt, int M, int use_grid) {
    PackedScene ps;
    auto t0 = std::chrono::steady_clock::now();
    std::vector<PrimBox> prims = make_prim_boxes(sp, S, pl, P);
    if (use_grid < 0) {
        // measured on B200: the walk needs half the steps of the tree walk and still renders the headline frame
        // in 300 ms against 175 ms, so the grid is opt-in (RTB_GRID=1) and the default keeps everything in the tree
        use_grid = 0;
        if (const char* e = getenv("RTB_GRID")) use_grid = atoi(e) ? 1 : 0;
    }
    std::vector<char> in_grid;
    std::vector<int> grid_ids;
    if (use_grid && select_grid_spheres(sp, S, in_grid)) {
        std::vector<PrimBox> rest;
        rest.reserve(prims.size());
        for (const PrimBox& b : prims) {
            if (b.type == 0 && in_grid[b.index]) grid_ids.push_back(b.index);
            else rest.push_back(b);
        }
        ps.bvh = build_wide_bvh(rest, kMaxLeafPrims);
        build_grid(prims, grid_ids, (int)ps.bvh.sphere_order.size(), ps.grid);  // prims[i] is sphere i's box for i < S
        // bounds of what the tree holds (hoisted primitives excluded: they are tested before anything else)
        for (int a = 0; a < 3; ++a) { ps.grid.tree_lo[a] = 3.0e38f; ps.grid.tree_hi[a] = -3.0e38f; }
        if (!ps.bvh.nodes.empty()) {
            const WideNode& root = ps.bvh.nodes[0];
            for (int a = 0; a < 3; ++a)
                for (int c = 0; c < 4; ++c) {
                    ps.grid.tree_lo[a] = std::fmin(ps.grid.tree_lo[a], root.lo[a][c]);
                    ps.grid.tree_hi[a] = std::fmax(ps.grid.tree_hi[a], root.hi[a][c]);
                }
        }
    } else {
        ps.bvh = build_wide_bvh(prims, kMaxLeafPrims);
    }
    ps.build_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    const WideBvh& bvh = ps.bvh;
    std::vector<int> order = bvh.sphere_order;
    order.insert(order.end(), grid_ids.begin(), grid_ids.end());
    ps.spheres.resize(order.size() * 4);
    ps.sphere_meta.resize(order.size() * 2);
    for (size_t k = 0; k < order.size(); ++k) {
        const SphereIn& s = sp[order[k]];
        float* o = &ps.spheres[4 * k];
        o[0] = s.c[0]; o[1] = s.c[1]; o[2] = s.c[2]; o[3] = s.r;
        ps.sphere_meta[2 * k] = s.mat;
        ps.sphere_meta[2 * k + 1] = order[k];
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
