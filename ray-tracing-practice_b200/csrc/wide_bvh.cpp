#include <cstdlib>
// wide_bvh.cpp -- binned-SAH build + collapse to 128-byte 4-wide nodes (see wide_bvh.h).
// Stands where the reference's build_bvh / build_bvh_recursive stand (include/bvh_builder.h:52-120); nothing of
// their median-split algorithm is kept -- the tree quality (Q4 in SURVEY.md: ~100 node visits per ray) is what
// this replaces.
#include "wide_bvh.h"

#include <algorithm>
#include <array>
#include <functional>
#include <atomic>
#include <cmath>
#include <cstring>
#include <deque>
#include <limits>
#include <thread>

namespace rtb {
namespace {

constexpr int kBins = 16;
constexpr float kCostNode = 1.0f;  // one wide-node visit (8 x LDG.128 + 4 slab tests)
static float cost_prim_value() {
    const char* e = getenv("RTB_COST_PRIM");  // tuning knob for profiling runs
    return e ? (float)atof(e) : 0.5f;
}
static const float kCostPrim = cost_prim_value();  // one leaf primitive test, in units of a wide-node visit

struct Box3 {
    float lo[3], hi[3];
    void reset() {
        for (int a = 0; a < 3; ++a) { lo[a] = std::numeric_limits<float>::infinity(); hi[a] = -lo[a]; }
    }
    void grow(const float* l, const float* h) {
        for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], l[a]); hi[a] = std::max(hi[a], h[a]); }
    }
    void grow(const Box3& b) { grow(b.lo, b.hi); }
    float area() const {
        float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        if (dx < 0 || dy < 0 || dz < 0) return 0.f;
        return 2.f * (dx * dy + dy * dz + dz * dx);
    }
};

struct BinNode {
    Box3 box;
    int left = -1, right = -1;  // children (binary), -1 for leaves
    int first = 0, count = 0;   // range in ids[] for leaves
    int type = -1;              // leaf primitive type
};

struct Builder {
    const std::vector<PrimBox>& prims;
    int max_leaf;
    std::vector<int> ids;
    std::vector<float> cent;  // 3 per prim
    std::vector<BinNode> nodes;
    std::atomic<int> next_node{0};
    std::atomic<int> live_threads{0};
    int max_threads;

    Builder(const std::vector<PrimBox>& p, int ml) : prims(p), max_leaf(ml) {
        ids.resize(p.size());
        cent.resize(p.size() * 3);
        for (size_t i = 0; i < p.size(); ++i) {
            ids[i] = (int)i;
            for (int a = 0; a < 3; ++a) cent[3 * i + a] = 0.5f * (p[i].lo[a] + p[i].hi[a]);
        }
        nodes.resize(p.empty() ? 1 : 2 * p.size());
        max_threads = (int)std::thread::hardware_concurrency();
        if (max_threads < 1) max_threads = 1;
    }

    int alloc() { return next_node.fetch_add(1); }

    bool homogeneous(int b, int e) const {
        int t = prims[ids[b]].type;
        for (int i = b + 1; i < e; ++i)
            if (prims[ids[i]].type != t) return false;
        return true;
    }

    void build(int me, int b, int e, int depth) {
        BinNode& n = nodes[me];
        n.box.reset();
        Box3 cb;
        cb.reset();
        for (int i = b; i < e; ++i) {
            const PrimBox& p = prims[ids[i]];
            n.box.grow(p.lo, p.hi);
            const float* c = &cent[3 * (size_t)ids[i]];
            cb.grow(c, c);
        }
        int cnt = e - b;
        bool can_leaf = cnt <= max_leaf && homogeneous(b, e);
        if (cnt == 1) { make_leaf(n, b, e); return; }

        // binned SAH over the three axes
        float best_cost = std::numeric_limits<float>::infinity();
        int best_axis = -1, best_split = -1;
        float parent_area = std::max(n.box.area(), 1e-30f);
        for (int a = 0; a < 3; ++a) {
            float ext = cb.hi[a] - cb.lo[a];
            if (!(ext > 0.f)) continue;
            Box3 bb[kBins];
            int bc[kBins];
            for (int k = 0; k < kBins; ++k) { bb[k].reset(); bc[k] = 0; }
            float scale = kBins * (1.f - 1e-6f) / ext;
            for (int i = b; i < e; ++i) {
                int k = (int)((cent[3 * (size_t)ids[i] + a] - cb.lo[a]) * scale);
                k = std::min(std::max(k, 0), kBins - 1);
                bb[k].grow(prims[ids[i]].lo, prims[ids[i]].hi);
                bc[k]++;
            }
            float right_area[kBins];
            int right_cnt[kBins];
            Box3 acc;
            acc.reset();
            int c = 0;
            for (int k = kBins - 1; k > 0; --k) {
                acc.grow(bb[k]);
                c += bc[k];
                right_area[k] = acc.area();
                right_cnt[k] = c;
            }
            acc.reset();
            c = 0;
            for (int k = 0; k < kBins - 1; ++k) {
                acc.grow(bb[k]);
                c += bc[k];
                if (c == 0 || right_cnt[k + 1] == 0) continue;
                // leaves hold up to max_leaf primitives: charge whole leaf tests, which is what traversal pays
                float cost = kCostNode * 0.5f + kCostPrim * (acc.area() * c + right_area[k + 1] * right_cnt[k + 1]) / parent_area;
                if (cost < best_cost) { best_cost = cost; best_axis = a; best_split = k; }
            }
        }
        if (can_leaf && (best_axis < 0 || kCostPrim * cnt <= best_cost)) { make_leaf(n, b, e); return; }

        int mid;
        if (best_axis >= 0) {
            float ext = cb.hi[best_axis] - cb.lo[best_axis];
            float scale = kBins * (1.f - 1e-6f) / ext;
            float lo = cb.lo[best_axis];
            int axis = best_axis, split = best_split;
            auto it = std::partition(ids.begin() + b, ids.begin() + e, [&](int id) {
                int k = (int)((cent[3 * (size_t)id + axis] - lo) * scale);
                k = std::min(std::max(k, 0), kBins - 1);
                return k <= split;
            });
            mid = (int)(it - ids.begin());
        } else {
            // coincident centroids: split mixed ranges by type, else in halves
            if (!homogeneous(b, e)) {
                auto it = std::partition(ids.begin() + b, ids.begin() + e, [&](int id) { return prims[id].type == 0; });
                mid = (int)(it - ids.begin());
            } else {
                mid = (b + e) / 2;
            }
        }
        if (mid == b || mid == e) mid = (b + e) / 2;

        int l = alloc(), r = alloc();
        nodes[me].left = l;
        nodes[me].right = r;
        bool spawn = cnt > 32768 && live_threads.load() < max_threads;
        if (spawn) {
            live_threads++;
            std::thread t([this, l, b, mid, depth]() { build(l, b, mid, depth + 1); live_threads--; });
            build(r, mid, e, depth + 1);
            t.join();
        } else {
            build(l, b, mid, depth + 1);
            build(r, mid, e, depth + 1);
        }
    }

    void make_leaf(BinNode& n, int b, int e) {
        n.left = n.right = -1;
        n.first = b;
        n.count = e - b;
        n.type = prims[ids[b]].type;
    }
};

}  // namespace

// indices (into all_prims) of the primitives to hoist: box area >= kHoistAreaFraction of the scene box's, largest first
static std::vector<int> select_hoisted(const std::vector<PrimBox>& all_prims) {
    std::vector<int> pick;
    if (const char* e = getenv("RTB_HOIST")) if (atoi(e) == 0) return pick;
    Box3 scene;
    scene.reset();
    for (const PrimBox& p : all_prims) scene.grow(p.lo, p.hi);
    const float sa = scene.area();
    if (!(sa > 0.f)) return pick;
    std::vector<std::pair<float, int>> cand;
    for (size_t i = 0; i < all_prims.size(); ++i) {
        Box3 b;
        b.reset();
        b.grow(all_prims[i].lo, all_prims[i].hi);
        if (b.area() >= kHoistAreaFraction * sa) cand.emplace_back(-b.area(), (int)i);
    }
    std::sort(cand.begin(), cand.end());
    for (size_t k = 0; k < cand.size() && (int)k < kMaxHoisted; ++k) pick.push_back(cand[k].second);
    std::sort(pick.begin(), pick.end());
    return pick;
}

WideBvh build_wide_bvh(const std::vector<PrimBox>& all_prims, int max_leaf) {
    WideBvh out;
    max_leaf = std::min(std::max(max_leaf, 1), 4);
    const std::vector<int> hoist = select_hoisted(all_prims);
    std::vector<PrimBox> prims;  // what the tree is built over
    prims.reserve(all_prims.size());
    for (size_t i = 0, h = 0; i < all_prims.size(); ++i) {
        if (h < hoist.size() && hoist[h] == (int)i) { ++h; continue; }
        prims.push_back(all_prims[i]);
    }
    // hoisted primitives take the slots after the tree's in the re-ordered arrays
    auto append_hoisted = [&]() {
        for (int i : hoist) {
            const PrimBox& p = all_prims[i];
            std::vector<int>& ord = p.type == 0 ? out.sphere_order : out.plane_order;
            out.hoisted_refs.push_back(leaf_ref(p.type, (int)ord.size(), 1));
            ord.push_back(p.index);
        }
    };
    if (prims.empty()) {
        WideNode root;
        for (int a = 0; a < 3; ++a)
            for (int i = 0; i < 4; ++i) { root.lo[a][i] = std::numeric_limits<float>::infinity(); root.hi[a][i] = -root.lo[a][i]; }
        for (int i = 0; i < 4; ++i) { root.child[i] = kEmptyChild; root.pad[i] = 0; }
        out.nodes.push_back(root);
        out.depth = 1;
        append_hoisted();
        return out;
    }
    Builder B(prims, max_leaf);
    int root = B.alloc();
    B.build(root, 0, (int)prims.size(), 0);
    const std::vector<BinNode>& bn = B.nodes;

    // depth-first pass: place each leaf's primitives contiguously in the per-type order arrays
    std::vector<int> leaf_first(bn.size(), 0);
    {
        std::vector<int> st{root};
        while (!st.empty()) {
            int i = st.back();
            st.pop_back();
            const BinNode& n = bn[i];
            if (n.left < 0) {
                std::vector<int>& ord = n.type == 0 ? out.sphere_order : out.plane_order;
                leaf_first[i] = (int)ord.size();
                for (int k = 0; k < n.count; ++k) ord.push_back(prims[B.ids[n.first + k]].index);
                out.leaves++;
            } else {
                st.push_back(n.right);
                st.push_back(n.left);
            }
        }
    }

    // Which binary nodes become wide nodes: dynamic programme over the binary tree (after Ylitie, Karras, Laine, "Efficient
    // incoherent ray traversal on GPUs through compressed wide BVHs", HPG 2017, section 4.1).  With one primitive per leaf the
    // leaf term of the SAH is the same for every collapse, so the cost to minimise is the summed area of the wide nodes:
    //   c[n][0] = area(n) + min_k (c[l][k-1] + c[r][3-k]),  k = 1..3 roots from the left subtree, 4-k from the right
    //   c[n][i] = min(c[n][i-1], min_k (c[l][k-1] + c[r][i-k]))      subtree n represented by at most i+1 roots
    // Children have larger indices than their parents (alloc order), so a descending sweep is a post-order.
    static const bool use_dp = !(getenv("RTB_COLLAPSE") && atoi(getenv("RTB_COLLAPSE")) == 0);
    const int nbin = B.next_node.load();
    std::vector<std::array<float, 3>> cost(use_dp ? nbin : 0);
    std::vector<std::array<unsigned char, 3>> choice(use_dp ? nbin : 0);
    if (use_dp) {
        for (int n = nbin - 1; n >= 0; --n) {
            const BinNode& b = bn[n];
            if (b.left < 0) { cost[n] = {0.f, 0.f, 0.f}; choice[n] = {0, 0, 0}; continue; }
            const auto &cl = cost[b.left], &cr = cost[b.right];
            float best = std::numeric_limits<float>::infinity();
            int bk = 1;
            for (int k = 1; k <= 3; ++k) {
                float c = cl[k - 1] + cr[3 - k];
                if (c < best) { best = c; bk = k; }
            }
            cost[n][0] = b.box.area() + best;
            choice[n][0] = (unsigned char)bk;
            for (int i = 1; i < 3; ++i) {  // at most i+1 roots
                float bi = cost[n][i - 1];
                int ki = 0;
                for (int k = 1; k <= i; ++k) {
                    float c = cl[k - 1] + cr[i - k];
                    if (c < bi) { bi = c; ki = k; }
                }
                cost[n][i] = bi;
                choice[n][i] = (unsigned char)ki;
            }
        }
    }
    // the (at most `roots`) forest roots that represent binary subtree n
    std::function<void(int, int, int*, int&)> gather = [&](int n, int roots, int* kids, int& nk) {
        if (bn[n].left < 0 || roots <= 1) { kids[nk++] = n; return; }
        int k = choice[n][roots - 1];
        if (k == 0) { gather(n, roots - 1, kids, nk); return; }
        gather(bn[n].left, k, kids, nk);
        gather(bn[n].right, roots - k, kids, nk);
    };

    // breadth-first collapse
    struct Item { int bin; int wide; };
    std::deque<Item> q;
    out.nodes.emplace_back();
    q.push_back({root, 0});
    float root_area = std::max(bn[root].box.area(), 1e-30f);
    double sah = 0;
    while (!q.empty()) {
        Item it = q.front();
        q.pop_front();
        int kids[4];
        int nk = 0;
        if (bn[it.bin].left < 0) {
            kids[nk++] = it.bin;  // single-leaf scene
        } else if (use_dp) {
            gather(bn[it.bin].left, choice[it.bin][0], kids, nk);
            gather(bn[it.bin].right, 4 - choice[it.bin][0], kids, nk);
        } else {
            kids[nk++] = bn[it.bin].left;
            kids[nk++] = bn[it.bin].right;
            while (nk < 4) {
                int pick = -1;
                float pa = -1.f;
                for (int k = 0; k < nk; ++k)
                    if (bn[kids[k]].left >= 0 && bn[kids[k]].box.area() > pa) { pa = bn[kids[k]].box.area(); pick = k; }
                if (pick < 0) break;
                int open = kids[pick];
                kids[pick] = bn[open].left;
                kids[nk++] = bn[open].right;
            }
        }
        WideNode w;
        for (int i = 0; i < 4; ++i) {
            w.pad[i] = 0;
            if (i < nk) {
                const BinNode& c = bn[kids[i]];
                for (int a = 0; a < 3; ++a) { w.lo[a][i] = c.box.lo[a]; w.hi[a][i] = c.box.hi[a]; }
                if (c.left < 0) {
                    w.child[i] = leaf_ref(c.type, leaf_first[kids[i]], c.count);
                    sah += kCostPrim * c.count * c.box.area() / root_area;
                } else {
                    int idx = (int)out.nodes.size();
                    out.nodes.emplace_back();
                    w.child[i] = idx;
                    q.push_back({kids[i], idx});
                }
            } else {
                for (int a = 0; a < 3; ++a) { w.lo[a][i] = std::numeric_limits<float>::infinity(); w.hi[a][i] = -w.lo[a][i]; }
                w.child[i] = kEmptyChild;
            }
        }
        sah += kCostNode * bn[it.bin].box.area() / root_area;
        out.nodes[it.wide] = w;
    }
    out.sah_cost = (float)sah;

    // depth and worst-case stack need (children are after their parents in BFS order: sweep backwards)
    size_t nn = out.nodes.size();
    std::vector<int> dep(nn, 1), need(nn, 0);
    for (size_t i = nn; i-- > 0;) {
        const WideNode& w = out.nodes[i];
        int k = 0, dmax = 0, nmax = 0;
        for (int c = 0; c < 4; ++c) {
            if (w.child[c] == kEmptyChild) continue;
            ++k;
            if (w.child[c] >= 0) { dmax = std::max(dmax, dep[w.child[c]]); nmax = std::max(nmax, need[w.child[c]]); }
        }
        dep[i] = 1 + dmax;
        need[i] = (k > 0 ? k - 1 : 0) + nmax;
    }
    out.depth = dep[0];
    out.max_stack = need[0];
    append_hoisted();
    return out;
}

}  // namespace rtb

// ---------------------------------------------------------------------------------------------------------
#include "ref_layouts.h"
namespace rtb {

std::vector<PrimBox> make_prim_boxes(const SphereIn* sp, int S, const PlaneIn* pl, int P) {
    std::vector<PrimBox> prims;
    prims.reserve((size_t)S + P);
    for (int i = 0; i < S; ++i) {
        PrimBox b;
        float r = std::fabs(sp[i].r);
        float m = std::fmax(std::fabs(sp[i].c[0]), std::fmax(std::fabs(sp[i].c[1]), std::fabs(sp[i].c[2])));
        // the reference's float discriminant can accept rays that pass slightly outside the exact sphere
        float pad = 1e-3f * r + 1e-6f * (m + r);
        for (int a = 0; a < 3; ++a) { b.lo[a] = sp[i].c[a] - r - pad; b.hi[a] = sp[i].c[a] + r + pad; }
        b.type = 0;
        b.index = i;
        prims.push_back(b);
    }
    for (int i = 0; i < P; ++i) {
        PrimBox b;
        const PlaneIn& p = pl[i];
        float m = 0.f;
        for (int a = 0; a < 3; ++a) {
            float c0 = p.base[a], c1 = p.base[a] + p.u[a], c2 = p.base[a] + p.v[a], c3 = p.base[a] + p.u[a] + p.v[a];
            float lo = std::fmin(c0, std::fmin(c1, c2)), hi = std::fmax(c0, std::fmax(c1, c2));
            if (p.type != 2) { lo = std::fmin(lo, c3); hi = std::fmax(hi, c3); }
            b.lo[a] = lo;
            b.hi[a] = hi;
            m = std::fmax(m, std::fmax(std::fabs(lo), std::fabs(hi)));
        }
        float pad = 1e-4f + 4e-6f * m;  // flat boxes get the reference's own 1e-4 thickness (include/aabb.h:92-97) and more
        for (int a = 0; a < 3; ++a) { b.lo[a] -= pad; b.hi[a] += pad; }
        b.type = 1;
        b.index = i;
        prims.push_back(b);
    }
    return prims;
}

int validate_wide_bvh(const WideBvh& bvh, const std::vector<PrimBox>& prims) {
    size_t ns = 0, np = 0;
    for (const PrimBox& p : prims) (p.type == 0 ? ns : np)++;
    if (bvh.sphere_order.size() != ns || bvh.plane_order.size() != np) return 1;
    std::vector<char> seen_s(ns, 0), seen_p(np, 0), seen_n(bvh.nodes.size(), 0);
    // caller index -> prim box
    std::vector<const PrimBox*> sbox(ns, nullptr), pbox(np, nullptr);
    for (const PrimBox& p : prims) {
        if (p.type == 0) { if ((size_t)p.index >= ns) return 2; sbox[p.index] = &p; }
        else { if ((size_t)p.index >= np) return 2; pbox[p.index] = &p; }
    }
    if (bvh.nodes.empty()) return 3;
    seen_n[0] = 1;
    for (size_t i = 0; i < bvh.nodes.size(); ++i) {
        if (!seen_n[i]) return 4;  // unreachable node, or child index not after its parent
        const WideNode& w = bvh.nodes[i];
        for (int c = 0; c < 4; ++c) {
            int32_t ch = w.child[c];
            if (ch == kEmptyChild) continue;
            if (ch >= 0) {
                if ((size_t)ch >= bvh.nodes.size() || (size_t)ch <= i || seen_n[ch]) return 5;
                seen_n[ch] = 1;
                const WideNode& k = bvh.nodes[ch];
                for (int cc = 0; cc < 4; ++cc) {
                    if (k.child[cc] == kEmptyChild) continue;
                    for (int a = 0; a < 3; ++a)
                        if (k.lo[a][cc] < w.lo[a][c] || k.hi[a][cc] > w.hi[a][c]) return 6;  // child box escapes its parent
                }
            } else {
                uint32_t bits = (uint32_t)ch;
                int type = (bits >> 30) & 1, cnt = ((bits >> 28) & 3) + 1, first = bits & 0x0fffffff;
                const std::vector<int>& ord = type == 0 ? bvh.sphere_order : bvh.plane_order;
                std::vector<char>& seen = type == 0 ? seen_s : seen_p;
                if ((size_t)first + cnt > ord.size()) return 7;
                for (int k = 0; k < cnt; ++k) {
                    int orig = ord[first + k];
                    if (orig < 0 || (size_t)orig >= seen.size() || seen[orig]) return 8;
                    seen[orig] = 1;
                    const PrimBox* pb = type == 0 ? sbox[orig] : pbox[orig];
                    for (int a = 0; a < 3; ++a)
                        if (pb->lo[a] < w.lo[a][c] || pb->hi[a] > w.hi[a][c]) return 9;  // primitive escapes its leaf box
                }
            }
        }
    }
    for (int32_t ch : bvh.hoisted_refs) {
        uint32_t bits = (uint32_t)ch;
        int type = (bits >> 30) & 1, cnt = ((bits >> 28) & 3) + 1, first = bits & 0x0fffffff;
        const std::vector<int>& ord = type == 0 ? bvh.sphere_order : bvh.plane_order;
        std::vector<char>& seen = type == 0 ? seen_s : seen_p;
        if (ch >= 0 || cnt != 1 || (size_t)first >= ord.size()) return 11;
        int orig = ord[first];
        if (orig < 0 || (size_t)orig >= seen.size() || seen[orig]) return 12;
        seen[orig] = 1;
    }
    if ((int)bvh.hoisted_refs.size() > kMaxHoisted) return 13;
    for (char s : seen_s) if (!s) return 10;
    for (char s : seen_p) if (!s) return 10;
    return 0;
}

}  // namespace rtb
