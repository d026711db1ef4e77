// bvh_gpu.cu -- device BVH build + scene packing: the kernels around the per-element steps
// of lbvh.cuh.  Replaces build_bvh + the array uploads of create_scene (include/bvh_builder.h:99-120,
// src/main.cu:428-474) for callers that rebuild often: the caller's AoS arrays go up once and everything the render
// kernels read -- 128-byte wide nodes, leaf-ordered SoA spheres / planes -- is produced on the GPU.
// The only library call is cub's radix sort / prefix sum (CUDA toolkit); the tree construction is hand-written.
#include "bvh_gpu.h"

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/block/block_scan.cuh>

#include <algorithm>
#include <chrono>
#include <cstring>
#include <cstdlib>
#include <vector>

#include "lbvh.cuh"

namespace rtb {
namespace {

using lbvh::Box;
constexpr int kThreads = 256;
inline int blocks_for(int n) { return (n + kThreads - 1) / kThreads; }

// ---- 1. boxes + bounds (ordered-uint atomics): bounds12 = centre min xyz, centre max xyz, box min xyz, box max xyz
__global__ void k_boxes(const SphereIn* __restrict__ sp, int S, const PlaneIn* __restrict__ pl, int n, Box* __restrict__ prim_boxes,
                        unsigned int* __restrict__ bounds12) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned int mn[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu}, mx[6] = {0u, 0u, 0u, 0u, 0u, 0u};
    if (i < n) {
        Box b = lbvh::prim_box(sp, S, pl, i);
        prim_boxes[i] = b;
        for (int a = 0; a < 3; ++a) {
            unsigned int c = lbvh::float_to_ordered(__fmul_rn(0.5f, __fadd_rn(b.lo[a], b.hi[a])));
            mn[a] = c;
            mx[a] = c;
            mn[3 + a] = lbvh::float_to_ordered(b.lo[a]);
            mx[3 + a] = lbvh::float_to_ordered(b.hi[a]);
        }
    }
    for (int a = 0; a < 6; ++a) {
        mn[a] = __reduce_min_sync(0xffffffffu, mn[a]);
        mx[a] = __reduce_max_sync(0xffffffffu, mx[a]);
    }
    if ((threadIdx.x & 31) == 0) {
        for (int a = 0; a < 3; ++a) {
            atomicMin(&bounds12[a], mn[a]);
            atomicMax(&bounds12[3 + a], mx[a]);
            atomicMin(&bounds12[6 + a], mn[3 + a]);
            atomicMax(&bounds12[9 + a], mx[3 + a]);
        }
    }
}

// ---- 1b. hoist candidates: primitives whose box has a large share of the scene box's area (rare: the floor quad).  The
// host picks at most four of them (largest first) from this list.
__global__ void k_hoist(const Box* __restrict__ prim_boxes, int n, const unsigned int* __restrict__ bounds12, int* __restrict__ cand_count,
                        int2* __restrict__ cand) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Box sb;
    for (int a = 0; a < 3; ++a) {
        sb.lo[a] = lbvh::ordered_to_float(bounds12[6 + a]);
        sb.hi[a] = lbvh::ordered_to_float(bounds12[9 + a]);
    }
    const Box b = prim_boxes[i];
    if (lbvh::hoist_candidate(b, lbvh::box_area(sb))) {
        int k = atomicAdd(cand_count, 1);
        if (k < lbvh::kHoistCandCap) cand[k] = make_int2(i, __float_as_int(lbvh::box_area(b)));
    }
}

// ---- 2. Morton keys; hoisted primitives get the largest key, so the sort puts them behind the tree's primitives
__global__ void k_keys(const Box* __restrict__ prim_boxes, int n, const unsigned int* __restrict__ bounds12, int4 hoisted,
                       unsigned long long* __restrict__ keys, unsigned int* __restrict__ vals) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float cmin[3], cmax[3];
    for (int a = 0; a < 3; ++a) {
        cmin[a] = lbvh::ordered_to_float(bounds12[a]);
        cmax[a] = lbvh::ordered_to_float(bounds12[3 + a]);
    }
    const bool h = i == hoisted.x || i == hoisted.y || i == hoisted.z || i == hoisted.w;
    keys[i] = h ? ~0ull : lbvh::morton_key(prim_boxes[i], cmin, cmax);
    vals[i] = (unsigned int)i;
}

// ---- 3b. after the sort: 1 where the primitive at sorted position p is a sphere (scanned into sphere positions)
__global__ void k_flags(const unsigned int* __restrict__ vals, int n, int S, int* __restrict__ is_sphere) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) is_sphere[p] = vals[p] < (unsigned int)S;
}

// ---- 3c. leaf boxes, leaf references and the leaf-ordered SoA primitive streams (layouts of scene_pack.h)
__global__ void k_pack(const SphereIn* __restrict__ sp, int S, const PlaneIn* __restrict__ pl, const unsigned int* __restrict__ vals,
                       const int* __restrict__ sphere_pos, const Box* __restrict__ prim_boxes, int n, int n_tree, Box* __restrict__ boxes,
                       int* __restrict__ leaf_refs, float4* __restrict__ spheres, int2* __restrict__ sphere_meta,
                       float4* __restrict__ planes, float4* __restrict__ planes_fast) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    int id = (int)vals[p];
    if (p < n_tree) boxes[(n_tree - 1) + p] = prim_boxes[id];  // positions >= n_tree hold the hoisted primitives
    if (id < S) {
        int pos = sphere_pos[p];
        const SphereIn& s = sp[id];
        spheres[pos] = make_float4(s.c[0], s.c[1], s.c[2], s.r);
        sphere_meta[pos] = make_int2(s.mat, id);
        leaf_refs[p] = lbvh::make_leaf_ref(0, pos);
    } else {
        int pos = p - sphere_pos[p];
        const PlaneIn& q = pl[id - S];
        float4* o = planes + 5 * (size_t)pos;
        o[0] = make_float4(q.n[0], q.n[1], q.n[2], q.D);
        o[1] = make_float4(q.w[0], q.w[1], q.w[2], __int_as_float(q.type));
        o[2] = make_float4(q.u[0], q.u[1], q.u[2], __int_as_float(q.mat));
        o[3] = make_float4(q.v[0], q.v[1], q.v[2], __int_as_float(id - S));
        o[4] = make_float4(q.base[0], q.base[1], q.base[2], 0.f);
        float f[12];
        lbvh::plane_fast_record(q, f);
        float4* of = planes_fast + 3 * (size_t)pos;
        of[0] = make_float4(f[0], f[1], f[2], f[3]);
        of[1] = make_float4(f[4], f[5], f[6], f[7]);
        of[2] = make_float4(f[8], f[9], f[10], f[11]);
        leaf_refs[p] = lbvh::make_leaf_ref(1, pos);
    }
}

// ---- 4. binary radix tree
__global__ void k_radix(const unsigned long long* __restrict__ keys, int n, int2* __restrict__ children, int* __restrict__ parent) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int l, r;
    lbvh::radix_node((const uint64_t*)keys, n, i, l, r);
    children[i] = make_int2(l, r);
    parent[l] = i;
    parent[r] = i;
}

// ---- 5. bottom-up fit: the second thread to arrive at an inner node owns it and everything below it (lbvh.cuh, fit_node:
// tree rotation, box, collapse table)
__global__ void k_fit(int2* children, int* parent, int n, Box* boxes, lbvh::CollapseDp* dp, unsigned int* arrived) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n || n < 2) return;
    int cur = parent[(n - 1) + p];
    while (true) {
        __threadfence();
        if (atomicAdd(&arrived[cur], 1u) == 0u) return;
        lbvh::fit_node(children, parent, boxes, dp, n, cur);
        if (cur == 0) return;
        cur = __ldcg(parent + cur);
    }
}

// ---- 6. collapse, one breadth-first level per launch pair
__global__ void k_collapse_count(const int2* __restrict__ children, const lbvh::CollapseDp* __restrict__ dp, int n,
                                 const int* __restrict__ items, int count, int4* __restrict__ kids_out, int* __restrict__ inner_cnt) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= count) return;
    int kids[4];
    inner_cnt[k] = lbvh::collapse_pick_dp(children, dp, n, items[k], kids);
    kids_out[k] = make_int4(kids[0], kids[1], kids[2], kids[3]);
}
__global__ void k_collapse_write(const Box* __restrict__ boxes, const int* __restrict__ leaf_refs, int n,
                                 const int* __restrict__ items, const int4* __restrict__ kids_in,
                                 const int* __restrict__ inner_off, const int* __restrict__ inner_cnt, int count, int level_base,
                                 float* __restrict__ nodes, int* __restrict__ next_items, int* __restrict__ next_count,
                                 float prim_cost, double* __restrict__ sah) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    float cost = 0.f;
    if (k < count) {
        int4 kd = kids_in[k];
        int kids[4] = {kd.x, kd.y, kd.z, kd.w};
        lbvh::collapse_write(boxes, leaf_refs, n, kids, level_base + count, inner_off[k], nodes + 32 * (size_t)(level_base + k), next_items);
        if (k == count - 1) *next_count = inner_off[k] + inner_cnt[k];
        float root_area = fmaxf(lbvh::box_area(boxes[0]), 1e-30f);
        cost = lbvh::box_area(boxes[n == 1 ? 0 : items[k]]) / root_area;
        for (int c = 0; c < 4; ++c)
            if (kids[c] >= n - 1) cost += prim_cost * lbvh::box_area(boxes[kids[c]]) / root_area;
    }
    for (int o = 16; o > 0; o >>= 1) cost += __shfl_xor_sync(0xffffffffu, cost, o);
    if ((threadIdx.x & 31) == 0 && cost != 0.f) atomicAdd(sah, (double)cost);
}

// ---- 6b. the first levels of the collapse (while a level has at most 1024 nodes) in ONE block: the same steps as the
// count / scan / write launches above with a block scan in between, so the ~6 smallest levels cost one launch and one host
// round trip instead of three launches and a round trip each.  Node order and contents are identical.
constexpr int kTopThreads = 1024;
constexpr int kTopMaxLevels = 24;
struct TopLevels {
    int base, count, nlevels;
    int level_base[kTopMaxLevels], level_count[kTopMaxLevels];
};
__global__ void __launch_bounds__(kTopThreads) k_collapse_top(const int2* __restrict__ children, const lbvh::CollapseDp* __restrict__ dp,
                                                               const Box* __restrict__ boxes, const int* __restrict__ leaf_refs, int n,
                                                               int* items_a, int* items_b, float* nodes, float prim_cost, double* sah,
                                                               TopLevels* out) {
    typedef cub::BlockScan<int, kTopThreads> Scan;
    __shared__ typename Scan::TempStorage scan_tmp;
    int* cur = items_a;  // items_a[0] = 0 (the root) on entry; plain pointers: each level reads what the previous one wrote
    int* nxt = items_b;
    int base = 0, count = 1, lvl = 0;
    const int k = threadIdx.x;
    const float root_area = fmaxf(lbvh::box_area(boxes[0]), 1e-30f);
    while (count > 0 && count <= kTopThreads && lvl < kTopMaxLevels) {
        int kids[4] = {-1, -1, -1, -1};
        int inner = 0, item = 0;
        if (k < count) {
            item = cur[k];
            inner = lbvh::collapse_pick_dp(children, dp, n, item, kids);
        }
        int off = 0, total = 0;
        Scan(scan_tmp).ExclusiveSum(inner, off, total);
        float cost = 0.f;
        if (k < count) {
            lbvh::collapse_write(boxes, leaf_refs, n, kids, base + count, off, nodes + 32 * (size_t)(base + k), nxt);
            cost = lbvh::box_area(boxes[n == 1 ? 0 : item]) / root_area;
            for (int c = 0; c < 4; ++c)
                if (kids[c] >= n - 1) cost += prim_cost * lbvh::box_area(boxes[kids[c]]) / root_area;
        }
        for (int o = 16; o > 0; o >>= 1) cost += __shfl_xor_sync(0xffffffffu, cost, o);
        if ((k & 31) == 0 && cost != 0.f) atomicAdd(sah, (double)cost);
        if (k == 0) { out->level_base[lvl] = base; out->level_count[lvl] = count; }
        base += count;
        count = total;
        ++lvl;
        int* t = cur; cur = nxt; nxt = t;
        __syncthreads();  // the next level's items (global memory, written by this block) and scan_tmp are safe to reuse
    }
    if (k == 0) { out->base = base; out->count = count; out->nlevels = lvl; }
}

// ---- 7. depth / stack need of one level (children live in later levels, already final)
__global__ void k_depth(const float* __restrict__ nodes, int base, int count, int* dep, int* need) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= count) return;
    int d, s;
    lbvh::depth_need(nodes + 32 * (size_t)(base + k), dep, need, d, s);
    dep[base + k] = d;
    need[base + k] = s;
}

struct Arena {
    char* base = nullptr;
    size_t size = 0, used = 0;
    template <class T> T* take(size_t count) {
        size_t bytes = (count * sizeof(T) + 255) & ~(size_t)255;
        T* p = reinterpret_cast<T*>(base + used);
        used += bytes;
        return p;
    }
};
inline size_t pad256(size_t b) { return (b + 255) & ~(size_t)255; }

}  // namespace

static cudaError_t write_empty_root(void* d_nodes, cudaStream_t stream) {
    float h[32];
    for (int a = 0; a < 3; ++a)
        for (int i = 0; i < 4; ++i) { h[a * 4 + i] = __builtin_huge_valf(); h[12 + a * 4 + i] = -__builtin_huge_valf(); }
    int32_t* ch = reinterpret_cast<int32_t*>(h + 24);
    for (int i = 0; i < 4; ++i) { ch[i] = lbvh::kEmpty; ch[4 + i] = 0; }
    return cudaMemcpyAsync(d_nodes, h, 128, cudaMemcpyHostToDevice, stream);
}

#define BG_CUDA(expr)                     \
    do {                                  \
        cudaError_t e_ = (expr);          \
        if (e_ != cudaSuccess) {          \
            r.fail_line = __LINE__;       \
            if (arena.base) cudaFree(arena.base); \
            if (d_sp) cudaFree(d_sp);     \
            if (d_pl) cudaFree(d_pl);     \
            gpu_build_free(r);            \
            return e_;                    \
        }                                 \
    } while (0)

void gpu_build_free(GpuBuildResult& r) {
    void* bufs[] = {r.d_nodes, r.d_spheres, r.d_sphere_meta, r.d_planes, r.d_planes_fast};
    for (void* b : bufs)
        if (b) cudaFree(b);
    r.d_nodes = r.d_spheres = r.d_sphere_meta = r.d_planes = r.d_planes_fast = nullptr;
}

cudaError_t gpu_build_scene(const SphereIn* h_sp, int S, const PlaneIn* h_pl, int P, cudaStream_t stream, GpuBuildResult& r) {
    r = GpuBuildResult{};
    const int n = S + P;
    Arena arena;
    SphereIn* d_sp = nullptr;
    PlaneIn* d_pl = nullptr;
    auto t0 = std::chrono::steady_clock::now();

    // outputs (owned by the caller afterwards).  A wide node is rooted at a distinct binary inner node: <= max(n-1,1) nodes.
    const int max_nodes = n > 1 ? n - 1 : 1;
    BG_CUDA(cudaMalloc(&r.d_nodes, (size_t)max_nodes * 128));
    BG_CUDA(cudaMalloc(&r.d_spheres, S ? (size_t)S * 16 : 16));
    BG_CUDA(cudaMalloc(&r.d_sphere_meta, S ? (size_t)S * 8 : 16));
    BG_CUDA(cudaMalloc(&r.d_planes, P ? (size_t)P * 80 : 16));
    BG_CUDA(cudaMalloc(&r.d_planes_fast, P ? (size_t)P * 48 : 16));
    r.bytes_nodes = (size_t)max_nodes * 128;
    r.bytes_prims = (size_t)(S ? S * 24 : 32) + (size_t)(P ? P * 128 : 32);

    if (n == 0) {  // empty scene: one node whose four slots fail every slab test
        BG_CUDA(write_empty_root(r.d_nodes, stream));
        BG_CUDA(cudaStreamSynchronize(stream));
        r.num_nodes = 1;
        r.depth = 1;
        return cudaSuccess;
    }

    // the caller's arrays, as they are
    if (S) {
        BG_CUDA(cudaMalloc((void**)&d_sp, (size_t)S * sizeof(SphereIn)));
        BG_CUDA(cudaMemcpyAsync(d_sp, h_sp, (size_t)S * sizeof(SphereIn), cudaMemcpyHostToDevice, stream));
    }
    if (P) {
        BG_CUDA(cudaMalloc((void**)&d_pl, (size_t)P * sizeof(PlaneIn)));
        BG_CUDA(cudaMemcpyAsync(d_pl, h_pl, (size_t)P * sizeof(PlaneIn), cudaMemcpyHostToDevice, stream));
    }

    // one arena for every temporary
    size_t sort_tmp = 0, scan_tmp = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_tmp, (unsigned long long*)nullptr, (unsigned long long*)nullptr, (unsigned int*)nullptr,
                                    (unsigned int*)nullptr, n, 0, 64, stream);
    cub::DeviceScan::ExclusiveSum(nullptr, scan_tmp, (int*)nullptr, (int*)nullptr, n, stream);
    size_t need = pad256((size_t)n * sizeof(Box)) + 2 * pad256((size_t)n * 8) + 2 * pad256((size_t)n * 4) +  // prim boxes, keys x2, vals x2
                  pad256((size_t)(2 * n) * sizeof(Box)) + pad256((size_t)n * 8) + pad256((size_t)(2 * n) * 4) +  // boxes, children, parent
                  8 * pad256((size_t)n * 4) +       // arrived, is_sphere, sphere_pos, leaf_refs, items x2, inner_cnt, inner_off
                  2 * pad256((size_t)n * 16) +      // kids, collapse tables
                  2 * pad256((size_t)max_nodes * 4) +  // dep, need
                  pad256(sort_tmp) + pad256(scan_tmp) + 6 * 256 + pad256(sizeof(TopLevels)) + pad256(lbvh::kHoistCandCap * sizeof(int2));
    BG_CUDA(cudaMalloc((void**)&arena.base, need));
    arena.size = need;
    Box* prim_boxes = arena.take<Box>(n);
    unsigned long long* keys_a = arena.take<unsigned long long>(n);
    unsigned long long* keys_b = arena.take<unsigned long long>(n);
    unsigned int* vals_a = arena.take<unsigned int>(n);
    unsigned int* vals_b = arena.take<unsigned int>(n);
    Box* boxes = arena.take<Box>(2 * (size_t)n);
    int2* children = arena.take<int2>(n);
    int* parent = arena.take<int>(2 * (size_t)n);
    unsigned int* arrived = arena.take<unsigned int>(n);
    int* is_sphere = arena.take<int>(n);
    int* sphere_pos = arena.take<int>(n);
    int* leaf_refs = arena.take<int>(n);
    int* items_a = arena.take<int>(n);
    int* items_b = arena.take<int>(n);
    int* inner_cnt = arena.take<int>(n);
    int* inner_off = arena.take<int>(n);
    int4* kids = arena.take<int4>(n);
    lbvh::CollapseDp* dp = arena.take<lbvh::CollapseDp>(n);
    int* dep = arena.take<int>(max_nodes);
    int* need_s = arena.take<int>(max_nodes);
    void* d_sort_tmp = arena.take<char>(sort_tmp);
    void* d_scan_tmp = arena.take<char>(scan_tmp);
    unsigned int* bounds12 = arena.take<unsigned int>(12);
    int* cand_count = arena.take<int>(1);
    int2* cand = arena.take<int2>(lbvh::kHoistCandCap);
    int* next_count = arena.take<int>(1);
    TopLevels* top = arena.take<TopLevels>(1);
    double* sah = arena.take<double>(1);

    BG_CUDA(cudaStreamSynchronize(stream));
    auto t1 = std::chrono::steady_clock::now();

    const unsigned int b12[12] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
    BG_CUDA(cudaMemcpyAsync(bounds12, b12, sizeof b12, cudaMemcpyHostToDevice, stream));
    BG_CUDA(cudaMemsetAsync(arrived, 0, (size_t)n * 4, stream));
    BG_CUDA(cudaMemsetAsync(sah, 0, 8, stream));
    BG_CUDA(cudaMemsetAsync(cand_count, 0, 4, stream));
    k_boxes<<<blocks_for(n), kThreads, 0, stream>>>(d_sp, S, d_pl, n, prim_boxes, bounds12);

    // hoisted primitives: the device lists the candidates, the host picks at most four (largest area first, then lowest
    // index -- select_hoisted's rule in wide_bvh.cpp)
    int hoist[4] = {-1, -1, -1, -1};
    int nh = 0;
    {
        const char* e = getenv("RTB_HOIST");
        if (!(e && atoi(e) == 0)) {
            k_hoist<<<blocks_for(n), kThreads, 0, stream>>>(prim_boxes, n, bounds12, cand_count, cand);
            int h_count = 0;
            BG_CUDA(cudaMemcpyAsync(&h_count, cand_count, 4, cudaMemcpyDeviceToHost, stream));
            BG_CUDA(cudaStreamSynchronize(stream));
            if (h_count > 0 && h_count <= lbvh::kHoistCandCap) {
                std::vector<int2> hc((size_t)h_count);
                BG_CUDA(cudaMemcpy(hc.data(), cand, (size_t)h_count * sizeof(int2), cudaMemcpyDeviceToHost));
                std::vector<std::pair<float, int>> byarea;
                for (const int2& c : hc) {
                    float a;
                    memcpy(&a, &c.y, 4);
                    byarea.emplace_back(-a, c.x);
                }
                std::sort(byarea.begin(), byarea.end());
                for (size_t k = 0; k < byarea.size() && nh < lbvh::kMaxHoist; ++k) hoist[nh++] = byarea[k].second;
            }
        }
    }
    const int nt = n - nh;  // primitives in the tree
    k_keys<<<blocks_for(n), kThreads, 0, stream>>>(prim_boxes, n, bounds12, make_int4(hoist[0], hoist[1], hoist[2], hoist[3]), keys_a, vals_a);
    BG_CUDA(cub::DeviceRadixSort::SortPairs(d_sort_tmp, sort_tmp, keys_a, keys_b, vals_a, vals_b, n, 0, 64, stream));
    k_flags<<<blocks_for(n), kThreads, 0, stream>>>(vals_b, n, S, is_sphere);
    BG_CUDA(cub::DeviceScan::ExclusiveSum(d_scan_tmp, scan_tmp, is_sphere, sphere_pos, n, stream));
    k_pack<<<blocks_for(n), kThreads, 0, stream>>>(d_sp, S, d_pl, vals_b, sphere_pos, prim_boxes, n, nt, boxes, leaf_refs,
                                                  (float4*)r.d_spheres, (int2*)r.d_sphere_meta, (float4*)r.d_planes, (float4*)r.d_planes_fast);
    if (nt > 1) {
        k_radix<<<blocks_for(nt - 1), kThreads, 0, stream>>>(keys_b, nt, children, parent);
        k_fit<<<blocks_for(nt), kThreads, 0, stream>>>(children, parent, nt, boxes, dp, arrived);
    }
    BG_CUDA(cudaGetLastError());
    r.num_hoisted = nh;
    if (nh) BG_CUDA(cudaMemcpyAsync(r.hoisted, leaf_refs + nt, (size_t)nh * 4, cudaMemcpyDeviceToHost, stream));
    if (nt == 0) {  // everything hoisted: the tree is one node whose slots fail every slab test
        BG_CUDA(write_empty_root(r.d_nodes, stream));
        BG_CUDA(cudaStreamSynchronize(stream));
        r.num_nodes = 1;
        r.depth = 1;
        r.leaves = n;
        cudaFree(arena.base);
        if (d_sp) cudaFree(d_sp);
        if (d_pl) cudaFree(d_pl);
        return cudaSuccess;
    }

    // breadth-first collapse; the host only learns each level's size (4 bytes per level)
    const char* pc = getenv("RTB_COST_PRIM");
    const float prim_cost = pc ? (float)atof(pc) : 0.5f;
    std::vector<int> level_base, level_count;
    int root_item = 0;
    BG_CUDA(cudaMemcpyAsync(items_a, &root_item, 4, cudaMemcpyHostToDevice, stream));
    int* cur = items_a;
    int* nxt = items_b;
    int base = 0, count = 1;
    {   // the small top levels in one launch (k_collapse_top), one read-back
        k_collapse_top<<<1, kTopThreads, 0, stream>>>(children, dp, boxes, leaf_refs, nt, items_a, items_b, (float*)r.d_nodes, prim_cost, sah, top);
        TopLevels h_top;
        BG_CUDA(cudaMemcpyAsync(&h_top, top, sizeof h_top, cudaMemcpyDeviceToHost, stream));
        BG_CUDA(cudaStreamSynchronize(stream));
        for (int l = 0; l < h_top.nlevels; ++l) { level_base.push_back(h_top.level_base[l]); level_count.push_back(h_top.level_count[l]); }
        base = h_top.base;
        count = h_top.count;
        if (h_top.nlevels & 1) { cur = items_b; nxt = items_a; }
    }
    while (count > 0) {
        if (base + count > max_nodes) { r.fail_line = __LINE__; BG_CUDA(cudaErrorUnknown); }
        level_base.push_back(base);
        level_count.push_back(count);
        k_collapse_count<<<blocks_for(count), kThreads, 0, stream>>>(children, dp, nt, cur, count, kids, inner_cnt);
        BG_CUDA(cub::DeviceScan::ExclusiveSum(d_scan_tmp, scan_tmp, inner_cnt, inner_off, count, stream));
        k_collapse_write<<<blocks_for(count), kThreads, 0, stream>>>(boxes, leaf_refs, nt, cur, kids, inner_off, inner_cnt, count, base,
                                                                    (float*)r.d_nodes, nxt, next_count, prim_cost, sah);
        int h_next = 0;
        BG_CUDA(cudaMemcpyAsync(&h_next, next_count, 4, cudaMemcpyDeviceToHost, stream));
        BG_CUDA(cudaStreamSynchronize(stream));
        base += count;
        count = h_next;
        int* t = cur; cur = nxt; nxt = t;
    }
    r.num_nodes = base;
    r.levels = (int)level_base.size();
    for (int l = r.levels - 1; l >= 0; --l)
        k_depth<<<blocks_for(level_count[l]), kThreads, 0, stream>>>((const float*)r.d_nodes, level_base[l], level_count[l], dep, need_s);
    double h_sah = 0;
    BG_CUDA(cudaMemcpyAsync(&r.depth, dep, 4, cudaMemcpyDeviceToHost, stream));
    BG_CUDA(cudaMemcpyAsync(&r.max_stack, need_s, 4, cudaMemcpyDeviceToHost, stream));
    BG_CUDA(cudaMemcpyAsync(&h_sah, sah, 8, cudaMemcpyDeviceToHost, stream));
    BG_CUDA(cudaStreamSynchronize(stream));
    BG_CUDA(cudaGetLastError());
    auto t2 = std::chrono::steady_clock::now();
    r.sah_cost = (float)h_sah;
    r.leaves = n;
    r.upload_ms = std::chrono::duration<float, std::milli>(t1 - t0).count();
    r.build_ms = std::chrono::duration<float, std::milli>(t2 - t1).count();
    cudaFree(arena.base);
    if (d_sp) cudaFree(d_sp);
    if (d_pl) cudaFree(d_pl);
    return cudaSuccess;
}

}  // namespace rtb
