// wide_bvh.h -- host-side builder of the cache-line-aligned 4-wide BVH the render kernels traverse.
//
// Replaces the reference's host builder + node layout (include/bvh_builder.h:52-120: recursive median split,
// one primitive per leaf, 36-byte AoS BVHNode read with scalar loads, include/bvh.h:7-12) with
//   * a binned-SAH binary build (16 bins per axis, all three axes),
//   * a collapse to 4-wide nodes by repeatedly opening the child with the largest surface area,
//   * 128-byte nodes (one L1/L2 line) stored breadth-first so the first N nodes -- the top levels -- are one
//     contiguous block the kernel stages in shared memory,
//   * leaves of up to 4 same-type primitives, referencing contiguous ranges of re-ordered SoA buffers.
#pragma once
#include <cstdint>
#include <vector>
#include "leaf_size.h"

namespace rtb {

// One 128-byte wide node = 8 x float4 (LDG.128 / LDS.128 each).  Child i's box is
// (lo[0][i], lo[1][i], lo[2][i]) .. (hi[0][i], hi[1][i], hi[2][i]).  Empty slots have lo=+inf, hi=-inf.
struct alignas(128) WideNode {
    float lo[3][4];
    float hi[3][4];
    int32_t child[4];  // >=0: index of an inner WideNode; <0: leaf, see leaf_ref(); kEmpty for unused slots
    int32_t pad[4];
};
static_assert(sizeof(WideNode) == 128, "WideNode must be exactly one 128-byte line");

constexpr int32_t kEmptyChild = 0x7fffffff;
constexpr int kMaxHoisted = 4;
// a primitive is hoisted when its box has at least this fraction of the scene box's surface area
constexpr float kHoistAreaFraction = 0.25f;
// leaf reference: bit31 set | type (1 bit, 0 sphere / 1 plane) | count-1 (2 bits) | first (28 bits)
inline int32_t leaf_ref(int type, int first, int count) {
    return (int32_t)(0x80000000u | ((uint32_t)type << 30) | ((uint32_t)(count - 1) << 28) | (uint32_t)first);
}

struct PrimBox {
    float lo[3], hi[3];
    int type;   // 0 sphere, 1 plane
    int index;  // index into the caller's array
};

struct WideBvh {
    std::vector<WideNode> nodes;
    std::vector<int> sphere_order;  // new position -> caller's sphere index
    std::vector<int> plane_order;   // new position -> caller's plane index
    int depth = 0;                  // wide levels
    int max_stack = 0;              // worst-case traversal stack entries (near-first, push the rest)
    int leaves = 0;
    float sah_cost = 0.f;
    int32_t root_ref = 0;           // 0 (inner root) or a leaf ref when the whole scene fits in one leaf
    // Primitives that span most of the scene (the floor quad) are kept OUT of the tree: every ray tests them once, before
    // the traversal, while the warp is still converged (traverse.cuh, trav_begin).  Leaf references, at most kMaxHoisted.
    std::vector<int32_t> hoisted_refs;
};

struct SphereIn;
struct PlaneIn;
// Conservative boxes: a superset of everything the reference's leaf tests (include/sphere.h:24, include/plane.h:57)
// can accept, so culling never removes a hit the reference's linear scan would report.
std::vector<PrimBox> make_prim_boxes(const SphereIn* sp, int S, const PlaneIn* pl, int P);
// Structural check used by the CPU tests: boxes nest, every primitive is referenced exactly once, leaf ranges are
// type-homogeneous, BFS indices increase.  Returns 0 when sound, else a non-zero code.
int validate_wide_bvh(const WideBvh& bvh, const std::vector<PrimBox>& prims);

// Builds the wide BVH over `prims` (boxes must already be conservative).  max_leaf in [1,4].
WideBvh build_wide_bvh(const std::vector<PrimBox>& prims, int max_leaf = kMaxLeafPrims);

}  // namespace rtb
