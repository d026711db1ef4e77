// leaf_size.h -- primitives per BVH leaf, shared by the host builder (wide_bvh.h) and the device leaf step (traverse.cuh).
// Measured on B200 (100k spheres, 1080p, 256 spp): 1 beats 2..4 by 2-3 % -- a leaf test runs with ~3 of 32 lanes active,
// so a second primitive in a leaf costs the warp more than one more wide-node visit costs a lane.
#pragma once
#ifndef RTB_MAX_LEAF
#define RTB_MAX_LEAF 1
#endif
namespace rtb {
constexpr int kMaxLeafPrims = RTB_MAX_LEAF;  // 1..4
}
