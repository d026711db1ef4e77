// bvh_gpu.h -- host interface of the device BVH build (bvh_gpu.cu).
#pragma once
#include <cuda_runtime.h>

#include <cstddef>

#include "ref_layouts.h"

namespace rtb {

struct GpuBuildResult {
    // device buffers in the layouts of scene_pack.h / wide_bvh.h; owned by the caller once gpu_build_scene returns success
    void* d_nodes = nullptr;        // 128-byte wide nodes, breadth-first
    void* d_spheres = nullptr;      // float4 (cx,cy,cz,r), leaf (Morton) order
    void* d_sphere_meta = nullptr;  // int2 (material, caller's index)
    void* d_planes = nullptr;       // 5 x float4 per plane
    void* d_planes_fast = nullptr;  // 3 x float4 per plane: the fast-mode record (lbvh.cuh, plane_fast_record)
    size_t bytes_nodes = 0, bytes_prims = 0;
    int num_nodes = 0, depth = 0, max_stack = 0, leaves = 0, levels = 0;
    int num_hoisted = 0;            // primitives kept out of the tree (wide_bvh.h), tested once per ray before the traversal
    int hoisted[4] = {0, 0, 0, 0};  // their leaf references
    float sah_cost = 0.f;
    float upload_ms = 0.f;  // allocation + H2D of the caller's arrays
    float build_ms = 0.f;   // boxes, keys, sort, tree, fit, collapse, depth sweep (wall clock, stream idle on both sides)
    int fail_line = 0;
};

// Builds the wide BVH and the SoA primitive streams on the current device.  One primitive per leaf.
cudaError_t gpu_build_scene(const SphereIn* h_spheres, int S, const PlaneIn* h_planes, int P, cudaStream_t stream, GpuBuildResult& out);
void gpu_build_free(GpuBuildResult& r);

}  // namespace rtb
