// scene_pack.h -- host-side preparation of the device scene: wide BVH + leaf-ordered SoA primitive buffers.
// Replaces the layout half of create_scene's upload (src/main.cu:428-478): AoS SphereData/PlaneData/MaterialData
// become float4 streams the kernels read with 128-bit loads.
#pragma once
#include <vector>

#include "ref_layouts.h"
#include "wide_bvh.h"

namespace rtb {

struct PackedScene {
    WideBvh bvh;
    std::vector<float> spheres;      // 4 per sphere: cx,cy,cz,r                      (leaf order)
    std::vector<int> sphere_meta;    // 2 per sphere: material_idx, caller's index
    std::vector<float> planes;       // 20 per plane: (n,D) (w,type) (u,mat) (v,caller's index) (base,0)
    std::vector<float> planes_fast;  // 12 per plane: the fast-mode record of lbvh.cuh (plane_fast_record), same order
    std::vector<float> mats;         // 12 per material: (albedo,type) (emit,fuzz) (absorption,ir)
    float build_ms = 0.f;
};

PackedScene pack_scene(const SphereIn* sp, int S, const PlaneIn* pl, int P, const MatIn* mt, int M);

}  // namespace rtb
