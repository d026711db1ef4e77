// scene_pack.h -- host-side preparation of the device scene: wide BVH + leaf-ordered SoA primitive buffers.
// Replaces the layout half of create_scene's upload (src/main.cu:428-478): AoS SphereData/PlaneData/MaterialData
// become float4 streams the kernels read with 128-bit loads.
#pragma once
#include <vector>

#include "ref_layouts.h"
#include "wide_bvh.h"

namespace rtb {

// Uniform grid over a dense field of small spheres (the BASELINE sphere slabs): a ray that starts inside such a field pays
// the DEPTH of a tree before anything can be culled (every ancestor of the leaf it starts in contains its origin); a grid
// finds the cells along the ray by arithmetic and stops at the first cell boundary behind a hit.  Spheres in the grid are
// NOT in the tree; everything else (planes, outlying or large spheres) is.
struct GridAccel {
    bool enabled = false;
    float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0}, cw[3] = {1, 1, 1};
    int n[3] = {1, 1, 1};
    std::vector<unsigned int> cells;   // 2 per cell: (first ref, count)
    std::vector<int> refs;             // leaf references (wide_bvh.h leaf_ref) of the spheres overlapping the cell
    std::vector<int> sphere_ids;       // caller's indices of the grid's spheres; they follow the tree's in the SoA arrays
    float tree_lo[3] = {0, 0, 0}, tree_hi[3] = {0, 0, 0};  // bounds of everything in the tree (empty: lo > hi)
};

struct PackedScene {
    WideBvh bvh;
    GridAccel grid;
    std::vector<float> spheres;      // 4 per sphere: cx,cy,cz,r                      (leaf order)
    std::vector<int> sphere_meta;    // 2 per sphere: material_idx, caller's index
    std::vector<float> planes;       // 20 per plane: (n,D) (w,type) (u,mat) (v,caller's index) (base,0)
    std::vector<float> planes_fast;  // 12 per plane: the fast-mode record of lbvh.cuh (plane_fast_record), same order
    std::vector<float> mats;         // 12 per material: (albedo,type) (emit,fuzz) (absorption,ir)
    float build_ms = 0.f;
};

// use_grid: -1 = the RTB_GRID environment variable (default off: measured slower), 0 never, 1 where
// select_grid_spheres finds a field
PackedScene pack_scene(const SphereIn* sp, int S, const PlaneIn* pl, int P, const MatIn* mt, int M, int use_grid = 0);

}  // namespace rtb
