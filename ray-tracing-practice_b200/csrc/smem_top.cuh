// smem_top.cuh -- staging of the BVH's top levels (the first `top_nodes` wide nodes, breadth-first) in shared memory,
// plus the shared-memory carve-up every traversal kernel uses: [top nodes | stack distances | stack refs].
#pragma once
#include "traverse.cuh"

namespace rtb {

struct TraversalSmem {
    float4* top;
    float* st;
    int* sr;
};

__device__ __forceinline__ TraversalSmem carve_and_stage(const DevScene& sc, float4* smem4) {
    TraversalSmem t;
    t.top = smem4;
    t.st = reinterpret_cast<float*>(smem4 + sc.top_nodes * kTopStride);
    t.sr = reinterpret_cast<int*>(t.st + kSmemStack * blockDim.x);
    int n4 = sc.top_nodes * 8;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) {
        int node = i >> 3, k = i & 7;
        t.top[node * kTopStride + k] = __ldg(sc.nodes + i);  // 9 x float4 stride: 128-bit LDS of different nodes spread over banks
    }
    __syncthreads();
    return t;
}

}  // namespace rtb
