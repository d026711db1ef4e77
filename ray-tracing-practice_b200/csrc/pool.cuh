// pool.cuh -- the per-warp path pool of the megakernel: kPool path slots in shared memory per warp.
//
// A lane's registers hold only the TRAVERSAL state of the segment it is tracing (RayTrav).  Everything that belongs to
// the path -- origin/direction of the current segment, throughput, radiance, the pixel it samples, the pixel's
// running sum -- lives in a slot of the warp's pool.  With twice as many slots as lanes, a lane that finishes its
// traversal parks the hit in its slot and immediately picks up another slot whose ray is ready, and the expensive,
// divergent shading code runs only when 32 parked hits have gathered: one pass, one lane per hit, every lane busy.
// All bookkeeping is warp-local (ballot / popc ranks into two ring buffers of slot ids), so there is no inter-warp
// synchronisation anywhere in the kernel.
#pragma once
#include "path.cuh"

namespace rtb {

constexpr int kPool = 64;  // slots per warp (power of two)
enum PoolField {
    kFOx = 0, kFOy, kFOz, kFDx, kFDy, kFDz, kFBx, kFBy, kFBz, kFLx, kFLy, kFLz, kFAx, kFAy, kFAz,
    kFPix, kFSample, kFDepth, kFRng, kFHitT, kFHitRef, kFSpare, kFReadyQ, kFHitQ, kPoolFields
};
constexpr int kPoolWords = kPoolFields * kPool;  // 32-bit words per warp

struct WarpPool {
    float* f;  // SoA: field k of slot s at f[k * kPool + s]
    __device__ __forceinline__ float& fl(int k, int s) const { return f[k * kPool + s]; }
    __device__ __forceinline__ int& in(int k, int s) const { return reinterpret_cast<int*>(f)[k * kPool + s]; }
    __device__ __forceinline__ V3 ld3(int k, int s) const { return mk3(fl(k, s), fl(k + 1, s), fl(k + 2, s)); }
    __device__ __forceinline__ void st3(int k, int s, V3 v) const { fl(k, s) = v.x; fl(k + 1, s) = v.y; fl(k + 2, s) = v.z; }
};

// warp-uniform tile pool + ring-buffer cursors
struct WarpSched {
    int pool_tile, pool_blk, pool_next;
    bool exhausted;
    int ready_head, ready_cnt, hit_head, hit_cnt;
};

}  // namespace rtb
