// kernels.h -- host-visible launch interface of the sm_100a kernels (kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "path.cuh"

namespace rtb {

struct FrameParams {
    CameraPod cam;
    uint2 key;          // Philox key
    int tex_mode;       // 0 cpu semantics, 1 texture unit
    int tile_rank, tile_nranks;
    int tiles_x, tiles_y, my_tiles;
    int trace_min;                    // render_mega, time-sliced TRACE: a TRACE phase ends as soon as fewer than trace_min lanes are still
                                      // traversing and some lane waits to be shaded (1 = every lane runs to completion)
    int leaf_min;                     // megakernel: lanes that must hold a leaf before the primitive tests run (kernels.cu)
    int refill_min, park_min;         // pooled megakernel: lanes that must be out of rays before a refill; parked hits before a shading pass
    int regroup;                      // pooled megakernel: 1 = order the parked hits by material class before a shading pass (mega_pool.cu)
    float* fb;                        // W*H*3 sums
    // sample chunks (render_mega only): a work item is (8x4 pixel block, chunk of chunk_spp samples).  With nchunks > 1 a lane
    // adds its chunk's partial sum to the pixel's 64-bit FIXED-POINT accumulator acc[pixel][channel] (value * 2^32, one
    // RED.ADD.U64 per channel and item); integer addition is associative, so the frame does not depend on the order in
    // which the chunks arrive, and finalize_chunks converts the sums to float once.  nchunks == 1: the lane stores to fb.
    long long* acc;
    int chunk_spp, nchunks;
    int chunk_major;                  // item order: 0 = (block, chunk) chunk fastest; 1 = all blocks of chunk 0, then chunk 1, ...
    // Work queues (zeroed before launch): nqueues heads, kQueueStride words apart.  render_mega: queue q owns the 8x4 blocks
    // kb with kb % nqueues == q; an SM's warps draw from queue smid % nqueues and move on to the next queue when theirs runs
    // dry, so that the warps of one SM work on the SAME block (its sample chunks) at the same time and share its part of the
    // tree in L1.  nqueues == 1 is the single global queue (the other kernels use head 0 only).
    unsigned int* tile_counter;
    int nqueues;
    int queue_per_cta;                // 1: a warp's home queue is blockIdx.x % nqueues instead of smid % nqueues
    unsigned long long* counters;     // 6 x u64: segments,nodes,spheres,planes,hits,samples
};

struct LaunchCfg {
    int num_sms;
    int blocks_per_sm;
    int threads;
};

constexpr int kTile = 16;  // == RT_TILE
constexpr int kQueueStride = 16;  // 64 bytes between queue heads
constexpr int kMaxQueues = 1024;

// Interleaved-tile partition of a frame over `nranks` devices.  Tiles (row-major ids) are dealt in groups of nranks
// consecutive ids, one tile of every group to every rank; WHICH tile of group g a rank gets is rotated by a hash of g.
// (Plain t % nranks hands a rank the same tile COLUMNS in every row whenever the tiles per row are a multiple of nranks --
// 120 tiles at 1080p, 480 at 8K, for 2 / 4 / 8 ranks -- i.e. vertical stripes, and a rank's cost then follows whatever
// the picture holds in its stripes.)  k = the rank's k-th tile = group index; rank r's packed buffer holds its tiles in
// increasing k.  The last group may be short: a (k, rank) whose tile id is >= the tile count owns nothing.
#ifndef RTB_TILE_ROT
#define RTB_TILE_ROT 1
#endif
__host__ __device__ inline int tile_rot(unsigned int g, int nranks) {
#if RTB_TILE_ROT == 1
    // floor(frac(g * golden ratio) * nranks): the TOP bits of the 32-bit product are the Kronecker (golden-ratio) sequence,
    // which is low-discrepancy in g -- over ANY run of consecutive groups every rank gets every position of the group almost
    // equally often, so a cost that varies smoothly along the rows cancels instead of adding up as a random walk
    return nranks > 1 ? (int)(((unsigned long long)(g * 0x9E3779B9u) * (unsigned long long)nranks) >> 32) : 0;
#else
    return nranks > 1 ? (int)(((g * 0x9E3779B1u) >> 16) % (unsigned int)nranks) : 0;  // a hash of g: pseudo-random positions
#endif
}
__host__ __device__ inline int tile_of_rank(int k, int rank, int nranks) {
    int p = rank - tile_rot((unsigned int)k, nranks);
    if (p < 0) p += nranks;
    return k * nranks + p;
}
__host__ __device__ inline void tile_owner(int t, int nranks, int& rank, int& k) {
    k = t / nranks;
    rank = (t % nranks + tile_rot((unsigned int)k, nranks)) % nranks;
}
// the unit of work a warp fetches with one atomicAdd: an 8x4-pixel block (one pixel per lane) of a 16x16 tile.  Tiles
// stay the unit of the multi-GPU partition; blocks keep the end-of-frame tail short (8x more, 8x smaller work items).
constexpr int kBlockPixels = 32;
constexpr int kBlocksPerTile = kTile * kTile / kBlockPixels;
// render_mega's launch geometry: 4 CTAs x 224 threads per SM = 28 warps at 72 registers (no spills with the time-sliced TRACE
// loop).  Measured on the headline frame at trace_min 6: 224x4 180.1 ms, 128x7 180.5, 288x3 181.9 (all 72 registers);
// 256x4 191.0 and 192x5 193.6 (64 registers, ~120 B of spills); the round-1 geometry 256x3 (80 registers) 192.1.
#ifndef RTB_MEGA_THREADS
#define RTB_MEGA_THREADS 224
#endif
#ifndef RTB_MEGA_BLOCKS
#define RTB_MEGA_BLOCKS 4
#endif
constexpr int kMegaThreads = RTB_MEGA_THREADS;      // CTA size of render_mega
constexpr int kMegaBlocksPerSM = RTB_MEGA_BLOCKS;   // resident CTAs per SM it is compiled (register-capped) for
constexpr int kPoolThreads = 256, kPoolBlocksPerSM = 3;  // render_pool / wavefront kernels: the geometry they were measured with
constexpr int kTraceMinDefault = 6, kTraceMinSmallTree = 10;  // time-sliced TRACE: see FrameParams::trace_min
constexpr int kLeafMinDefault = 12;

// persistent megakernel.  rng: 0 Philox / 1 reference stream; exact: 0/1; count: 0/1
cudaError_t launch_render_mega(const DevScene& sc, const FrameParams& fp, int rng, int exact, int count, LaunchCfg cfg,
                               cudaStream_t stream);
// persistent megakernel with a per-warp path pool in shared memory (mega_pool.cu)
cudaError_t launch_render_pool(const DevScene& sc, const FrameParams& fp, int rng, int exact, int count, LaunchCfg cfg,
                               cudaStream_t stream);
size_t pool_smem_bytes(int top_nodes, int threads);
// wavefront variant (generate / extend / shade kernels over global queues); scratch is owned by the caller
struct WavefrontScratch {
    void* buf = nullptr;
    size_t bytes = 0;
};
cudaError_t launch_render_wavefront(const DevScene& sc, const FrameParams& fp, int rng, int exact, int count, LaunchCfg cfg,
                                    WavefrontScratch& scratch, cudaStream_t stream);
// exact = 1: the reference's leaf arithmetic; 0: the production (fast fp32 + SFU) arithmetic of render_mega
cudaError_t launch_trace_primary(const DevScene& sc, const float* d_rays6, int n, float tmin, float tmax, int exact, int* d_type,
                                 int* d_index, float* d_t, int num_sms, cudaStream_t stream);
// fb[pixel] = float(acc[pixel]) * 2^-32 for the pixels of this rank's tiles (render_mega with nchunks > 1)
cudaError_t launch_finalize_chunks(const FrameParams& fp, int num_sms, cudaStream_t stream);
constexpr float kAccScale = 4294967296.0f, kAccInvScale = 2.3283064365386963e-10f;  // 2^32, 2^-32
cudaError_t launch_tonemap(const float* d_fb, size_t n_values, int divisor, unsigned char* d_out, int num_sms, cudaStream_t stream);
cudaError_t launch_tiles_pack(const float* d_fb, int W, int H, int rank, int nranks, float* d_packed, int num_sms, cudaStream_t stream);
// skip_rank >= 0: leave that rank's tiles alone (they are already in d_fb: the caller rendered them there); rank_stride:
// floats between two ranks' packed buffers (0 = tiles_packed_floats, i.e. back to back)
cudaError_t launch_tiles_unpack(const float* d_packed_all, int W, int H, int nranks, int skip_rank, size_t rank_stride, float* d_fb,
                                int num_sms, cudaStream_t stream);
size_t tiles_packed_floats(int W, int H, int nranks);
size_t mega_smem_bytes(int top_nodes, int threads);

}  // namespace rtb
