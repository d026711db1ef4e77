// kernels.cu -- the sm_100a kernels of the per-pixel render path.
//
// render_mega replaces the reference's render_kernel + ray_color (src/camera.cu:17-34, 218-252).  The reference
// maps one thread to one pixel and runs its whole spp x depth loop nest there, so a warp lives as long as its
// slowest pixel and dead lanes idle through every bounce.  Here the kernel is PERSISTENT (one grid sized to the
// machine: SMs x resident CTAs) and a warp alternates between two phases, all 32 lanes in phase:
//   SHADE  every lane finishes the segment it just traced (miss -> background; hit -> emission + scatter, one Philox
//          block drawn at a single site).  A lane whose path ended starts the next sample of ITS pixel in the same
//          phase, so lanes never idle on path-length variance (62 % of paths end at the first vertex, the tail runs to
//          max_depth); a lane whose pixel is complete takes the next pixel of the warp's current work item.  Work
//          items are 8x4-pixel blocks of the 16x16 tiles, fetched with ONE atomicAdd per warp (lane 0) and a shuffle
//          broadcast; ballot/popc ranks hand the pixels to the lanes that need one.
//   TRACE  every lane steps the closest-hit traversal of its segment (4-wide BVH, near-first, per-thread stack in local
//          memory).  The phase is TIME-SLICED: it ends when fewer than trace_min (6) lanes are still traversing and some lane
//          waits to be shaded; the unfinished lanes keep ref / sp / closest / the hit so far (their stack is in local memory),
//          sit out the SHADE phase and resume next to the fresh rays of the others.  The constants derived from (o, d) are
//          recomputed for every lane after SHADE, so none of them is live across the shading code: 72 registers, no spills,
//          4 CTAs x 224 threads per SM.  Measured: 180.1 vs 189.9 ms on the headline frame, 267 vs 295 ms on
//          the reference's shipped scene, against "every lane runs to completion" at 3 x 256 threads and 80 registers.
// Measured alternatives that lost and why are in DESIGN.md 4.2: a shared-memory stack, warp votes that batch the leaf tests,
// pre-tracing the next camera ray in idle lanes, a shared-memory lane counter instead of the vote, and the pooled variant
// (mega_pool.cu).
// Per-pixel sums are accumulated in registers in sample order (the reference's summation order) and stored once,
// so the output is deterministic and, in parity mode, comparable bit for bit.
#include "kernels.h"
#ifndef RTB_LOOP_SHAPE
#define RTB_LOOP_SHAPE 4
#endif
#include "path.cuh"
#include "smem_top.cuh"

namespace rtb {

// TOP: the first sc.top_nodes wide nodes are read from shared memory (measured: no gain on B200 -- the L1 already holds
// the top levels -- so the default launch uses TOP = false and saves the branch and two registers in the node loop)
// GRID: the scene has a sphere grid (scene_pack.h): its own instantiation, so that scenes without one do not carry the
// grid walk in their traversal loop (measured: compiled in but unused it cost the headline frame 15 %)
template <int MODE, bool X, bool COUNT, bool TOP, bool GRID>
__global__ void __launch_bounds__(kMegaThreads, kMegaBlocksPerSM) render_mega(const __grid_constant__ DevScene sc,
                                                                                const __grid_constant__ FrameParams fp) {
    extern __shared__ float4 smem4[];
    const TraversalSmem ts = carve_and_stage(sc, smem4);
    const float4* s_top = ts.top;
    float2 lstack[kLocalStack];
    TravStack stk;
    stk.lo = lstack;
    stk.tid = threadIdx.x;
    stk.nt = kMegaThreads;  // launch geometry is fixed: constant strides, no rematerialised address math in push / pop
    if (TOP) {
        stk.st = ts.st; stk.sr = ts.sr;
    } else {  // nothing staged: the stack sits at the start of dynamic shared memory, at compile-time offsets
        stk.st = reinterpret_cast<float*>(smem4);
        stk.sr = reinterpret_cast<int*>(smem4) + kSmemStack * kMegaThreads;
    }

    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int W = fp.cam.W, H = fp.cam.H, spp = fp.cam.spp;
    const int nchunks = fp.nchunks;

    // warp-uniform tile pool
    int pool_x0 = 0, pool_y0 = 0, pool_chunk = 0, pool_next = kBlockPixels;
    bool exhausted = false;
    unsigned queue;  // the work queue this warp draws from: its SM's, until that one is dry
    if (fp.queue_per_cta) queue = blockIdx.x;
    else asm("mov.u32 %0, %%smid;" : "=r"(queue));
    queue %= (unsigned)fp.nqueues;

    // per-lane state
    RayTrav rt;
    trav_init<X>(rt, mk3(0, 0, 0), mk3(0, 0, 1), 0.001f, 1e30f);
    rt.ref = kRefNeed;
    bool has_pixel = false;
    int px = 0, py = 0, sample = 0, chunk = 0;
    PathState ps;
    ps.o = mk3(0, 0, 0); ps.d = mk3(0, 0, 1); ps.beta = mk3(1, 1, 1); ps.L = mk3(0, 0, 0); ps.depth = 0;
    V3 acc = mk3(0, 0, 0);
    Rng<MODE> rng;
    rng.attach(0u, 0u, fp.key);
    uint32_t pixel_base = 0;
    WorkCount wc = {0, 0, 0, 0, 0, 0};

    for (;;) {
        // ================= SHADE phase: every lane of the warp is between two segments (in phase)
        {
            // ---- SHADE 1: second half of the finished segments (miss -> background, else emission + scatter)
            bool start_ray = false;
            if (rt.ref == kRefDone) {
                rng.attach((uint32_t)sample, (uint32_t)(py * W + px), fp.key);
                if (finish_segment<MODE, X, COUNT>(sc, fp.cam, fp.tex_mode, trav_hit(rt), ps, rng, wc)) {
                    // sample done: pixel += radiance (src/camera.cu:30), next sample or pixel done
                    acc = add3<X>(acc, ps.L);
                    ++sample;
                    if (sample >= min(spp, (chunk + 1) * fp.chunk_spp)) {
                        const size_t pix = 3 * ((size_t)py * W + px);
                        if (nchunks > 1) {
                            // fixed point, 2^-32 units: the float -> integer conversion is exact for |v| < 2^31 (below 2^-9 to
                            // 2^-33 absolute), so the accumulator holds the exact sum of the chunks' fp32 sums in any order
                            unsigned long long* a = reinterpret_cast<unsigned long long*>(fp.acc) + pix;
                            atomicAdd(a, (unsigned long long)__float2ll_rn(acc.x * kAccScale));
                            atomicAdd(a + 1, (unsigned long long)__float2ll_rn(acc.y * kAccScale));
                            atomicAdd(a + 2, (unsigned long long)__float2ll_rn(acc.z * kAccScale));
                        } else {
                            float* out = fp.fb + pix;
                            out[0] = acc.x; out[1] = acc.y; out[2] = acc.z;
                        }
                        has_pixel = false;
                    }
                    rt.ref = kRefNeed;
                } else {
                    start_ray = true;
                }
            }
            // ---- SHADE 2: hand out pixels to lanes that have none (warp-aggregated fetch)
            unsigned need = __ballot_sync(full, rt.ref == kRefNeed && !has_pixel);
            while (need && !exhausted) {
                if (pool_next >= kBlockPixels) {
                    // item = (tile, 8x4 block, sample chunk), chunk fastest: the chunks of a block are in flight together and
                    // find the block's part of the tree in cache.  Queue q holds the blocks kb = j * nqueues + q.
                    const unsigned per_chunk = (unsigned)fp.my_tiles * kBlocksPerTile;
                    const unsigned nq = (unsigned)fp.nqueues;
                    unsigned k = 0xffffffffu;
                    if (lane == 0) {
                        for (unsigned tries = 0; tries < nq; ++tries) {
                            const unsigned mine = (per_chunk + nq - 1u - queue) / nq * (unsigned)nchunks;  // items of this queue
                            const unsigned got = atomicAdd(fp.tile_counter + queue * kQueueStride, 1u);
                            if (got < mine) { k = got; break; }
                            queue = queue + 1u == nq ? 0u : queue + 1u;   // dry: steal from the next queue (and stay there)
                        }
                    }
                    k = __shfl_sync(full, k, 0);
                    queue = __shfl_sync(full, queue, 0);
                    if (k == 0xffffffffu) { exhausted = true; break; }
                    const unsigned kb = fp.chunk_major ? (k % ((per_chunk + nq - 1u - queue) / nq)) * nq + queue
                                                       : (k / (unsigned)nchunks) * nq + queue;
                    const int tile = tile_of_rank((int)(kb / kBlocksPerTile), fp.tile_rank, fp.tile_nranks);  // >= tile count: no pixels
                    const int blk = (int)(kb % kBlocksPerTile);  // 8 blocks of 8x4 pixels inside the 16x16 tile
                    pool_x0 = (tile % fp.tiles_x) * kTile + (blk & 1) * 8;
                    pool_y0 = (tile / fp.tiles_x) * kTile + (blk >> 1) * 4;
                    pool_chunk = (int)(fp.chunk_major ? k / ((per_chunk + nq - 1u - queue) / nq) : k % (unsigned)nchunks);
                    pool_next = 0;
                }
                int avail = kBlockPixels - pool_next;
                int rank = __popc(need & ((1u << lane) - 1u));
                if (rt.ref == kRefNeed && !has_pixel && rank < avail) {
                    int q = pool_next + rank;
                    int x = pool_x0 + (q & 7);
                    int y = pool_y0 + (q >> 3);
                    if (x < W && y < H && (spp <= 0 || fp.cam.depth <= 0)) {
                        float* out = fp.fb + 3 * ((size_t)y * W + x);  // no samples / no segments: the pixel is black
                        out[0] = 0.f; out[1] = 0.f; out[2] = 0.f;
                    } else if (x < W && y < H) {
                        has_pixel = true;
                        px = x; py = y;
                        chunk = pool_chunk;
                        sample = pool_chunk * fp.chunk_spp;
                        acc = mk3(0, 0, 0);
                        pixel_base = pixel_seed_base(x, y, W);
                    }
                }
                int n_need = __popc(need);
                pool_next += n_need < avail ? n_need : avail;
                need = __ballot_sync(full, rt.ref == kRefNeed && !has_pixel);
            }
            // ---- SHADE 3: start the next sample of the lane's pixel; lanes left without a pixel retire
            if (rt.ref == kRefNeed) {
                if (has_pixel) {
                    begin_path<MODE, X>(fp.cam, px, py, sample, pixel_base, fp.key, rng, ps);
                    if (COUNT) wc.samples++;
                    start_ray = true;
                } else {
                    rt.ref = kRefIdle;
                }
            }
#ifndef RTB_SPLIT_SETUP
#define RTB_SPLIT_SETUP 0
#endif
#if RTB_LOOP_SHAPE == 4 || RTB_SPLIT_SETUP
            // time-sliced TRACE: a lane may arrive here in the middle of a traversal (it keeps ref, sp, closest and the hit so
            // far; its stack is in local memory).  The constants derived from (o, d) are recomputed for EVERY lane -- the same
            // values for a resumed ray -- so none of them is live across the shading code above.
            trav_setup<X>(rt, ps.o, ps.d, 0.001f);
            if (start_ray) {
                trav_reset(rt, 1e30f);  // Interval(0.001f, 1e30f), src/camera.cu:226
                for (int k = 0; k < sc.num_hoisted; ++k) trav_leaf<X, COUNT>(sc, rt, wc, sc.hoisted[k]);
                if (GRID) trav_grid_begin<X>(sc, rt);
            }
#else
            if (start_ray) trav_begin<X, COUNT>(sc, rt, ps.o, ps.d, 0.001f, 1e30f, wc);  // Interval(0.001f, 1e30f), src/camera.cu:226
#endif
        }
        if (__all_sync(full, rt.ref == kRefIdle)) break;  // no pixels left for any lane
        // ================= TRACE phase: every lane runs its segment's closest-hit traversal to completion (traverse.cuh).
        // One iteration = (leaf test + pop)? then (wide node)?: a lane whose leaf test pops a wide node visits it in the same
        // iteration, together with the lanes that were at wide nodes anyway.  Measured on B200: warp votes
        // that batch the leaf tests, lanes waiting for each other, and postponed leaves (below) all cost more than they save.
#if RTB_LOOP_SHAPE == 1
        while (ref_is_active(rt.ref)) {
            if (GRID && ref_is_grid(rt.ref)) trav_grid_step<X, COUNT>(sc, rt, wc);
            if (ref_is_leaf(rt.ref)) trav_leaf_pop<X, COUNT>(sc, rt, stk, wc);
            if (ref_is_inner(rt.ref)) {
                if (trav_inner<X, COUNT, TOP>(sc, s_top, stk, rt, wc)) trav_pop(rt, stk);
            }
        }
#elif RTB_LOOP_SHAPE == 4
        // time-sliced: the phase ends as soon as fewer than trace_min lanes are still traversing and some lane waits to be
        // shaded.  The unfinished lanes sit out the SHADE phase (predicated off) and resume here, next to the fresh rays of
        // the others -- the traversal runs with more lanes per instruction, the shading code runs more often.
        for (;;) {
            const unsigned am = __ballot_sync(full, ref_is_active(rt.ref));
            if (am == 0u) break;
            if (__popc(am) < fp.trace_min && __any_sync(full, rt.ref == kRefDone)) break;
            if (GRID && ref_is_grid(rt.ref)) trav_grid_step<X, COUNT>(sc, rt, wc);
            if (ref_is_leaf(rt.ref)) trav_leaf_pop<X, COUNT>(sc, rt, stk, wc);
            if (ref_is_inner(rt.ref)) {
                if (trav_inner<X, COUNT, TOP>(sc, s_top, stk, rt, wc)) trav_pop(rt, stk);
            }
        }
#endif
    }

    // ---- work counters: warp reduce, one atomic per warp and counter
    unsigned int vals[6] = {wc.segments, wc.nodes, wc.spheres, wc.planes, wc.hits, wc.samples};
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        if (!COUNT && i != 0) continue;
        unsigned long long v = vals[i];
        for (int s = 16; s > 0; s >>= 1) v += __shfl_down_sync(full, v, s);
        if (lane == 0 && v) atomicAdd(fp.counters + i, v);
    }
}

size_t mega_smem_bytes(int top_nodes, int threads) {
    return (size_t)top_nodes * kTopStride * sizeof(float4) + (size_t)kSmemStack * threads * (sizeof(float) + sizeof(int));
}

template <int MODE, bool X, bool COUNT, bool TOP, bool GRID>
static cudaError_t launch_mega_tt(const DevScene& sc, const FrameParams& fp, LaunchCfg cfg, cudaStream_t stream) {
    size_t smem = mega_smem_bytes(sc.top_nodes, cfg.threads);
    cudaError_t e = cudaFuncSetAttribute(render_mega<MODE, X, COUNT, TOP, GRID>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    render_mega<MODE, X, COUNT, TOP, GRID><<<cfg.num_sms * cfg.blocks_per_sm, cfg.threads, smem, stream>>>(sc, fp);
    return cudaGetLastError();
}
template <int MODE, bool X, bool COUNT>
static cudaError_t launch_mega_t(const DevScene& sc, const FrameParams& fp, LaunchCfg cfg, cudaStream_t stream) {
    if (sc.grid_on) {  // nothing is staged in shared memory on this path
        DevScene g = sc;
        g.top_nodes = 0;
        return launch_mega_tt<MODE, X, COUNT, false, true>(g, fp, cfg, stream);
    }
    return sc.top_nodes > 0 ? launch_mega_tt<MODE, X, COUNT, true, false>(sc, fp, cfg, stream)
                            : launch_mega_tt<MODE, X, COUNT, false, false>(sc, fp, cfg, stream);
}

cudaError_t launch_render_mega(const DevScene& sc, const FrameParams& fp, int rng, int exact, int count, LaunchCfg cfg,
                               cudaStream_t stream) {
    int sel = (rng ? 4 : 0) | (exact ? 2 : 0) | (count ? 1 : 0);
    switch (sel) {
        case 0: return launch_mega_t<kRngPhilox, false, false>(sc, fp, cfg, stream);
        case 1: return launch_mega_t<kRngPhilox, false, true>(sc, fp, cfg, stream);
        case 2: return launch_mega_t<kRngPhilox, true, false>(sc, fp, cfg, stream);
        case 3: return launch_mega_t<kRngPhilox, true, true>(sc, fp, cfg, stream);
        case 4: return launch_mega_t<kRngReference, false, false>(sc, fp, cfg, stream);
        case 5: return launch_mega_t<kRngReference, false, true>(sc, fp, cfg, stream);
        case 6: return launch_mega_t<kRngReference, true, false>(sc, fp, cfg, stream);
        default: return launch_mega_t<kRngReference, true, true>(sc, fp, cfg, stream);
    }
}

// ---------------------------------------------------------------- primary-ray closest hit (parity entry point)
// X = true: the exact-math leaf tests (the reference's operation order, include/sphere.h:24-47, include/plane.h:57-96).
// X = false: the PRODUCTION arithmetic -- the very functions render_mega<.., false, ..> is built from (trav_begin with the
// hoisted primitives, trav_leaf_pop -> sphere_t<false> / plane_t<false>, trav_inner<false> with the un-widened slab test, the
// SFU approximations as compiled for the device), driven by the megakernel's own loop shape, so the ids / t this entry
// point returns are the ids / t the benchmarked kernel shades.
template <bool X>
__global__ void __launch_bounds__(256, 2) trace_primary_kernel(const __grid_constant__ DevScene sc, const float* __restrict__ rays6,
                                                               int n, float tmin, float tmax, int* __restrict__ type,
                                                               int* __restrict__ index, float* __restrict__ t) {
    extern __shared__ float4 smem4[];
    const TraversalSmem ts = carve_and_stage(sc, smem4);
    const float4* s_top = ts.top;
    float* s_st = ts.st;
    int* s_sr = ts.sr;
    WorkCount wc = {0, 0, 0, 0, 0, 0};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float* r = rays6 + 6 * (size_t)i;
        HitInfo hit;
        if (X) {
            trace_closest<true, false>(sc, s_top, s_st, s_sr, mk3(r[0], r[1], r[2]), mk3(r[3], r[4], r[5]), tmin, tmax, hit, wc);
        } else {
            float2 lstack[kLocalStack];
            TravStack stk;
            stk.st = s_st; stk.sr = s_sr; stk.lo = lstack;
            stk.tid = threadIdx.x; stk.nt = blockDim.x;
            RayTrav rt;
            trav_begin<false, false>(sc, rt, mk3(r[0], r[1], r[2]), mk3(r[3], r[4], r[5]), tmin, tmax, wc);
            while (ref_is_active(rt.ref)) {  // render_mega's TRACE loop body
                if (ref_is_grid(rt.ref)) trav_grid_step<false, false>(sc, rt, wc);
                if (ref_is_leaf(rt.ref)) trav_leaf_pop<false, false>(sc, rt, stk, wc);
                if (ref_is_inner(rt.ref)) {
                    if (trav_inner<false, false, true>(sc, s_top, stk, rt, wc)) trav_pop(rt, stk);
                }
            }
            hit = trav_hit(rt);
            if (rt.htype >= 0)  // fast mode does not carry the caller's index through the traversal
                hit.orig = rt.htype == 0 ? __ldg(sc.sphere_meta + rt.hslot).y : __float_as_int(__ldg(sc.planes + 5 * (size_t)rt.hslot + 3).w);
        }
        type[i] = hit.type;
        index[i] = hit.orig;
        t[i] = hit.type < 0 ? 0.0f : hit.t;
    }
}

template <bool X>
static cudaError_t launch_trace_primary_t(const DevScene& sc, const float* d_rays6, int n, float tmin, float tmax, int* d_type,
                                          int* d_index, float* d_t, int num_sms, cudaStream_t stream) {
    const int threads = 256;
    size_t smem = mega_smem_bytes(sc.top_nodes, threads);
    cudaError_t e = cudaFuncSetAttribute(trace_primary_kernel<X>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int blocks = (n + threads - 1) / threads;
    if (blocks > num_sms * 2 * 4) blocks = num_sms * 2 * 4;
    if (blocks < 1) blocks = 1;
    trace_primary_kernel<X><<<blocks, threads, smem, stream>>>(sc, d_rays6, n, tmin, tmax, d_type, d_index, d_t);
    return cudaGetLastError();
}
cudaError_t launch_trace_primary(const DevScene& sc, const float* d_rays6, int n, float tmin, float tmax, int exact, int* d_type,
                                 int* d_index, float* d_t, int num_sms, cudaStream_t stream) {
    return exact ? launch_trace_primary_t<true>(sc, d_rays6, n, tmin, tmax, d_type, d_index, d_t, num_sms, stream)
                 : launch_trace_primary_t<false>(sc, d_rays6, n, tmin, tmax, d_type, d_index, d_t, num_sms, stream);
}

// ---------------------------------------------------------------- sample chunks -> framebuffer (render_mega with nchunks > 1)
// fb = float(acc) * 2^-32 for the pixels of this rank's tiles (the other pixels of fb are not this rank's to write).
// A tile row is 48 consecutive values: one thread per value, 3 x 16 threads per row.
__global__ void finalize_chunks_kernel(const __grid_constant__ FrameParams fp) {
    const int W = fp.cam.W, H = fp.cam.H;
    const int ntiles = fp.tiles_x * fp.tiles_y;
    for (int k = blockIdx.x; k < fp.my_tiles; k += gridDim.x) {
        const int t = tile_of_rank(k, fp.tile_rank, fp.tile_nranks);
        if (t >= ntiles) continue;
        const int x0 = (t % fp.tiles_x) * kTile, y0 = (t / fp.tiles_x) * kTile;
        const int nx = min(kTile, W - x0) * 3;  // values per row of this tile
        for (int q = threadIdx.x; q < kTile * kTile * 3; q += blockDim.x) {
            const int row = q / (kTile * 3), c = q % (kTile * 3);
            const int y = y0 + row;
            if (y < H && c < nx) {
                const size_t i = 3 * ((size_t)y * W + x0) + c;
                fp.fb[i] = __ll2float_rn(__ldcs(fp.acc + i)) * kAccInvScale;
            }
        }
    }
}
cudaError_t launch_finalize_chunks(const FrameParams& fp, int num_sms, cudaStream_t stream) {
    finalize_chunks_kernel<<<num_sms * 8, 256, 0, stream>>>(fp);
    return cudaGetLastError();
}

// ---------------------------------------------------------------- output stage (ISaver::writeColor arithmetic)
// src/camera.cu:136-148: c/=N (multiply by the double reciprocal narrowed), sqrt, clamp [0,0.999], *256, truncate
__global__ void tonemap_kernel(const float* __restrict__ fb, size_t n, float inv_div, unsigned char* __restrict__ out) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        float v = __fsqrt_rn(__fmul_rn(inv_div, fb[i]));
        v = v < 0.0f ? 0.0f : (v > 0.999f ? 0.999f : v);
        out[i] = (unsigned char)(int)__fmul_rn(256.0f, v);
    }
}
cudaError_t launch_tonemap(const float* d_fb, size_t n_values, int divisor, unsigned char* d_out, int num_sms, cudaStream_t stream) {
    float inv = (float)(1.0 / (double)(float)divisor);
    int blocks = (int)((n_values + 255) / 256);
    if (blocks > num_sms * 8) blocks = num_sms * 8;
    if (blocks < 1) blocks = 1;
    tonemap_kernel<<<blocks, 256, 0, stream>>>(d_fb, n_values, inv, d_out);
    return cudaGetLastError();
}

// ---------------------------------------------------------------- tile pack / unpack for the multi-GPU gather
// rank r's packed buffer holds its tiles in increasing k (kernels.h: tile_of_rank), 16x16x3 floats each.
size_t tiles_packed_floats(int W, int H, int nranks) {
    size_t tiles = (size_t)((W + kTile - 1) / kTile) * ((H + kTile - 1) / kTile);
    size_t per = (tiles + nranks - 1) / nranks;
    return per * kTile * kTile * 3;
}
__global__ void tiles_pack_kernel(const float* __restrict__ fb, int W, int H, int rank, int nranks, float* __restrict__ packed) {
    const int tiles_x = (W + kTile - 1) / kTile, ntiles = tiles_x * ((H + kTile - 1) / kTile);
    const int my = (ntiles + nranks - 1) / nranks;
    for (int k = blockIdx.x; k < my; k += gridDim.x) {
        const int t = tile_of_rank(k, rank, nranks);
        const int x = (t % tiles_x) * kTile + (threadIdx.x & 15), y = (t / tiles_x) * kTile + (threadIdx.x >> 4);
        float* dst = packed + ((size_t)k * kTile * kTile + threadIdx.x) * 3;
        if (t < ntiles && x < W && y < H) {
            const float* src = fb + 3 * ((size_t)y * W + x);
            dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2];
        } else {
            dst[0] = dst[1] = dst[2] = 0.0f;
        }
    }
}
__global__ void tiles_unpack_kernel(const float* __restrict__ all, int W, int H, int nranks, int skip_rank, size_t per_rank,
                                    float* __restrict__ fb) {
    const int tiles_x = (W + kTile - 1) / kTile, ntiles = tiles_x * ((H + kTile - 1) / kTile);
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
        int r, k;
        tile_owner(t, nranks, r, k);
        if (r == skip_rank) continue;
        const int x = (t % tiles_x) * kTile + (threadIdx.x & 15), y = (t / tiles_x) * kTile + (threadIdx.x >> 4);
        if (x < W && y < H) {
            const float* src = all + (size_t)r * per_rank + ((size_t)k * kTile * kTile + threadIdx.x) * 3;
            float* dst = fb + 3 * ((size_t)y * W + x);
            dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2];
        }
    }
}
cudaError_t launch_tiles_pack(const float* d_fb, int W, int H, int rank, int nranks, float* d_packed, int num_sms, cudaStream_t stream) {
    tiles_pack_kernel<<<num_sms * 4, kTile * kTile, 0, stream>>>(d_fb, W, H, rank, nranks, d_packed);
    return cudaGetLastError();
}
cudaError_t launch_tiles_unpack(const float* d_packed_all, int W, int H, int nranks, int skip_rank, size_t rank_stride, float* d_fb,
                                int num_sms, cudaStream_t stream) {
    if (rank_stride == 0) rank_stride = tiles_packed_floats(W, H, nranks);
    tiles_unpack_kernel<<<num_sms * 4, kTile * kTile, 0, stream>>>(d_packed_all, W, H, nranks, skip_rank, rank_stride, d_fb);
    return cudaGetLastError();
}

}  // namespace rtb
