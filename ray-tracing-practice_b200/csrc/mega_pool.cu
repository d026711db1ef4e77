// mega_pool.cu -- render_pool: the persistent megakernel with a per-warp path pool (pool.cuh).
//
// Replaces the reference's render_kernel + ray_color (src/camera.cu:17-34, 218-252).  Each warp owns kPool path
// slots in shared memory and schedules three kinds of work over them, choosing by ballot/popc each iteration:
//   REFILL  lanes whose traversal finished park (t, primitive) in their slot, append the slot to the warp's hit
//           queue and take the next slot from the ready queue (6 floats + trav_begin);
//   SHADE   when 32 hits are parked: one lane per hit runs the second half of the segment (emission, scatter, next
//           sample, next pixel) -- the divergent code always runs with a full warp;
//   LEAF / INNER traversal steps (traverse.cuh) for the lanes that hold a leaf / an inner node.
// A slot owns one pixel at a time and runs its samples in order, so per-pixel sums are formed in sample order (the
// reference's summation order) and stored once: deterministic, and bit-comparable in parity mode.
#include "kernels.h"
#include "pool.cuh"
#include "smem_top.cuh"

namespace rtb {

// warp-aggregated pixel handout: lanes with `want` get the next pixel of the warp's 16x16 tile (one atomicAdd per
// tile by lane 0).  Returns true with (x, y) for lanes that received a pixel.
__device__ __forceinline__ bool take_pixel(const FrameParams& fp, WarpSched& ws, bool want, int lane, int& x_out, int& y_out) {
    const unsigned full = 0xffffffffu;
    const int W = fp.cam.W, H = fp.cam.H;
    bool got = false;
    unsigned need = __ballot_sync(full, want);
    while (need && !ws.exhausted) {
        if (ws.pool_next >= kBlockPixels) {
            unsigned k = 0;
            if (lane == 0) k = atomicAdd(fp.tile_counter, 1u);
            k = __shfl_sync(full, k, 0);
            if ((int)(k / kBlocksPerTile) >= fp.my_tiles) { ws.exhausted = true; break; }
            ws.pool_tile = tile_of_rank((int)(k / kBlocksPerTile), fp.tile_rank, fp.tile_nranks);
            ws.pool_blk = (int)(k % kBlocksPerTile);
            ws.pool_next = 0;
        }
        int avail = kBlockPixels - ws.pool_next;
        int rank = __popc(need & ((1u << lane) - 1u));
        if (want && !got && rank < avail) {
            int q = ws.pool_next + rank;
            int blk = ws.pool_blk;  // 8 blocks of 8x4 pixels inside the 16x16 tile
            int x = (ws.pool_tile % fp.tiles_x) * kTile + (blk & 1) * 8 + (q & 7);
            int y = (ws.pool_tile / fp.tiles_x) * kTile + (blk >> 1) * 4 + (q >> 3);
            if (x < W && y < H && (fp.cam.spp <= 0 || fp.cam.depth <= 0)) {
                float* out = fp.fb + 3 * ((size_t)y * W + x);  // no samples / no segments: the pixel is black
                out[0] = 0.f; out[1] = 0.f; out[2] = 0.f;
            } else if (x < W && y < H) {
                got = true;
                x_out = x; y_out = y;
            }
        }
        int n_need = __popc(need);
        ws.pool_next += n_need < avail ? n_need : avail;
        need = __ballot_sync(full, want && !got);
    }
    return got;
}

// One shading pass: lane `lane` (if act) works on slot s.  fresh: the slot has no hit to shade, it only needs a pixel.
// Slots that end with a ray to trace are appended to the ready queue.
template <int MODE, bool X, bool COUNT>
__device__ __forceinline__ void shade_pass(const DevScene& sc, const FrameParams& fp, const WarpPool& wp, WarpSched& ws, bool act, int s,
                                           bool fresh, int lane, WorkCount& wc) {
    const unsigned full = 0xffffffffu;
    const int W = fp.cam.W, spp = fp.cam.spp;
    bool produce = false, needpix = act && fresh, newsample = false;
    PathState ps;
    Rng<MODE> rng;
    rng.attach(0u, 0u, fp.key);
    V3 acc = mk3(0, 0, 0);
    int px = 0, py = 0, sample = 0;
    if (act && !fresh) {
        int pix = wp.in(kFPix, s);
        px = pix & 0xffff; py = (int)((unsigned)pix >> 16);
        sample = wp.in(kFSample, s);
        ps.o = wp.ld3(kFOx, s); ps.d = wp.ld3(kFDx, s); ps.beta = wp.ld3(kFBx, s); ps.L = wp.ld3(kFLx, s);
        ps.depth = wp.in(kFDepth, s);
        HitInfo hit;
        hit.t = wp.fl(kFHitT, s);
        int href = wp.in(kFHitRef, s);
        hit.type = href < 0 ? -1 : (href >> 30);
        hit.slot = href & 0x3fffffff;
        hit.orig = -1;
        rng.attach((uint32_t)sample, (uint32_t)(py * W + px), fp.key);
        rng.restore((uint32_t)wp.in(kFRng, s));
        if (!finish_segment<MODE, X, COUNT>(sc, fp.cam, fp.tex_mode, hit, ps, rng, wc)) {
            wp.st3(kFOx, s, ps.o); wp.st3(kFDx, s, ps.d); wp.st3(kFBx, s, ps.beta); wp.st3(kFLx, s, ps.L);
            wp.in(kFDepth, s) = ps.depth;
            if (MODE == kRngReference) wp.in(kFRng, s) = (int)rng.save();
            produce = true;
        } else {
            // sample done: pixel += radiance (src/camera.cu:30), next sample or pixel done
            acc = add3<X>(wp.ld3(kFAx, s), ps.L);
            ++sample;
            if (sample >= spp) {
                float* out = fp.fb + 3 * ((size_t)py * W + px);
                out[0] = acc.x; out[1] = acc.y; out[2] = acc.z;
                needpix = true;
            } else {
                newsample = true;
            }
        }
    }
    if (__any_sync(full, needpix)) {
        if (take_pixel(fp, ws, needpix, lane, px, py)) {
            acc = mk3(0, 0, 0);
            sample = 0;
            newsample = true;
        }
    }
    if (newsample) {
        begin_path<MODE, X>(fp.cam, px, py, sample, pixel_seed_base(px, py, W), fp.key, rng, ps);
        if (COUNT) wc.samples++;
        wp.st3(kFOx, s, ps.o); wp.st3(kFDx, s, ps.d); wp.st3(kFBx, s, ps.beta); wp.st3(kFLx, s, ps.L);
        wp.st3(kFAx, s, acc);
        wp.in(kFDepth, s) = 0;
        wp.in(kFPix, s) = px | (py << 16);
        wp.in(kFSample, s) = sample;
        if (MODE == kRngReference) wp.in(kFRng, s) = (int)rng.save();
        produce = true;
    }
    unsigned pm = __ballot_sync(full, produce);
    if (produce) wp.in(kFReadyQ, (ws.ready_head + ws.ready_cnt + __popc(pm & ((1u << lane) - 1u))) & (kPool - 1)) = s;
    ws.ready_cnt += __popc(pm);
    __syncwarp();
}

// north_star subsystem 3, "regroup by material after a bounce" (the divergent switch it targets is material_scatter's,
// include/materials.h:70-140): when more hits are parked than one shading pass takes, the hit queue (<= kPool = 64 entries,
// two per lane) is stably ordered by the class of what was hit -- miss, LAMBERTIAN, METAL, DIELECTRIC, DIFFUSE_LIGHT -- with
// one ballot-ranked counting sort, so that the 32 hits of a pass fall into as few branches of shade_vertex as possible.
// Which hits share a pass changes nothing in the image: every slot's sample stream is its own.
__device__ __forceinline__ void regroup_hits_by_material(const DevScene& sc, const WarpPool& wp, WarpSched& ws, int lane) {
    const unsigned full = 0xffffffffu, lt = (1u << lane) - 1u;
    int e[2], c[2], r[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int q = lane + 32 * h;
        e[h] = q < ws.hit_cnt ? wp.in(kFHitQ, (ws.hit_head + q) & (kPool - 1)) : -1;
        c[h] = 7;
        if (e[h] >= 0) {
            const int href = wp.in(kFHitRef, e[h]);
            c[h] = 0;
            if (href >= 0) {
                const int slot = href & 0x3fffffff;
                const int mat = (href >> 30) == 0 ? __ldg(sc.sphere_meta + slot).x : __float_as_int(__ldg(sc.planes + 5 * (size_t)slot + 2).w);
                c[h] = 1 + __float_as_int(__ldg(sc.mats + 3 * (size_t)mat).w);
            }
        }
        r[h] = 0;
    }
    int base = 0;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        const unsigned m0 = __ballot_sync(full, c[0] == k), m1 = __ballot_sync(full, c[1] == k);
        if (c[0] == k) r[0] = base + __popc(m0 & lt);
        if (c[1] == k) r[1] = base + __popc(m0) + __popc(m1 & lt);
        base += __popc(m0) + __popc(m1);
    }
    __syncwarp();
#pragma unroll
    for (int h = 0; h < 2; ++h)
        if (e[h] >= 0) wp.in(kFHitQ, (ws.hit_head + r[h]) & (kPool - 1)) = e[h];
    __syncwarp();
}

template <int MODE, bool X, bool COUNT>
__global__ void __launch_bounds__(kPoolThreads, kPoolBlocksPerSM) render_pool(const __grid_constant__ DevScene sc,
                                                                                const __grid_constant__ FrameParams fp) {
    extern __shared__ float4 smem4[];
    const TraversalSmem ts = carve_and_stage(sc, smem4);
    const float4* s_top = ts.top;
    float2 lstack[kLocalStack];
    TravStack stk;
    stk.st = ts.st; stk.sr = ts.sr; stk.lo = lstack;
    stk.tid = threadIdx.x; stk.nt = blockDim.x;

    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    WarpPool wp;
    wp.f = reinterpret_cast<float*>(ts.sr + kSmemStack * blockDim.x) + (threadIdx.x >> 5) * kPoolWords;
    const int refill_min = fp.refill_min, leaf_min = fp.leaf_min, park_min = fp.park_min;

    WarpSched ws;
    ws.pool_tile = 0; ws.pool_blk = 0; ws.pool_next = kBlockPixels; ws.exhausted = false;
    ws.ready_head = 0; ws.ready_cnt = 0; ws.hit_head = 0; ws.hit_cnt = 0;
    WorkCount wc = {0, 0, 0, 0, 0, 0};

    RayTrav rt;
    trav_init<X>(rt, mk3(0, 0, 0), mk3(0, 0, 1), 0.001f, 1e30f);
    rt.ref = kRefIdle;
    int slot = 0;

    // every slot starts by asking for a pixel
#pragma unroll 1
    for (int k = 0; k < kPool; k += 32) shade_pass<MODE, X, COUNT>(sc, fp, wp, ws, true, k + lane, true, lane, wc);

    for (;;) {
        const unsigned finm = __ballot_sync(full, rt.ref == kRefDone);
        const unsigned idlem = __ballot_sync(full, rt.ref == kRefIdle);
        const unsigned travm = ~(finm | idlem);
        // ---- REFILL: park finished traversals, hand ready slots to lanes without a ray
        if ((finm != 0u && (__popc(finm) >= refill_min || travm == 0u)) || (idlem != 0u && ws.ready_cnt > 0)) {
            if (rt.ref == kRefDone) {
                wp.fl(kFHitT, slot) = rt.closest;
                wp.in(kFHitRef, slot) = rt.htype < 0 ? -1 : ((rt.htype << 30) | rt.hslot);
                wp.in(kFHitQ, (ws.hit_head + ws.hit_cnt + __popc(finm & lt_mask)) & (kPool - 1)) = slot;
                rt.ref = kRefIdle;
            }
            ws.hit_cnt += __popc(finm);
            __syncwarp();
            const unsigned want = finm | idlem;
            const int n = min(__popc(want), ws.ready_cnt);
            const int r = __popc(want & lt_mask);
            if (rt.ref == kRefIdle && r < n) {
                slot = wp.in(kFReadyQ, (ws.ready_head + r) & (kPool - 1));
                trav_begin<X, COUNT>(sc, rt, wp.ld3(kFOx, slot), wp.ld3(kFDx, slot), 0.001f, 1e30f, wc);  // Interval(0.001f, 1e30f), src/camera.cu:226
            }
            ws.ready_head = (ws.ready_head + n) & (kPool - 1);
            ws.ready_cnt -= n;
        }
        // ---- SHADE: a full warp of parked hits, or whatever is parked when lanes starve
        const unsigned starve = __ballot_sync(full, rt.ref == kRefIdle);
        if (ws.hit_cnt >= park_min || (ws.hit_cnt > 0 && ws.ready_cnt == 0 && starve != 0u)) {
            if (fp.regroup && ws.hit_cnt > 32) regroup_hits_by_material(sc, wp, ws, lane);
            const int n = min(ws.hit_cnt, 32);
            const int s = wp.in(kFHitQ, (ws.hit_head + lane) & (kPool - 1));
            ws.hit_head = (ws.hit_head + n) & (kPool - 1);
            ws.hit_cnt -= n;
            shade_pass<MODE, X, COUNT>(sc, fp, wp, ws, lane < n, s, false, lane, wc);
            continue;  // refill before stepping
        }
        if (starve == full && ws.hit_cnt == 0 && ws.ready_cnt == 0) break;  // nothing parked, nothing ready, nobody tracing
        // ---- GRID: one step of the sphere-grid walk (scenes with a grid, scene_pack.h)
        if (ref_is_grid(rt.ref)) trav_grid_step<X, COUNT>(sc, rt, wc);
        // ---- LEAF: primitive tests, when enough lanes hold a leaf or no lane has an inner node to visit
        const unsigned lmask = __ballot_sync(full, ref_is_leaf(rt.ref));
        const unsigned imask = __ballot_sync(full, ref_is_inner(rt.ref));
        if (lmask != 0u && (__popc(lmask) >= leaf_min || imask == 0u)) {
            if (ref_is_leaf(rt.ref)) {
                trav_leaf<X, COUNT>(sc, rt, wc);
                trav_pop(rt, stk);
            }
        }
        // ---- INNER: one wide node for every lane that holds one
        if (ref_is_inner(rt.ref)) {
            if (trav_inner<X, COUNT>(sc, s_top, stk, rt, wc)) trav_pop(rt, stk);
        }
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

size_t pool_smem_bytes(int top_nodes, int threads) {
    return mega_smem_bytes(top_nodes, threads) + (size_t)(threads / 32) * kPoolWords * sizeof(float);
}

template <int MODE, bool X, bool COUNT>
static cudaError_t launch_pool_t(const DevScene& sc, const FrameParams& fp, LaunchCfg cfg, cudaStream_t stream) {
    size_t smem = pool_smem_bytes(sc.top_nodes, cfg.threads);
    cudaError_t e = cudaFuncSetAttribute(render_pool<MODE, X, COUNT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    render_pool<MODE, X, COUNT><<<cfg.num_sms * cfg.blocks_per_sm, cfg.threads, smem, stream>>>(sc, fp);
    return cudaGetLastError();
}

cudaError_t launch_render_pool(const DevScene& sc, const FrameParams& fp, int rng, int exact, int count, LaunchCfg cfg,
                               cudaStream_t stream) {
    int sel = (rng ? 4 : 0) | (exact ? 2 : 0) | (count ? 1 : 0);
    switch (sel) {
        case 0: return launch_pool_t<kRngPhilox, false, false>(sc, fp, cfg, stream);
        case 1: return launch_pool_t<kRngPhilox, false, true>(sc, fp, cfg, stream);
        case 2: return launch_pool_t<kRngPhilox, true, false>(sc, fp, cfg, stream);
        case 3: return launch_pool_t<kRngPhilox, true, true>(sc, fp, cfg, stream);
        case 4: return launch_pool_t<kRngReference, false, false>(sc, fp, cfg, stream);
        case 5: return launch_pool_t<kRngReference, false, true>(sc, fp, cfg, stream);
        case 6: return launch_pool_t<kRngReference, true, false>(sc, fp, cfg, stream);
        default: return launch_pool_t<kRngReference, true, true>(sc, fp, cfg, stream);
    }
}

}  // namespace rtb
