// wavefront.cu -- wavefront variant of the render path, kept to be measured against the megakernel (north_star (3)).
//
// One PASS renders K consecutive samples of every pixel this rank owns.  Paths live in global memory (SoA); each
// bounce is two kernels over a compact queue of live path ids:
//     wf_generate : camera rays for (pixel, sample) -> state, queue[0]
//     wf_extend   : closest hit for every queued path                     (traverse.cuh, same code as the megakernel)
//     wf_shade    : emission + scatter; survivors are appended to the next bounce's queue with a warp-aggregated
//                   atomic (ballot / popc compaction), finished paths leave their radiance in the state
//     wf_resolve  : pixel += sample radiances in sample order (the reference's summation order, src/camera.cu:30)
// Streams are keyed by (pixel, sample, bounce), so the image is bit-identical to the megakernel's.
// No host synchronisation inside a frame: queue lengths stay on the device, every bounce is launched.
#include "kernels.h"
#include "smem_top.cuh"

namespace rtb {
namespace {

struct WfState {
    float* o[3];
    float* d[3];
    float* beta[3];
    float* L[3];
    float* hit_t;
    int* hit_ref;          // type<<30 | slot ; -1 = miss
    uint32_t* seed;        // reference stream only
    int* queue[2];
    unsigned int* qcount;  // [max_depth + 2]
    int slots;             // pixel slots of this rank = my_tiles * 256
    int K;                 // samples per pass
};

__device__ __forceinline__ bool slot_pixel(const FrameParams& fp, int slot, int& x, int& y) {
    int tile = tile_of_rank(slot >> 8, fp.tile_rank, fp.tile_nranks), q = slot & 255, blk = q >> 5;
    x = (tile % fp.tiles_x) * kTile + (blk & 1) * 8 + (q & 7);
    y = (tile / fp.tiles_x) * kTile + (blk >> 1) * 4 + ((q >> 3) & 3);
    return x < fp.cam.W && y < fp.cam.H;
}

__device__ __forceinline__ void queue_push(bool want, int value, int* queue, unsigned int* count) {
    unsigned m = __ballot_sync(0xffffffffu, want);
    if (!m) return;
    int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
    unsigned base = 0;
    if (lane == leader) base = atomicAdd(count, (unsigned)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (want) queue[base + __popc(m & ((1u << lane) - 1u))] = value;
}

__device__ __forceinline__ void flush_counts(const WorkCount& wc, unsigned long long* counters, bool all) {
    unsigned int vals[6] = {wc.segments, wc.nodes, wc.spheres, wc.planes, wc.hits, wc.samples};
    int lane = threadIdx.x & 31;
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        if (!all && i != 0) continue;
        unsigned long long v = vals[i];
        for (int s = 16; s > 0; s >>= 1) v += __shfl_down_sync(0xffffffffu, v, s);
        if (lane == 0 && v) atomicAdd(counters + i, v);
    }
}

template <int MODE, bool X, bool COUNT>
__global__ void __launch_bounds__(256) wf_generate(const __grid_constant__ FrameParams fp, const __grid_constant__ WfState st, int s0) {
    WorkCount wc = {0, 0, 0, 0, 0, 0};
    int total = st.slots * st.K;
    int rounds = (total + gridDim.x * blockDim.x - 1) / (gridDim.x * blockDim.x);
    for (int r = 0; r < rounds; ++r) {
        int id = (r * gridDim.x + blockIdx.x) * blockDim.x + threadIdx.x;
        bool ok = false;
        if (id < total) {
            int slot = id % st.slots, k = id / st.slots, x, y;
            if (s0 + k < fp.cam.spp && slot_pixel(fp, slot, x, y)) {
                Rng<MODE> rng;
                PathState ps;
                begin_path<MODE, X>(fp.cam, x, y, s0 + k, pixel_seed_base(x, y, fp.cam.W), fp.key, rng, ps);
                st.o[0][id] = ps.o.x; st.o[1][id] = ps.o.y; st.o[2][id] = ps.o.z;
                st.d[0][id] = ps.d.x; st.d[1][id] = ps.d.y; st.d[2][id] = ps.d.z;
                st.beta[0][id] = 1.f; st.beta[1][id] = 1.f; st.beta[2][id] = 1.f;
                st.L[0][id] = 0.f; st.L[1][id] = 0.f; st.L[2][id] = 0.f;
                if constexpr (MODE == kRngReference) st.seed[id] = rng.s;
                ok = fp.cam.depth > 0;
                if (COUNT) wc.samples++;
            }
        }
        queue_push(ok, id, st.queue[0], st.qcount);
    }
    if (COUNT) flush_counts(wc, fp.counters, true);
}

template <bool X, bool COUNT>
__global__ void __launch_bounds__(256, 2) wf_extend(const __grid_constant__ DevScene sc, const __grid_constant__ FrameParams fp,
                                                    const __grid_constant__ WfState st, int bounce) {
    extern __shared__ float4 smem4[];
    const TraversalSmem ts = carve_and_stage(sc, smem4);
    const int* queue = st.queue[bounce & 1];
    const unsigned n = st.qcount[bounce];
    WorkCount wc = {0, 0, 0, 0, 0, 0};
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        int id = queue[i];
        HitInfo hit;
        trace_closest<X, COUNT>(sc, ts.top, ts.st, ts.sr, mk3(st.o[0][id], st.o[1][id], st.o[2][id]),
                                mk3(st.d[0][id], st.d[1][id], st.d[2][id]), 0.001f, 1e30f, hit, wc);
        st.hit_t[id] = hit.t;
        st.hit_ref[id] = hit.type < 0 ? -1 : ((hit.type << 30) | hit.slot);
    }
    if (COUNT) flush_counts(wc, fp.counters, true);
}

template <int MODE, bool X, bool COUNT>
__global__ void __launch_bounds__(256) wf_shade(const __grid_constant__ DevScene sc, const __grid_constant__ FrameParams fp,
                                                const __grid_constant__ WfState st, int bounce, int s0) {
    const int* queue = st.queue[bounce & 1];
    int* next = st.queue[(bounce + 1) & 1];
    const unsigned n = st.qcount[bounce];
    WorkCount wc = {0, 0, 0, 0, 0, 0};
    unsigned rounds = (n + gridDim.x * blockDim.x - 1) / (gridDim.x * blockDim.x);
    for (unsigned r = 0; r < rounds; ++r) {
        unsigned i = (r * gridDim.x + blockIdx.x) * blockDim.x + threadIdx.x;
        bool alive = false;
        int id = 0;
        if (i < n) {
            id = queue[i];
            PathState ps;
            ps.o = mk3(st.o[0][id], st.o[1][id], st.o[2][id]);
            ps.d = mk3(st.d[0][id], st.d[1][id], st.d[2][id]);
            ps.beta = mk3(st.beta[0][id], st.beta[1][id], st.beta[2][id]);
            ps.L = mk3(st.L[0][id], st.L[1][id], st.L[2][id]);
            ps.depth = bounce;
            HitInfo hit;
            int ref = st.hit_ref[id];
            hit.t = st.hit_t[id];
            hit.type = ref < 0 ? -1 : (ref >> 30) & 1;
            hit.slot = ref & 0x3fffffff;
            hit.orig = -1;
            Rng<MODE> rng;
            if constexpr (MODE == kRngReference) {
                rng.s = st.seed[id];
            } else {
                int slot = id % st.slots, k = id / st.slots, x, y;
                slot_pixel(fp, slot, x, y);
                rng.attach((uint32_t)(s0 + k), (uint32_t)(y * fp.cam.W + x), fp.key);
            }
            alive = !finish_segment<MODE, X, COUNT>(sc, fp.cam, fp.tex_mode, hit, ps, rng, wc);
            st.L[0][id] = ps.L.x; st.L[1][id] = ps.L.y; st.L[2][id] = ps.L.z;
            if (alive) {
                st.o[0][id] = ps.o.x; st.o[1][id] = ps.o.y; st.o[2][id] = ps.o.z;
                st.d[0][id] = ps.d.x; st.d[1][id] = ps.d.y; st.d[2][id] = ps.d.z;
                st.beta[0][id] = ps.beta.x; st.beta[1][id] = ps.beta.y; st.beta[2][id] = ps.beta.z;
                if constexpr (MODE == kRngReference) st.seed[id] = rng.s;
            }
        }
        queue_push(alive, id, next, st.qcount + bounce + 1);
    }
    flush_counts(wc, fp.counters, COUNT);
}

template <bool X>
__global__ void __launch_bounds__(256) wf_resolve(const __grid_constant__ FrameParams fp, const __grid_constant__ WfState st, int s0) {
    for (int slot = blockIdx.x * blockDim.x + threadIdx.x; slot < st.slots; slot += gridDim.x * blockDim.x) {
        int x, y;
        if (!slot_pixel(fp, slot, x, y)) continue;
        float* out = fp.fb + 3 * ((size_t)y * fp.cam.W + x);
        V3 acc = s0 == 0 ? mk3(0, 0, 0) : mk3(out[0], out[1], out[2]);
        for (int k = 0; k < st.K && s0 + k < fp.cam.spp; ++k) {
            int id = k * st.slots + slot;
            acc = add3<X>(acc, mk3(st.L[0][id], st.L[1][id], st.L[2][id]));
        }
        out[0] = acc.x; out[1] = acc.y; out[2] = acc.z;
    }
}

template <int MODE, bool X, bool COUNT>
cudaError_t run(const DevScene& sc, const FrameParams& fp, LaunchCfg cfg, const WfState& st, cudaStream_t stream) {
    const int threads = 256;
    size_t smem = mega_smem_bytes(sc.top_nodes, threads);
    cudaError_t e = cudaFuncSetAttribute(wf_extend<X, COUNT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int grid_wide = cfg.num_sms * 8, grid_ext = cfg.num_sms * 2 * 2;
    if (fp.cam.spp <= 0) {  // nothing to sample: the frame is all zeros
        wf_resolve<X><<<grid_wide, threads, 0, stream>>>(fp, st, 0);
        return cudaGetLastError();
    }
    for (int s0 = 0; s0 < fp.cam.spp; s0 += st.K) {
        e = cudaMemsetAsync(st.qcount, 0, sizeof(unsigned int) * (fp.cam.depth + 2), stream);
        if (e != cudaSuccess) return e;
        wf_generate<MODE, X, COUNT><<<grid_wide, threads, 0, stream>>>(fp, st, s0);
        for (int b = 0; b < fp.cam.depth; ++b) {
            wf_extend<X, COUNT><<<grid_ext, threads, smem, stream>>>(sc, fp, st, b);
            wf_shade<MODE, X, COUNT><<<grid_wide, threads, 0, stream>>>(sc, fp, st, b, s0);
        }
        wf_resolve<X><<<grid_wide, threads, 0, stream>>>(fp, st, s0);
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

}  // namespace

cudaError_t launch_render_wavefront(const DevScene& sc, const FrameParams& fp, int rng, int exact, int count, LaunchCfg cfg,
                                    WavefrontScratch& scratch, cudaStream_t stream) {
    WfState st;
    st.slots = fp.my_tiles * kTile * kTile;
    if (st.slots <= 0) return cudaSuccess;
    long long budget = 16ll << 20;  // paths in flight per pass
    long long K = budget / st.slots;
    if (K < 1) K = 1;
    if (K > (fp.cam.spp > 0 ? fp.cam.spp : 1)) K = fp.cam.spp > 0 ? fp.cam.spp : 1;
    st.K = (int)K;
    size_t paths = (size_t)st.slots * st.K;
    size_t words = paths * (12 + 2 + 1 + 2) + (size_t)fp.cam.depth + 2 + 64;
    size_t bytes = words * 4;
    if (scratch.bytes < bytes) {
        if (scratch.buf) cudaFree(scratch.buf);
        scratch.buf = nullptr;
        scratch.bytes = 0;
        cudaError_t e = cudaMalloc(&scratch.buf, bytes);
        if (e != cudaSuccess) return e;
        scratch.bytes = bytes;
    }
    float* p = (float*)scratch.buf;
    for (int a = 0; a < 3; ++a) { st.o[a] = p; p += paths; }
    for (int a = 0; a < 3; ++a) { st.d[a] = p; p += paths; }
    for (int a = 0; a < 3; ++a) { st.beta[a] = p; p += paths; }
    for (int a = 0; a < 3; ++a) { st.L[a] = p; p += paths; }
    st.hit_t = p; p += paths;
    st.hit_ref = (int*)p; p += paths;
    st.seed = (uint32_t*)p; p += paths;
    st.queue[0] = (int*)p; p += paths;
    st.queue[1] = (int*)p; p += paths;
    st.qcount = (unsigned int*)p;
    int sel = (rng ? 4 : 0) | (exact ? 2 : 0) | (count ? 1 : 0);
    switch (sel) {
        case 0: return run<kRngPhilox, false, false>(sc, fp, cfg, st, stream);
        case 1: return run<kRngPhilox, false, true>(sc, fp, cfg, st, stream);
        case 2: return run<kRngPhilox, true, false>(sc, fp, cfg, st, stream);
        case 3: return run<kRngPhilox, true, true>(sc, fp, cfg, st, stream);
        case 4: return run<kRngReference, false, false>(sc, fp, cfg, st, stream);
        case 5: return run<kRngReference, false, true>(sc, fp, cfg, st, stream);
        case 6: return run<kRngReference, true, false>(sc, fp, cfg, st, stream);
        default: return run<kRngReference, true, true>(sc, fp, cfg, st, stream);
    }
}

}  // namespace rtb
