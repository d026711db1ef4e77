// rt_api.cu -- the C-ABI declared in include/rt_b200.h: scene upload, frame launch, parity trace, output stage.
// Owns every device allocation behind the opaque rt_scene handle (the reference leaves them to the caller and
// leaks the BVH node array, src/main.cu:453-474 vs :322-344).  No CPU fallback: every compute entry point needs
// a CUDA device and fails with RT_ERR_NO_DEVICE / RT_ERR_CUDA otherwise.
#include <cuda_runtime.h>

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "bvh_gpu.h"
#include "rt_scene_impl.h"
#include "scene_pack.h"
#include "wide_bvh.h"

using namespace rtb;

namespace rtb {
std::string& last_error() {
    thread_local std::string e;
    return e;
}
}  // namespace rtb

static int check_device() {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) {
        cudaGetLastError();
        return fail(RT_ERR_NO_DEVICE, std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0"));
    }
    return RT_OK;
}

extern "C" {

const char* rt_last_error(void) { return last_error().c_str(); }
const char* rt_version(void) { return "rt_b200 0.1 (sm_100a)"; }
int rt_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int rt_set_device(int device) {
    int rc = check_device();
    if (rc) return rc;
    RT_CUDA(cudaSetDevice(device));
    return RT_OK;
}

static int upload(void** dst, const void* src, size_t bytes, unsigned long long& total) {
    *dst = nullptr;
    size_t alloc = bytes ? bytes : 16;
    RT_CUDA(cudaMalloc(dst, alloc));
    if (bytes) RT_CUDA(cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice));
    total += alloc;
    return RT_OK;
}

int rt_scene_create(const void* spheres, int S, const void* planes, int P, const void* materials, int M, const float* tex_rgba,
                    int tw, int th, int tex_mat, rt_scene** out) {
    return rt_scene_create_ex(spheres, S, planes, P, materials, M, tex_rgba, tw, th, tex_mat, nullptr, out);
}

int rt_scene_create_ex(const void* spheres, int S, const void* planes, int P, const void* materials, int M, const float* tex_rgba,
                       int tw, int th, int tex_mat, const rt_scene_opts* sopts, rt_scene** out) {
    if (!out) return fail(RT_ERR_INVALID, "rt_scene_create: out is NULL");
    int builder = RT_BVH_HOST_SAH;
    if (const char* e = getenv("RTB_BVH_BUILDER")) builder = (!strcmp(e, "gpu") || !strcmp(e, "1")) ? RT_BVH_GPU_LBVH : RT_BVH_HOST_SAH;
    if (sopts && sopts->struct_size >= (int)(2 * sizeof(int))) builder = sopts->bvh_builder;
    if (builder != RT_BVH_HOST_SAH && builder != RT_BVH_GPU_LBVH) return fail(RT_ERR_INVALID, "rt_scene_create_ex: unknown bvh_builder");
    *out = nullptr;
    if (S < 0 || P < 0 || M <= 0) return fail(RT_ERR_INVALID, "rt_scene_create: negative primitive count or no materials");
    if ((S && !spheres) || (P && !planes) || !materials) return fail(RT_ERR_INVALID, "rt_scene_create: NULL array");
    if ((long long)S >= (1 << 28) || (long long)P >= (1 << 28)) return fail(RT_ERR_INVALID, "rt_scene_create: too many primitives");
    const SphereIn* sp = (const SphereIn*)spheres;
    const PlaneIn* pl = (const PlaneIn*)planes;
    const MatIn* mt = (const MatIn*)materials;
    for (int i = 0; i < S; ++i)
        if (sp[i].mat < 0 || sp[i].mat >= M) return fail(RT_ERR_INVALID, "rt_scene_create: sphere material_idx out of range");
    for (int i = 0; i < P; ++i) {
        if (pl[i].mat < 0 || pl[i].mat >= M) return fail(RT_ERR_INVALID, "rt_scene_create: plane material_idx out of range");
        if (pl[i].type < 0 || pl[i].type > 2) return fail(RT_ERR_INVALID, "rt_scene_create: plane type must be QUAD/ELLIPSE/TRIANGLE");
    }
    for (int i = 0; i < M; ++i)
        if (mt[i].type < 0 || mt[i].type > 3) return fail(RT_ERR_INVALID, "rt_scene_create: unknown material type");
    if (tex_rgba && (tw <= 0 || th <= 0 || tex_mat < 0 || tex_mat >= M))
        return fail(RT_ERR_INVALID, "rt_scene_create: bad texture size or material index");
    int rc = check_device();
    if (rc) return rc;

    auto t1 = std::chrono::steady_clock::now();
    rt_scene* s = new rt_scene();
    auto bail = [&](int code) { rt_scene_destroy(s); return code; };
    cudaGetDevice(&s->device);
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, s->device) == cudaSuccess) s->num_sms = prop.multiProcessorCount;
    unsigned long long total = 0;
    struct { int nodes, depth, max_stack, leaves; float sah, build_ms; } tree{};
    float gpu_upload_ms = 0.f;
    if (builder == RT_BVH_GPU_LBVH) {
        // ---- device build (bvh_gpu.cu): the caller's arrays go up as they are; nodes + SoA streams are produced on the GPU
        GpuBuildResult g;
        cudaError_t ge = gpu_build_scene(sp, S, pl, P, nullptr, g);
        if (ge != cudaSuccess) {
            cudaGetLastError();
            char b[256];
            snprintf(b, sizeof b, "CUDA error = %u (%s) in the device BVH build (bvh_gpu.cu:%d)", (unsigned)ge, cudaGetErrorString(ge), g.fail_line);
            fail(RT_ERR_CUDA, b);
            return bail(RT_ERR_CUDA);
        }
        s->d_nodes = g.d_nodes;
        s->d_spheres = g.d_spheres;
        s->d_sphere_meta = g.d_sphere_meta;
        s->d_planes = g.d_planes;
        s->d_planes_fast = g.d_planes_fast;
        total += g.bytes_nodes + g.bytes_prims;
        tree = {g.num_nodes, g.depth, g.max_stack, g.leaves, g.sah_cost, g.build_ms};
        gpu_upload_ms = g.upload_ms;
        s->dev.num_hoisted = g.num_hoisted;
        for (int k = 0; k < g.num_hoisted; ++k) s->dev.hoisted[k] = g.hoisted[k];
        PackedScene tmp = pack_scene(nullptr, 0, nullptr, 0, mt, M);  // material records only
        if ((rc = upload(&s->d_mats, tmp.mats.data(), tmp.mats.size() * sizeof(float), total))) return bail(rc);
    } else {
        // ---- wide BVH over conservative boxes + SoA primitive buffers in leaf order (host, scene_pack.cpp)
        PackedScene pk = pack_scene(sp, S, pl, P, mt, M, -1);
        const WideBvh& bvh = pk.bvh;
        tree = {(int)bvh.nodes.size(), bvh.depth, bvh.max_stack, bvh.leaves, bvh.sah_cost, pk.build_ms};
        s->dev.num_hoisted = (int)bvh.hoisted_refs.size();
        for (int k = 0; k < s->dev.num_hoisted; ++k) s->dev.hoisted[k] = bvh.hoisted_refs[k];
        if ((rc = upload(&s->d_nodes, bvh.nodes.data(), bvh.nodes.size() * sizeof(WideNode), total))) return bail(rc);
        if ((rc = upload(&s->d_spheres, pk.spheres.data(), pk.spheres.size() * sizeof(float), total))) return bail(rc);
        if ((rc = upload(&s->d_sphere_meta, pk.sphere_meta.data(), pk.sphere_meta.size() * sizeof(int), total))) return bail(rc);
        if ((rc = upload(&s->d_planes, pk.planes.data(), pk.planes.size() * sizeof(float), total))) return bail(rc);
        if ((rc = upload(&s->d_planes_fast, pk.planes_fast.data(), pk.planes_fast.size() * sizeof(float), total))) return bail(rc);
        if ((rc = upload(&s->d_mats, pk.mats.data(), pk.mats.size() * sizeof(float), total))) return bail(rc);
        if (pk.grid.enabled) {
            const GridAccel& g = pk.grid;
            if ((rc = upload(&s->d_grid_cells, g.cells.data(), g.cells.size() * sizeof(unsigned int), total))) return bail(rc);
            if ((rc = upload(&s->d_grid_refs, g.refs.data(), g.refs.size() * sizeof(int), total))) return bail(rc);
            s->dev.grid_on = 1;
            s->dev.grid_cells = (const uint2*)s->d_grid_cells;
            s->dev.grid_refs = (const int*)s->d_grid_refs;
            for (int a = 0; a < 3; ++a) {
                s->dev.grid_lo[a] = g.lo[a]; s->dev.grid_hi[a] = g.hi[a]; s->dev.grid_cw[a] = g.cw[a];
                s->dev.grid_icw[a] = 1.0f / g.cw[a]; s->dev.grid_n[a] = g.n[a];
                s->dev.tree_lo[a] = g.tree_lo[a]; s->dev.tree_hi[a] = g.tree_hi[a];
            }
            s->stats.grid_spheres = (int)g.sphere_ids.size();
            s->stats.grid_cells = g.n[0] * g.n[1] * g.n[2];
        }
    }
    if (tree.max_stack > kSmemStack + kLocalStack) {
        fail(RT_ERR_UNSUPPORTED, "rt_scene_create: BVH needs a deeper traversal stack than the kernels provide");
        return bail(RT_ERR_UNSUPPORTED);
    }
    s->dev.tex_mat = -1;
    if (tex_rgba) {
        size_t bytes = (size_t)tw * th * 4 * sizeof(float);
        if ((rc = upload(&s->d_tex, tex_rgba, bytes, total))) return bail(rc);
        // texture object with the reference's descriptor (src/main.cu:24-46): float4 array, wrap, linear, normalised
        cudaChannelFormatDesc cd = cudaCreateChannelDesc(32, 32, 32, 32, cudaChannelFormatKindFloat);
        auto chk = [&](cudaError_t e, const char* what) {
            if (e == cudaSuccess) return 0;
            cudaGetLastError();
            return fail(RT_ERR_CUDA, std::string("CUDA error (") + cudaGetErrorString(e) + ") in " + what);
        };
        if ((rc = chk(cudaMallocArray(&s->tex_array, &cd, tw, th), "cudaMallocArray"))) return bail(rc);
        size_t pitch = (size_t)tw * 4 * sizeof(float);
        if ((rc = chk(cudaMemcpy2DToArray(s->tex_array, 0, 0, tex_rgba, pitch, pitch, th, cudaMemcpyHostToDevice), "cudaMemcpy2DToArray")))
            return bail(rc);
        cudaResourceDesc rd;
        memset(&rd, 0, sizeof rd);
        rd.resType = cudaResourceTypeArray;
        rd.res.array.array = s->tex_array;
        cudaTextureDesc td;
        memset(&td, 0, sizeof td);
        td.addressMode[0] = cudaAddressModeWrap;
        td.addressMode[1] = cudaAddressModeWrap;
        td.filterMode = cudaFilterModeLinear;
        td.readMode = cudaReadModeElementType;
        td.normalizedCoords = 1;
        if ((rc = chk(cudaCreateTextureObject(&s->tex_obj, &rd, &td, nullptr), "cudaCreateTextureObject"))) return bail(rc);
        total += bytes;
        s->dev.tex_mat = tex_mat;
        s->dev.tex_w = tw;
        s->dev.tex_h = th;
    }
    auto chk2 = [&](cudaError_t e) { if (e != cudaSuccess) { cudaGetLastError(); fail(RT_ERR_CUDA, cudaGetErrorString(e)); return 1; } return 0; };
    if (chk2(cudaMalloc((void**)&s->d_tile_counter, kMaxQueues * kQueueStride * sizeof(unsigned int))) || chk2(cudaMalloc((void**)&s->d_counters, 64)) ||
        chk2(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking)) || chk2(cudaEventCreate(&s->ev0)) ||
        chk2(cudaEventCreate(&s->ev1)) || chk2(cudaEventCreate(&s->ev2)))
        return bail(RT_ERR_CUDA);
    auto t2 = std::chrono::steady_clock::now();

    s->dev.nodes = (const float4*)s->d_nodes;
    s->dev.spheres = (const float4*)s->d_spheres;
    s->dev.sphere_meta = (const int2*)s->d_sphere_meta;
    s->dev.planes = (const float4*)s->d_planes;
    s->dev.planes_fast = (const float4*)s->d_planes_fast;
    s->dev.mats = (const float4*)s->d_mats;
    s->dev.tex = (const float4*)s->d_tex;
    s->dev.tex_obj = s->tex_obj;
    s->dev.num_nodes = tree.nodes;
    s->default_top = s->dev.num_nodes < 64 ? s->dev.num_nodes : 64;
    s->dev.top_nodes = s->default_top;

    s->stats.num_spheres = S;
    s->stats.num_planes = P;
    s->stats.num_materials = M;
    s->stats.wide_nodes = tree.nodes;
    s->stats.wide_depth = tree.depth;
    s->stats.max_stack = tree.max_stack;
    s->stats.leaves = tree.leaves;
    s->stats.sah_cost = tree.sah;
    s->stats.build_ms = tree.build_ms;
    // everything of scene creation that is not the tree build: allocations, H2D of arrays / texture, stream + events
    s->stats.upload_ms = std::chrono::duration<float, std::milli>(t2 - t1).count() - tree.build_ms;
    s->stats.bvh_builder = builder;
    s->stats.input_upload_ms = gpu_upload_ms;
    s->stats.num_hoisted = s->dev.num_hoisted;
    s->stats.device_bytes = total;
    *out = s;
    return RT_OK;
}

int rt_scene_rebuild(rt_scene* s, const void* spheres, int S, const void* planes, int P) {
    if (!s) return fail(RT_ERR_INVALID, "rt_scene_rebuild: scene is NULL");
    if (S < 0 || P < 0 || (S && !spheres) || (P && !planes)) return fail(RT_ERR_INVALID, "rt_scene_rebuild: bad array");
    if ((long long)S >= (1 << 28) || (long long)P >= (1 << 28)) return fail(RT_ERR_INVALID, "rt_scene_rebuild: too many primitives");
    const SphereIn* sp = (const SphereIn*)spheres;
    const PlaneIn* pl = (const PlaneIn*)planes;
    const int M = s->stats.num_materials;
    for (int i = 0; i < S; ++i)
        if (sp[i].mat < 0 || sp[i].mat >= M) return fail(RT_ERR_INVALID, "rt_scene_rebuild: sphere material_idx out of range");
    for (int i = 0; i < P; ++i) {
        if (pl[i].mat < 0 || pl[i].mat >= M) return fail(RT_ERR_INVALID, "rt_scene_rebuild: plane material_idx out of range");
        if (pl[i].type < 0 || pl[i].type > 2) return fail(RT_ERR_INVALID, "rt_scene_rebuild: plane type must be QUAD/ELLIPSE/TRIANGLE");
    }
    if (s->pending) return fail(RT_ERR_INVALID, "rt_scene_rebuild: a frame is in flight (call rt_render_finish first)");
    DeviceGuard guard(s->device);
    RT_CUDA(cudaStreamSynchronize(s->stream));
    GpuBuildResult g;
    cudaError_t ge = gpu_build_scene(sp, S, pl, P, nullptr, g);
    if (ge != cudaSuccess) {
        cudaGetLastError();
        char b[256];
        snprintf(b, sizeof b, "CUDA error = %u (%s) in the device BVH build (bvh_gpu.cu:%d)", (unsigned)ge, cudaGetErrorString(ge), g.fail_line);
        return fail(RT_ERR_CUDA, b);  // the scene keeps its previous geometry
    }
    if (g.max_stack > kSmemStack + kLocalStack) {
        gpu_build_free(g);
        return fail(RT_ERR_UNSUPPORTED, "rt_scene_rebuild: BVH needs a deeper traversal stack than the kernels provide");
    }
    void* old[] = {s->d_nodes, s->d_spheres, s->d_sphere_meta, s->d_planes, s->d_planes_fast, s->d_grid_cells, s->d_grid_refs};
    s->d_grid_cells = s->d_grid_refs = nullptr;  // the device builder puts everything in the tree
    s->dev.grid_on = 0;
    s->stats.grid_spheres = s->stats.grid_cells = 0;
    for (void* b : old)
        if (b) cudaFree(b);
    s->d_nodes = g.d_nodes;
    s->d_spheres = g.d_spheres;
    s->d_sphere_meta = g.d_sphere_meta;
    s->d_planes = g.d_planes;
    s->d_planes_fast = g.d_planes_fast;
    s->dev.planes_fast = (const float4*)s->d_planes_fast;
    s->dev.nodes = (const float4*)s->d_nodes;
    s->dev.spheres = (const float4*)s->d_spheres;
    s->dev.sphere_meta = (const int2*)s->d_sphere_meta;
    s->dev.planes = (const float4*)s->d_planes;
    s->dev.num_nodes = g.num_nodes;
    s->default_top = g.num_nodes < 64 ? g.num_nodes : 64;
    s->dev.top_nodes = s->default_top;
    s->dev.num_hoisted = g.num_hoisted;
    for (int k = 0; k < g.num_hoisted; ++k) s->dev.hoisted[k] = g.hoisted[k];
    s->stats.num_spheres = S;
    s->stats.num_planes = P;
    s->stats.wide_nodes = g.num_nodes;
    s->stats.wide_depth = g.depth;
    s->stats.max_stack = g.max_stack;
    s->stats.leaves = g.leaves;
    s->stats.sah_cost = g.sah_cost;
    s->stats.build_ms = g.build_ms;
    s->stats.input_upload_ms = g.upload_ms;
    s->stats.bvh_builder = RT_BVH_GPU_LBVH;
    s->stats.num_hoisted = g.num_hoisted;
    return RT_OK;
}

void rt_scene_destroy(rt_scene* s) {
    if (!s) return;
    DeviceGuard guard(s->device);
    if (s->stream) cudaStreamSynchronize(s->stream);
    if (s->tex_obj) cudaDestroyTextureObject(s->tex_obj);
    if (s->tex_array) cudaFreeArray(s->tex_array);
    void* bufs[] = {s->d_nodes, s->d_spheres, s->d_sphere_meta, s->d_planes, s->d_planes_fast, s->d_mats, s->d_tex, s->d_fb, s->d_u8, s->d_grid_cells, s->d_grid_refs,
                    s->d_tile_counter, s->d_counters, s->d_trace, s->wf.buf, s->d_acc, s->d_packed};
    for (void* b : bufs)
        if (b) cudaFree(b);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    if (s->ev2) cudaEventDestroy(s->ev2);
    if (s->stream) cudaStreamDestroy(s->stream);
    cudaGetLastError();
    delete s;
}

int rt_scene_get_stats(const rt_scene* s, rt_scene_stats* out) {
    if (!s || !out) return fail(RT_ERR_INVALID, "rt_scene_get_stats: NULL argument");
    *out = s->stats;
    return RT_OK;
}

}  // extern "C"

namespace rtb {

static void default_opts(rt_render_opts& d) {
    memset(&d, 0, sizeof d);
    d.tex_mode = RT_TEX_GPU;
    d.top_nodes_in_smem = -1;
    d.tile_nranks = 1;
    d.sample_chunk = 0;
}

// struct_size says how much of the struct the caller knows: fields beyond it (a caller compiled against an older header)
// and everything when it is <= 0 (a zero-initialised struct) take the defaults
rt_render_opts resolve_opts(const rt_render_opts* in) {
    rt_render_opts o;
    default_opts(o);
    if (in && in->struct_size > 0) {
        size_t n = (size_t)in->struct_size < sizeof o ? (size_t)in->struct_size : sizeof o;
        rt_render_opts d = o;
        memcpy(&o, in, n);
        if (n < sizeof o) memcpy((char*)&o + n, (char*)&d + n, sizeof o - n);
    }
    if (o.tile_nranks < 1) o.tile_nranks = 1;
    return o;
}

int ensure_framebuffer(rt_scene* s, size_t n) {
    if (s->fb_floats < n) {
        if (s->d_fb) cudaFree(s->d_fb);
        s->d_fb = nullptr;
        s->fb_floats = 0;
        RT_CUDA(cudaMalloc((void**)&s->d_fb, n * sizeof(float)));
        s->fb_floats = n;
    }
    return RT_OK;
}

int launch_frame(rt_scene* s, const void* cam76, const rt_render_opts* opts_in, float* d_fb, cudaStream_t stream) {
    if (!s || !cam76 || !d_fb) return fail(RT_ERR_INVALID, "rt_render: NULL argument");
    rt_render_opts o = resolve_opts(opts_in);
    FrameParams fp;
    memset(&fp, 0, sizeof fp);
    memcpy(&fp.cam, cam76, sizeof(CameraPod));
    if (fp.cam.W <= 0 || fp.cam.H <= 0 || fp.cam.spp < 0 || fp.cam.depth < 0) return fail(RT_ERR_INVALID, "rt_render: bad CameraData");
    if (o.tile_rank < 0 || o.tile_rank >= o.tile_nranks) return fail(RT_ERR_INVALID, "rt_render: tile_rank out of range");
    if (o.kernel != RT_KERNEL_MEGA && o.kernel != RT_KERNEL_WAVEFRONT && o.kernel != RT_KERNEL_POOL) return fail(RT_ERR_INVALID, "rt_render: unknown kernel");
    fp.key = make_uint2((unsigned)(o.seed & 0xffffffffu), (unsigned)(o.seed >> 32));
    fp.tex_mode = o.tex_mode;
    fp.tile_rank = o.tile_rank;
    fp.tile_nranks = o.tile_nranks;
    fp.tiles_x = (fp.cam.W + kTile - 1) / kTile;
    fp.tiles_y = (fp.cam.H + kTile - 1) / kTile;
    int tiles = fp.tiles_x * fp.tiles_y;
    fp.my_tiles = (tiles + o.tile_nranks - 1) / o.tile_nranks;  // groups of the interleave; a rank's slot in the last one may be empty
    fp.fb = d_fb;
    fp.tile_counter = s->d_tile_counter;
    fp.nqueues = s->num_sms < kMaxQueues ? s->num_sms : kMaxQueues;  // one queue per SM (render_mega)
    if (const char* e = getenv("RTB_QUEUES")) fp.nqueues = atoi(e);
    fp.queue_per_cta = 0;
    if (const char* e = getenv("RTB_QUEUE_PER_CTA")) fp.queue_per_cta = atoi(e) ? 1 : 0;
    if (fp.nqueues < 1) fp.nqueues = 1;
    if (fp.nqueues > kMaxQueues) fp.nqueues = kMaxQueues;
    fp.counters = s->d_counters;
    // sample chunks: work items of a few samples (the tail is what limits strong scaling over GPUs, and short items keep a
    // warp's lanes on the same few pixels and in step: measured, items of 171 samples render the headline frame 6 % slower
    // than items of 2).  Only the production stream is chunked: parity mode keeps the reference's per-pixel summation order.
    fp.acc = nullptr;
    fp.chunk_spp = fp.cam.spp > 0 ? fp.cam.spp : 1;
    fp.nchunks = 1;
    fp.chunk_major = 0;
    if (const char* e = getenv("RTB_CHUNK_MAJOR")) fp.chunk_major = atoi(e) ? 1 : 0;
    int want_chunk = o.sample_chunk;
    if (o.rng != RT_RNG_PHILOX) want_chunk = -1;
    if (want_chunk == 0) {
        want_chunk = (o.rng == RT_RNG_PHILOX && o.kernel == RT_KERNEL_MEGA && fp.cam.spp >= 32) ? 2 : -1;
        if (const char* e = getenv("RTB_SAMPLE_CHUNK")) want_chunk = atoi(e) > 0 ? atoi(e) : -1;
    }
    if (want_chunk > 0 && o.kernel == RT_KERNEL_MEGA && fp.cam.spp > want_chunk && fp.cam.depth > 0) {
        // the chunks of a pixel meet in a 64-bit fixed-point accumulator plane (2 x the framebuffer's bytes, L2 resident at
        // 1080p); if it cannot be allocated the frame renders unsplit -- a longer tail, the same samples
        const size_t need = (size_t)fp.cam.W * fp.cam.H * 3;
        if (s->acc_values < need) {
            if (s->d_acc) cudaFree(s->d_acc);
            s->d_acc = nullptr;
            s->acc_values = 0;
            if (cudaMalloc((void**)&s->d_acc, need * sizeof(long long)) == cudaSuccess) s->acc_values = need;
            else cudaGetLastError();
        }
        if (s->d_acc) {
            // at most kMaxChunksPerPixel items per pixel: beyond that the accumulator's atomics cost more than the shorter items
            // gain (measured on the shipped scene at 2500 spp: 1250 chunks of 2 render 7 % slower than 125 of 20)
            const int kMaxChunksPerPixel = 128;
            if (o.sample_chunk <= 0 && (fp.cam.spp + want_chunk - 1) / want_chunk > kMaxChunksPerPixel)
                want_chunk = (fp.cam.spp + kMaxChunksPerPixel - 1) / kMaxChunksPerPixel;
            fp.chunk_spp = want_chunk;
            fp.nchunks = (fp.cam.spp + want_chunk - 1) / want_chunk;
            fp.acc = s->d_acc;
        }
    }
    DevScene dev = s->dev;
    // megakernel default: no staging (measured no gain); the other kernels keep the 64-node default
    int top = o.top_nodes_in_smem < 0 ? (o.kernel == RT_KERNEL_MEGA ? 0 : s->default_top) : o.top_nodes_in_smem;
    if (top > dev.num_nodes) top = dev.num_nodes;
    if (top > 1024) top = 1024;
    dev.top_nodes = top;
    LaunchCfg cfg{s->num_sms, kMegaBlocksPerSM, kMegaThreads};
    if (o.kernel != RT_KERNEL_MEGA) cfg = LaunchCfg{s->num_sms, kPoolBlocksPerSM, kPoolThreads};
    // scheduling thresholds of the megakernels; the RTB_* variables are tuning knobs for profiling runs
    fp.leaf_min = kLeafMinDefault;
    if (const char* e = getenv("RTB_LEAF_MIN")) fp.leaf_min = atoi(e);
    fp.refill_min = 32;
    fp.park_min = 32;
    if (const char* e = getenv("RTB_REFILL_MIN")) fp.refill_min = atoi(e);
    if (const char* e = getenv("RTB_PARK_MIN")) fp.park_min = atoi(e);
    // small trees (the reference's shipped scene: 86 nodes, 3.6 visits per ray) finish a traversal in a few steps, so a TRACE
    // phase is worth ending earlier: measured 255.9 ms at 10 vs 260.7 at 6 on the shipped scene, 179.0 vs 174.9 on the headline
    fp.trace_min = s->dev.num_nodes < 1024 ? kTraceMinSmallTree : kTraceMinDefault;
    if (const char* e = getenv("RTB_TRACE_MIN")) fp.trace_min = atoi(e);
    fp.regroup = 0;
    if (const char* e = getenv("RTB_POOL_REGROUP")) fp.regroup = atoi(e) ? 1 : 0;
    if (fp.refill_min < 1) fp.refill_min = 1;
    if (fp.park_min < 1) fp.park_min = 1;
    RT_CUDA(cudaMemsetAsync(s->d_tile_counter, 0, kMaxQueues * kQueueStride * sizeof(unsigned int), stream));
    RT_CUDA(cudaMemsetAsync(s->d_counters, 0, 64, stream));
    RT_CUDA(cudaEventRecord(s->ev0, stream));
    if (fp.acc) RT_CUDA(cudaMemsetAsync(fp.acc, 0, (size_t)fp.cam.W * fp.cam.H * 3 * sizeof(long long), stream));  // inside the timed region
    if (o.kernel == RT_KERNEL_MEGA) RT_CUDA(launch_render_mega(dev, fp, o.rng, o.exact_math, o.count_work, cfg, stream));
    else if (o.kernel == RT_KERNEL_POOL) RT_CUDA(launch_render_pool(dev, fp, o.rng, o.exact_math, o.count_work, cfg, stream));
    else RT_CUDA(launch_render_wavefront(dev, fp, o.rng, o.exact_math, o.count_work, cfg, s->wf, stream));
    if (fp.nchunks > 1) RT_CUDA(launch_finalize_chunks(fp, s->num_sms, stream));
    RT_CUDA(cudaEventRecord(s->ev1, stream));
    s->pending = true;
    s->pending_stream = stream;
    return RT_OK;
}

}  // namespace rtb

extern "C" {

int rt_render_finish(rt_scene* s, unsigned long long* segs, float* kernel_ms, rt_work_counters* counters) {
    if (!s) return fail(RT_ERR_INVALID, "rt_render_finish: NULL scene");
    if (!s->pending) return fail(RT_ERR_INVALID, "rt_render_finish: no frame in flight");
    DeviceGuard guard(s->device);
    RT_CUDA(cudaEventSynchronize(s->ev1));
    s->pending = false;
    unsigned long long c[8];
    RT_CUDA(cudaMemcpyAsync(c, s->d_counters, 48, cudaMemcpyDeviceToHost, s->pending_stream));
    RT_CUDA(cudaStreamSynchronize(s->pending_stream));
    if (segs) *segs = c[0];
    if (kernel_ms) RT_CUDA(cudaEventElapsedTime(kernel_ms, s->ev0, s->ev1));
    if (counters) {
        counters->segments = c[0]; counters->node_visits = c[1]; counters->sphere_tests = c[2];
        counters->plane_tests = c[3]; counters->shaded_hits = c[4]; counters->samples = c[5];
    }
    return RT_OK;
}

int rt_render_device(rt_scene* s, const void* cam76, const rt_render_opts* opts, void* d_fb, void* cuda_stream) {
    if (!s) return fail(RT_ERR_INVALID, "rt_render_device: NULL scene");
    DeviceGuard guard(s->device);
    return launch_frame(s, cam76, opts, (float*)d_fb, (cudaStream_t)cuda_stream);
}

int rt_render(rt_scene* s, const void* cam76, const rt_render_opts* opts, float* fb, unsigned long long* segs, float* kernel_ms,
              rt_work_counters* counters) {
    if (!s || !cam76 || !fb) return fail(RT_ERR_INVALID, "rt_render: NULL argument");
    CameraPod cam;
    memcpy(&cam, cam76, sizeof cam);
    if (cam.W <= 0 || cam.H <= 0) return fail(RT_ERR_INVALID, "rt_render: bad CameraData");
    DeviceGuard guard(s->device);
    size_t n = (size_t)cam.W * cam.H * 3;
    int rc = ensure_framebuffer(s, n);
    if (rc) return rc;
    rt_render_opts o = resolve_opts(opts);
    if (o.tile_nranks > 1) RT_CUDA(cudaMemsetAsync(s->d_fb, 0, n * sizeof(float), s->stream));
    rc = launch_frame(s, cam76, opts, s->d_fb, s->stream);
    if (rc) return rc;
    RT_CUDA(cudaMemcpyAsync(fb, s->d_fb, n * sizeof(float), cudaMemcpyDeviceToHost, s->stream));
    rc = rt_render_finish(s, segs, kernel_ms, counters);
    if (rc) return rc;
    RT_CUDA(cudaStreamSynchronize(s->stream));
    return RT_OK;
}

// render + device-side output stage: the frame never leaves the GPU as floats (3 B/pixel D2H instead of 12)
int rt_render_u8(rt_scene* s, const void* cam76, const rt_render_opts* opts, int divisor, unsigned char* rgb,
                 unsigned long long* segs, float* kernel_ms) {
    if (!s || !cam76 || !rgb || divisor <= 0) return fail(RT_ERR_INVALID, "rt_render_u8: bad argument");
    CameraPod cam;
    memcpy(&cam, cam76, sizeof cam);
    if (cam.W <= 0 || cam.H <= 0) return fail(RT_ERR_INVALID, "rt_render_u8: bad CameraData");
    DeviceGuard guard(s->device);
    size_t n = (size_t)cam.W * cam.H * 3;
    int rc = ensure_framebuffer(s, n);
    if (rc) return rc;
    rt_render_opts o = resolve_opts(opts);
    if (o.tile_nranks > 1) RT_CUDA(cudaMemsetAsync(s->d_fb, 0, n * sizeof(float), s->stream));
    rc = launch_frame(s, cam76, opts, s->d_fb, s->stream);
    if (rc) return rc;
    rc = rt_render_finish(s, segs, kernel_ms, nullptr);
    if (rc) return rc;
    return rt_tonemap_u8(s, nullptr, cam.W, cam.H, divisor, rgb);
}

int rt_trace_primary_ex(rt_scene* s, const float* rays6, int n, float tmin, float tmax, const rt_trace_opts* topts, int* type,
                        int* index, float* t) {
    if (!s || (n > 0 && (!rays6 || !type || !index || !t)) || n < 0) return fail(RT_ERR_INVALID, "rt_trace_primary: bad argument");
    int exact = 1;
    if (topts && topts->struct_size >= (int)(2 * sizeof(int))) exact = topts->exact_math ? 1 : 0;
    if (n == 0) return RT_OK;
    DeviceGuard guard(s->device);
    size_t need = (size_t)n * (6 * 4 + 4 + 4 + 4);
    if (s->trace_bytes < need) {
        if (s->d_trace) cudaFree(s->d_trace);
        s->d_trace = nullptr;
        s->trace_bytes = 0;
        RT_CUDA(cudaMalloc(&s->d_trace, need));
        s->trace_bytes = need;
    }
    float* d_rays = (float*)s->d_trace;
    int* d_type = (int*)(d_rays + 6 * (size_t)n);
    int* d_index = d_type + n;
    float* d_t = (float*)(d_index + n);
    RT_CUDA(cudaMemcpyAsync(d_rays, rays6, (size_t)n * 24, cudaMemcpyHostToDevice, s->stream));
    DevScene dev = s->dev;
    if (!exact) dev.top_nodes = 0;  // as the production megakernel launches: every node through the L1 (__ldg), nothing staged
    RT_CUDA(launch_trace_primary(dev, d_rays, n, tmin, tmax, exact, d_type, d_index, d_t, s->num_sms, s->stream));
    RT_CUDA(cudaMemcpyAsync(type, d_type, (size_t)n * 4, cudaMemcpyDeviceToHost, s->stream));
    RT_CUDA(cudaMemcpyAsync(index, d_index, (size_t)n * 4, cudaMemcpyDeviceToHost, s->stream));
    RT_CUDA(cudaMemcpyAsync(t, d_t, (size_t)n * 4, cudaMemcpyDeviceToHost, s->stream));
    RT_CUDA(cudaStreamSynchronize(s->stream));
    return RT_OK;
}

int rt_trace_primary(rt_scene* s, const float* rays6, int n, float tmin, float tmax, int* type, int* index, float* t) {
    return rt_trace_primary_ex(s, rays6, n, tmin, tmax, nullptr, type, index, t);
}

int rt_tonemap_u8(rt_scene* s, const void* d_fb, int W, int H, int divisor, unsigned char* rgb) {
    if (!s || !rgb || W <= 0 || H <= 0 || divisor <= 0) return fail(RT_ERR_INVALID, "rt_tonemap_u8: bad argument");
    const float* src = d_fb ? (const float*)d_fb : s->d_fb;
    if (!src) return fail(RT_ERR_INVALID, "rt_tonemap_u8: no framebuffer");
    DeviceGuard guard(s->device);
    size_t n = (size_t)W * H * 3;
    if (s->u8_bytes < n) {
        if (s->d_u8) cudaFree(s->d_u8);
        s->d_u8 = nullptr;
        s->u8_bytes = 0;
        RT_CUDA(cudaMalloc((void**)&s->d_u8, n));
        s->u8_bytes = n;
    }
    RT_CUDA(launch_tonemap(src, n, divisor, s->d_u8, s->num_sms, s->stream));
    RT_CUDA(cudaMemcpyAsync(rgb, s->d_u8, n, cudaMemcpyDeviceToHost, s->stream));
    RT_CUDA(cudaStreamSynchronize(s->stream));
    return RT_OK;
}

unsigned long long rt_tiles_packed_floats(int W, int H, int nranks) { return tiles_packed_floats(W, H, nranks < 1 ? 1 : nranks); }
int rt_tiles_pack(rt_scene* s, const void* d_fb, int W, int H, int rank, int nranks, void* d_packed, void* stream) {
    if (!s || !d_fb || !d_packed || nranks < 1 || rank < 0 || rank >= nranks) return fail(RT_ERR_INVALID, "rt_tiles_pack: bad argument");
    DeviceGuard guard(s->device);
    RT_CUDA(launch_tiles_pack((const float*)d_fb, W, H, rank, nranks, (float*)d_packed, s->num_sms, (cudaStream_t)stream));
    return RT_OK;
}
int rt_tiles_unpack(rt_scene* s, const void* d_all, int W, int H, int nranks, void* d_fb, void* stream) {
    if (!s || !d_fb || !d_all || nranks < 1) return fail(RT_ERR_INVALID, "rt_tiles_unpack: bad argument");
    DeviceGuard guard(s->device);
    RT_CUDA(launch_tiles_unpack((const float*)d_all, W, H, nranks, -1, 0, (float*)d_fb, s->num_sms, (cudaStream_t)stream));
    return RT_OK;
}
void rt_tile_owner(int tile, int nranks, int* rank, int* k) {
    int r = 0, kk = tile;
    if (nranks > 1) tile_owner(tile, nranks, r, kk);
    if (rank) *rank = r;
    if (k) *k = kk;
}

}  // extern "C"
