/* rt_b200.h -- C-ABI of the B200-native per-pixel render path (librt_b200.so).
 *
 * Drop-in boundary for zloyaloha/ray-tracing-practice.  The reference has no FFI; its seam is the C++ call
 *     gpu_render(SceneParams&, SceneData&)            include/camera.cuh:114, src/camera.cu:290
 *       -> Camera::render(color* d_fb)                src/camera.cu:198
 *       -> render_kernel<<<grid,16x16>>>(color* fb)   src/camera.cu:17
 * with its inputs in the host arrays create_scene fills (src/main.cu:346-497).  Every entry point below takes
 * exactly those arrays, in the reference's own struct layouts, as plain pointers + sizes:
 *     SphereData   32 B  {float center[3] @0; float radius @12; int material_idx @16}          include/sphere.h:8
 *     PlaneData    80 B  {int type @0; float D @4; int material_idx @8; float w[3] @12, u[3] @24, v[3] @36,
 *                         base[3] @48, normal[3] @60}                                          include/plane.h:9
 *     MaterialData 64 B  {int type @0; float fuzz @4, ir @8; float absorption[3] @12, albedo[3] @24,
 *                         emit[3] @36; u64 tex_obj @48; u64 cpu_tex @56 (both ignored here)}   include/materials.h:53
 *     CameraData   76 B  {float origin[3], pixel00_loc[3], pixel_delta_u[3], pixel_delta_v[3], background[3];
 *                         int image_width, image_height, samples_per_pixel, max_depth}         include/camera.cuh:86
 * The library owns all device memory behind an opaque handle; the caller owns every host buffer.
 * All functions return RT_OK (0) or an RT_ERR_* code; rt_last_error() gives the message (thread-local).
 * There is NO CPU fallback: without a CUDA device every compute entry point fails with RT_ERR_NO_DEVICE.
 */
#ifndef RT_B200_H
#define RT_B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rt_scene rt_scene;

enum {
    RT_OK = 0,
    RT_ERR_INVALID = 1,     /* bad argument (null pointer, negative count, material index out of range ...) */
    RT_ERR_CUDA = 2,        /* a CUDA runtime call failed; message carries file:line and the CUDA error string */
    RT_ERR_NO_DEVICE = 3,   /* no usable CUDA device (the reference would exit(99), include/camera.cuh:20-29) */
    RT_ERR_UNSUPPORTED = 4
};

enum { RT_KERNEL_MEGA = 0, RT_KERNEL_WAVEFRONT = 1, RT_KERNEL_POOL = 2 };
enum { RT_RNG_PHILOX = 0,      /* counter-based Philox4x32-10 keyed on (seed; pixel, sample, bounce) -- production */
       RT_RNG_REFERENCE = 1 }; /* the reference's stateful wang_hash stream (include/random_utils.h) incl. its
                                  rejection loops -- parity mode: reproduces the reference host render per sample */
enum { RT_TEX_CPU = 0,         /* tex2D_cpu semantics: V flipped, no half texel (include/materials.h:20-51) */
       RT_TEX_GPU = 1 };       /* CUDA texture unit as the reference sets it up (src/main.cu:37-43) -- default */

typedef struct rt_render_opts {
    int struct_size;            /* = sizeof(rt_render_opts) as the caller compiled it: fields beyond it take their defaults
                                   (<= 0: every field does) */
    int kernel;                 /* RT_KERNEL_* */
    int rng;                    /* RT_RNG_* */
    int exact_math;             /* 1: leaf tests + shading in the reference host build's exact operation order
                                   (no FMA contraction, double where the reference widens); 0: fast fp32 */
    int tex_mode;               /* RT_TEX_* */
    unsigned long long seed;    /* Philox key (ignored by RT_RNG_REFERENCE, which is keyed by pixel like the reference) */
    int tile_rank, tile_nranks; /* interleaved-tile partition: this call renders tiles t with t % nranks == rank.
                                   nranks <= 1 renders everything.  Pixels of other ranks are left untouched. */
    int count_work;             /* 1: also count node visits / leaf tests / shaded hits (slower; for the roofline) */
    int top_nodes_in_smem;      /* -1 default (megakernel: 0, others: 64); 0 off; N>0: stage the first N wide nodes in shared memory */
    int sample_chunk;           /* megakernel work-item length in samples.  A pixel's chunks are summed in a 64-bit fixed-point
                                   accumulator (2^-32 units, 24 bytes per pixel of scratch): integer adds commute, so the frame is
                                   deterministic whatever order the chunks finish in, and equal for every tile partition.
                                   0 auto (2 for the Philox stream when spp >= 32); -1 never split a pixel (always the case for
                                   RT_RNG_REFERENCE: the reference's summation order) */
} rt_render_opts;

typedef struct rt_work_counters {
    unsigned long long segments;      /* path segments = hit_scene calls (the "rays" of the Mrays/s metric) */
    unsigned long long node_visits;   /* wide nodes fetched */
    unsigned long long sphere_tests;
    unsigned long long plane_tests;
    unsigned long long shaded_hits;
    unsigned long long samples;
} rt_work_counters;

typedef struct rt_scene_stats {
    int num_spheres, num_planes, num_materials;
    int wide_nodes, wide_depth, max_stack, leaves;
    float sah_cost, build_ms, upload_ms;
    unsigned long long device_bytes;
    int bvh_builder;            /* RT_BVH_* the tree was built with */
    float input_upload_ms;      /* RT_BVH_GPU_LBVH: allocation + H2D of the caller's arrays (part of upload_ms) */
    int num_hoisted;            /* primitives kept out of the tree because they span most of the scene (the floor quad):
                                   tested once per ray before the traversal; at most 4 */
    int grid_spheres;           /* spheres held by the uniform grid instead of the tree (0: no grid, the default).  With
                                   RTB_GRID=1 in the environment the host builder puts a dense field of many similar small
                                   spheres (>= 4096 and at least half of the scene's) into a grid; measured slower than the
                                   tree on B200, kept as the measured alternative */
    int grid_cells;             /* cells of that grid */
} rt_scene_stats;

/* Which builder replaces build_bvh (include/bvh_builder.h:99-120). */
enum { RT_BVH_HOST_SAH = 0,    /* host, binned SAH, multi-threaded: best trees (default) */
       RT_BVH_GPU_LBVH = 1 };  /* device: Morton sort + binary radix tree + bottom-up fit + wide collapse, all on the GPU
                                  (SURVEY 8 f1): ~100x faster to build, ~1.3x more node visits per ray */
typedef struct rt_scene_opts {
    int struct_size;            /* = sizeof(rt_scene_opts) */
    int bvh_builder;            /* RT_BVH_* */
} rt_scene_opts;

/* Replaces create_scene's upload + build_bvh (src/main.cu:428-478, include/bvh_builder.h:99): copies the three
 * arrays, builds the wide BVH and SoA primitive buffers, uploads.  tex_rgba: tw*th RGBA float texels exactly as
 * stbi_loadf(...,4) returns them (src/main.cu:18), applied to materials[tex_material_idx]; NULL/-1 for none. */
int rt_scene_create(const void* spheres, int num_spheres, const void* planes, int num_planes, const void* materials,
                    int num_materials, const float* tex_rgba, int tex_w, int tex_h, int tex_material_idx, rt_scene** out);
/* Same with options (opts may be NULL = rt_scene_create).  With RT_BVH_GPU_LBVH the caller's arrays are uploaded as they
 * are and the tree + SoA buffers are produced by kernels (csrc/bvh_gpu.cu); results are identical in exact mode (the closest
 * hit does not depend on the tree), the environment variable RTB_BVH_BUILDER=gpu selects it where opts is NULL. */
int rt_scene_create_ex(const void* spheres, int num_spheres, const void* planes, int num_planes, const void* materials,
                       int num_materials, const float* tex_rgba, int tex_w, int tex_h, int tex_material_idx,
                       const rt_scene_opts* opts, rt_scene** out);
/* New geometry for an existing scene (materials and texture stay): rebuilds the tree and the SoA buffers with the DEVICE
 * builder and swaps them in -- what an animated scene calls once per frame (the reference's scene is static; its build_bvh,
 * include/bvh_builder.h:99, runs once).  Primitive counts may change; material indices must fit the scene's materials.
 * On failure the scene keeps its previous geometry.  Not to be called while a rt_render_device frame is in flight. */
int rt_scene_rebuild(rt_scene* scene, const void* spheres, int num_spheres, const void* planes, int num_planes);
/* Replaces destroy_scene_arrays + destroy_all_materials (src/main.cu:310-344). */
void rt_scene_destroy(rt_scene* scene);
int rt_scene_get_stats(const rt_scene* scene, rt_scene_stats* out);

/* Replaces Camera::render up to and including its D2H copy (src/camera.cu:198-209): renders one frame and
 * writes fb_sum_rgb[(j*W+i)*3+c] = SUM over samples of radiance (the reference's framebuffer contract,
 * src/camera.cu:33) into HOST memory.  opts may be NULL (defaults: megakernel, Philox seed 0, fast math, GPU
 * texture semantics).  ray_segments / kernel_ms / counters may be NULL. */
int rt_render(rt_scene* scene, const void* camera_data76, const rt_render_opts* opts, float* fb_sum_rgb,
              unsigned long long* ray_segments, float* kernel_ms, rt_work_counters* counters);
/* Same, framebuffer stays in DEVICE memory (d_fb_sum_rgb: W*H*3 floats) and the launch is asynchronous on
 * `cuda_stream` (a cudaStream_t, may be NULL).  Counters are fetched by rt_render_finish. */
int rt_render_device(rt_scene* scene, const void* camera_data76, const rt_render_opts* opts, void* d_fb_sum_rgb,
                     void* cuda_stream);
int rt_render_finish(rt_scene* scene, unsigned long long* ray_segments, float* kernel_ms, rt_work_counters* counters);

/* Closest hit with primitive ids for caller-supplied rays (host arrays; rays6 = N x {ox,oy,oz,dx,dy,dz}).
 * prim_type: 0 sphere / 1 plane / -1 miss (BVHNode::type convention, include/bvh.h:11); prim_index: index into
 * the caller's original array; t: hit parameter in units of the UNNORMALISED direction, like HitRecord::t.
 * Replaces hit_scene (include/scene.h:23-35) for rays the caller supplies.  rt_trace_primary runs the exact-math leaf
 * tests (parity target: the reference's leaf tests over all primitives, bit for bit). */
int rt_trace_primary(rt_scene* scene, const float* rays6, int num_rays, float tmin, float tmax, int* prim_type,
                     int* prim_index, float* t);
typedef struct rt_trace_opts {
    int struct_size;            /* = sizeof(rt_trace_opts) */
    int exact_math;             /* 1: as rt_trace_primary.  0: the PRODUCTION arithmetic -- the traversal and leaf code the default
                                   render (rt_render_opts.exact_math = 0) is compiled from: fast fp32 with FMA contraction, the
                                   two-dot-product plane record, the un-widened slab test, SFU reciprocal / sqrt -- so ids and t
                                   are the ones the benchmarked kernel shades */
} rt_trace_opts;
/* opts may be NULL (= rt_trace_primary). */
int rt_trace_primary_ex(rt_scene* scene, const float* rays6, int num_rays, float tmin, float tmax, const rt_trace_opts* opts,
                        int* prim_type, int* prim_index, float* t);

/* ---- allocation-free variant.  NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernel, same
 * arguments and same results as rt_trace_primary_ex, but the device staging area for the rays and the results is a
 * file-scope `static __device__` array (csrc/rt_noalloc.cu) instead of a cudaMalloc'ed buffer: the call performs no
 * cudaMalloc / cudaFree and calls no allocating library routine.  Limits: num_rays <= RT_NOALLOC_MAX_RAYS (more is rejected
 * with RT_ERR_INVALID and a message naming the limit); the array is RT_NOALLOC_MAX_RAYS x 36 bytes = 9 MiB per GPU that
 * loads the library, held until the process exits; every call on a GPU shares it, so calls that overlap on one GPU race and
 * must be serialised by the caller.  The array keeps the last call's contents; the call overwrites every slot it reads, so
 * nothing has to be cleared. */
#define RT_NOALLOC_MAX_RAYS 262144
int rt_trace_primary_noalloc(rt_scene* scene, const float* rays6, int num_rays, float tmin, float tmax, const rt_trace_opts* opts,
                             int* prim_type, int* prim_index, float* t);
int rt_noalloc_max_rays(void); /* = RT_NOALLOC_MAX_RAYS of the library as built */

/* Device-side output stage replacing ISaver::writeColor's arithmetic (src/camera.cu:136-148): byte =
 * (unsigned char)(256*clamp(sqrt(sum/divisor),0,0.999)).  d_fb_sum_rgb is a device pointer, rgb_u8 a host one. */
int rt_tonemap_u8(rt_scene* scene, const void* d_fb_sum_rgb, int width, int height, int divisor, unsigned char* rgb_u8);

/* rt_render followed by the output stage without the float download: replaces Camera::render + the per-pixel
 * saver->writeColor loop (src/camera.cu:198-215).  rgb_u8 is a HOST buffer of W*H*3 bytes, row-major. */
int rt_render_u8(rt_scene* scene, const void* camera_data76, const rt_render_opts* opts, int divisor, unsigned char* rgb_u8,
                 unsigned long long* ray_segments, float* kernel_ms);

/* Tile partition of the multi-GPU frame.  Tiles are RT_TILE x RT_TILE pixels with row-major ids; they are dealt to the
 * ranks in groups of nranks consecutive ids, one tile of each group per rank, the position inside group g rotated by a
 * hash of g (so that a rank does not own the same tile columns in every row).  rt_tile_owner gives the owning rank of a
 * tile and k = its position in that rank's packed buffer (16x16x3 floats per tile, increasing k). */
#define RT_TILE 16
void rt_tile_owner(int tile, int nranks, int* rank, int* k);
int rt_tiles_pack(rt_scene* scene, const void* d_fb_sum_rgb, int width, int height, int rank, int nranks, void* d_packed,
                  void* cuda_stream);
int rt_tiles_unpack(rt_scene* scene, const void* d_packed_all, int width, int height, int nranks, void* d_fb_sum_rgb,
                    void* cuda_stream);
unsigned long long rt_tiles_packed_floats(int width, int height, int nranks); /* per-rank packed buffer length */

/* ---- one frame on N GPUs (the reference is single-GPU: gpu_render -> Camera::render, src/camera.cu:290-349, :198-216).
 * One rank per GPU -- a process (torchrun, MPI) or a thread of one process.  Every rank holds the whole scene (its own
 * rt_scene on its own device) and calls rt_render_tiles with the same camera and options: each renders its interleaved
 * tiles, the packed tiles go to rank 0 with grouped ncclSend / ncclRecv over NVLink, rank 0 assembles the frame on its
 * device and downloads it once.  NCCL is bound at run time: the libnccl.so.2 the process already holds (e.g. the one torch
 * loaded), else $RTB_NCCL_LIB, else the system library; without one nranks > 1 fails with RT_ERR_UNSUPPORTED.  A process that
 * also uses torch should import it BEFORE the first rt_comm_* call, so that both share one NCCL. */
typedef struct rt_comm rt_comm;
#define RT_COMM_ID_BYTES 128
/* rank 0: an ncclUniqueId (RT_COMM_ID_BYTES bytes) to hand to every rank by the launcher's own means (MPI_Bcast, a
 * torch.distributed broadcast, a file) */
int rt_comm_unique_id(void* id_out);
/* every rank, on its device (the calling thread's current device; see rt_set_device): ncclCommInitRank.  nranks == 1 needs no id. */
int rt_comm_init_rank(const void* id, int rank, int nranks, rt_comm** out);
/* one process driving ndev devices (devices NULL: 0..ndev-1): ncclCommInitAll; out receives ndev communicators, rank i on
 * devices[i].  rt_render_tiles must then be called from ndev threads, one per rank. */
int rt_comm_init_all(int ndev, const int* devices, rt_comm** out);
void rt_comm_destroy(rt_comm* comm);
int rt_comm_rank(const rt_comm* comm);
int rt_comm_size(const rt_comm* comm);
typedef struct rt_tiles_info {
    int struct_size;                     /* = sizeof(rt_tiles_info), set by the caller */
    int rank, nranks;
    unsigned long long segments_rank;    /* path segments this rank traced */
    unsigned long long segments_total;   /* rank 0: over all ranks (the counts travel with the tiles); 0 elsewhere */
    float kernel_ms;                     /* this rank's render kernel (+ chunk sum) */
    float device_ms;                     /* this rank's stream from the start of the render to the end of its send (rank 0:
                                            of the gather and the unpack) -- CUDA events */
    unsigned long long gathered_bytes;   /* rank 0: bytes received over NCCL */
    unsigned long long d2h_bytes;        /* rank 0: bytes downloaded into fb_sum_rgb */
} rt_tiles_info;
/* comm NULL or of size 1: the whole frame on this device.  fb_sum_rgb: HOST buffer of W*H*3 floats on rank 0 (the
 * reference's framebuffer contract, as rt_render); NULL there leaves the assembled frame in the scene's device framebuffer
 * (rt_tonemap_u8(scene, NULL, ...) reads it); ignored on the other ranks.  opts->tile_rank / tile_nranks are taken from comm. */
int rt_render_tiles(rt_scene* scene, rt_comm* comm, const void* camera_data76, const rt_render_opts* opts, float* fb_sum_rgb,
                    rt_tiles_info* info);

const char* rt_last_error(void);
int rt_device_count(void);
/* Selects the CUDA device of the CALLING THREAD for subsequent rt_scene_create / rt_comm_init_rank calls.  A scene lives
 * on the device it was created on, and every entry point that takes a scene switches to that device for the duration of
 * the call, so a handle may be used from any thread.  The reference is single-GPU (it never calls cudaSetDevice). */
int rt_set_device(int device);
const char* rt_version(void);

#ifdef __cplusplus
}
#endif
#endif
