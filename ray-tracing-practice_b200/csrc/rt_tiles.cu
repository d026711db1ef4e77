// rt_tiles.cu -- ONE frame on N GPUs behind the C-ABI: rt_comm_* and rt_render_tiles (include/rt_b200.h).
//
// The reference renders a frame on one device (gpu_render -> Camera::render, src/camera.cu:290-349, :198-216).  Here every
// rank (one per GPU: a process under torchrun / MPI, or a thread of one process) holds the whole scene, renders the
// interleaved 16x16 tiles that are its share (kernels.h: tile_of_rank), packs them, and the packed tiles travel to rank 0
// with grouped ncclSend / ncclRecv over NVLink; rank 0 scatters them into its framebuffer (its own tiles are already there)
// and downloads the assembled frame ONCE.  Pixels are independent and the scene is read-only, so this gather is the only
// exchange of the path -- there is no compute step to fuse it with (W*H*12/N bytes per rank against a whole frame of path
// tracing), it is one collective per frame on the scene's stream, behind the render kernel.
//
// NCCL is bound at run time (dlopen of libnccl.so.2: inside a torchrun rank that is the copy torch already loaded, in
// rt_main the system library), so librt_b200.so has no link-time dependency on it and single-GPU callers never touch it.
#include <dlfcn.h>

#include <cstring>
#include <mutex>
#include <vector>

#include "rt_scene_impl.h"

using namespace rtb;

namespace {

// the slice of nccl.h this file uses (ABI-stable since NCCL 2.7: ncclSend / ncclRecv)
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;     // ncclSuccess = 0
constexpr int kNcclFloat = 7; // ncclFloat32
static_assert(sizeof(ncclUniqueId) == RT_COMM_ID_BYTES, "RT_COMM_ID_BYTES must be NCCL_UNIQUE_ID_BYTES");

struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string why;
};

NcclApi& nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        // a copy the process already holds (torch's bundled one, when torch was imported first) is the one to share; else
        // RTB_NCCL_LIB, else the system library.  Loaded RTLD_LOCAL: nothing else should resolve against our choice.
        const char* names[] = {getenv("RTB_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
        api.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
        for (const char* n : names) {
            if (api.lib) break;
            if (!n || !*n) continue;
            api.lib = dlopen(n, RTLD_NOW | RTLD_LOCAL);
            if (!api.lib) api.why = dlerror();
        }
        if (!api.lib) return;
        auto sym = [&](const char* n) {
            void* p = dlsym(api.lib, n);
            if (!p) { api.why = std::string("missing symbol ") + n; }
            return p;
        };
        api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
        api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
        api.CommInitAll = (decltype(api.CommInitAll))sym("ncclCommInitAll");
        api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
        api.Send = (decltype(api.Send))sym("ncclSend");
        api.Recv = (decltype(api.Recv))sym("ncclRecv");
        api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
        api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
        api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
        if (!api.GetUniqueId || !api.CommInitRank || !api.CommInitAll || !api.CommDestroy || !api.Send || !api.Recv ||
            !api.GroupStart || !api.GroupEnd || !api.GetErrorString) {
            dlclose(api.lib);
            api.lib = nullptr;
        }
    });
    return api;
}

int need_nccl() {
    if (nccl().lib) return RT_OK;
    return fail(RT_ERR_UNSUPPORTED, "NCCL is not available (dlopen libnccl.so.2: " + nccl().why + ")");
}

#define RT_NCCL(expr)                                                                                   \
    do {                                                                                                \
        ncclResult_t r_ = (expr);                                                                       \
        if (r_ != 0) {                                                                                  \
            char b_[384];                                                                               \
            snprintf(b_, sizeof b_, "NCCL error = %d (%s) at %s:%d '%s'", (int)r_, nccl().GetErrorString(r_), __FILE__, __LINE__, #expr); \
            return fail(RT_ERR_CUDA, b_);                                                               \
        }                                                                                               \
    } while (0)

constexpr size_t kTailFloats = 4;  // per-rank trailer of the packed buffer: the rank's segment count (u64) + padding to 16 B

// copies the render kernel's segment counter behind the rank's packed tiles, so it travels with them
__global__ void stash_segments_kernel(const unsigned long long* __restrict__ counters, float* __restrict__ tail) {
    reinterpret_cast<unsigned long long*>(tail)[0] = counters[0];
    reinterpret_cast<unsigned long long*>(tail)[1] = 0ull;
}

}  // namespace

struct rt_comm {
    ncclComm_t comm = nullptr;
    int rank = 0, nranks = 1, device = 0;
};

extern "C" {

int rt_comm_unique_id(void* id_out) {
    if (!id_out) return fail(RT_ERR_INVALID, "rt_comm_unique_id: NULL argument");
    int rc = need_nccl();
    if (rc) return rc;
    ncclUniqueId id;
    RT_NCCL(nccl().GetUniqueId(&id));
    memcpy(id_out, &id, sizeof id);
    return RT_OK;
}

int rt_comm_init_rank(const void* id_in, int rank, int nranks, rt_comm** out) {
    if (!out) return fail(RT_ERR_INVALID, "rt_comm_init_rank: out is NULL");
    *out = nullptr;
    if (nranks < 1 || rank < 0 || rank >= nranks) return fail(RT_ERR_INVALID, "rt_comm_init_rank: rank out of range");
    rt_comm* c = new rt_comm();
    c->rank = rank;
    c->nranks = nranks;
    if (cudaGetDevice(&c->device) != cudaSuccess) { cudaGetLastError(); delete c; return fail(RT_ERR_NO_DEVICE, "rt_comm_init_rank: no CUDA device"); }
    if (nranks > 1) {
        if (!id_in) { delete c; return fail(RT_ERR_INVALID, "rt_comm_init_rank: NULL id"); }
        int rc = need_nccl();
        if (rc) { delete c; return rc; }
        ncclUniqueId id;
        memcpy(&id, id_in, sizeof id);
        ncclResult_t r = nccl().CommInitRank(&c->comm, nranks, id, rank);
        if (r != 0) {
            delete c;
            return fail(RT_ERR_CUDA, std::string("ncclCommInitRank failed: ") + nccl().GetErrorString(r));
        }
    }
    *out = c;
    return RT_OK;
}

int rt_comm_init_all(int ndev, const int* devices, rt_comm** out) {
    if (!out || ndev < 1) return fail(RT_ERR_INVALID, "rt_comm_init_all: bad argument");
    std::vector<int> devs(ndev);
    for (int i = 0; i < ndev; ++i) devs[i] = devices ? devices[i] : i;
    std::vector<ncclComm_t> comms(ndev, nullptr);
    if (ndev > 1) {
        int rc = need_nccl();
        if (rc) return rc;
        RT_NCCL(nccl().CommInitAll(comms.data(), ndev, devs.data()));
    }
    for (int i = 0; i < ndev; ++i) {
        out[i] = new rt_comm();
        out[i]->comm = comms[i];
        out[i]->rank = i;
        out[i]->nranks = ndev;
        out[i]->device = devs[i];
    }
    return RT_OK;
}

void rt_comm_destroy(rt_comm* c) {
    if (!c) return;
    if (c->comm && nccl().lib) {
        DeviceGuard guard(c->device);
        nccl().CommDestroy(c->comm);
    }
    delete c;
}

int rt_comm_rank(const rt_comm* c) { return c ? c->rank : 0; }
int rt_comm_size(const rt_comm* c) { return c ? c->nranks : 1; }

int rt_render_tiles(rt_scene* s, rt_comm* c, const void* cam76, const rt_render_opts* opts_in, float* fb, rt_tiles_info* info) {
    if (!s || !cam76) return fail(RT_ERR_INVALID, "rt_render_tiles: NULL argument");
    const int rank = c ? c->rank : 0, nranks = c ? c->nranks : 1;
    if (c && c->device != s->device) return fail(RT_ERR_INVALID, "rt_render_tiles: the communicator and the scene live on different devices");
    CameraPod cam;
    memcpy(&cam, cam76, sizeof cam);
    if (cam.W <= 0 || cam.H <= 0) return fail(RT_ERR_INVALID, "rt_render_tiles: bad CameraData");
    DeviceGuard guard(s->device);
    const size_t n = (size_t)cam.W * cam.H * 3;
    int rc = ensure_framebuffer(s, n);
    if (rc) return rc;
    rt_render_opts o = resolve_opts(opts_in);
    o.struct_size = (int)sizeof o;
    o.tile_rank = rank;
    o.tile_nranks = nranks;
    cudaStream_t st = s->stream;
    rc = launch_frame(s, cam76, &o, s->d_fb, st);  // records ev0 / ev1 around render (+ chunk sum)
    if (rc) return rc;
    const size_t per = tiles_packed_floats(cam.W, cam.H, nranks), stride = per + kTailFloats;
    unsigned long long seg_rank = 0, seg_total = 0, gathered = 0;
    std::vector<unsigned long long> tails;
    if (nranks > 1) {
        const size_t need = rank == 0 ? stride * nranks : stride;
        if (s->packed_floats < need) {
            if (s->d_packed) cudaFree(s->d_packed);
            s->d_packed = nullptr;
            s->packed_floats = 0;
            RT_CUDA(cudaMalloc((void**)&s->d_packed, need * sizeof(float)));
            s->packed_floats = need;
        }
        float* mine = s->d_packed;  // rank 0: slot 0 of the gather buffer (only its trailer is used)
        if (rank != 0) RT_CUDA(launch_tiles_pack(s->d_fb, cam.W, cam.H, rank, nranks, mine, s->num_sms, st));
        stash_segments_kernel<<<1, 1, 0, st>>>(s->d_counters, mine + per);
        RT_CUDA(cudaGetLastError());
        if (rank == 0) {
            RT_NCCL(nccl().GroupStart());
            for (int r = 1; r < nranks; ++r) RT_NCCL(nccl().Recv(s->d_packed + (size_t)r * stride, stride, kNcclFloat, r, c->comm, st));
            RT_NCCL(nccl().GroupEnd());
            RT_CUDA(launch_tiles_unpack(s->d_packed, cam.W, cam.H, nranks, 0, stride, s->d_fb, s->num_sms, st));  // rank 0's tiles are in place
            gathered = (unsigned long long)(nranks - 1) * stride * sizeof(float);
        } else {
            RT_NCCL(nccl().Send(mine, stride, kNcclFloat, 0, c->comm, st));
        }
    }
    RT_CUDA(cudaEventRecord(s->ev2, st));
    unsigned long long d2h = 0;
    if (rank == 0 && fb) {
        RT_CUDA(cudaMemcpyAsync(fb, s->d_fb, n * sizeof(float), cudaMemcpyDeviceToHost, st));
        d2h += n * sizeof(float);
    }
    if (rank == 0 && nranks > 1) {
        tails.resize(nranks);
        RT_CUDA(cudaMemcpy2DAsync(tails.data(), sizeof(unsigned long long), s->d_packed + per, stride * sizeof(float),
                                  sizeof(unsigned long long), nranks, cudaMemcpyDeviceToHost, st));
    }
    float kms = 0.f;
    rc = rt_render_finish(s, &seg_rank, &kms, nullptr);  // waits for ev1, reads the counters
    if (rc) return rc;
    RT_CUDA(cudaStreamSynchronize(st));
    seg_total = seg_rank;
    if (rank == 0 && nranks > 1) {
        seg_total = 0;
        for (unsigned long long v : tails) seg_total += v;
    }
    if (info && info->struct_size >= (int)sizeof(int)) {
        rt_tiles_info out;
        memset(&out, 0, sizeof out);
        out.rank = rank;
        out.nranks = nranks;
        out.segments_rank = seg_rank;
        out.segments_total = rank == 0 ? seg_total : 0;
        out.kernel_ms = kms;
        RT_CUDA(cudaEventElapsedTime(&out.device_ms, s->ev0, s->ev2));
        out.gathered_bytes = gathered;
        out.d2h_bytes = d2h;
        size_t m = (size_t)info->struct_size < sizeof out ? (size_t)info->struct_size : sizeof out;
        out.struct_size = info->struct_size;
        memcpy(info, &out, m);
    }
    return RT_OK;
}

}  // extern "C"
