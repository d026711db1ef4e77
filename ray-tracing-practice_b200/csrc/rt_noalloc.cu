// rt_noalloc.cu -- rt_trace_primary_noalloc: the closest-hit entry point for callers that must not allocate (code under CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same kernel as rt_trace_primary_ex
// (trace_primary_kernel, kernels.cu); the device staging area rt_trace_primary_ex takes from cudaMalloc is a file-scope
// __device__ array here, so the call allocates nothing -- no cudaMalloc / cudaFree, no allocating library routine.
//
// The array exists once per GPU that loads the library (RT_NOALLOC_MAX_RAYS x 36 bytes = 9 MiB, held until the process
// exits) and EVERY call on that GPU shares it: calls that overlap on one GPU (two host threads, two scenes) race on it and
// must be serialised by the caller.  It keeps the previous call's contents; nothing in it has to start at zero -- the ray
// slots [0, n) are overwritten by the upload and the kernel writes type / index / t of every ray in [0, n) before they are
// read back -- so no clear is needed and none is done.
//
// In host code `g_noalloc_scratch` is only the host-side stand-in of the symbol, not a device address: the device address
// is taken with cudaGetSymbolAddress, once per device, with that device current (scratch_of_device), and that pointer is
// what the kernel gets.
#include <cuda_runtime.h>

#include <mutex>

#include "rt_scene_impl.h"

using namespace rtb;

namespace {

constexpr size_t kBytesPerRay = 6 * sizeof(float) + sizeof(int) + sizeof(int) + sizeof(float);  // rays6 in; type, index, t out
constexpr int kMaxDevices = 64;

static __device__ __align__(256) unsigned char g_noalloc_scratch[(size_t)RT_NOALLOC_MAX_RAYS * kBytesPerRay];

std::mutex g_mu;
void* g_scratch_ptr[kMaxDevices];  // device address of g_noalloc_scratch on device i (0: not looked up yet)

// device address of the scratch array on `device`, which must be the calling thread's current device
int scratch_of_device(int device, void** out) {
    if (device < 0 || device >= kMaxDevices) return fail(RT_ERR_UNSUPPORTED, "rt_trace_primary_noalloc: device index beyond 63");
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_ptr[device]) {
        void* p = nullptr;
        RT_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) return fail(RT_ERR_CUDA, "rt_trace_primary_noalloc: cudaGetSymbolAddress returned a null address");
        g_scratch_ptr[device] = p;
    }
    *out = g_scratch_ptr[device];
    return RT_OK;
}

}  // namespace

extern "C" {

int rt_noalloc_max_rays(void) { return RT_NOALLOC_MAX_RAYS; }

int rt_trace_primary_noalloc(rt_scene* s, const float* rays6, int n, float tmin, float tmax, const rt_trace_opts* topts, int* type,
                             int* index, float* t) {
    if (!s || (n > 0 && (!rays6 || !type || !index || !t)) || n < 0) return fail(RT_ERR_INVALID, "rt_trace_primary_noalloc: bad argument");
    if (n > RT_NOALLOC_MAX_RAYS) {
        char b[200];
        snprintf(b, sizeof b, "rt_trace_primary_noalloc: %d rays exceed the static scratch of RT_NOALLOC_MAX_RAYS = %d rays "
                              "(split the batch, or use rt_trace_primary_ex, which allocates)", n, RT_NOALLOC_MAX_RAYS);
        return fail(RT_ERR_INVALID, b);
    }
    int exact = 1;
    if (topts && topts->struct_size >= (int)(2 * sizeof(int))) exact = topts->exact_math ? 1 : 0;
    if (n == 0) return RT_OK;
    DeviceGuard guard(s->device);
    void* base = nullptr;
    int rc = scratch_of_device(s->device, &base);
    if (rc) return rc;
    float* d_rays = (float*)base;
    int* d_type = (int*)(d_rays + 6 * (size_t)n);
    int* d_index = d_type + n;
    float* d_t = (float*)(d_index + n);
    RT_CUDA(cudaMemcpyAsync(d_rays, rays6, (size_t)n * 24, cudaMemcpyHostToDevice, s->stream));
    DevScene dev = s->dev;
    if (!exact) dev.top_nodes = 0;  // as rt_trace_primary_ex: the production megakernel's node path
    RT_CUDA(launch_trace_primary(dev, d_rays, n, tmin, tmax, exact, d_type, d_index, d_t, s->num_sms, s->stream));
    RT_CUDA(cudaMemcpyAsync(type, d_type, (size_t)n * 4, cudaMemcpyDeviceToHost, s->stream));
    RT_CUDA(cudaMemcpyAsync(index, d_index, (size_t)n * 4, cudaMemcpyDeviceToHost, s->stream));
    RT_CUDA(cudaMemcpyAsync(t, d_t, (size_t)n * 4, cudaMemcpyDeviceToHost, s->stream));
    RT_CUDA(cudaStreamSynchronize(s->stream));
    return RT_OK;
}

}  // extern "C"
