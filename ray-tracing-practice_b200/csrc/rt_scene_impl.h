// rt_scene_impl.h -- what the translation units behind the C-ABI (rt_api.cu, rt_tiles.cu) share: the opaque scene handle,
// the thread-local error string, the CUDA error macro and the device guard.  Not part of the public interface.
#pragma once
#include <cuda_runtime.h>

#include <cstdio>
#include <string>

#include "../../include/rt_b200.h"
#include "kernels.h"

namespace rtb {
std::string& last_error();
inline int fail(int code, const std::string& msg) {
    last_error() = msg;
    return code;
}
}  // namespace rtb

#define RT_CUDA(expr)                                                                                              \
    do {                                                                                                           \
        cudaError_t e_ = (expr);                                                                                   \
        if (e_ != cudaSuccess) {                                                                                   \
            char b_[512];                                                                                          \
            snprintf(b_, sizeof b_, "CUDA error = %u (%s) at %s:%d '%s'", (unsigned)e_, cudaGetErrorString(e_), __FILE__, \
                     __LINE__, #expr);                                                                             \
            cudaGetLastError();                                                                                    \
            return rtb::fail(RT_ERR_CUDA, b_);                                                                     \
        }                                                                                                          \
    } while (0)

namespace rtb {
// A scene lives on the device it was created on; every entry point that takes one switches the calling thread to that device
// for the duration of the call (and back), so a handle can be used from any thread whatever its current device is.
struct DeviceGuard {
    int prev = -1;
    bool switched = false;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) == cudaSuccess && prev != dev && cudaSetDevice(dev) == cudaSuccess) switched = true;
    }
    ~DeviceGuard() {
        if (switched) cudaSetDevice(prev);
    }
    DeviceGuard(const DeviceGuard&) = delete;
    DeviceGuard& operator=(const DeviceGuard&) = delete;
};


// launches one frame's kernels on `stream` (render + chunk sum); rt_render_finish collects counters and timing
int launch_frame(rt_scene* s, const void* cam76, const rt_render_opts* opts_in, float* d_fb, cudaStream_t stream);
rt_render_opts resolve_opts(const rt_render_opts* in);
int ensure_framebuffer(rt_scene* s, size_t floats);
}  // namespace rtb

struct rt_scene {
    int device = 0;
    int num_sms = 148;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    rt_scene_stats stats{};
    rtb::DevScene dev{};
    int default_top = 0;
    // owned device memory
    void* d_nodes = nullptr;
    void* d_spheres = nullptr;
    void* d_sphere_meta = nullptr;
    void* d_planes = nullptr;
    void* d_planes_fast = nullptr;
    void* d_mats = nullptr;
    void* d_tex = nullptr;
    void* d_grid_cells = nullptr;
    void* d_grid_refs = nullptr;
    cudaArray_t tex_array = nullptr;
    cudaTextureObject_t tex_obj = 0;
    float* d_fb = nullptr;
    size_t fb_floats = 0;
    unsigned char* d_u8 = nullptr;
    size_t u8_bytes = 0;
    unsigned int* d_tile_counter = nullptr;
    unsigned long long* d_counters = nullptr;
    void* d_trace = nullptr;
    size_t trace_bytes = 0;
    rtb::WavefrontScratch wf;
    long long* d_acc = nullptr;  // fixed-point accumulators of the sample chunks (render_mega), W*H*3 values
    size_t acc_values = 0;
    float* d_packed = nullptr; // rt_render_tiles: nranks packed tile buffers on rank 0 (own slot unused), one elsewhere
    size_t packed_floats = 0;
    cudaEvent_t ev2 = nullptr; // end of the gather + unpack of rt_render_tiles
    bool pending = false;
    cudaStream_t pending_stream = nullptr;
};

