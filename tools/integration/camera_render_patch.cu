// The patch a maintainer of zloyaloha/ray-tracing-practice would apply to render through librt_b200.so: Camera::render's body
// (src/camera.cu:198-216: render_kernel<<<>>> + cudaMemcpy + saver loop) and the one-time setup at the top of gpu_render
// (src/camera.cu:291).  It compiles against the reference's own headers: tests/test_host.py builds it with
//   nvcc -std=c++17 -c -I/root/reference/include -Iinclude tools/integration/camera_render_patch.cu
// wherever /root/reference exists.  Everything else of the reference -- read_scene_params, create_scene, the camera path,
// build_camera_data, the savers, main -- stays as it is; main calls create_scene(..., use_gpu=false) for the GPU mode too,
// because the library takes the HOST arrays and does the upload itself.
#include <cstdlib>
#include <iostream>
#include <vector>

#include "camera.cuh"   // the reference's header, unmodified
#include "rt_b200.h"

static rt_scene* g_rt_scene = nullptr;

static void rt_check(int rc) {  // keeps checkCudaErrors' contract (include/camera.cuh:20-29): print, exit(99)
    if (rc != RT_OK) {
        std::cerr << "rt_b200: " << rt_last_error() << "\n";
        std::exit(99);
    }
}

// replaces `cudaMemcpyToSymbol(d_scene_data_const, &host_scene, ...)` at the top of gpu_render (src/camera.cu:291).
// host_scene is what create_scene(..., use_gpu=false) leaves behind (src/main.cu:479-496): host pointers, and the floor
// texture as the CpuTexture of material 0 (src/main.cu:354-362).
void rt_b200_setup(const SceneData& host_scene) {
    const CpuTexture* tex = host_scene.num_materials > 0 ? host_scene.d_materials[0].cpu_tex : nullptr;
    rt_check(rt_scene_create(host_scene.d_spheres, host_scene.num_spheres, host_scene.d_planes, host_scene.num_planes,
                             host_scene.d_materials, host_scene.num_materials, tex ? tex->data : nullptr, tex ? tex->width : 0,
                             tex ? tex->height : 0, tex ? 0 : -1, &g_rt_scene));
}

void rt_b200_teardown() {  // next to cudaFree(d_fb) at the end of gpu_render (src/camera.cu:348)
    rt_scene_destroy(g_rt_scene);
    g_rt_scene = nullptr;
}

// replaces the body of Camera::render (src/camera.cu:198-216); d_fb is no longer used (the library owns device memory)
void Camera::render(color* /*d_fb*/) const {
    CameraData cam = build_camera_data();  // src/camera.cu:171-196, unchanged
    std::vector<color> host_fb(static_cast<std::size_t>(imageWidth) * imageHeight);
    static_assert(sizeof(color) == 3 * sizeof(float), "color is three floats: the framebuffer contract of rt_render");
    // opts = NULL: megakernel, Philox seed 0, fast fp32, texture-unit semantics.  For a regression run against the reference's
    // own --cpu frames: rt_render_opts o = {sizeof o}; o.rng = RT_RNG_REFERENCE; o.exact_math = 1; o.tex_mode = RT_TEX_CPU;
    rt_check(rt_render(g_rt_scene, &cam, nullptr, reinterpret_cast<float*>(host_fb.data()), nullptr, nullptr, nullptr));
    for (int j = 0; j < imageHeight; ++j)
        for (int i = 0; i < imageWidth; ++i) saver->writeColor(host_fb[static_cast<std::size_t>(j) * imageWidth + i]);
}
