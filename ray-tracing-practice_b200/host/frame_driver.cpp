// frame_driver.cpp -- gpu_render: the reference's frame loop (src/camera.cu:290-349) over the C-ABI.
// Per frame: camera path -> rt_render (kernel + D2H of the sums) -> output stage -> file; stdout line
// "n\tms\ttotal_rays" with total_rays = W*H*spp as the reference prints it (primary samples only).
// The ms printed is, like the reference's, the whole frame (render + download + encode + write).
#include <chrono>
#include <cstdio>
#include <ostream>

#include "../../include/rt_b200.h"
#include "scene_host.h"

namespace rtb {

int gpu_render(const SceneParams& p, const HostScene& scene, int format, std::ostream& log) {
    rt_scene* h = nullptr;
    const float* tex = scene.texels.empty() ? nullptr : scene.texels.data();
    int rc = rt_scene_create(scene.spheres.data(), (int)scene.spheres.size(), scene.planes.data(), (int)scene.planes.size(),
                             scene.materials.data(), (int)scene.materials.size(), tex, scene.tex_w, scene.tex_h,
                             tex ? scene.floor_material : -1, &h);
    if (rc != RT_OK) {
        std::fprintf(stderr, "%s\n", rt_last_error());
        return 99;  // check_cuda's exit code (include/camera.cuh:20-29)
    }
    std::vector<float> fb((size_t)p.width * p.height * 3);
    std::vector<unsigned char> rgb(fb.size());
    for (int n = 0; n < p.num_frames; ++n) {
        char filename[256];
        std::snprintf(filename, sizeof filename, p.output_path.c_str(), n);
        CameraImage cam = camera_for_frame(p, n);
        auto t0 = std::chrono::steady_clock::now();
        rc = rt_render(h, &cam, nullptr, fb.data(), nullptr, nullptr, nullptr);
        if (rc != RT_OK) {
            std::fprintf(stderr, "%s\n", rt_last_error());
            rt_scene_destroy(h);
            return 99;
        }
        // the reference hands sqrt_rays_per_pixel (not spp) to its saver (src/camera.cu:300)
        rc = rt_tonemap_u8(h, nullptr, p.width, p.height, p.render.sqrt_rays_per_pixel, rgb.data());
        if (rc != RT_OK) {
            std::fprintf(stderr, "%s\n", rt_last_error());
            rt_scene_destroy(h);
            return 99;
        }
        if (format == 1) save_png(filename, p.width, p.height, rgb.data());
        else if (format == 2) save_ppm_text(filename, p.width, p.height, rgb.data());
        else save_binary(filename, p.width, p.height, rgb.data());  // open failures are silent, as in the reference
        float ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
        long long total_rays = (long long)p.width * p.height * p.render.sqrt_rays_per_pixel * p.render.sqrt_rays_per_pixel;
        log << n << "\t" << ms << "\t" << total_rays << "\n";
    }
    rt_scene_destroy(h);
    return 0;
}

}  // namespace rtb
