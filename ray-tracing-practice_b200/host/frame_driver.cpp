// frame_driver.cpp -- gpu_render: the reference's frame loop (src/camera.cu:290-349) over the C-ABI, pipelined.
// Per frame: camera path -> rt_render_u8 (kernel + device-side output stage + 3 B/pixel download) -> a writer thread
// encodes and writes the file while the NEXT frame renders (the reference does render, download, per-pixel writeColor
// and file I/O strictly one after the other inside its timed region).  stdout line "n\tms\ttotal_rays" with
// total_rays = W*H*spp as the reference prints it (primary samples only); ms = wall time from the start of
// the frame until its image is handed to the writer (render + output stage + download).
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <mutex>
#include <ostream>
#include <string>
#include <thread>
#include <vector>

#include "../../include/rt_b200.h"
#include "scene_host.h"

namespace rtb {

namespace {
// one image in flight to the writer, one being filled: classic double buffer
struct Writer {
    std::mutex m;
    std::condition_variable cv;
    std::vector<unsigned char> img;
    std::string path;
    int w = 0, h = 0, format = 0;
    bool full = false, stop = false;
    std::thread th;
    Writer() : th([this] { run(); }) {}
    void run() {
        for (;;) {
            std::vector<unsigned char> local;
            std::string p;
            int lw, lh, lf;
            {
                std::unique_lock<std::mutex> lk(m);
                cv.wait(lk, [this] { return full || stop; });
                if (!full) return;
                local.swap(img);
                p = path; lw = w; lh = h; lf = format;
                full = false;
            }
            cv.notify_all();
            if (lf == 1) save_png(p.c_str(), lw, lh, local.data());
            else if (lf == 2) save_ppm_text(p.c_str(), lw, lh, local.data());
            else save_binary(p.c_str(), lw, lh, local.data());  // open failures are silent, as in the reference
        }
    }
    void submit(std::vector<unsigned char>& rgb, const char* p, int W, int H, int fmt) {
        std::unique_lock<std::mutex> lk(m);
        cv.wait(lk, [this] { return !full; });
        img.swap(rgb);
        path = p; w = W; h = H; format = fmt;
        full = true;
        lk.unlock();
        cv.notify_all();
    }
    ~Writer() {
        {
            std::unique_lock<std::mutex> lk(m);
            cv.wait(lk, [this] { return !full; });
            stop = true;
        }
        cv.notify_all();
        th.join();
    }
};
}  // namespace

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
    const size_t bytes = (size_t)p.width * p.height * 3;
    int status = 0;
    {
        Writer writer;
        std::vector<unsigned char> rgb;
        for (int n = 0; n < p.num_frames; ++n) {
            char filename[256];
            std::snprintf(filename, sizeof filename, p.output_path.c_str(), n);
            CameraImage cam = camera_for_frame(p, n);
            auto t0 = std::chrono::steady_clock::now();
            rgb.resize(bytes);
            // the reference hands sqrt_rays_per_pixel (not spp) to its saver (src/camera.cu:300)
            rc = rt_render_u8(h, &cam, nullptr, p.render.sqrt_rays_per_pixel, rgb.data(), nullptr, nullptr);
            if (rc != RT_OK) {
                std::fprintf(stderr, "%s\n", rt_last_error());
                status = 99;
                break;
            }
            writer.submit(rgb, filename, p.width, p.height, format);
            float ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
            long long total_rays = (long long)p.width * p.height * p.render.sqrt_rays_per_pixel * p.render.sqrt_rays_per_pixel;
            log << n << "\t" << ms << "\t" << total_rays << "\n";
        }
    }  // ~Writer: the last file is on disk before we return
    rt_scene_destroy(h);
    return status;
}

}  // namespace rtb
