// frame_driver.cpp -- gpu_render: the reference's frame loop (src/camera.cu:290-349) over the C-ABI, pipelined.
// Per frame: camera path -> rt_render_u8 (kernel + device-side output stage + 3 B/pixel download) -> a writer thread
// encodes and writes the file while the NEXT frame renders (the reference does render, download, per-pixel writeColor
// and file I/O strictly one after the other inside its timed region; DriverOpts::pipeline = false does the
// same here).  stdout line "n\tms\ttotal_rays" with total_rays = W*H*spp as the reference prints it (primary samples
// only); ms = wall time from the start of the frame until its image is handed to the writer (render + output
// stage + download), or until the file is written when the pipeline is off.
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <memory>
#include <mutex>
#include <algorithm>
#include <cstdlib>
#include <ostream>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

#include "../../include/rt_b200.h"
#include "scene_host.h"

namespace rtb {

static void save_frame(int format, const char* path, int W, int H, const unsigned char* rgb);

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
            save_frame(lf, p.c_str(), lw, lh, local.data());
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

// the reference hands its config's output_path to sprintf as the FORMAT (src/camera.cu:318): accept what that is meant for --
// literal text, "%%", and at most one int conversion -- and refuse anything else instead of reading stray varargs
static bool output_pattern_ok(const std::string& p) {
    int conversions = 0;
    for (size_t i = 0; i < p.size(); ++i) {
        if (p[i] != '%') continue;
        if (i + 1 < p.size() && p[i + 1] == '%') { ++i; continue; }
        size_t j = i + 1;
        while (j < p.size() && (p[j] == '0' || p[j] == '-' || p[j] == '+' || p[j] == ' ')) ++j;
        while (j < p.size() && p[j] >= '0' && p[j] <= '9') ++j;
        if (j >= p.size() || (p[j] != 'd' && p[j] != 'i' && p[j] != 'u' && p[j] != 'x' && p[j] != 'X')) return false;
        ++conversions;
        i = j;
    }
    return conversions <= 1;
}

// prints the per-frame lines in frame order as soon as the next expected frame is complete (the reference streams one line
// per frame, src/camera.cu:344-346): immediately with one worker, through this cursor with several
struct OrderedLog {
    std::mutex m;
    std::ostream& out;
    std::vector<std::string> lines;
    std::vector<char> done;
    size_t next = 0;
    OrderedLog(std::ostream& o, int n) : out(o), lines((size_t)std::max(n, 0)), done((size_t)std::max(n, 0), 0) {}
    void put(int n, std::string line) {
        std::lock_guard<std::mutex> lk(m);
        lines[(size_t)n] = std::move(line);
        done[(size_t)n] = 1;
        while (next < done.size() && done[next]) {
            out << lines[next];
            lines[next].clear();
            ++next;
        }
        out.flush();
    }
};

static std::string frame_line(const SceneParams& p, int n, float ms) {
    long long total_rays = (long long)p.width * p.height * p.render.sqrt_rays_per_pixel * p.render.sqrt_rays_per_pixel;
    std::ostringstream os;
    os << n << "\t" << ms << "\t" << total_rays << "\n";
    return os.str();
}

static void save_frame(int format, const char* path, int W, int H, const unsigned char* rgb) {
    if (format == 1) save_png(path, W, H, rgb);
    else if (format == 2) save_ppm_text(path, W, H, rgb);
    else save_binary(path, W, H, rgb);  // open failures are silent, as in the reference
}

static rt_render_opts driver_render_opts(const DriverOpts& d) {
    rt_render_opts o;
    std::memset(&o, 0, sizeof o);
    o.struct_size = (int)sizeof o;
    o.tex_mode = RT_TEX_GPU;
    o.top_nodes_in_smem = -1;
    o.tile_nranks = 1;
    if (d.parity) {  // the reference program's own arithmetic and random stream: frames byte-identical to its --cpu run
        o.rng = RT_RNG_REFERENCE;
        o.exact_math = 1;
        o.tex_mode = RT_TEX_CPU;
    }
    return o;
}

static rt_scene* create_device_scene(const HostScene& scene) {
    rt_scene* h = nullptr;
    const float* tex = scene.texels.empty() ? nullptr : scene.texels.data();
    int rc = rt_scene_create(scene.spheres.data(), (int)scene.spheres.size(), scene.planes.data(), (int)scene.planes.size(),
                             scene.materials.data(), (int)scene.materials.size(), tex, scene.tex_w, scene.tex_h,
                             tex ? scene.floor_material : -1, &h);
    if (rc != RT_OK) {
        std::fprintf(stderr, "%s\n", rt_last_error());
        return nullptr;
    }
    return h;
}

// One device's share of the frame loop: frames first, first+stride, ... on the CURRENT device (the caller selected it).
static int render_frames_on_device(const SceneParams& p, const HostScene& scene, const DriverOpts& d, int first, int stride,
                                   OrderedLog& log) {
    rt_scene* h = create_device_scene(scene);
    if (!h) return 99;  // check_cuda's exit code (include/camera.cuh:20-29)
    const size_t bytes = (size_t)p.width * p.height * 3;
    const rt_render_opts ro = driver_render_opts(d);
    int status = 0;
    {
        std::unique_ptr<Writer> writer(d.pipeline ? new Writer() : nullptr);
        std::vector<unsigned char> rgb;
        for (int n = first; n < p.num_frames; n += stride) {
            char filename[512];
            std::snprintf(filename, sizeof filename, p.output_path.c_str(), n);
            CameraImage cam = camera_for_frame(p, n);
            auto t0 = std::chrono::steady_clock::now();
            rgb.resize(bytes);
            // the reference hands sqrt_rays_per_pixel (not spp) to its saver (src/camera.cu:300)
            int rc = rt_render_u8(h, &cam, &ro, p.render.sqrt_rays_per_pixel, rgb.data(), nullptr, nullptr);
            if (rc != RT_OK) {
                std::fprintf(stderr, "%s\n", rt_last_error());
                status = 99;
                break;
            }
            if (writer) writer->submit(rgb, filename, p.width, p.height, d.format);
            else save_frame(d.format, filename, p.width, p.height, rgb.data());  // inside the frame's time, like the reference (Q9)
            float ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
            log.put(n, frame_line(p, n, ms));
        }
    }  // ~Writer: the last file is on disk before we return
    rt_scene_destroy(h);
    return status;
}

// --split tiles: every frame is rendered by ALL devices together (rt_render_tiles: interleaved tiles, NCCL gather to rank 0);
// rank 0 runs the output stage and hands the image to the writer.  One thread per rank, in lock step through the collective.
static int render_frames_tiled(const SceneParams& p, const HostScene& scene, const DriverOpts& d, int ndev, OrderedLog& log) {
    std::vector<rt_comm*> comms((size_t)ndev, nullptr);
    if (rt_comm_init_all(ndev, nullptr, comms.data()) != RT_OK) {
        std::fprintf(stderr, "%s\n", rt_last_error());
        return 99;
    }
    std::vector<int> status((size_t)ndev, 0);
    std::vector<std::thread> th;
    const rt_render_opts ro = driver_render_opts(d);
    for (int g = 0; g < ndev; ++g)
        th.emplace_back([&, g] {
            if (rt_set_device(g) != RT_OK) { std::fprintf(stderr, "%s\n", rt_last_error()); status[g] = 99; }
            rt_scene* h = status[g] ? nullptr : create_device_scene(scene);
            if (!h) status[g] = 99;
            // a rank that failed to set up still has to take part in every collective, or the others would wait forever:
            // nothing sensible to do but stop the whole run
            if (status[g]) { std::fprintf(stderr, "rt_b200: rank %d could not be set up\n", g); std::exit(99); }
            std::unique_ptr<Writer> writer(g == 0 && d.pipeline ? new Writer() : nullptr);
            std::vector<unsigned char> rgb;
            const size_t bytes = (size_t)p.width * p.height * 3;
            for (int n = 0; n < p.num_frames; ++n) {
                char filename[512];
                std::snprintf(filename, sizeof filename, p.output_path.c_str(), n);
                CameraImage cam = camera_for_frame(p, n);
                auto t0 = std::chrono::steady_clock::now();
                int rc = rt_render_tiles(h, comms[g], &cam, &ro, nullptr, nullptr);  // assembled frame stays on rank 0's device
                if (rc == RT_OK && g == 0) {
                    rgb.resize(bytes);
                    rc = rt_tonemap_u8(h, nullptr, p.width, p.height, p.render.sqrt_rays_per_pixel, rgb.data());
                }
                if (rc != RT_OK) { std::fprintf(stderr, "%s\n", rt_last_error()); std::exit(99); }
                if (g == 0) {
                    if (writer) writer->submit(rgb, filename, p.width, p.height, d.format);
                    else save_frame(d.format, filename, p.width, p.height, rgb.data());
                    float ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
                    log.put(n, frame_line(p, n, ms));
                }
            }
            writer.reset();
            rt_scene_destroy(h);
        });
    for (auto& t : th) t.join();
    for (rt_comm* c : comms) rt_comm_destroy(c);
    for (int s : status)
        if (s) return s;
    return 0;
}

// Frames are independent (src/camera.cu:297-347 renders them one after the other): with num_gpus > 1 device g renders
// frames g, g+num_gpus, ... from its own copy of the scene, each with its own writer thread (or, with split_tiles, all
// devices share every frame); the stdout lines come out in frame order as frames complete.
// num_gpus <= 0: RTB_GPUS from the environment, else 1.
int gpu_render(const SceneParams& p, const HostScene& scene, const DriverOpts& d_in, std::ostream& log_out) {
    DriverOpts d = d_in;
    if (d.num_gpus <= 0) {
        const char* e = std::getenv("RTB_GPUS");
        d.num_gpus = e ? std::atoi(e) : 1;
    }
    if (!output_pattern_ok(p.output_path)) {
        std::fprintf(stderr, "rt_b200: output path '%s' is used as a printf format for the frame number: it may hold text, %%%% and at most one %%d\n",
                     p.output_path.c_str());
        return 1;
    }
    const int have = rt_device_count();
    if (have < 1) {
        std::fprintf(stderr, "rt_b200: no CUDA device\n");
        return 99;
    }
    OrderedLog log(log_out, p.num_frames);
    if (d.split_tiles && d.num_gpus > 1) {
        if (d.num_gpus > have) {
            std::fprintf(stderr, "rt_b200: --split tiles needs one device per rank (%d asked, %d present)\n", d.num_gpus, have);
            return 1;
        }
        return render_frames_tiled(p, scene, d, d.num_gpus, log);
    }
    // more workers than devices is allowed (worker g uses device g % have): it only costs memory, and it lets a one-GPU box
    // run the same threaded path
    int workers = std::max(1, std::min(d.num_gpus, std::max(p.num_frames, 1)));
    std::vector<int> status((size_t)workers, 0);
    if (workers == 1) {
        status[0] = render_frames_on_device(p, scene, d, 0, 1, log);
    } else {
        std::vector<std::thread> th;
        for (int g = 0; g < workers; ++g)
            th.emplace_back([&, g] {
                if (rt_set_device(g % have) != RT_OK) { std::fprintf(stderr, "%s\n", rt_last_error()); status[g] = 99; return; }
                status[g] = render_frames_on_device(p, scene, d, g, workers, log);
            });
        for (auto& t : th) t.join();
    }
    for (int s : status)
        if (s) return s;
    return 0;
}

int gpu_render(const SceneParams& p, const HostScene& scene, int format, std::ostream& log, int num_gpus) {
    DriverOpts d;
    d.format = format;
    d.num_gpus = num_gpus;
    return gpu_render(p, scene, d, log);
}

}  // namespace rtb
