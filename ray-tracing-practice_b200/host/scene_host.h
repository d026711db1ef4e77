// scene_host.h -- host-side mirror of the reference's surface around the render path (C++; the reference is C++).
//
//   SceneParams & friends     <- include/scene_params.h:8-58 (same names, same fields)
//   read_scene_params         <- src/main.cu:499-550 (stdin token order preserved exactly)
//   default_config_text       <- print_default_config, src/main.cu:552-570
//   create_scene              <- src/main.cu:346-426 (material / primitive order preserved: indices are parity-relevant)
//   camera_for_frame          <- the per-frame camera path + Camera::build_camera_data, src/camera.cu:297-324, 171-196
//   gpu_render                <- src/camera.cu:290-349, driving the C-ABI in include/rt_b200.h instead of render_kernel
//   BinarySaver/PNG/P3 output <- src/camera.cu:52-153
//   make_synthetic_spheres    -- N-sphere extension used by BASELINE.json configs 2-5
#pragma once
#include <cstdint>
#include <istream>
#include <string>
#include <vector>

#include "../csrc/ref_layouts.h"

namespace rtb {

struct Vec3f {
    float e[3];
};

struct CameraPathParams {
    float rc0, zc0, phic0;
    float Arc, Azc;
    float wrc, wzc, wc;
    float prc, pzc;
    float rn0, zn0, phin0;
    float Arn, Azn;
    float wrn, wzn, wn;
    float prn, pzn;
};
struct BodyParams {
    Vec3f center, col;
    float radius, reflection_coeff, transparency_coeff;
    int lights_on_edge;
};
struct FloorParams {
    Vec3f corners[4];
    std::string texture_path;
    Vec3f tint;
    float reflection_coeff;
};
struct LightSourceParams {
    Vec3f position, col;
};
struct RenderParams {
    int max_depth, sqrt_rays_per_pixel;
};
struct SceneParams {
    int num_frames = 0;
    std::string output_path;
    int width = 0, height = 0;
    float fov_degrees = 0;
    CameraPathParams camera_path{};
    std::vector<BodyParams> bodies;
    FloorParams floor{};
    std::vector<LightSourceParams> lights;
    RenderParams render{};
};

// The arrays create_scene hands to the render path, in the reference's own struct layouts.
struct HostScene {
    std::vector<SphereIn> spheres;
    std::vector<PlaneIn> planes;
    std::vector<MatIn> materials;
    int floor_material = -1;        // material that receives the floor texture (index 0 in create_scene)
    std::vector<float> texels;      // RGBA float texels as stbi_loadf(...,4) returns them; empty = untextured
    int tex_w = 0, tex_h = 0;
};

struct CameraImage {  // CameraData, include/camera.cuh:86-96 (76 bytes)
    float origin[3], pixel00_loc[3], pixel_delta_u[3], pixel_delta_v[3], background[3];
    int image_width, image_height, samples_per_pixel, max_depth;
};
static_assert(sizeof(CameraImage) == 76, "CameraData layout");

SceneParams read_scene_params(std::istream& in);
std::string default_config_text();
// load_texture=false leaves HostScene::texels empty (the reference prints and continues when the file is missing)
bool create_scene(const SceneParams& p, HostScene& out, bool load_texture, std::string* err);
PlaneIn make_plane(const Vec3f& base, const Vec3f& u, const Vec3f& v, int material_idx, int type);
CameraImage build_camera_data(int height, int width, const Vec3f& lookfrom, const Vec3f& lookat, float vfov, int spp, int max_depth);
CameraImage camera_for_frame(const SceneParams& p, int frame);
// appends N seeded random spheres + a 64-entry material palette (SURVEY.md section 8d, config 2-4 generator)
void make_synthetic_spheres(HostScene& scene, int n, uint32_t seed);

// savers: byte = (unsigned char)(256*clamp(sqrt(sum/divisor),0,0.999)) (src/camera.cu:136-148)
void tonemap_host(const float* fb_sum, size_t n_values, int divisor, unsigned char* out);
bool save_binary(const std::string& path, int W, int H, const unsigned char* rgb);   // BinarySaver, src/camera.cu:128-153
bool save_ppm_text(const std::string& path, int W, int H, const unsigned char* rgb); // FileSaver, src/camera.cu:58-74
bool save_png(const std::string& path, int W, int H, const unsigned char* rgb);      // PNGSaver, src/camera.cu:94-126
bool load_texels(const std::string& path, std::vector<float>& rgba, int& w, int& h); // stbi_loadf(...,4), src/main.cu:18

// gpu_render (src/camera.cu:290-349): every frame through the C-ABI; prints "n\tms\ttotal_rays" per frame.
// format: 0 binary (the reference's default), 1 png, 2 text ppm.  Returns 0, or 99 like check_cuda's exit code.
// num_gpus > 1 distributes the (independent) frames over that many devices of the box, frame n on device n % num_gpus;
// 0 = take RTB_GPUS from the environment, default 1.
int gpu_render(const SceneParams& p, const HostScene& scene, int format, std::ostream& log, int num_gpus = 0);
struct DriverOpts {
    int format = 0;            // 0 binary, 1 png, 2 text ppm
    int num_gpus = 0;          // 0: RTB_GPUS or 1
    bool parity = false;       // RT_RNG_REFERENCE + exact math + CPU texture semantics: the reference program's own frames, byte for byte
    bool pipeline = true;      // false: write each file before the next frame starts (the reference's strictly serial loop)
    bool split_tiles = false;  // with num_gpus > 1: all devices render EVERY frame together (rt_render_tiles) instead of a frame each
};
int gpu_render(const SceneParams& p, const HostScene& scene, const DriverOpts& opts, std::ostream& log);

}  // namespace rtb
