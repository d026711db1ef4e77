// host_api.cpp -- C exports of the host mirror (scene_host.h) so tests / bench.py can drive it through ctypes.
// Pure host code: callable without a GPU.
#include <cstring>
#include <sstream>

#include "../csrc/scene_pack.h"
#include "scene_host.h"

using namespace rtb;

namespace {
struct HostHandle {
    SceneParams params;
    HostScene scene;
    std::string err;
};
}  // namespace

extern "C" {

// parse a config text (stdin token stream of the reference) and assemble the scene.  texture_override: NULL keeps the
// path of the config, "" skips the texture, anything else replaces the path.
int rth_scene_from_config(const char* text, const char* texture_override, void** out) {
    HostHandle* h = new HostHandle();
    std::istringstream in(text);
    h->params = read_scene_params(in);
    bool load = true;
    if (texture_override) {
        if (!*texture_override) load = false;
        else h->params.floor.texture_path = texture_override;
    }
    if (!create_scene(h->params, h->scene, load, &h->err)) { *out = h; return 1; }
    *out = h;
    return 0;
}
int rth_scene_empty(void** out) { *out = new HostHandle(); return 0; }
void rth_scene_destroy(void* h) { delete (HostHandle*)h; }
const char* rth_scene_error(void* h) { return ((HostHandle*)h)->err.c_str(); }
void rth_scene_counts(void* hh, int* out5) {
    HostHandle* h = (HostHandle*)hh;
    out5[0] = (int)h->scene.spheres.size(); out5[1] = (int)h->scene.planes.size(); out5[2] = (int)h->scene.materials.size();
    out5[3] = h->scene.tex_w; out5[4] = h->scene.tex_h;
}
void rth_scene_copy(void* hh, void* sp, void* pl, void* mt, float* tex) {
    HostHandle* h = (HostHandle*)hh;
    if (sp) std::memcpy(sp, h->scene.spheres.data(), h->scene.spheres.size() * sizeof(SphereIn));
    if (pl) std::memcpy(pl, h->scene.planes.data(), h->scene.planes.size() * sizeof(PlaneIn));
    if (mt) std::memcpy(mt, h->scene.materials.data(), h->scene.materials.size() * sizeof(MatIn));
    if (tex && !h->scene.texels.empty()) std::memcpy(tex, h->scene.texels.data(), h->scene.texels.size() * sizeof(float));
}
void rth_scene_add_spheres(void* hh, int n, unsigned seed) { make_synthetic_spheres(((HostHandle*)hh)->scene, n, seed); }
// W,H,max_depth,sqrt_spp,num_frames + fov
void rth_scene_params(void* hh, int* out5, float* fov) {
    HostHandle* h = (HostHandle*)hh;
    out5[0] = h->params.width; out5[1] = h->params.height; out5[2] = h->params.render.max_depth;
    out5[3] = h->params.render.sqrt_rays_per_pixel; out5[4] = h->params.num_frames;
    *fov = h->params.fov_degrees;
}
void rth_camera_for_frame(void* hh, int frame, void* out76) {
    CameraImage c = camera_for_frame(((HostHandle*)hh)->params, frame);
    std::memcpy(out76, &c, 76);
}
void rth_build_camera(int H, int W, const float* from3, const float* at3, float vfov, int spp, int depth, void* out76) {
    Vec3f f{{from3[0], from3[1], from3[2]}}, a{{at3[0], at3[1], at3[2]}};
    CameraImage c = build_camera_data(H, W, f, a, vfov, spp, depth);
    std::memcpy(out76, &c, 76);
}
void rth_make_planes(const float* base_u_v9, const int* mat, const int* type, int n, void* out80) {
    PlaneIn* o = (PlaneIn*)out80;
    for (int i = 0; i < n; ++i) {
        const float* p = base_u_v9 + 9 * (size_t)i;
        o[i] = make_plane(Vec3f{{p[0], p[1], p[2]}}, Vec3f{{p[3], p[4], p[5]}}, Vec3f{{p[6], p[7], p[8]}}, mat[i], type[i]);
    }
}
const char* rth_default_config(void) {
    static std::string s = default_config_text();
    return s.c_str();
}
void rth_tonemap(const float* fb, unsigned long long n, int divisor, unsigned char* out) { tonemap_host(fb, (size_t)n, divisor, out); }
int rth_save(const char* path, int W, int H, const unsigned char* rgb, int format) {
    bool ok = format == 1 ? save_png(path, W, H, rgb) : format == 2 ? save_ppm_text(path, W, H, rgb) : save_binary(path, W, H, rgb);
    return ok ? 0 : 1;
}
int rth_load_texels(const char* path, float* out, int* w, int* h) {  // out may be NULL to query the size
    std::vector<float> t;
    if (!load_texels(path, t, *w, *h)) return 1;
    if (out) std::memcpy(out, t.data(), t.size() * sizeof(float));
    return 0;
}
// wide-BVH statistics + structural validation for host-side tests: out = nodes, depth, max_stack, leaves
int rth_bvh_check(const void* sp, int S, const void* pl, int P, int* out4, float* sah, float* build_ms) {
    PackedScene pk = pack_scene((const SphereIn*)sp, S, (const PlaneIn*)pl, P, nullptr, 0);
    out4[0] = (int)pk.bvh.nodes.size(); out4[1] = pk.bvh.depth; out4[2] = pk.bvh.max_stack; out4[3] = pk.bvh.leaves;
    if (sah) *sah = pk.bvh.sah_cost;
    if (build_ms) *build_ms = pk.build_ms;
    std::vector<PrimBox> prims = make_prim_boxes((const SphereIn*)sp, S, (const PlaneIn*)pl, P);
    return validate_wide_bvh(pk.bvh, prims);
}
// gpu_render over the C-ABI for a parsed config (needs a GPU).  Captures the per-frame log lines.
int rth_gpu_render(void* hh, int format, char* log, int log_cap) {
    HostHandle* h = (HostHandle*)hh;
    std::ostringstream os;
    int rc = gpu_render(h->params, h->scene, format, os);
    if (log && log_cap > 0) { std::strncpy(log, os.str().c_str(), log_cap - 1); log[log_cap - 1] = 0; }
    return rc;
}
// flags: 1 parity, 2 no pipeline, 4 split tiles
int rth_gpu_render_ex(void* hh, int format, int flags, int num_gpus, char* log, int log_cap) {
    HostHandle* h = (HostHandle*)hh;
    std::ostringstream os;
    DriverOpts d;
    d.format = format;
    d.num_gpus = num_gpus;
    d.parity = flags & 1;
    d.pipeline = !(flags & 2);
    d.split_tiles = flags & 4;
    int rc = gpu_render(h->params, h->scene, d, os);
    if (log && log_cap > 0) { std::strncpy(log, os.str().c_str(), log_cap - 1); log[log_cap - 1] = 0; }
    return rc;
}
void rth_set_output(void* hh, const char* pattern, int num_frames, int W, int H, int max_depth, int sqrt_spp) {
    HostHandle* h = (HostHandle*)hh;
    if (pattern) h->params.output_path = pattern;
    if (num_frames > 0) h->params.num_frames = num_frames;
    if (W > 0) h->params.width = W;
    if (H > 0) h->params.height = H;
    if (max_depth > 0) h->params.render.max_depth = max_depth;
    if (sqrt_spp > 0) h->params.render.sqrt_rays_per_pixel = sqrt_spp;
}

}  // extern "C"
