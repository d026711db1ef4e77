// oracle/ref_harness.cu -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// Builds the UNMODIFIED reference (zloyaloha/ray-tracing-practice) into
// oracle/_ref/libref_rt.so by #including its two translation units from where
// they lie (-I$(REF) -I$(REF)/include); no reference source is copied into this
// repository.  This file only adds an extern "C" surface around the reference's
// own functions so that tests / bench.py's cpu_baseline + `--impl reference`
// legs can drive them through ctypes:
//   * known-answer entry points around wang_hash / random_float / hit_sphere /
//     hit_plane / AABB::hit / reflectance / material_scatter / tex2D_cpu
//   * scene assembly through the reference's read_scene_params + create_scene
//     (src/main.cu:346-550) or from caller arrays + the reference's build_bvh
//   * closest-hit with primitive ids: brute force over the reference's leaf
//     tests (the commented-out linear path, include/scene.h:37-51) and an
//     id-recording walk of the reference's BVHNode array that is cross-checked
//     against the real hit_bvh (include/bvh.h:19-65) on every ray
//   * the CPU frame: Camera::render_cpu / ray_color_host (src/camera.cu:36-50,
//     254-288), row-parallel over std::thread for the timed baseline
//   * the reference CUDA render_kernel (src/camera.cu:17-34) fed the same arrays
//     (GPU baseline, only callable on a box with a GPU).
#include <atomic>
#include <cstring>
#include <sstream>
#include <thread>

#define main ref_main_unused_
#include "src/main.cu"
#undef main
#include "src/camera.cu"

namespace {

struct RefScene {
    SceneData sd{};
    std::vector<MaterialData> mats;
    std::vector<SphereData> sph;
    std::vector<PlaneData> pl;
    std::vector<BVHNode> nodes;
    std::vector<BVHTree> trees;
    SceneParams params;
    bool has_params = false;
    std::vector<float> tex_copy;  // owned texels when built from arrays
    CpuTexture* own_tex = nullptr;
    int tex_mat = -1;
    // device mirror for the reference CUDA kernel
    bool gpu_ready = false;
    SceneData dsd{};
    BVHNode* d_nodes = nullptr;
    cudaTextureObject_t d_tex = 0;
    color* d_fb = nullptr;
    size_t d_fb_pixels = 0;
};

Ray make_ray(const float* r6) { return Ray(point3(r6[0], r6[1], r6[2]), vec3(r6[3], r6[4], r6[5])); }

void put_rec(const HitRecord& rec, float* out) {
    out[0] = rec.t;
    out[1] = rec.normal.x();
    out[2] = rec.normal.y();
    out[3] = rec.normal.z();
    out[4] = rec.u;
    out[5] = rec.v;
    out[6] = rec.front_face ? 1.0f : 0.0f;
    out[7] = (float)rec.material_idx;
    out[8] = rec.point.x();
    out[9] = rec.point.y();
    out[10] = rec.point.z();
}

int bvh_depth(const std::vector<BVHNode>& n, int i) {
    if (n.empty()) return 0;
    if (n[i].left < 0) return 1;
    int a = bvh_depth(n, n[i].left), b = bvh_depth(n, n[i].right);
    return 1 + (a > b ? a : b);
}

// segment-counting transcription of ray_color_host's loop (src/camera.cu:254-288);
// calls the reference's own hit_scene / material_scatter.  ref_selfcheck_color()
// proves it returns the identical colour as the unmodified function.
color ray_color_counted(Ray r, unsigned int& seed, const SceneData& sd, const CameraData& cam, unsigned long long& segs) {
    color total(0, 0, 0), beta(1, 1, 1);
    Ray cur = r;
    for (int depth = 0; depth < cam.max_depth; depth++) {
        HitRecord rec;
        segs++;
        if (!hit_scene(cur, Interval(0.001f, 1e30f), rec, sd)) {
            total += beta * cam.background;
            break;
        }
        MaterialData m = sd.d_materials[rec.material_idx];
        if (m.cpu_tex != nullptr) m.albedo = m.albedo * tex2D_cpu(m.cpu_tex, rec.u, rec.v);
        total += beta * material_emit(m);
        Ray sc;
        color att;
        if (!material_scatter(cur, rec, att, sc, seed, m)) break;
        beta *= att;
        cur = sc;
    }
    return total;
}

}  // namespace

extern "C" {

// ---------------------------------------------------------------- KATs
unsigned int ref_wang_hash(unsigned int s) { return wang_hash(s); }
float ref_random_float(unsigned int* seed) { return random_float(*seed); }
float ref_reflectance(float c, float r) { return reflectance(c, r); }
void ref_random_in_unit_sphere(unsigned int* seed, float* out3) {
    vec3 v = random_in_unit_sphere(*seed);
    out3[0] = v.x(); out3[1] = v.y(); out3[2] = v.z();
}
void ref_random_in_hemisphere(const float* n3, unsigned int* seed, float* out3) {
    vec3 v = random_in_hemisphere(vec3(n3[0], n3[1], n3[2]), *seed);
    out3[0] = v.x(); out3[1] = v.y(); out3[2] = v.z();
}
int ref_hit_sphere(const float* ray6, float tmin, float tmax, const float* cr4, float* rec11) {
    SphereData s(point3(cr4[0], cr4[1], cr4[2]), cr4[3], 0);
    HitRecord rec{};
    bool h = hit_sphere(make_ray(ray6), Interval(tmin, tmax), rec, s);
    if (h) put_rec(rec, rec11);
    return h;
}
// plane given as base,u,v (9 floats) + type; also returns the ctor's derived D,w,normal (7 floats)
int ref_hit_plane(const float* ray6, float tmin, float tmax, const float* buv9, int type, float* rec11, float* derived7) {
    PlaneData p(point3(buv9[0], buv9[1], buv9[2]), vec3(buv9[3], buv9[4], buv9[5]), vec3(buv9[6], buv9[7], buv9[8]), 0,
                (PlaneType)type);
    if (derived7) {
        derived7[0] = p.D;
        derived7[1] = p.w.x(); derived7[2] = p.w.y(); derived7[3] = p.w.z();
        derived7[4] = p.normal.x(); derived7[5] = p.normal.y(); derived7[6] = p.normal.z();
    }
    HitRecord rec{};
    bool h = hit_plane(make_ray(ray6), Interval(tmin, tmax), rec, p);
    if (h) put_rec(rec, rec11);
    return h;
}
int ref_aabb_hit(const float* lo3, const float* hi3, const float* ray6, float tmin, float tmax) {
    AABB b(point3(lo3[0], lo3[1], lo3[2]), point3(hi3[0], hi3[1], hi3[2]));
    return b.hit(make_ray(ray6), Interval(tmin, tmax));
}
// material passed as the reference's 64-byte MaterialData image; rec: point3,normal3,t,front,u,v
int ref_material_scatter(const float* ray6, const float* point3_, const float* normal3, float t, int front,
                         const void* mat64, unsigned int* seed, float* att3, float* out_ray6) {
    MaterialData m;
    std::memcpy((void*)&m, mat64, sizeof(MaterialData));
    HitRecord rec{};
    rec.point = point3(point3_[0], point3_[1], point3_[2]);
    rec.normal = vec3(normal3[0], normal3[1], normal3[2]);
    rec.t = t;
    rec.front_face = front != 0;
    color att(0, 0, 0);
    Ray sc;
    bool alive = material_scatter(make_ray(ray6), rec, att, sc, *seed, m);
    att3[0] = att.x(); att3[1] = att.y(); att3[2] = att.z();
    point3 o = sc.origin();
    vec3 d = sc.direction();
    out_ray6[0] = o.x(); out_ray6[1] = o.y(); out_ray6[2] = o.z();
    out_ray6[3] = d.x(); out_ray6[4] = d.y(); out_ray6[5] = d.z();
    return alive;
}
void ref_tex2d_cpu(const float* rgba, int w, int h, float u, float v, float* out3) {
    CpuTexture t{const_cast<float*>(rgba), w, h};
    color c = tex2D_cpu(&t, u, v);
    out3[0] = c.x(); out3[1] = c.y(); out3[2] = c.z();
}
int ref_struct_sizes(int* out) {
    int k = 0;
    out[k++] = sizeof(vec3); out[k++] = sizeof(Ray); out[k++] = sizeof(Interval); out[k++] = sizeof(AABB);
    out[k++] = sizeof(BVHNode); out[k++] = sizeof(BVHTree); out[k++] = sizeof(SphereData); out[k++] = sizeof(PlaneData);
    out[k++] = sizeof(MaterialData); out[k++] = sizeof(HitRecord); out[k++] = sizeof(CameraData); out[k++] = sizeof(SceneData);
    return k;
}
// stbi_loadf exactly as load_texture_cpu does (src/main.cu:52-60); caller frees with ref_free_texels
float* ref_load_texels(const char* path, int* w, int* h) {
    int ch;
    return stbi_loadf(path, w, h, &ch, 4);
}
void ref_free_texels(float* p) { if (p) stbi_image_free(p); }

// ---------------------------------------------------------------- scenes
int ref_scene_from_config(const char* cfg_text, const char* texture_override, void** out) {
    auto* s = new RefScene();
    std::istringstream in(cfg_text);
    s->params = read_scene_params(in);
    s->has_params = true;
    if (texture_override) s->params.floor.texture_path = texture_override;
    if (s->params.lights.empty()) { delete s; return 1; }  // create_scene reads lights[0] (src/main.cu:365)
    create_scene(s->params, s->sd, s->sph, s->mats, s->pl, s->nodes, s->trees, false);
    *out = s;
    return 0;
}

int ref_scene_from_arrays(const void* spheres, int S, const void* planes, int P, const void* mats, int M, const float* tex_rgba,
                          int tw, int th, int tex_mat, void** out) {
    auto* s = new RefScene();
    const SphereData* sp = (const SphereData*)spheres;
    const PlaneData* pp = (const PlaneData*)planes;
    s->sph.assign(sp, sp + S);
    s->pl.assign(pp, pp + P);
    s->mats.resize(M);
    std::memcpy((void*)s->mats.data(), mats, sizeof(MaterialData) * (size_t)M);
    for (auto& m : s->mats) { m.tex_obj = 0; m.cpu_tex = nullptr; }
    if (tex_rgba && tex_mat >= 0 && tex_mat < M) {
        s->tex_copy.assign(tex_rgba, tex_rgba + (size_t)tw * th * 4);
        s->own_tex = new CpuTexture{s->tex_copy.data(), tw, th};
        s->mats[tex_mat].cpu_tex = s->own_tex;
        s->tex_mat = tex_mat;
    }
    // host wiring exactly as create_scene's !use_gpu branch (src/main.cu:479-496)
    s->sd.d_spheres = s->sph.data();
    s->sd.d_planes = s->pl.data();
    s->sd.d_materials = s->mats.data();
    s->sd.num_spheres = S;
    s->sd.num_planes = P;
    s->sd.num_materials = M;
    s->nodes = build_bvh(s->sph, s->pl);
    BVHTree tree;
    tree.nodes = s->nodes.data();
    tree.num_nodes = (int)s->nodes.size();
    s->trees.push_back(tree);
    s->sd.d_bvh_trees = s->trees.data();
    s->sd.num_bvh_trees = 1;
    *out = s;
    return 0;
}

void ref_scene_counts(void* h, int* out6) {
    auto* s = (RefScene*)h;
    out6[0] = (int)s->sph.size();
    out6[1] = (int)s->pl.size();
    out6[2] = (int)s->mats.size();
    out6[3] = (int)s->nodes.size();
    out6[4] = bvh_depth(s->nodes, 0);
    int tex = -1;
    for (size_t i = 0; i < s->mats.size(); ++i)
        if (s->mats[i].cpu_tex) tex = (int)i;
    out6[5] = tex;
}

void ref_scene_copy(void* h, void* spheres, void* planes, void* mats, void* nodes) {
    auto* s = (RefScene*)h;
    if (spheres) std::memcpy(spheres, (const void*)s->sph.data(), s->sph.size() * sizeof(SphereData));
    if (planes) std::memcpy(planes, (const void*)s->pl.data(), s->pl.size() * sizeof(PlaneData));
    if (mats) {
        std::memcpy(mats, (const void*)s->mats.data(), s->mats.size() * sizeof(MaterialData));
        // pointers are meaningless outside this process image: zero them in the copy
        for (size_t i = 0; i < s->mats.size(); ++i) std::memset((char*)mats + i * sizeof(MaterialData) + 48, 0, 16);
    }
    if (nodes) std::memcpy(nodes, (const void*)s->nodes.data(), s->nodes.size() * sizeof(BVHNode));
}

// texture of the scene (if any): returns pointer to RGBA float texels owned by the scene
const float* ref_scene_texture(void* h, int* w, int* hgt) {
    auto* s = (RefScene*)h;
    for (auto& m : s->mats)
        if (m.cpu_tex) { *w = m.cpu_tex->width; *hgt = m.cpu_tex->height; return m.cpu_tex->data; }
    *w = *hgt = 0;
    return nullptr;
}

void ref_scene_params_render(void* h, int* out) {  // W,H,max_depth,sqrt_spp,num_frames ; fov via float slot
    auto* s = (RefScene*)h;
    out[0] = s->params.width; out[1] = s->params.height; out[2] = s->params.render.max_depth;
    out[3] = s->params.render.sqrt_rays_per_pixel; out[4] = s->params.num_frames;
    float f = s->params.fov_degrees;
    std::memcpy(&out[5], &f, 4);
}

void ref_scene_destroy(void* h);

// CameraData exactly as cpu_render/gpu_render derive it per frame (src/camera.cu:297-324)
void ref_build_camera(int H, int W, const float* from3, const float* at3, float vfov, int spp, int max_depth, void* out76) {
    Camera cam(H, W, nullptr, point3(from3[0], from3[1], from3[2]), point3(at3[0], at3[1], at3[2]));
    cam.vfov = vfov;
    cam.samplesPerPixel = spp;
    cam.maxDepth = max_depth;
    cam.background_color = color(0, 0, 0);
    CameraData d = cam.build_camera_data();
    std::memcpy(out76, &d, sizeof(CameraData));
}

// whole reference CPU program path for a config: cpu_render writes its frames to params.output_path
int ref_cpu_render_config(void* h) {
    auto* s = (RefScene*)h;
    if (!s->has_params) return 1;
    cpu_render(s->params, s->sd);
    return 0;
}

// ---------------------------------------------------------------- closest hit with ids
// mode 0: brute force (spheres then planes, include/scene.h:37-51 order); mode 1: id-recording BVH walk.
// returns number of rays where the id-recording walk disagreed with the real hit_bvh (must be 0).
int ref_trace(void* h, const float* rays6, int N, int mode, float tmin, float tmax, int* type, int* index, float* rec11) {
    auto* s = (RefScene*)h;
    std::atomic<int> bad{0};
    int nth = (int)std::thread::hardware_concurrency();
    if (nth < 1) nth = 1;
    if (N < 4096) nth = 1;
    auto work = [&](int t0) {
        for (int k = t0; k < N; k += nth) {
            Ray r = make_ray(rays6 + 6 * (size_t)k);
            HitRecord best{};
            int bt = -1, bi = -1;
            float closest = tmax;
            if (mode == 0) {
                HitRecord tmp;
                for (int i = 0; i < (int)s->sph.size(); ++i)
                    if (hit_sphere(r, Interval(tmin, closest), tmp, s->sph[i])) { closest = tmp.t; best = tmp; bt = 0; bi = i; }
                for (int i = 0; i < (int)s->pl.size(); ++i)
                    if (hit_plane(r, Interval(tmin, closest), tmp, s->pl[i])) { closest = tmp.t; best = tmp; bt = 1; bi = i; }
            } else {
                // same visiting discipline as hit_bvh (include/bvh.h:28-62) with (type,right) recorded at leaves;
                // child order: left first (the reference's own order is undefined).
                const BVHNode* nodes = s->nodes.data();
                int nn = (int)s->nodes.size();
                int stack[64];
                int sp = 0;
                if (nn > 0) stack[sp++] = 0;
                while (sp > 0) {
                    int ni = stack[--sp];
                    const BVHNode& node = nodes[ni];
                    if (!node.box.hit(r, Interval(tmin, closest))) continue;
                    if (node.left < 0) {
                        HitRecord tmp;
                        bool hh = false;
                        if (node.type == 0) hh = hit_sphere(r, Interval(tmin, closest), tmp, s->sph[node.right]);
                        else if (node.type == 1) hh = hit_plane(r, Interval(tmin, closest), tmp, s->pl[node.right]);
                        if (hh) { closest = tmp.t; best = tmp; bt = node.type; bi = node.right; }
                    } else {
                        stack[sp++] = node.right;
                        stack[sp++] = node.left;
                    }
                }
                HitRecord chk{};
                bool hh = hit_bvh(r, Interval(tmin, tmax), chk, s->nodes.data(), nn, s->sph.data(), s->pl.data());
                if (hh != (bt >= 0) || (hh && chk.t != best.t)) bad++;
            }
            type[k] = bt;
            index[k] = bi;
            if (rec11) {
                if (bt >= 0) put_rec(best, rec11 + 11 * (size_t)k);
                else std::memset(rec11 + 11 * (size_t)k, 0, 44);
            }
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nth; ++t) th.emplace_back(work, t);
    work(0);
    for (auto& t : th) t.join();
    return bad.load();
}

// ---------------------------------------------------------------- CPU frame
// Pixels (i,j) with i%xstride==x0 && j%ystride==y0 are rendered exactly as Camera::render_cpu
// does (src/camera.cu:36-50): base=wang_hash(i*W+j), seed=wang_hash(base+s), get_ray,
// ray_color_host, sum.  fb is W*H*3 floats (untouched pixels left as they are).  Rows are
// distributed over `nthreads` std::threads (the reference itself is single-threaded).
// counted!=0 uses the segment-counting transcription and returns path segments in *segments.
int ref_render_cpu(void* h, const void* cam76, int nthreads, int x0, int xstride, int y0, int ystride, float* fb,
                   unsigned long long* segments, int counted) {
    auto* s = (RefScene*)h;
    CameraData cam;
    std::memcpy(&cam, cam76, sizeof(CameraData));
    if (nthreads < 1) nthreads = 1;
    std::atomic<int> next_row{0};
    std::atomic<unsigned long long> segs{0};
    int W = cam.image_width, H = cam.image_height;
    auto work = [&]() {
        unsigned long long local = 0;
        for (;;) {
            int j = next_row.fetch_add(1);
            if (j >= H) break;
            if (j % ystride != y0) continue;
            for (int i = x0; i < W; i += xstride) {
                color px(0, 0, 0);
                unsigned int base = wang_hash((unsigned int)((size_t)i * cam.image_width + j));
                for (int sidx = 0; sidx < cam.samples_per_pixel; ++sidx) {
                    unsigned int seed = wang_hash(base + sidx);
                    Ray r = cam.get_ray(i, j, seed);
                    if (counted) px += ray_color_counted(r, seed, s->sd, cam, local);
                    else px += ray_color_host(r, seed, s->sd, cam);
                }
                float* o = fb + 3 * ((size_t)j * W + i);
                o[0] = px.x(); o[1] = px.y(); o[2] = px.z();
            }
        }
        segs += local;
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nthreads; ++t) th.emplace_back(work);
    work();
    for (auto& t : th) t.join();
    if (segments) *segments = segs.load();
    return 0;
}

// The unmodified single-threaded Camera::render_cpu over the whole frame.
int ref_render_cpu_unmodified(void* h, int H, int W, const float* from3, const float* at3, float vfov, int spp, int max_depth,
                              float* fb) {
    auto* s = (RefScene*)h;
    Camera cam(H, W, nullptr, point3(from3[0], from3[1], from3[2]), point3(at3[0], at3[1], at3[2]));
    cam.vfov = vfov;
    cam.samplesPerPixel = spp;
    cam.maxDepth = max_depth;
    cam.background_color = color(0, 0, 0);
    std::vector<color> out((size_t)W * H);
    cam.render_cpu(s->sd, out);
    std::memcpy(fb, (const void*)out.data(), out.size() * sizeof(color));
    return 0;
}

// ---------------------------------------------------------------- reference CUDA kernel (GPU box only)
static int ref_gpu_prepare(RefScene* s) {
    if (s->gpu_ready) return 0;
    // upload as create_scene's use_gpu branch does (src/main.cu:428-478)
    SceneData d{};
    std::vector<MaterialData> mats = s->mats;
    for (auto& m : mats) {
        if (m.cpu_tex) {
            // texture object with the reference's descriptor (src/main.cu:24-46): float4 array, wrap, linear, normalised
            cudaChannelFormatDesc cd = cudaCreateChannelDesc(32, 32, 32, 32, cudaChannelFormatKindFloat);
            cudaArray_t arr;
            if (cudaMallocArray(&arr, &cd, m.cpu_tex->width, m.cpu_tex->height) != cudaSuccess) return 2;
            size_t pitch = (size_t)m.cpu_tex->width * 4 * sizeof(float);
            if (cudaMemcpy2DToArray(arr, 0, 0, m.cpu_tex->data, pitch, pitch, m.cpu_tex->height, cudaMemcpyHostToDevice) != cudaSuccess)
                return 2;
            cudaResourceDesc rd;
            std::memset(&rd, 0, sizeof(rd));
            rd.resType = cudaResourceTypeArray;
            rd.res.array.array = arr;
            cudaTextureDesc td;
            std::memset(&td, 0, sizeof(td));
            td.addressMode[0] = cudaAddressModeWrap;
            td.addressMode[1] = cudaAddressModeWrap;
            td.filterMode = cudaFilterModeLinear;
            td.readMode = cudaReadModeElementType;
            td.normalizedCoords = 1;
            if (cudaCreateTextureObject(&s->d_tex, &rd, &td, NULL) != cudaSuccess) return 2;
            m.tex_obj = s->d_tex;
        }
        m.cpu_tex = nullptr;
    }
    auto up = [](const void* src, size_t bytes, void** dst) -> int {
        *dst = nullptr;
        if (bytes == 0) return 0;
        if (cudaMalloc(dst, bytes) != cudaSuccess) return 2;
        return cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice) == cudaSuccess ? 0 : 2;
    };
    if (up((const void*)s->sph.data(), s->sph.size() * sizeof(SphereData), (void**)&d.d_spheres)) return 2;
    if (up((const void*)s->pl.data(), s->pl.size() * sizeof(PlaneData), (void**)&d.d_planes)) return 2;
    if (up((const void*)mats.data(), mats.size() * sizeof(MaterialData), (void**)&d.d_materials)) return 2;
    if (up((const void*)s->nodes.data(), s->nodes.size() * sizeof(BVHNode), (void**)&s->d_nodes)) return 2;
    BVHTree tree;
    tree.nodes = s->d_nodes;
    tree.num_nodes = (int)s->nodes.size();
    if (up(&tree, sizeof(BVHTree), (void**)&d.d_bvh_trees)) return 2;
    d.num_bvh_trees = 1;
    d.num_spheres = 0;
    d.num_planes = 0;
    d.num_materials = (int)mats.size();
    s->dsd = d;
    s->gpu_ready = true;
    return 0;
}

// Launches the reference's own render_kernel (src/camera.cu:17-34, grid/block as Camera::render :201-204)
// `reps` times and returns the mean kernel-only ms (event pair around the launch) + the last framebuffer.
int ref_gpu_render(void* h, const void* cam76, int reps, float* fb, float* kernel_ms) {
    auto* s = (RefScene*)h;
    int rc = ref_gpu_prepare(s);
    if (rc) return rc;
    CameraData cam;
    std::memcpy(&cam, cam76, sizeof(CameraData));
    size_t npx = (size_t)cam.image_width * cam.image_height;
    if (s->d_fb_pixels < npx) {
        if (s->d_fb) cudaFree(s->d_fb);
        if (cudaMalloc((void**)&s->d_fb, npx * sizeof(color)) != cudaSuccess) return 2;
        s->d_fb_pixels = npx;
    }
    if (cudaMemcpyToSymbol(d_scene_data_const, &s->dsd, sizeof(SceneData)) != cudaSuccess) return 3;
    if (cudaMemcpyToSymbol(d_cam_data_const, &cam, sizeof(CameraData)) != cudaSuccess) return 3;
    dim3 threads(16, 16);
    dim3 blocks((cam.image_width + 15) / 16, (cam.image_height + 15) / 16);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float total = 0;
    for (int r = 0; r < reps; ++r) {
        cudaEventRecord(e0);
        render_kernel<<<blocks, threads>>>(s->d_fb);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) return 4;
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        total += ms;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (cudaGetLastError() != cudaSuccess) return 4;
    if (fb && cudaMemcpy(fb, s->d_fb, npx * sizeof(color), cudaMemcpyDeviceToHost) != cudaSuccess) return 5;
    if (kernel_ms) *kernel_ms = total / (reps > 0 ? reps : 1);
    return 0;
}

// The reference PROGRAM's GPU path for a config scene: its own gpu_render loop (src/camera.cu:290-349) -- per frame the camera
// of the path, Camera::render (render_kernel + cudaDeviceSynchronize + D2H + the saver's per-pixel writes, :198-215) inside
// the reference's own event pair -- writing its frames to params.output_path.  The "n\tms\ttotal_rays" lines it prints to
// std::cout are captured into `out` (the f3 timing baseline).
int ref_gpu_render_config(void* h, char* out, int cap) {
    auto* s = (RefScene*)h;
    if (!s->has_params) return 1;
    int rc = ref_gpu_prepare(s);
    if (rc) return rc;
    std::ostringstream os;
    std::streambuf* old = std::cout.rdbuf(os.rdbuf());
    gpu_render(s->params, s->dsd);
    std::cout.rdbuf(old);
    if (out && cap > 0) snprintf(out, (size_t)cap, "%s", os.str().c_str());
    return cudaGetLastError() == cudaSuccess ? 0 : 4;
}

void ref_scene_destroy(void* h) {
    auto* s = (RefScene*)h;
    if (!s) return;
    if (s->gpu_ready) {
        cudaFree(s->dsd.d_spheres);
        cudaFree(s->dsd.d_planes);
        cudaFree(s->dsd.d_materials);
        cudaFree(s->dsd.d_bvh_trees);
        cudaFree(s->d_nodes);
        if (s->d_tex) destroy_texture_resources(s->d_tex);
        if (s->d_fb) cudaFree(s->d_fb);
    }
    for (auto& m : s->mats) {
        if (m.cpu_tex && m.cpu_tex != s->own_tex) {
            stbi_image_free(m.cpu_tex->data);
            delete m.cpu_tex;
        }
        m.cpu_tex = nullptr;
    }
    delete s->own_tex;
    delete s;
}

// proves the counting transcription equals the unmodified ray_color_host on n random samples of a frame
int ref_selfcheck_color(void* h, const void* cam76, int n) {
    auto* s = (RefScene*)h;
    CameraData cam;
    std::memcpy(&cam, cam76, sizeof(CameraData));
    int bad = 0;
    unsigned int pick = 12345u;
    for (int k = 0; k < n; ++k) {
        pick = wang_hash(pick);
        int i = pick % cam.image_width;
        pick = wang_hash(pick);
        int j = pick % cam.image_height;
        unsigned int s1 = wang_hash(wang_hash(i * cam.image_width + j) + k), s2 = s1;
        Ray r1 = cam.get_ray(i, j, s1), r2 = cam.get_ray(i, j, s2);
        unsigned long long segs = 0;
        color a = ray_color_host(r1, s1, s->sd, cam);
        color b = ray_color_counted(r2, s2, s->sd, cam, segs);
        if (std::memcmp(&a, &b, sizeof(color)) != 0 || s1 != s2) bad++;
    }
    return bad;
}

}  // extern "C"
