// scene_host.cpp -- config parser, scene assembly, camera path (see scene_host.h for the reference map).
// Arithmetic follows the reference host code operation by operation (float ops, vec3 division through a double
// reciprocal, include/vec3.h:97) because the arrays built here are compared bit for bit with the reference's
// create_scene output (tests/golden/shipped_scene.npz).
#include "scene_host.h"

#include <cmath>
#include <cstring>
#include <random>
#include <sstream>

namespace rtb {
namespace {

using V = Vec3f;
inline V mk(float x, float y, float z) { return V{{x, y, z}}; }
inline V operator+(const V& a, const V& b) { return mk(a.e[0] + b.e[0], a.e[1] + b.e[1], a.e[2] + b.e[2]); }
inline V operator-(const V& a, const V& b) { return mk(a.e[0] - b.e[0], a.e[1] - b.e[1], a.e[2] - b.e[2]); }
inline V operator-(const V& a) { return mk(-a.e[0], -a.e[1], -a.e[2]); }
inline V operator*(float t, const V& v) { return mk(t * v.e[0], t * v.e[1], t * v.e[2]); }
inline V operator*(const V& v, float t) { return t * v; }
inline V operator/(const V& v, float t) { return (float)(1.0 / (double)t) * v; }
inline float dot(const V& a, const V& b) { return a.e[0] * b.e[0] + a.e[1] * b.e[1] + a.e[2] * b.e[2]; }
inline V cross(const V& a, const V& b) {
    return mk(a.e[1] * b.e[2] - a.e[2] * b.e[1], a.e[2] * b.e[0] - a.e[0] * b.e[2], a.e[0] * b.e[1] - a.e[1] * b.e[0]);
}
inline float length(const V& a) { return sqrtf(dot(a, a)); }
inline V unit(const V& a) { return a / length(a); }
inline void put(float* d, const V& v) { d[0] = v.e[0]; d[1] = v.e[1]; d[2] = v.e[2]; }

MatIn blank_material(int type) {
    MatIn m;
    std::memset(&m, 0, sizeof m);
    m.type = type;
    return m;
}

// One glass solid = faces + a thin metal strip along every edge + `lights_per_edge` small emissive spheres on each
// strip (add_octahedron / add_cube / add_dodecahedron, src/main.cu:62-308).  The three generators differ only in
// their vertex / face / edge tables, the face-plane kind and the emission order, which `Solid` captures.
struct Solid {
    const float (*verts)[3];
    int n_verts;
    float face_distance_factor;  // distance centre->face as a multiple of r
    bool zero_scale_when_small;  // cube: light_scale=0 if the face is closer than the light radius (src/main.cu:79-83)
};

struct SolidBuilder {
    HostScene& sc;
    V center;
    float r;
    int body_mat, border_mat, light_mat, lights_per_edge;
    std::vector<V> vtx, vtx_light;
    float light_radius;

    void place(const Solid& s, float dist_to_face) {
        light_radius = r / 100 * 2;
        float light_scale = 1.0f;
        if (dist_to_face > light_radius) light_scale = (dist_to_face - light_radius) / dist_to_face;
        else if (s.zero_scale_when_small) light_scale = 0.0f;
        vtx.resize(s.n_verts);
        vtx_light.resize(s.n_verts);
        for (int i = 0; i < s.n_verts; ++i) {
            V dir = unit(mk(s.verts[i][0], s.verts[i][1], s.verts[i][2]));
            vtx[i] = center + dir * r;
            vtx_light[i] = center + dir * (r * light_scale);
        }
    }
    void face(const V& a, const V& b, const V& c, int type) { sc.planes.push_back(make_plane(a, b - a, c - a, body_mat, type)); }
    void edge(int i0, int i1) {
        V start = vtx_light[i0], end = vtx_light[i1];
        V edge_vec = end - start;
        V mid = (start + end) * 0.5f;
        V radial = unit(mid - center);
        V tangent = unit(cross(edge_vec, radial));
        float width = r * 0.05f;
        V base = start - tangent * (width * 0.5f);
        sc.planes.push_back(make_plane(base, edge_vec, tangent * width, border_mat, 0));
        for (int k = 0; k < lights_per_edge; ++k) {
            float t = (k + 0.5f) / lights_per_edge;
            V pos = (1.0f - t) * start + t * end;
            SphereIn s;
            std::memset(&s, 0, sizeof s);
            put(s.c, pos);
            s.r = light_radius;
            s.mat = light_mat;
            sc.spheres.push_back(s);
        }
    }
};

const float kOctaVerts[6][3] = {{0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}, {1, 0, 0}, {-1, 0, 0}};
const int kOctaFaces[8][3] = {{0, 2, 4}, {0, 4, 3}, {0, 3, 5}, {0, 5, 2}, {1, 4, 2}, {1, 3, 4}, {1, 5, 3}, {1, 2, 5}};
const int kOctaEdges[12][2] = {{0, 2}, {0, 4}, {0, 3}, {0, 5}, {1, 2}, {1, 4}, {1, 3}, {1, 5}, {2, 4}, {4, 3}, {3, 5}, {5, 2}};

const float kCubeVerts[8][3] = {{-1, -1, -1}, {1, -1, -1}, {1, 1, -1}, {-1, 1, -1}, {-1, -1, 1}, {1, -1, 1}, {1, 1, 1}, {-1, 1, 1}};
const int kCubeEdges[12][2] = {{0, 1}, {1, 5}, {5, 4}, {4, 0}, {3, 2}, {2, 6}, {6, 7}, {7, 3}, {0, 3}, {1, 2}, {5, 6}, {4, 7}};
const int kCubeFaces[6][4] = {{4, 5, 6, 7}, {1, 0, 3, 2}, {5, 1, 2, 6}, {4, 7, 3, 0}, {7, 6, 2, 3}, {0, 1, 5, 4}};

constexpr float kPhi = 1.61803398875f;
constexpr float kInvPhi = 1.0f / kPhi;
const float kDodeVerts[20][3] = {{1, 1, 1},          {1, 1, -1},          {1, -1, 1},          {1, -1, -1},         {-1, 1, 1},
                                 {-1, 1, -1},        {-1, -1, 1},         {-1, -1, -1},        {0, kPhi, kInvPhi},  {0, kPhi, -kInvPhi},
                                 {0, -kPhi, kInvPhi}, {0, -kPhi, -kInvPhi}, {kInvPhi, 0, kPhi},  {kInvPhi, 0, -kPhi}, {-kInvPhi, 0, kPhi},
                                 {-kInvPhi, 0, -kPhi}, {kPhi, kInvPhi, 0},  {kPhi, -kInvPhi, 0}, {-kPhi, kInvPhi, 0}, {-kPhi, -kInvPhi, 0}};
const int kDodeFaces[12][5] = {{12, 2, 17, 16, 0}, {8, 4, 14, 12, 0},  {16, 1, 9, 8, 0},   {17, 3, 13, 1, 16}, {13, 15, 5, 9, 1},  {14, 6, 10, 2, 12},
                               {10, 11, 3, 17, 2}, {3, 11, 7, 15, 13}, {18, 19, 6, 14, 4}, {9, 5, 18, 4, 8},   {7, 11, 10, 6, 19}, {5, 15, 7, 19, 18}};

void add_octahedron(SolidBuilder& b) {  // src/main.cu:248-308: faces first, then strips
    Solid s{kOctaVerts, 6, 0.f, false};
    b.place(s, b.r * 0.57735026919f);
    for (auto& f : kOctaFaces) b.face(b.vtx[f[0]], b.vtx[f[1]], b.vtx[f[2]], 2);
    for (auto& e : kOctaEdges) b.edge(e[0], e[1]);
}
void add_cube(SolidBuilder& b) {  // src/main.cu:62-132: strips first, then quads (u = B-A, v = D-A)
    Solid s{kCubeVerts, 8, 0.f, true};
    b.place(s, b.r / std::sqrt(3.0f));
    for (auto& e : kCubeEdges) b.edge(e[0], e[1]);
    for (auto& f : kCubeFaces) b.face(b.vtx[f[0]], b.vtx[f[1]], b.vtx[f[3]], 0);
}
void add_dodecahedron(SolidBuilder& b) {  // src/main.cu:137-233: per face a 3-triangle fan, then its not-yet-seen edges
    Solid s{kDodeVerts, 20, 0.f, false};
    b.place(s, b.r * 0.79465447229f);
    bool seen[20][20];
    std::memset(seen, 0, sizeof seen);
    for (auto& f : kDodeFaces) {
        for (int k = 1; k <= 3; ++k) b.face(b.vtx[f[0]], b.vtx[f[k]], b.vtx[f[k + 1]], 2);
        for (int i = 0; i < 5; ++i) {
            int a = f[i], c = f[(i + 1) % 5];
            int lo = a < c ? a : c, hi = a < c ? c : a;
            if (seen[lo][hi]) continue;
            seen[lo][hi] = true;
            b.edge(lo, hi);
        }
    }
}

}  // namespace

// PlaneData's host constructor (include/plane.h:19-28)
PlaneIn make_plane(const Vec3f& base, const Vec3f& u, const Vec3f& v, int material_idx, int type) {
    PlaneIn p;
    std::memset(&p, 0, sizeof p);
    p.type = type;
    p.mat = material_idx;
    V n = cross(u, v);
    V normal = unit(n);
    p.D = dot(normal, base);
    put(p.w, n / dot(n, n));
    put(p.u, u);
    put(p.v, v);
    put(p.base, base);
    put(p.n, normal);
    return p;
}

SceneParams read_scene_params(std::istream& in) {
    SceneParams p;
    in >> p.num_frames >> p.output_path >> p.width >> p.height >> p.fov_degrees;
    CameraPathParams& c = p.camera_path;
    in >> c.rc0 >> c.zc0 >> c.phic0 >> c.Arc >> c.Azc >> c.wrc >> c.wzc >> c.wc >> c.prc >> c.pzc;
    in >> c.rn0 >> c.zn0 >> c.phin0 >> c.Arn >> c.Azn >> c.wrn >> c.wzn >> c.wn >> c.prn >> c.pzn;
    p.bodies.resize(3);  // fixed by the format (src/main.cu:517)
    for (BodyParams& b : p.bodies) {
        in >> b.center.e[0] >> b.center.e[1] >> b.center.e[2] >> b.col.e[0] >> b.col.e[1] >> b.col.e[2];
        in >> b.radius >> b.reflection_coeff >> b.transparency_coeff >> b.lights_on_edge;
    }
    for (V& corner : p.floor.corners) in >> corner.e[0] >> corner.e[1] >> corner.e[2];
    in >> p.floor.texture_path >> p.floor.tint.e[0] >> p.floor.tint.e[1] >> p.floor.tint.e[2] >> p.floor.reflection_coeff;
    int n_lights = 0;
    in >> n_lights;
    if (n_lights > 4) n_lights = 4;  // src/main.cu:538-540
    if (n_lights < 0) n_lights = 0;
    p.lights.resize(n_lights);
    for (LightSourceParams& l : p.lights)
        in >> l.position.e[0] >> l.position.e[1] >> l.position.e[2] >> l.col.e[0] >> l.col.e[1] >> l.col.e[2];
    in >> p.render.max_depth >> p.render.sqrt_rays_per_pixel;
    return p;
}

std::string default_config_text() {
    return "100\n"
           "/home/zloyaloha/development/ray-tracing-practice/images/render_%d.png\n"
           "1080 720 50\n"
           "15.0 4.5 3.14159    0.0 4.5    0.0 1.0 1.0    0.0 -1.57\n"
           "0.0 4.5 0.0    0.0 4.5    0.0 1.0 0.0    0.0 -1.57\n"
           "0.0 0.0 3.0     0.3 0.0 0.0     3.0     1.5     0.1     3\n"
           "4 0.0 6.0     0.0 0.3 0.0     3.0     1.2     0.1     2\n"
           "8 0.0 9.0     0.0 0.0 0.3     3.0     1     0.1     1\n"
           "-15.0 -15.0 -1.0      -15.0 15.0 -1.0       15.0 15.0 -1.0        15.0 -15.0 -1.0 ../floor2.jpg\n"
           "1.0 1.0 1.0\n"
           "0.3\n"
           "4\n"
           "-15.0 -15.0 1  10.0 10.0 10.0\n"
           "-15.0 15.0 1   10.0 10.0 10.0\n"
           "15.0 15.0 1    10.0 10.0 10.0\n"
           "15.0 -15.0 1   10.0 10.0 10.0\n"
           "50 50\n";
}

bool create_scene(const SceneParams& p, HostScene& out, bool load_texture, std::string* err) {
    if (p.lights.empty()) {  // the reference reads lights[0] unconditionally (src/main.cu:365)
        if (err) *err = "create_scene: the scene needs at least one light";
        return false;
    }
    if (p.bodies.size() != 3) {
        if (err) *err = "create_scene: exactly 3 bodies expected";
        return false;
    }
    out = HostScene();
    // [0] floor: METAL, albedo = tint, fuzz = reflection coefficient, optional texture (src/main.cu:349-361)
    MatIn floor_mat = blank_material(1);
    put(floor_mat.albedo, p.floor.tint);
    floor_mat.fuzz = p.floor.reflection_coeff;
    out.materials.push_back(floor_mat);
    out.floor_material = 0;
    if (load_texture && !p.floor.texture_path.empty()) {
        if (!load_texels(p.floor.texture_path, out.texels, out.tex_w, out.tex_h)) {
            out.texels.clear();
            out.tex_w = out.tex_h = 0;
            if (err) *err = "Failed to load texture: " + p.floor.texture_path;  // non-fatal, like src/main.cu:19-22
        }
    }
    // [1] edge lights: DIFFUSE_LIGHT, emit = lights[0].col * 0.1 (src/main.cu:363-367)
    MatIn edge_light = blank_material(3);
    put(edge_light.emit, p.lights[0].col * 0.1f);
    out.materials.push_back(edge_light);
    const int edge_light_idx = 1;

    for (size_t i = 0; i < p.bodies.size(); ++i) {
        const BodyParams& body = p.bodies[i];
        // [2+2i] glass: DIELECTRIC, ir = 1 + reflection, Beer-Lambert absorption from colour / transparency (:375-383)
        MatIn glass = blank_material(2);
        glass.ir = 1.0f + body.reflection_coeff;
        float strength = (1.0f - body.transparency_coeff) * 0.5f;
        glass.absorb[0] = strength * (1.0f - body.col.e[0]);
        glass.absorb[1] = strength * (1.0f - body.col.e[1]);
        glass.absorb[2] = strength * (1.0f - body.col.e[2]);
        out.materials.push_back(glass);
        int body_idx = (int)out.materials.size() - 1;
        // [3+2i] edge strips: METAL albedo .5 fuzz .6 (:385-410)
        MatIn border = blank_material(1);
        border.albedo[0] = border.albedo[1] = border.albedo[2] = 0.5f;
        border.fuzz = 0.6f;
        out.materials.push_back(border);
        int border_idx = (int)out.materials.size() - 1;
        SolidBuilder sb{out, body.center, body.radius, body_idx, border_idx, edge_light_idx, body.lights_on_edge, {}, {}, 0.f};
        if (i == 0) add_octahedron(sb);
        else if (i == 1) add_cube(sb);
        else add_dodecahedron(sb);
    }
    // floor quad (:413-415)
    out.planes.push_back(make_plane(p.floor.corners[0], p.floor.corners[1] - p.floor.corners[0],
                                    p.floor.corners[3] - p.floor.corners[0], 0, 0));
    // one emissive material + unit sphere per light (:417-426)
    for (const LightSourceParams& l : p.lights) {
        MatIn m = blank_material(3);
        put(m.emit, l.col);
        out.materials.push_back(m);
        SphereIn s;
        std::memset(&s, 0, sizeof s);
        put(s.c, l.position);
        s.r = 1.0f;
        s.mat = (int)out.materials.size() - 1;
        out.spheres.push_back(s);
    }
    return true;
}

// Camera::build_camera_data (src/camera.cu:171-196); vup = (0,0,1), focal length 1
CameraImage build_camera_data(int height, int width, const Vec3f& from, const Vec3f& at, float vfov, int spp, int max_depth) {
    const float pi_f = 3.1415926535897932385f;
    float theta = (float)((double)(vfov * pi_f) / 180.0);
    float h = tanf(theta / 2);
    float viewport_height = (float)(2.0 * (double)h);
    float viewport_width = viewport_height * ((float)width / height);
    V w = unit(from - at);
    V u = unit(cross(mk(0, 0, 1), w));
    V v = cross(w, u);
    V horizontal = viewport_width * u, vertical = viewport_height * v;
    CameraImage c;
    put(c.origin, from);
    V du = horizontal / (float)width, dv = (-vertical) / (float)height;
    put(c.pixel_delta_u, du);
    put(c.pixel_delta_v, dv);
    V upper_left = from - w - horizontal / 2.0f + vertical / 2.0f;
    put(c.pixel00_loc, upper_left + 0.5f * (du + dv));
    c.background[0] = c.background[1] = c.background[2] = 0.0f;  // forced black, src/camera.cu:322
    c.image_width = width;
    c.image_height = height;
    c.samples_per_pixel = spp;
    c.max_depth = max_depth;
    return c;
}

// cylindrical sine paths of camera and look-at point (src/camera.cu:301-315)
CameraImage camera_for_frame(const SceneParams& p, int n) {
    const CameraPathParams& c = p.camera_path;
    float t = (float)((double)(((float)n / p.num_frames) * 2.0f) * M_PI);
    float r_c = c.rc0 + c.Arc * sinf(c.wrc * t + c.prc);
    float z_c = c.zc0 + c.Azc * sinf(c.wzc * t + c.pzc);
    float phi_c = c.phic0 + c.wc * t;
    V from = mk(r_c * cosf(phi_c), r_c * sinf(phi_c), z_c);
    float r_n = c.rn0 + c.Arn * sinf(c.wrn * t + c.prn);
    float z_n = c.zn0 + c.Azn * sinf(c.wzn * t + c.pzn);
    float phi_n = c.phin0 + c.wn * t;
    V at = mk(r_n * cosf(phi_n), r_n * sinf(phi_n), z_n);
    int s = p.render.sqrt_rays_per_pixel;
    return build_camera_data(p.height, p.width, from, at, p.fov_degrees, s * s, p.render.max_depth);
}

// N seeded random spheres resting on the floor slab: centres uniform in
// [-15,15]^2 x [-1+r, -1+5r], radius r*U(.5,1), r = 0.45*30/sqrt(N); 64-entry palette appended to the materials:
// 38 LAMBERTIAN (albedo U^3), 13 METAL (albedo .5+.5U, fuzz .3U), 10 DIELECTRIC (ir 1.5, no absorption),
// 3 DIFFUSE_LIGHT (emit 4) -- emitters are needed because the background is black (Q10).
void make_synthetic_spheres(HostScene& scene, int n, uint32_t seed) {
    std::mt19937 gen(seed);
    auto U = [&]() { return (float)(gen() >> 8) * (1.0f / 16777216.0f); };
    int base = (int)scene.materials.size();
    for (int k = 0; k < 64; ++k) {
        MatIn m;
        if (k < 38) {
            m = blank_material(0);
            for (int a = 0; a < 3; ++a) m.albedo[a] = U();
        } else if (k < 51) {
            m = blank_material(1);
            for (int a = 0; a < 3; ++a) m.albedo[a] = 0.5f + 0.5f * U();
            m.fuzz = 0.3f * U();
        } else if (k < 61) {
            m = blank_material(2);
            m.ir = 1.5f;
        } else {
            m = blank_material(3);
            m.emit[0] = m.emit[1] = m.emit[2] = 4.0f;
        }
        scene.materials.push_back(m);
    }
    float r = 0.45f * 30.0f / sqrtf((float)(n > 0 ? n : 1));
    scene.spheres.reserve(scene.spheres.size() + n);
    for (int i = 0; i < n; ++i) {
        SphereIn s;
        std::memset(&s, 0, sizeof s);
        s.c[0] = -15.0f + 30.0f * U();
        s.c[1] = -15.0f + 30.0f * U();
        s.c[2] = (-1.0f + r) + 4.0f * r * U();
        s.r = r * (0.5f + 0.5f * U());
        s.mat = base + (int)(gen() % 64u);
        scene.spheres.push_back(s);
    }
}

void tonemap_host(const float* fb, size_t n, int divisor, unsigned char* out) {
    float inv = (float)(1.0 / (double)(float)divisor);
    for (size_t i = 0; i < n; ++i) {
        float v = sqrtf(inv * fb[i]);
        v = v < 0.0f ? 0.0f : (v > 0.999f ? 0.999f : v);
        out[i] = (unsigned char)(256 * v);
    }
}

}  // namespace rtb
