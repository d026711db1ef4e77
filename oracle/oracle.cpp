// oracle/oracle.cpp -- TEST INFRASTRUCTURE ONLY.  CPU restatement of the per-pixel render path of
// zloyaloha/ray-tracing-practice (the reference).  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load this; the product never does.
//
// Parity status: PINNED.  Every function below is checked (tests/test_oracle_pins.py) against
//   (1) the known-answer vectors SURVEY.md section 8c extracted from the reference, and
//   (2) bit-exact outputs of the UNMODIFIED reference compiled here (oracle/_ref/libref_rt.so, built by
//       oracle/Makefile from /root/reference), committed as fixtures under tests/golden/ by
//       tests/golden/make_golden.py so the pins also hold on the GPU box where /root/reference is absent.
//
// Written from the reference's behaviour, not its text: own types, own control flow; each function cites the
// reference file:line whose arithmetic (operation order, float/double mixing, rounding points) it restates.
// Compile with -ffp-contract=off: the reference host build (x86-64, no -march) has no FMA contraction.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

namespace orc {

struct V3 {
    float x, y, z;
};
static inline V3 mk(float x, float y, float z) { return V3{x, y, z}; }
static inline V3 add(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline V3 sub(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline V3 mul(V3 a, V3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline V3 scale(float t, V3 v) { return mk(t * v.x, t * v.y, t * v.z); }
static inline V3 neg(V3 a) { return mk(-a.x, -a.y, -a.z); }
// include/vec3.h:99 -- (v0*w0 + v1*w1) + v2*w2, all float
static inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// include/vec3.h:101-103
static inline V3 cross(V3 a, V3 b) { return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
static inline float len2(V3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
static inline float len(V3 a) { return sqrtf(len2(a)); }
// include/vec3.h:97 -- division is a multiply by the DOUBLE reciprocal narrowed to float
static inline V3 divf(V3 v, float t) { return scale((float)(1.0 / (double)t), v); }
static inline V3 unit(V3 v) { return divf(v, len(v)); }  // include/vec3.h:104
// include/vec3.h:63
static inline V3 reflect(V3 v, V3 n) { return sub(v, scale(2 * dot(v, n), n)); }
// include/vec3.h:65-70
static inline V3 refract(V3 v, V3 n, float eta) {
    float c = fminf(dot(neg(v), n), 1.0f);
    V3 perp = scale(eta, add(v, scale(c, n)));
    float k = -sqrtf(fabsf((float)(1.0 - (double)len2(perp))));
    return add(perp, scale(k, n));
}

// ---------------------------------------------------------------- RNG (include/random_utils.h:7-42)
static inline uint32_t wang(uint32_t s) {
    s = (s ^ 61u) ^ (s >> 16);
    s *= 9u;
    s ^= s >> 4;
    s *= 0x27d4eb2du;
    s ^= s >> 15;
    return s;
}
static inline float rnd(uint32_t& s) {
    s = wang(s);
    return (float)s / 4294967296.0f;
}
static inline float rnd_pm1(uint32_t& s) { return -1.0f + (1.0f - -1.0f) * rnd(s); }
// include/random_utils.h:25-32.  The three draws are constructor ARGUMENTS; the reference host build
// (g++) evaluates them right to left, so the first draw lands in z.  Pinned by the golden fixture.
static inline V3 ball(uint32_t& s) {
    for (;;) {
        float z = rnd_pm1(s);
        float y = rnd_pm1(s);
        float x = rnd_pm1(s);
        V3 c = mk(x, y, z);
        if (len2(c) < 1.0f) return c;
    }
}
static inline V3 hemi(V3 n, uint32_t& s) {  // :36-42
    V3 d = unit(ball(s));
    return dot(d, n) > 0.0f ? d : neg(d);
}

struct Hit {
    V3 p, n;
    float t;
    bool front;
    int mat;
    float u, v;
};
static inline void face(Hit& h, V3 dir, V3 outward) {  // include/hittable_object.h:17-20
    h.front = dot(dir, outward) < 0;
    h.n = h.front ? outward : neg(outward);
}

// Reference struct images (sizes/offsets: SURVEY.md section 8 header, checked against _ref)
struct SphereRec {  // 32 B
    float c[3], r;
    int mat;
    int pad[3];
};
struct PlaneRec {  // 80 B
    int type;
    float D;
    int mat;
    float w[3], u[3], v[3], base[3], n[3];
    int pad[2];
};
struct MatRec {  // 64 B
    int type;
    float fuzz, ir;
    float absorb[3], albedo[3], emit[3];
    uint64_t tex_obj, cpu_tex;
};
struct NodeRec {  // 36 B
    float b[6];   // xmin,xmax,ymin,ymax,zmin,zmax
    int left, right, type;
};
static_assert(sizeof(SphereRec) == 32 && sizeof(PlaneRec) == 80 && sizeof(MatRec) == 64 && sizeof(NodeRec) == 36, "layout");
static inline V3 v3(const float* f) { return mk(f[0], f[1], f[2]); }

// include/sphere.h:16-22
static inline void sphere_uv(V3 p, float& u, float& v) {
    float theta = acosf(p.y);
    float phi = (float)((double)atan2f(-p.z, p.x) + M_PI);
    u = (float)((double)phi / (2 * M_PI));
    v = (float)((double)theta / M_PI);
}
// include/sphere.h:24-53
static inline bool hit_sphere(V3 o, V3 d, float tmin, float tmax, const SphereRec& s, Hit& h) {
    V3 c = v3(s.c);
    V3 oc = sub(o, c);
    float a = len2(d);
    float hb = dot(oc, d);
    float cc = len2(oc) - s.r * s.r;
    float D = hb * hb - a * cc;
    if (D < 0) return false;
    double sq = (double)sqrtf(D);
    double root = ((double)(-hb) - sq) / (double)a;
    float rf = (float)root;
    if (!(tmin <= rf && rf <= tmax)) {
        root = ((double)(-hb) + sq) / (double)a;
        rf = (float)root;
        if (!(tmin <= rf && rf <= tmax)) return false;
    }
    h.t = rf;
    h.p = add(o, scale(h.t, d));
    V3 outward = divf(sub(h.p, c), s.r);
    face(h, d, outward);
    h.mat = s.mat;
    sphere_uv(outward, h.u, h.v);
    return true;
}
// include/plane.h:19-28 (host constructor)
static inline void plane_derive(PlaneRec& p) {
    V3 u = v3(p.u), v = v3(p.v), b = v3(p.base);
    V3 n = cross(u, v);
    V3 nn = unit(n);
    p.n[0] = nn.x; p.n[1] = nn.y; p.n[2] = nn.z;
    p.D = dot(nn, b);
    V3 w = divf(n, dot(n, n));
    p.w[0] = w.x; p.w[1] = w.y; p.w[2] = w.z;
}
// include/plane.h:30-96
static inline bool hit_plane(V3 o, V3 d, float tmin, float tmax, const PlaneRec& p, Hit& h) {
    V3 n = v3(p.n);
    double denom = (double)dot(n, d);
    if ((double)fabsf((float)denom) < 1e-8) return false;
    double root = (double)(p.D - dot(n, o)) / denom;
    float rf = (float)root;
    if (!(tmin <= rf && rf <= tmax)) return false;
    V3 ip = add(o, scale(rf, d));
    V3 ph = sub(ip, v3(p.base));
    float alpha = dot(v3(p.w), cross(ph, v3(p.v)));
    float beta = dot(v3(p.w), cross(v3(p.u), ph));
    double a = alpha, b = beta;
    switch (p.type) {
        case 0:  // QUAD :30-38
            if (!(0.0f <= alpha && alpha <= 1.0f) || !(0.0f <= beta && beta <= 1.0f)) return false;
            break;
        case 1:  // ELLIPSE :40-46
            if ((double)(powf((float)(a - 0.5), 2) + powf((float)(b - 0.5), 2)) > 0.25) return false;
            break;
        case 2:  // TRIANGLE :48-55 (sum formed in double)
            if (a < 0 || b < 0 || (a + b) > 1) return false;
            break;
        default:
            break;
    }
    if (p.type >= 0 && p.type <= 2) { h.u = alpha; h.v = beta; }
    h.t = rf;
    h.p = ip;
    face(h, d, n);
    h.mat = p.mat;
    return true;
}
// include/aabb.h:42-65 -- per-axis float reciprocal, swap on negative, early out on max<=min
static inline bool hit_box(const float* b, V3 o, V3 d, float tmin, float tmax) {
    const float oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z};
    for (int a = 0; a < 3; ++a) {
        float inv = 1 / dd[a];
        float t1 = (b[2 * a] - oo[a]) * inv;
        float t2 = (b[2 * a + 1] - oo[a]) * inv;
        if (inv < 0) { float t = t1; t1 = t2; t2 = t; }
        if (t1 > tmin) tmin = t1;
        if (t2 < tmax) tmax = t2;
        if (tmax <= tmin) return false;
    }
    return true;
}

// ---------------------------------------------------------------- materials (include/materials.h:53-142)
enum { LAMBERTIAN = 0, METAL = 1, DIELECTRIC = 2, DIFFUSE_LIGHT = 3 };

static inline float schlick(float cosine, float ref_idx) {  // :64-68
    float r0 = (1 - ref_idx) / (1 + ref_idx);
    r0 = r0 * r0;
    return r0 + (1 - r0) * powf((1 - cosine), 5);
}
// :116-118 -- `exp` on a float argument resolves to the float overload in the reference host build
// (pinned by the golden scatter fixture)
static inline float exp_ref(float x) { return expf(x); }

// :70-140.  albedo is passed separately because ray_color modulates it by the texture first.
static inline bool scatter(V3 ro, V3 rd, const Hit& h, const MatRec& m, V3 albedo, uint32_t& seed, V3& att, V3& so, V3& sd) {
    switch (m.type) {
        case LAMBERTIAN: {
            V3 dir = hemi(h.n, seed);
            if (fabsf(dir.x) < 1e-8f && fabsf(dir.y) < 1e-8f && fabsf(dir.z) < 1e-8f) dir = h.n;
            so = h.p; sd = dir; att = albedo;
            return true;
        }
        case METAL: {
            if (rnd(seed) < 0.8f) {
                V3 refl = reflect(unit(rd), h.n);
                V3 b = ball(seed);
                so = h.p;
                sd = add(refl, scale(m.fuzz, b));
                att = albedo;
                return dot(sd, h.n) > 0;
            }
            V3 dir = hemi(h.n, seed);
            if (fabsf(dir.x) < 1e-8f && fabsf(dir.y) < 1e-8f && fabsf(dir.z) < 1e-8f) dir = h.n;
            so = h.p; sd = dir; att = albedo;
            return true;
        }
        case DIELECTRIC: {
            att = mk(1.0f, 1.0f, 1.0f);
            float ratio = h.front ? (float)(1.0 / (double)m.ir) : m.ir;
            V3 ud = unit(rd);
            float cos_t = fminf(dot(neg(ud), h.n), 1.0f);
            float sin_t = sqrtf((float)(1.0 - (double)(cos_t * cos_t)));
            bool tir = (double)(ratio * sin_t) > 1.0;
            V3 dir;
            if (tir || schlick(cos_t, ratio) > rnd(seed)) dir = reflect(ud, h.n);
            else dir = refract(ud, h.n, ratio);
            float dist = len(sub(h.p, ro));
            V3 tr = mk(exp_ref(-m.absorb[0] * dist), exp_ref(-m.absorb[1] * dist), exp_ref(-m.absorb[2] * dist));
            if (!h.front) att = mul(att, tr);
            float p = fmaxf(att.x, fmaxf(att.y, att.z));
            if (rnd(seed) > p) return false;
            att = scale((float)(1.0 / (double)p), att);
            float sgn = dot(dir, h.n) > 0 ? 1.0f : -1.0f;
            so = add(h.p, scale(sgn, scale(1e-4f, h.n)));
            sd = dir;
            return true;
        }
        default:
            return false;
    }
}

// ---------------------------------------------------------------- floor texture
// mode 0: the reference's CPU lookup tex2D_cpu (include/materials.h:20-51): wrap by -floor, V FLIPPED
//         (py=(1-v)*H), no half-texel offset, full-precision lerp.  Out-of-range x0/y0 (u or 1-v rounding to
//         exactly 1.0) index past the row/array in the reference (UB); here they wrap -- documented deviation.
// mode 1: the reference's GPU lookup (src/main.cu:37-43 descriptor; CUDA linear filtering, programming guide
//         "Texture Fetching"): wrap, normalised, NO flip, xB = u*W-0.5, weights quantised to 8 fractional bits.
static inline V3 texel(const float* t, int w, int x, int y) {
    const float* p = t + 4 * ((size_t)y * w + x);
    return mk(p[0], p[1], p[2]);
}
static inline V3 tex_lookup(const float* t, int w, int hgt, float u, float v, int mode) {
    if (mode == 0) {
        u = u - floorf(u);
        v = v - floorf(v);
        float px = u * w;
        float py = (1.0f - v) * hgt;
        int x0 = (int)px, y0 = (int)py;
        int x1 = (x0 + 1) % w, y1 = (y0 + 1) % hgt;
        float dx = px - x0, dy = py - y0;
        x0 %= w; y0 %= hgt;
        V3 c00 = texel(t, w, x0, y0), c10 = texel(t, w, x1, y0), c01 = texel(t, w, x0, y1), c11 = texel(t, w, x1, y1);
        V3 top = add(scale(1.0f - dx, c00), scale(dx, c10));
        V3 bot = add(scale(1.0f - dx, c01), scale(dx, c11));
        return add(scale(1.0f - dy, top), scale(dy, bot));
    }
    float fu = u - floorf(u), fv = v - floorf(v);
    float xb = fu * w - 0.5f, yb = fv * hgt - 0.5f;
    float fx = floorf(xb), fy = floorf(yb);
    float ax = floorf((xb - fx) * 256.0f + 0.5f) / 256.0f;
    float ay = floorf((yb - fy) * 256.0f + 0.5f) / 256.0f;
    int x0 = (((int)fx % w) + w) % w, y0 = (((int)fy % hgt) + hgt) % hgt;
    int x1 = (x0 + 1) % w, y1 = (y0 + 1) % hgt;
    V3 c00 = texel(t, w, x0, y0), c10 = texel(t, w, x1, y0), c01 = texel(t, w, x0, y1), c11 = texel(t, w, x1, y1);
    V3 r;
    r.x = (1 - ax) * (1 - ay) * c00.x + ax * (1 - ay) * c10.x + (1 - ax) * ay * c01.x + ax * ay * c11.x;
    r.y = (1 - ax) * (1 - ay) * c00.y + ax * (1 - ay) * c10.y + (1 - ax) * ay * c01.y + ax * ay * c11.y;
    r.z = (1 - ax) * (1 - ay) * c00.z + ax * (1 - ay) * c10.z + (1 - ax) * ay * c01.z + ax * ay * c11.z;
    return r;
}

// ---------------------------------------------------------------- reference BVH (include/aabb.h, bvh_builder.h)
struct Box {
    float b[6];
};
static inline void min_extent(float* lo, float* hi) {  // aabb.h:92-97 + interval.h:27-30
    if ((double)(*hi - *lo) < 0.0001) {
        float pad = 0.0001f / 2;
        *lo = *lo - pad;
        *hi = *hi + pad;
    }
}
static inline Box box_pts(V3 p, V3 q) {  // aabb.h:21-26
    Box r;
    r.b[0] = p.x <= q.x ? p.x : q.x; r.b[1] = p.x <= q.x ? q.x : p.x;
    r.b[2] = p.y <= q.y ? p.y : q.y; r.b[3] = p.y <= q.y ? q.y : p.y;
    r.b[4] = p.z <= q.z ? p.z : q.z; r.b[5] = p.z <= q.z ? q.z : p.z;
    for (int a = 0; a < 3; ++a) min_extent(&r.b[2 * a], &r.b[2 * a + 1]);
    return r;
}
static inline Box box_join(const Box& a, const Box& c) {  // aabb.h:28-33
    Box r;
    for (int k = 0; k < 3; ++k) {
        r.b[2 * k] = fminf(a.b[2 * k], c.b[2 * k]);
        r.b[2 * k + 1] = fmaxf(a.b[2 * k + 1], c.b[2 * k + 1]);
        min_extent(&r.b[2 * k], &r.b[2 * k + 1]);
    }
    return r;
}
struct Prim {
    Box box;
    int type, index;
    float c[3];
};
static Box sphere_box(const SphereRec& s) {  // bvh_builder.h:17-20
    V3 r = mk(s.r, s.r, s.r), c = v3(s.c);
    return box_pts(sub(c, r), add(c, r));
}
static Box plane_box(const PlaneRec& p) {  // bvh_builder.h:22-50
    V3 p0 = v3(p.base), p1 = add(p0, v3(p.u)), p2 = add(p0, v3(p.v)), p3 = add(add(p0, v3(p.u)), v3(p.v));
    float lo[3] = {fminf(p0.x, fminf(p1.x, p2.x)), fminf(p0.y, fminf(p1.y, p2.y)), fminf(p0.z, fminf(p1.z, p2.z))};
    float hi[3] = {fmaxf(p0.x, fmaxf(p1.x, p2.x)), fmaxf(p0.y, fmaxf(p1.y, p2.y)), fmaxf(p0.z, fmaxf(p1.z, p2.z))};
    if (p.type == 0 || p.type == 1) {
        lo[0] = fminf(lo[0], p3.x); hi[0] = fmaxf(hi[0], p3.x);
        lo[1] = fminf(lo[1], p3.y); hi[1] = fmaxf(hi[1], p3.y);
        lo[2] = fminf(lo[2], p3.z); hi[2] = fmaxf(hi[2], p3.z);
    }
    Box r = box_pts(mk(lo[0], lo[1], lo[2]), mk(hi[0], hi[1], hi[2]));
    for (int a = 0; a < 3; ++a) min_extent(&r.b[2 * a], &r.b[2 * a + 1]);  // pad(), aabb.h:35-40
    return r;
}
static int build_rec(std::vector<Prim>& pr, int s, int e, std::vector<NodeRec>& out) {  // bvh_builder.h:52-97
    int me = (int)out.size();
    out.push_back(NodeRec{});
    Box bb = pr[s].box;
    for (int i = s + 1; i < e; ++i) bb = box_join(bb, pr[i].box);
    std::memcpy(out[me].b, bb.b, sizeof(bb.b));
    if (e - s == 1) {
        out[me].left = -1;
        out[me].right = pr[s].index;
        out[me].type = pr[s].type;
        return me;
    }
    V3 c0 = v3(pr[s].c);
    Box cb = box_pts(c0, c0);
    for (int i = s + 1; i < e; ++i) {
        V3 c = v3(pr[i].c);
        cb = box_join(cb, box_pts(c, c));
    }
    int axis = 0;
    float ext = cb.b[1] - cb.b[0];
    if (cb.b[3] - cb.b[2] > ext) { axis = 1; ext = cb.b[3] - cb.b[2]; }
    if (cb.b[5] - cb.b[4] > ext) axis = 2;
    int mid = (s + e) / 2;
    std::nth_element(pr.begin() + s, pr.begin() + mid, pr.begin() + e,
                     [axis](const Prim& a, const Prim& b) { return a.c[axis] < b.c[axis]; });
    int l = build_rec(pr, s, mid, out);
    int r = build_rec(pr, mid, e, out);
    out[me].left = l;
    out[me].right = r;
    out[me].type = -1;
    return me;
}
static std::vector<NodeRec> build_ref_bvh(const std::vector<SphereRec>& sp, const std::vector<PlaneRec>& pl) {  // :99-120
    std::vector<Prim> pr;
    pr.reserve(sp.size() + pl.size());
    for (size_t i = 0; i < sp.size(); ++i) {
        Prim p;
        p.box = sphere_box(sp[i]);
        p.type = 0; p.index = (int)i;
        p.c[0] = sp[i].c[0]; p.c[1] = sp[i].c[1]; p.c[2] = sp[i].c[2];
        pr.push_back(p);
    }
    for (size_t i = 0; i < pl.size(); ++i) {
        Prim p;
        p.box = plane_box(pl[i]);
        p.type = 1; p.index = (int)i;
        V3 c = add(v3(pl[i].base), scale(0.5f, add(v3(pl[i].u), v3(pl[i].v))));
        p.c[0] = c.x; p.c[1] = c.y; p.c[2] = c.z;
        pr.push_back(p);
    }
    std::vector<NodeRec> out;
    if (!pr.empty()) build_rec(pr, 0, (int)pr.size(), out);
    return out;
}

// ---------------------------------------------------------------- scene + closest hit
struct Scene {
    std::vector<SphereRec> sph;
    std::vector<PlaneRec> pl;
    std::vector<MatRec> mats;
    std::vector<NodeRec> nodes;
    std::vector<float> tex;
    int tw = 0, th = 0, tex_mat = -1;
};

// mode 0: every primitive in array order (spheres, then planes) -- the linear path the reference keeps
//         commented out at include/scene.h:37-51; mode 1: walk of the reference-layout BVH with the visiting
//         discipline of hit_bvh (include/bvh.h:19-65), left child first.
static bool closest_hit(const Scene& sc, V3 o, V3 d, float tmin, float tmax, int mode, Hit& best, int& bt, int& bi) {
    float closest = tmax;
    bt = bi = -1;
    Hit tmp;
    if (mode == 0) {
        for (int i = 0; i < (int)sc.sph.size(); ++i)
            if (hit_sphere(o, d, tmin, closest, sc.sph[i], tmp)) { closest = tmp.t; best = tmp; bt = 0; bi = i; }
        for (int i = 0; i < (int)sc.pl.size(); ++i)
            if (hit_plane(o, d, tmin, closest, sc.pl[i], tmp)) { closest = tmp.t; best = tmp; bt = 1; bi = i; }
        return bt >= 0;
    }
    int stack[64];
    int sp = 0;
    if (!sc.nodes.empty()) stack[sp++] = 0;
    while (sp > 0) {
        const NodeRec& n = sc.nodes[stack[--sp]];
        if (!hit_box(n.b, o, d, tmin, closest)) continue;
        if (n.left < 0) {
            bool h = false;
            if (n.type == 0) h = hit_sphere(o, d, tmin, closest, sc.sph[n.right], tmp);
            else if (n.type == 1) h = hit_plane(o, d, tmin, closest, sc.pl[n.right], tmp);
            if (h) { closest = tmp.t; best = tmp; bt = n.type; bi = n.right; }
        } else if (sp + 2 <= 64) {
            stack[sp++] = n.right;
            stack[sp++] = n.left;
        }
    }
    return bt >= 0;
}

// ---------------------------------------------------------------- camera (src/camera.cu:171-196, camera.cuh:86-110)
struct Cam {  // 76-byte image of the reference's CameraData
    float origin[3], p00[3], du[3], dv[3], bg[3];
    int W, H, spp, depth;
};
static_assert(sizeof(Cam) == 76, "CameraData layout");
static inline void put3(float* o, V3 v) { o[0] = v.x; o[1] = v.y; o[2] = v.z; }

static Cam build_camera(int H, int W, V3 from, V3 at, float vfov, int spp, int depth) {
    const float pi_f = 3.1415926535897932385f;
    float theta = (float)((double)(vfov * pi_f) / 180.0);
    float h = tanf(theta / 2);  // `tan` on a float resolves to the float overload in the reference host build (pinned)
    float vh = (float)(2.0 * (double)h);
    float vw = vh * ((float)W / H);
    V3 w = unit(sub(from, at));
    V3 u = unit(cross(mk(0, 0, 1), w));
    V3 v = cross(w, u);
    V3 hor = scale(vw, u), ver = scale(vh, v);
    Cam c;
    put3(c.origin, from);
    V3 du = divf(hor, (float)W), dv = divf(neg(ver), (float)H);
    put3(c.du, du);
    put3(c.dv, dv);
    V3 ul = add(sub(sub(from, w), divf(hor, 2.0f)), divf(ver, 2.0f));
    put3(c.p00, add(ul, scale(0.5f, add(du, dv))));
    c.bg[0] = c.bg[1] = c.bg[2] = 0;
    c.W = W; c.H = H; c.spp = spp; c.depth = depth;
    return c;
}
static inline void camera_ray(const Cam& c, int i, int j, uint32_t& seed, V3& o, V3& d) {
    V3 pc = add(add(v3(c.p00), scale((float)i, v3(c.du))), scale((float)j, v3(c.dv)));
    float ox = rnd(seed) - 0.5f;
    float oy = rnd(seed) - 0.5f;
    V3 ps = add(add(pc, scale(ox, v3(c.du))), scale(oy, v3(c.dv)));
    o = v3(c.origin);
    d = sub(ps, o);
}

// ---------------------------------------------------------------- path integrator (src/camera.cu:254-288)
static V3 radiance(const Scene& sc, const Cam& cam, V3 o, V3 d, uint32_t& seed, int hit_mode, int tex_mode,
                   unsigned long long& segs) {
    V3 total = mk(0, 0, 0), beta = mk(1, 1, 1);
    for (int depth = 0; depth < cam.depth; ++depth) {
        Hit h;
        int bt, bi;
        ++segs;
        if (!closest_hit(sc, o, d, 0.001f, 1e30f, hit_mode, h, bt, bi)) {
            total = add(total, mul(beta, v3(cam.bg)));
            break;
        }
        const MatRec& m = sc.mats[h.mat];
        V3 albedo = v3(m.albedo);
        if (h.mat == sc.tex_mat && !sc.tex.empty()) {
            V3 t = tex_lookup(sc.tex.data(), sc.tw, sc.th, h.u, h.v, tex_mode);
            // CPU path: albedo * tex (camera.cu:267); GPU path: tex * albedo (camera.cu:235) -- commutative per lane
            albedo = mul(albedo, t);
        }
        total = add(total, mul(beta, v3(m.emit)));
        V3 att, so, sd;
        if (!scatter(o, d, h, m, albedo, seed, att, so, sd)) break;
        beta = mul(beta, att);
        o = so;
        d = sd;
    }
    return total;
}

}  // namespace orc

using namespace orc;
extern "C" {

unsigned int orc_wang_hash(unsigned int s) { return wang(s); }
float orc_random_float(unsigned int* s) { return rnd(*s); }
float orc_reflectance(float c, float r) { return schlick(c, r); }
void orc_random_in_unit_sphere(unsigned int* s, float* o) { put3(o, ball(*s)); }
void orc_random_in_hemisphere(const float* n, unsigned int* s, float* o) { put3(o, hemi(v3(n), *s)); }

static void put_hit(const Hit& h, float* o) {
    o[0] = h.t; o[1] = h.n.x; o[2] = h.n.y; o[3] = h.n.z; o[4] = h.u; o[5] = h.v;
    o[6] = h.front ? 1.0f : 0.0f; o[7] = (float)h.mat; o[8] = h.p.x; o[9] = h.p.y; o[10] = h.p.z;
}
int orc_hit_sphere(const float* ray6, float tmin, float tmax, const float* cr4, float* rec11) {
    SphereRec s{};
    s.c[0] = cr4[0]; s.c[1] = cr4[1]; s.c[2] = cr4[2]; s.r = cr4[3];
    Hit h{};
    bool ok = hit_sphere(v3(ray6), v3(ray6 + 3), tmin, tmax, s, h);
    if (ok) put_hit(h, rec11);
    return ok;
}
int orc_hit_plane(const float* ray6, float tmin, float tmax, const float* buv9, int type, float* rec11, float* derived7) {
    PlaneRec p{};
    p.type = type;
    std::memcpy(p.base, buv9, 12); std::memcpy(p.u, buv9 + 3, 12); std::memcpy(p.v, buv9 + 6, 12);
    plane_derive(p);
    if (derived7) { derived7[0] = p.D; std::memcpy(derived7 + 1, p.w, 12); std::memcpy(derived7 + 4, p.n, 12); }
    Hit h{};
    bool ok = hit_plane(v3(ray6), v3(ray6 + 3), tmin, tmax, p, h);
    if (ok) put_hit(h, rec11);
    return ok;
}
// fills D, w, normal of n plane records (80 B each) from base,u,v like the reference host ctor
void orc_plane_derive(void* planes80, int n) {
    PlaneRec* p = (PlaneRec*)planes80;
    for (int i = 0; i < n; ++i) plane_derive(p[i]);
}
int orc_aabb_hit(const float* lo, const float* hi, const float* ray6, float tmin, float tmax) {
    Box b = box_pts(v3(lo), v3(hi));
    return hit_box(b.b, v3(ray6), v3(ray6 + 3), tmin, tmax);
}
int orc_material_scatter(const float* ray6, const float* p3, const float* n3, float t, int front, const void* mat64,
                         unsigned int* seed, float* att3, float* out6) {
    MatRec m;
    std::memcpy(&m, mat64, 64);
    Hit h{};
    h.p = v3(p3); h.n = v3(n3); h.t = t; h.front = front != 0;
    V3 att = mk(0, 0, 0), so = mk(0, 0, 0), sd = mk(0, 0, 0);
    bool alive = scatter(v3(ray6), v3(ray6 + 3), h, m, v3(m.albedo), *seed, att, so, sd);
    put3(att3, att); put3(out6, so); put3(out6 + 3, sd);
    return alive;
}
void orc_tex2d(const float* rgba, int w, int hgt, float u, float v, int mode, float* out3) {
    put3(out3, tex_lookup(rgba, w, hgt, u, v, mode));
}
void orc_build_camera(int H, int W, const float* from3, const float* at3, float vfov, int spp, int depth, void* out76) {
    Cam c = build_camera(H, W, v3(from3), v3(at3), vfov, spp, depth);
    std::memcpy(out76, &c, 76);
}

int orc_scene_create(const void* spheres, int S, const void* planes, int P, const void* mats, int M, const float* tex, int tw,
                     int th, int tex_mat, void** out) {
    Scene* s = new Scene();
    s->sph.assign((const SphereRec*)spheres, (const SphereRec*)spheres + S);
    s->pl.assign((const PlaneRec*)planes, (const PlaneRec*)planes + P);
    s->mats.assign((const MatRec*)mats, (const MatRec*)mats + M);
    if (tex && tex_mat >= 0) {
        s->tex.assign(tex, tex + (size_t)tw * th * 4);
        s->tw = tw; s->th = th; s->tex_mat = tex_mat;
    }
    s->nodes = build_ref_bvh(s->sph, s->pl);
    *out = s;
    return 0;
}
void orc_scene_destroy(void* h) { delete (Scene*)h; }
static int depth_of(const std::vector<NodeRec>& n, int i) {
    if (n.empty()) return 0;
    if (n[i].left < 0) return 1;
    return 1 + std::max(depth_of(n, n[i].left), depth_of(n, n[i].right));
}
void orc_scene_counts(void* h, int* out6) {
    Scene* s = (Scene*)h;
    out6[0] = (int)s->sph.size(); out6[1] = (int)s->pl.size(); out6[2] = (int)s->mats.size();
    out6[3] = (int)s->nodes.size(); out6[4] = depth_of(s->nodes, 0); out6[5] = s->tex_mat;
}
void orc_scene_nodes(void* h, void* nodes36) {
    Scene* s = (Scene*)h;
    std::memcpy(nodes36, s->nodes.data(), s->nodes.size() * sizeof(NodeRec));
}

static int n_threads(int want) {
    int hw = (int)std::thread::hardware_concurrency();
    if (hw < 1) hw = 1;
    return want <= 0 ? hw : want;
}

int orc_trace(void* hs, const float* rays6, int N, int mode, float tmin, float tmax, int* type, int* index, float* rec11) {
    Scene* s = (Scene*)hs;
    int nth = N < 4096 ? 1 : n_threads(0);
    auto work = [&](int t0) {
        for (int k = t0; k < N; k += nth) {
            Hit h{};
            int bt, bi;
            bool ok = closest_hit(*s, v3(rays6 + 6 * (size_t)k), v3(rays6 + 6 * (size_t)k + 3), tmin, tmax, mode, h, bt, bi);
            type[k] = bt; index[k] = bi;
            if (rec11) {
                if (ok) put_hit(h, rec11 + 11 * (size_t)k);
                else std::memset(rec11 + 11 * (size_t)k, 0, 44);
            }
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nth; ++t) th.emplace_back(work, t);
    work(0);
    for (auto& t : th) t.join();
    return 0;
}

// Frame as Camera::render_cpu (src/camera.cu:36-50): per pixel base=wang(i*W+j) [i = x !], per sample
// seed=wang(base+s); pixel value = SUM over samples.  Only pixels with i%xs==x0 && j%ys==y0 are written.
int orc_render(void* hs, const void* cam76, int nthreads, int x0, int xs, int y0, int ys, int hit_mode, int tex_mode, float* fb,
               unsigned long long* segments) {
    Scene* s = (Scene*)hs;
    Cam cam;
    std::memcpy(&cam, cam76, 76);
    int nth = n_threads(nthreads);
    std::atomic<int> row{0};
    std::atomic<unsigned long long> total{0};
    auto work = [&]() {
        unsigned long long segs = 0;
        for (;;) {
            int j = row.fetch_add(1);
            if (j >= cam.H) break;
            if (j % ys != y0) continue;
            for (int i = x0; i < cam.W; i += xs) {
                V3 px = mk(0, 0, 0);
                uint32_t base = wang((uint32_t)((size_t)i * cam.W + j));
                for (int k = 0; k < cam.spp; ++k) {
                    uint32_t seed = wang(base + k);
                    V3 o, d;
                    camera_ray(cam, i, j, seed, o, d);
                    px = add(px, radiance(*s, cam, o, d, seed, hit_mode, tex_mode, segs));
                }
                put3(fb + 3 * ((size_t)j * cam.W + i), px);
            }
        }
        total += segs;
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nth; ++t) th.emplace_back(work);
    work();
    for (auto& t : th) t.join();
    if (segments) *segments = total.load();
    return 0;
}

}  // extern "C"
