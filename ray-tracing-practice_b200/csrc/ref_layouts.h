// ref_layouts.h -- byte images of the reference's host PODs as they cross the C-ABI (include/rt_b200.h).
// Offsets measured against the compiled reference (oracle/_ref, ref_struct_sizes) in tests/test_oracle_pins.py and tests/test_host.py (struct sizes and field offsets).
#pragma once
namespace rtb {
struct SphereIn {  // SphereData, include/sphere.h:8-14 (32 B, align 16)
    float c[3], r;
    int mat;
    int pad[3];
};
struct PlaneIn {  // PlaneData, include/plane.h:9-17 (80 B, align 16)
    int type;     // 0 QUAD, 1 ELLIPSE, 2 TRIANGLE
    float D;
    int mat;
    float w[3], u[3], v[3], base[3], n[3];
    int pad[2];
};
struct MatIn {  // MaterialData, include/materials.h:53-62 (64 B)
    int type;   // 0 LAMBERTIAN, 1 METAL, 2 DIELECTRIC, 3 DIFFUSE_LIGHT
    float fuzz, ir;
    float absorb[3], albedo[3], emit[3];
    unsigned long long tex_obj, cpu_tex;
};
static_assert(sizeof(SphereIn) == 32 && sizeof(PlaneIn) == 80 && sizeof(MatIn) == 64, "reference layouts");
}  // namespace rtb
