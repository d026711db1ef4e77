"""Pins the CPU oracle (oracle/oracle.cpp) to the reference: SURVEY.md section 8c known answers, golden fixtures produced
by the unmodified reference (tests/golden/make_golden.py), and -- where oracle/_ref exists -- the reference itself."""
import numpy as np
import pytest
from conftest import bits, golden


def test_survey_known_answers(orc):
    # SURVEY.md section 8c
    assert [orc.wang_hash(x) for x in (0, 1, 2, 12345, 0xFFFFFFFF)] == [3232319850, 663891101, 3329832309, 232713235, 1895078355]
    vals, seed = orc.random_floats(1, 4)
    assert np.allclose(vals, [0.154574186, 0.404735804, 0.186604708, 0.746444583], rtol=0, atol=1e-9)
    assert seed == 3205955024
    assert orc.wang_hash(0) == 3232319850 and orc.wang_hash(3232319850 + 0) == 3075307816  # pixel (0,0) W=200, sample 0
    h, rec = orc.hit_sphere([0, 0, 0, 0, 0, -1], 0.001, 1e30, [0, 0, -5, 1])
    assert h and rec[0] == 4 and tuple(rec[1:4]) == (0, 0, 1) and rec[6] == 1
    assert abs(rec[4] - 0.249999985) < 1e-8 and rec[5] == 0.5
    h, rec, der = orc.hit_plane([.25, .5, 0, 0, 0, -1], 0.001, 1e30, [-1, -1, -3, 2, 0, 0, 0, 2, 0], 0)
    assert h and rec[0] == 3 and rec[4] == 0.625 and rec[5] == 0.75 and der[0] == -3 and tuple(der[1:4]) == (0, 0, .25)
    assert orc.aabb_hit([-1, -1, -6], [1, 1, -4], [.25, .5, 0, 0, 0, -1], 0.001, 1e30)
    assert not orc.aabb_hit([-1, -1, -6], [1, 1, -4], [2, 0, 0, 0, 0, -1], 0.001, 1e30)
    assert abs(orc.reflectance(0.5, 1 / 1.5) - 0.0699999928) < 1e-9


def test_rng_and_sampling_vs_reference_fixture(orc):
    k = golden("kat.npz")
    for i, s in enumerate(k["rng_seeds"]):
        s = int(s)
        assert orc.wang_hash(s) == int(k["wang"][i])
        assert (bits(orc.random_floats(s, 4)[0]) == bits(k["floats"][i])).all()
        b, s2 = orc.random_in_unit_sphere(s)
        assert s2 == int(k["ball_seed_after"][i]) and (bits(b) == bits(k["ball"][i])).all()
        hm, s3 = orc.random_in_hemisphere(k["hemi_n"][i], s)
        assert s3 == int(k["hemi_seed_after"][i]) and (bits(hm) == bits(k["hemi"][i])).all()
    for (c, r), want in zip(k["refl_in"], k["refl_out"]):
        assert bits(orc.reflectance(float(c), float(r))) == bits(want)


def test_leaf_tests_vs_reference_fixture(orc):
    k = golden("kat.npz")
    for ray, cr, hit, rec in zip(k["sphere_ray"], k["sphere_cr"], k["sphere_hit"], k["sphere_rec"]):
        h, r = orc.hit_sphere(ray, 0.001, 1e30, cr)
        assert h == bool(hit) and (not h or (bits(r) == bits(rec)).all())
    for ray, buv, t, hit, rec, der in zip(k["plane_ray"], k["plane_buv"], k["plane_type"], k["plane_hit"], k["plane_rec"],
                                          k["plane_derived"]):
        h, r, d = orc.hit_plane(ray, 0.001, 1e30, buv, int(t))
        assert h == bool(hit) and (bits(d) == bits(der)).all() and (not h or (bits(r) == bits(rec)).all())
    for lo, hi, ray, hit in zip(k["box_lo"], k["box_hi"], k["box_ray"], k["box_hit"]):
        assert orc.aabb_hit(lo, hi, ray, 0.001, 50.0) == bool(hit)
    assert k["sphere_hit"].sum() > 50 and k["plane_hit"].sum() > 50 and 0 < k["box_hit"].sum() < len(k["box_hit"])


def test_scatter_camera_texture_vs_reference_fixture(orc):
    k = golden("kat.npz")
    alive_count = 0
    for i in range(len(k["sc_ray"])):
        a = orc.material_scatter(k["sc_ray"][i], k["sc_p"][i], k["sc_n"][i], float(k["sc_t"][i]), int(k["sc_front"][i]),
                                 k["sc_mat"][i:i + 1], int(k["sc_seed"][i]))
        assert a[0] == bool(k["sc_alive"][i]) and a[3] == int(k["sc_seed_after"][i])
        assert (bits(a[1]) == bits(k["sc_att"][i])).all()
        if a[0]:
            alive_count += 1
            assert (bits(a[2]) == bits(k["sc_out"][i])).all()
    assert alive_count > 100
    for inp, want in zip(k["cam_in"], k["cam_out"]):
        cam = orc.build_camera(int(inp[8]), int(inp[7]), inp[0:3], inp[3:6], float(np.float32(inp[6])), 16, 8)
        assert cam.tobytes() == want.tobytes()
    for (u, v), want in zip(k["tex_uv"], k["tex_out"]):
        assert (bits(orc.tex2d(k["tex"], float(u), float(v), 0)) == bits(want)).all()


def test_layouts_match_reference(orc):
    from oracle import oracle_py as O
    k = golden("kat.npz")
    # vec3 Ray Interval AABB BVHNode BVHTree SphereData PlaneData MaterialData HitRecord CameraData SceneData
    assert list(k["struct_sizes"]) == [12, 24, 8, 24, 36, 16, 32, 80, 64, 44, 76, 64]
    assert O.SPHERE.itemsize == 32 and O.PLANE.itemsize == 80 and O.MATERIAL.itemsize == 64 and O.BVHNODE.itemsize == 36
    assert O.CAMERA.itemsize == 76


def test_reference_bvh_and_frame_vs_fixture(orc):
    s = golden("shipped_scene.npz")
    sc = orc.scene(s["spheres"], s["planes"], s["materials"])
    c = sc.counts()
    assert (c["spheres"], c["planes"], c["materials"], c["nodes"], c["depth"]) == (94, 105, 12, 397, 9)  # SURVEY 8c scene KAT
    assert sc.nodes().tobytes() == s["nodes"].tobytes()
    g = golden("test_config_frame.npz")
    sc = orc.scene(g["spheres"], g["planes"], g["materials"])
    fb, segs = sc.render(g["camera"], hit_mode=1)
    assert segs == int(g["segments"][0])
    assert fb.tobytes() == g["fb_sum"].tobytes()
    from rtp_b200 import pixel_center_rays
    rays = pixel_center_rays(g["camera"])
    for mode, tag in ((0, "brute"), (1, "bvh")):
        ty, ix, rec = sc.trace(rays, mode)
        assert (ty == g["type_" + tag]).all() and (ix == g["index_" + tag]).all()
        assert (bits(rec[:, 0]) == bits(g["t_" + tag])).all()


def test_textured_synthetic_frame_vs_fixture(orc, rt):
    g = golden("synthetic2k_textured.npz")
    hs = rt.synthetic_scene(2000, W=96, H=54, sqrt_spp=2, depth=8, seed=1234, texture="")
    sp, pl, mt, _ = hs.arrays()
    sc = orc.scene(sp, pl, mt, g["tex"], 0)
    fb, segs = sc.render(g["camera"], hit_mode=1, tex_mode=0)
    assert segs == int(g["segments"][0]) and fb.tobytes() == g["fb_sum"].tobytes()
    rays = rt.pixel_center_rays(g["trace_camera"])
    ty, ix, rec = sc.trace(rays, 0)
    assert (ty == g["type_brute"]).all() and (ix == g["index_brute"]).all() and (bits(rec[:, 0]) == bits(g["t_brute"])).all()
    ty, ix, rec = sc.trace(rays, 1)
    assert (ty == g["type_bvh"]).all() and (ix == g["index_bvh"]).all()


def test_textured_spheres_frame_vs_fixture(orc):
    """get_sphere_uv -> tex2D_cpu (include/sphere.h:16-22, include/materials.h:20-51): the fixture's textured material sits on
    spheres (tests/golden/make_golden_r02.py); the restatement reproduces the reference's frame and the (u, v) of its hits"""
    g = golden("textured_spheres.npz")
    sc = orc.scene(g["spheres"], g["planes"], g["materials"], g["tex"], int(g["tex_mat"][0]))
    fb, segs = sc.render(g["camera"], hit_mode=1, tex_mode=0)
    assert segs == int(g["segments"][0]) and fb.tobytes() == g["fb_sum"].tobytes()
    import rtp_b200 as rt
    ty, ix, rec = sc.trace(rt.pixel_center_rays(g["trace_camera"]), 0)
    assert (ty == g["type_brute"]).all() and (ix == g["index_brute"]).all()
    assert (bits(rec) == bits(g["rec_brute"])).all()   # t, normal, u, v, front, material, point


def test_oracle_vs_live_reference(orc, ref):
    """where the compiled reference is present: fresh random comparison, bit for bit"""
    rng = np.random.default_rng(7)
    for _ in range(300):
        o = rng.uniform(-15, 15, 3); c = rng.uniform(-15, 15, 3); r = rng.uniform(0.01, 3)
        d = (c + rng.normal(size=3) * r * 0.7 - o) * rng.uniform(0.1, 2)
        ray = np.concatenate([o, d]); cr = np.concatenate([c, [r]])
        a, b = orc.hit_sphere(ray, 0.001, 1e30, cr), ref.hit_sphere(ray, 0.001, 1e30, cr)
        assert a[0] == b[0] and (bits(a[1]) == bits(b[1])).all()
    cfg = open("/root/reference/config.txt").read().replace("../floor2.jpg", "/root/reference/floor.jpg")
    rs = ref.scene_from_config(cfg)
    sp, pl, mt, nd = rs.arrays()
    tex = rs.texture()
    sc = orc.scene(sp, pl, mt, tex, 0)
    assert sc.nodes().tobytes() == nd.tobytes()
    cam = ref.build_camera(54, 96, [-15, 0, 4.5], [0, 0, 4.5], 50.0, 4, 10)
    fa, sa = sc.render(cam, hit_mode=1, tex_mode=0)
    fb, sb = rs.render(cam)
    assert sa == sb and fa.tobytes() == fb.tobytes()
    assert ref.lib.ref_selfcheck_color(rs.h, cam.ctypes.data, 500) == 0
