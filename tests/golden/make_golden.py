#!/usr/bin/env python
"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libref_rt.so, built by oracle/Makefile from
/root/reference).  Run in the build container (where /root/reference exists):

    python tests/golden/make_golden.py

The fixtures pin the oracle (tests/test_oracle_pins.py), the host mirror (tests/test_host.py) and -- on the GPU box,
where /root/reference is absent -- the CUDA path (tests/test_gpu_parity.py).  Everything here is produced by calling the
reference's own functions; nothing is computed by this repository's code except the random inputs.
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle_py as O  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
REF_DIR = "/root/reference"


def procedural_texture(w=64, h=48):
    """Small deterministic RGBA float texture (stands in for floor.jpg, which cannot travel to the GPU box)."""
    y, x = np.mgrid[0:h, 0:w].astype(np.float32)
    t = np.zeros((h, w, 4), np.float32)
    t[..., 0] = 0.15 + 0.8 * ((x // 8 + y // 8) % 2)
    t[..., 1] = 0.1 + 0.85 * x / (w - 1)
    t[..., 2] = 0.1 + 0.85 * y / (h - 1)
    t[..., 3] = 1.0
    return t


def zero_pad(a):
    """copy with struct padding zeroed so fixtures are deterministic"""
    out = np.zeros(a.shape, a.dtype)
    for n in a.dtype.names:
        out[n] = a[n]
    return out


def main():
    O.build()
    ref = O.Ref()
    rng = np.random.default_rng(20261104)

    # ---------------------------------------------------------------- 1. function-level known answers
    k = {}
    seeds = np.array([0, 1, 2, 12345, 0xFFFFFFFF] + list(rng.integers(0, 2 ** 32, 59)), np.uint64)
    k["rng_seeds"] = seeds
    k["wang"] = np.array([ref.wang_hash(int(s)) for s in seeds], np.uint64)
    k["floats"] = np.array([ref.random_floats(int(s), 4)[0] for s in seeds], np.float32)
    balls, ball_seeds, hemis, hemi_n, hemi_seeds = [], [], [], [], []
    for s in seeds:
        b, s2 = ref.random_in_unit_sphere(int(s))
        balls.append(b); ball_seeds.append(s2)
        n = rng.normal(size=3).astype(np.float32)
        hm, s3 = ref.random_in_hemisphere(n, int(s))
        hemis.append(hm); hemi_n.append(n); hemi_seeds.append(s3)
    k["ball"] = np.array(balls, np.float32); k["ball_seed_after"] = np.array(ball_seeds, np.uint64)
    k["hemi"] = np.array(hemis, np.float32); k["hemi_n"] = np.array(hemi_n, np.float32)
    k["hemi_seed_after"] = np.array(hemi_seeds, np.uint64)
    cs = rng.uniform(0, 1, 64).astype(np.float32); ri = rng.uniform(0.3, 2.5, 64).astype(np.float32)
    k["refl_in"] = np.stack([cs, ri], 1)
    k["refl_out"] = np.array([ref.reflectance(float(a), float(b)) for a, b in zip(cs, ri)], np.float32)

    N = 400
    sray, scr, shit, srec = [], [], [], []
    for i in range(N):
        o = rng.uniform(-15, 15, 3); c = rng.uniform(-15, 15, 3); r = rng.uniform(0.01, 3)
        tgt = c + rng.normal(size=3) * r * 0.7
        d = (tgt - o) * rng.uniform(0.1, 2)
        if i % 10 == 0:
            o = c + rng.normal(size=3) * r * 0.3  # origin inside: far root
        ray = np.concatenate([o, d]).astype(np.float32); cr = np.concatenate([c, [r]]).astype(np.float32)
        h, rec = ref.hit_sphere(ray, 0.001, 1e30, cr)
        sray.append(ray); scr.append(cr); shit.append(h); srec.append(rec)
    k["sphere_ray"] = np.array(sray); k["sphere_cr"] = np.array(scr); k["sphere_hit"] = np.array(shit)
    k["sphere_rec"] = np.array(srec)
    pray, pbuv, ptype, phit, prec, pder = [], [], [], [], [], []
    for i in range(N):
        base = rng.uniform(-5, 5, 3); u = rng.normal(size=3) * 2; v = rng.normal(size=3) * 2
        o = rng.uniform(-15, 15, 3); tgt = base + u * rng.uniform(-.2, 1.2) + v * rng.uniform(-.2, 1.2)
        d = (tgt - o) * rng.uniform(0.1, 2)
        ray = np.concatenate([o, d]).astype(np.float32); buv = np.concatenate([base, u, v]).astype(np.float32)
        t = i % 3
        h, rec, der = ref.hit_plane(ray, 0.001, 1e30, buv, t)
        pray.append(ray); pbuv.append(buv); ptype.append(t); phit.append(h); prec.append(rec); pder.append(der)
    k["plane_ray"] = np.array(pray); k["plane_buv"] = np.array(pbuv); k["plane_type"] = np.array(ptype, np.int32)
    k["plane_hit"] = np.array(phit); k["plane_rec"] = np.array(prec); k["plane_derived"] = np.array(pder)
    blo, bhi, bray, bhit = [], [], [], []
    for i in range(N):
        lo = rng.uniform(-5, 5, 3); hi = lo + rng.uniform(0, 3, 3) * (rng.random(3) > 0.1)
        o = rng.uniform(-15, 15, 3); d = (lo + (hi - lo) * rng.uniform(-.3, 1.3, 3) - o)
        if i % 9 == 0:
            d[int(rng.integers(0, 3))] = 0
        ray = np.concatenate([o, d]).astype(np.float32)
        blo.append(lo.astype(np.float32)); bhi.append(hi.astype(np.float32)); bray.append(ray)
        bhit.append(ref.aabb_hit(lo, hi, ray, 0.001, 50.0))
    k["box_lo"] = np.array(blo); k["box_hi"] = np.array(bhi); k["box_ray"] = np.array(bray); k["box_hit"] = np.array(bhit)
    # KATs quoted in SURVEY.md section 8c
    k["kat_sphere"] = ref.hit_sphere([0, 0, 0, 0, 0, -1], 0.001, 1e30, [0, 0, -5, 1])[1]
    hp = ref.hit_plane([.25, .5, 0, 0, 0, -1], 0.001, 1e30, [-1, -1, -3, 2, 0, 0, 0, 2, 0], 0)
    k["kat_plane_rec"] = hp[1]; k["kat_plane_derived"] = hp[2]
    mats = np.zeros(N, O.MATERIAL)
    mray, mp, mn, mt_, mfront, mseed, malive, matt, mout, mseed2 = [], [], [], [], [], [], [], [], [], []
    for i in range(N):
        m = mats[i:i + 1]
        m["type"] = i % 4; m["fuzz"] = rng.uniform(0, 1); m["ir"] = rng.uniform(1.0, 2.5)
        m["absorption"] = rng.uniform(0, 0.5, 3); m["albedo"] = rng.uniform(0, 1, 3); m["emit"] = rng.uniform(0, 4, 3)
        n = rng.normal(size=3); n /= np.linalg.norm(n); d = rng.normal(size=3) * rng.uniform(.2, 3)
        if d @ n > 0:
            n = -n
        p = rng.uniform(-5, 5, 3); t = rng.uniform(0.1, 5); o = p - d * t
        ray = np.concatenate([o, d]).astype(np.float32); front = int(rng.integers(0, 2)); s = int(rng.integers(0, 2 ** 32))
        alive, att, out, s2 = ref.material_scatter(ray, p, n, t, front, m, s)
        mray.append(ray); mp.append(p.astype(np.float32)); mn.append(n.astype(np.float32)); mt_.append(t)
        mfront.append(front); mseed.append(s); malive.append(alive); matt.append(att); mout.append(out); mseed2.append(s2)
    k["sc_mat"] = mats; k["sc_ray"] = np.array(mray); k["sc_p"] = np.array(mp); k["sc_n"] = np.array(mn)
    k["sc_t"] = np.array(mt_, np.float32); k["sc_front"] = np.array(mfront, np.int32); k["sc_seed"] = np.array(mseed, np.uint64)
    k["sc_alive"] = np.array(malive); k["sc_att"] = np.array(matt); k["sc_out"] = np.array(mout)
    k["sc_seed_after"] = np.array(mseed2, np.uint64)
    cams_in, cams_out = [], []
    for i in range(64):
        f = rng.uniform(-20, 20, 3).astype(np.float32); at = rng.uniform(-5, 5, 3).astype(np.float32)
        fov = np.float32(rng.uniform(20, 110)); W = int(rng.integers(16, 4000)); H = int(rng.integers(16, 3000))
        cams_in.append(np.concatenate([f, at, [fov, W, H]]).astype(np.float64))
        cams_out.append(ref.build_camera(H, W, f, at, float(fov), 16, 8)[0])
    k["cam_in"] = np.array(cams_in); k["cam_out"] = np.array(cams_out, O.CAMERA)
    tex = procedural_texture()
    uv = rng.uniform(-2, 2, (200, 2)).astype(np.float32)
    uv = uv[(uv[:, 1] != np.floor(uv[:, 1]))]
    k["tex"] = tex; k["tex_uv"] = uv
    k["tex_out"] = np.array([ref.tex2d(tex, float(a), float(b)) for a, b in uv], np.float32)
    k["struct_sizes"] = np.array(ref.struct_sizes(), np.int32)
    np.savez_compressed(os.path.join(OUT, "kat.npz"), **k)

    # ---------------------------------------------------------------- 2. shipped scene (config.txt, paths fixed: Q6)
    cfg = open(os.path.join(REF_DIR, "config.txt")).read().replace("../floor2.jpg", os.path.join(REF_DIR, "floor.jpg"))
    rs = ref.scene_from_config(cfg)
    sp, pl, mt, nd = rs.arrays()
    texf = rs.texture()
    s = dict(spheres=zero_pad(sp), planes=zero_pad(pl), materials=zero_pad(mt), counts=np.array(list(rs.counts().values())),
             nodes=zero_pad(nd), tex_shape=np.array(texf.shape), tex_checksum=np.array([np.float64(texf.astype(np.float64).sum())]),
             tex_corner=texf[:4, :4].copy(), tex_probe=texf[::133, ::200].copy())
    np.savez_compressed(os.path.join(OUT, "shipped_scene.npz"), **s)

    # ---------------------------------------------------------------- 3. test-config frame (create_test_config.py, untextured)
    tcfg = subprocess.run([sys.executable, os.path.join(REF_DIR, "create_test_config.py")], capture_output=True, text=True).stdout
    ts = ref.scene_from_config(tcfg)  # ../floor2.jpg is missing -> untextured, exactly like an unmodified run
    tsp, tpl, tmt, tnd = ts.arrays()
    p = ts.params()
    import rtp_b200 as rt
    hs = rt.HostScene(tcfg, texture="")
    cam = hs.camera(0)  # host mirror camera; cross-checked against the reference's own frame below
    fb_un = np.zeros((p["H"], p["W"], 3), np.float32)
    c0 = cam[0]
    # the unmodified single-threaded Camera::render_cpu, given lookfrom/lookat of the static camera
    lookfrom = np.array([15.0 * np.cos(np.float32(3.14159)), 15.0 * np.sin(np.float32(3.14159)), 4.5], np.float32)
    ref.lib.ref_render_cpu_unmodified(ts.h, p["H"], p["W"], lookfrom.ctypes.data, np.array([0, 0, 4.5], np.float32).ctypes.data,
                                      p["fov"], p["sqrt_spp"] ** 2, p["max_depth"], fb_un.ctypes.data)
    fb, segs = ts.render(cam)
    assert fb.tobytes() == fb_un.tobytes(), "host-mirror camera differs from the reference's frame-0 camera"
    rays = rt.pixel_center_rays(cam)
    ty0, ix0, rec0 = ts.trace(rays, 0)
    ty1, ix1, rec1 = ts.trace(rays, 1)
    np.savez_compressed(os.path.join(OUT, "test_config_frame.npz"), config=np.array(tcfg), spheres=zero_pad(tsp),
                        planes=zero_pad(tpl), materials=zero_pad(tmt), camera=cam, fb_sum=fb, segments=np.array([segs]),
                        type_brute=ty0.astype(np.int8), index_brute=ix0.astype(np.int16), t_brute=rec0[:, 0].copy(),
                        type_bvh=ty1.astype(np.int8), index_bvh=ix1.astype(np.int16), t_bvh=rec1[:, 0].copy())

    # ---------------------------------------------------------------- 4. the reference program end to end: 3 frames on the
    # moving camera path, written by the reference's own cpu_render + BinarySaver
    with tempfile.TemporaryDirectory() as td:
        mcfg = cfg.replace("1080 720 50", "48 32 50", 1)
        lines = mcfg.split("\n")
        lines[0] = "3"
        lines[1] = os.path.join(td, "f_%d.bin")
        lines[-2 if lines[-1] == "" else -1] = "4 2"
        mcfg = "\n".join(lines)
        ms = ref.scene_from_config(mcfg, texture_override="/nonexistent.jpg")
        assert ref.lib.ref_cpu_render_config(ms.h) == 0
        frames = [np.fromfile(os.path.join(td, "f_%d.bin" % i), np.uint8) for i in range(3)]
        mcfg_portable = mcfg.replace(td, "OUTDIR").replace(os.path.join(REF_DIR, "floor.jpg"), "/nonexistent.jpg")
    np.savez_compressed(os.path.join(OUT, "program_frames.npz"), config=np.array(mcfg_portable), frames=np.array(frames))

    # ---------------------------------------------------------------- 5. textured + synthetic-sphere scene (small)
    hs2 = rt.synthetic_scene(2000, W=96, H=54, sqrt_spp=2, depth=8, seed=1234, texture="")
    ssp, spl, smt, _ = hs2.arrays()
    cam2 = hs2.camera(0)
    rsyn = ref.scene(ssp, spl, smt, tex, 0)
    fb2, segs2 = rsyn.render(cam2)
    cam2b = rt.build_camera(90, 160, [-15, 0, 4.5], [0, 0, 4.5], 90.0, 1, 1)
    rays2 = rt.pixel_center_rays(cam2b)
    a0 = rsyn.trace(rays2, 0); a1 = rsyn.trace(rays2, 1)
    np.savez_compressed(os.path.join(OUT, "synthetic2k_textured.npz"), n_spheres=np.array([2000]), seed=np.array([1234]),
                        sphere_probe=zero_pad(ssp[::97]), material_probe=zero_pad(smt[::7]), tex=tex, camera=cam2, fb_sum=fb2,
                        segments=np.array([segs2]), trace_camera=cam2b, type_brute=a0[0].astype(np.int8), index_brute=a0[1],
                        t_brute=a0[2][:, 0].copy(), type_bvh=a1[0].astype(np.int8), index_bvh=a1[1], t_bvh=a1[2][:, 0].copy(),
                        counts=np.array(list(rsyn.counts().values())))
    for f in sorted(os.listdir(OUT)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
