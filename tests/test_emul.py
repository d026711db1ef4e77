"""CPU check of the DEVICE code's logic: the product's device headers (traverse.cuh / shade.cuh / path.cuh) compiled for
the host by tests/emul (test infrastructure, not a product path) against the oracle.  Exact-math results must be bit
identical; the GPU run of the same comparisons is tests/test_gpu_parity.py."""
import numpy as np
import pytest
from conftest import bits, golden


def test_exact_trace_matches_oracle_bit_for_bit(emul, EmuSceneCls, orc, rt):
    g = golden("test_config_frame.npz")
    es = EmuSceneCls(emul, g["spheres"], g["planes"], g["materials"])
    rays = rt.pixel_center_rays(g["camera"])
    ty, ix, t = es.trace(rays, exact=1)
    assert (ty == g["type_brute"]).all() and (ix == g["index_brute"]).all()
    hit = ty >= 0
    assert (bits(t[hit]) == bits(g["t_brute"][hit])).all()
    # fast-math traversal: same ids on this scene, t within fp32 noise
    ty2, ix2, t2 = es.trace(rays, exact=0)
    assert ((ty2 != ty) | (ix2 != ix)).mean() < 1e-3
    same = hit & (ty2 == ty) & (ix2 == ix)
    assert np.max(np.abs(t2[same] - t[same]) / t[same]) < 2e-4


def test_exact_trace_on_synthetic_spheres(emul, EmuSceneCls, orc, rt):
    g = golden("synthetic2k_textured.npz")
    hs = rt.synthetic_scene(2000, seed=1234)
    sp, pl, mt, _ = hs.arrays()
    es = EmuSceneCls(emul, sp, pl, mt)
    rays = rt.pixel_center_rays(g["trace_camera"])
    ty, ix, t = es.trace(rays, exact=1)
    mism = (ty != g["type_brute"]) | (ix != g["index_brute"])
    assert mism.sum() == 0, mism.sum()
    hit = ty >= 0
    assert hit.mean() > 0.3 and (bits(t[hit]) == bits(g["t_brute"][hit])).all()


def test_reference_stream_render_is_bit_identical(emul, EmuSceneCls, rt):
    """RT_RNG_REFERENCE + exact math + CPU texture semantics reproduces the reference host render per pixel"""
    g = golden("test_config_frame.npz")
    es = EmuSceneCls(emul, g["spheres"], g["planes"], g["materials"])
    fb, cnt = es.render(g["camera"], rng=1, exact=1, tex_mode=0)
    assert cnt[0] == int(g["segments"][0])
    assert fb.tobytes() == g["fb_sum"].tobytes()
    g = golden("synthetic2k_textured.npz")
    sp, pl, mt, _ = rt.synthetic_scene(2000, W=96, H=54, sqrt_spp=2, depth=8, seed=1234).arrays()
    es = EmuSceneCls(emul, sp, pl, mt, g["tex"], 0)
    fb, cnt = es.render(g["camera"], rng=1, exact=1, tex_mode=0)
    assert cnt[0] == int(g["segments"][0]) and fb.tobytes() == g["fb_sum"].tobytes()


def test_textured_spheres_render_is_bit_identical(emul, EmuSceneCls):
    """sphere_uv (shade.cuh) feeding the CPU-semantics texture lookup: the device code's exact policy, run on the host,
    reproduces the reference frame of a scene whose textured material sits on spheres"""
    g = golden("textured_spheres.npz")
    es = EmuSceneCls(emul, g["spheres"], g["planes"], g["materials"], g["tex"], int(g["tex_mat"][0]))
    fb, cnt = es.render(g["camera"], rng=1, exact=1, tex_mode=0)
    assert cnt[0] == int(g["segments"][0])
    same = (bits(fb) == bits(g["fb_sum"])).all(-1)
    assert same.mean() == 1.0, same.mean()


def test_philox_estimator_agrees_in_expectation(emul, EmuSceneCls, orc, rt):
    """production stream (Philox, direct sampling, fast math): same estimator => same image in expectation"""
    hs = rt.synthetic_scene(500, W=64, H=36, sqrt_spp=12, depth=8, seed=5)
    sp, pl, mt, _ = hs.arrays()
    cam = hs.camera(0)
    es = EmuSceneCls(emul, sp, pl, mt)
    fa, ca = es.render(cam, rng=0, exact=0, seed=1)
    fb_, cb = es.render(cam, rng=0, exact=0, seed=2)
    fo, so = orc.scene(sp, pl, mt).render(cam, hit_mode=1)
    spp = 144
    ma, mb, mo = fa.mean() / spp, fb_.mean() / spp, fo.mean() / spp
    assert abs(ma - mo) / mo < 0.05 and abs(mb - mo) / mo < 0.05
    assert abs(ca[0] - so) / so < 0.02  # path-length distribution agrees
    assert fa.tobytes() != fb_.tobytes()
    again, _ = es.render(cam, rng=0, exact=0, seed=1)
    assert again.tobytes() == fa.tobytes()  # counter-based: deterministic per (seed, pixel, sample, bounce)


# ---- device BVH build (csrc/lbvh.cuh + bvh_gpu.cu): the per-element steps run on the host

def _lbvh_check(emul, sp, pl):
    sp, pl = np.ascontiguousarray(sp), np.ascontiguousarray(pl)
    out = np.zeros(4, np.int32)
    rc = emul.emu_lbvh_check(sp.ctypes.data if len(sp) else None, len(sp), pl.ctypes.data if len(pl) else None, len(pl), out.ctypes.data)
    return rc, out


@pytest.mark.parametrize("n", [0, 1, 2, 3, 7, 500, 20000])
def test_lbvh_tree_is_sound(emul, rt, n):
    """boxes equal the host builder's bit for bit; every primitive is referenced once; boxes nest; BFS order; stack need <= 64"""
    hs = rt.synthetic_scene(n, seed=11)
    sp, pl, mt, _ = hs.arrays()
    rc, out = _lbvh_check(emul, sp, pl)
    assert rc == 0, rc
    assert out[3] == len(sp) + len(pl) and out[2] <= 64 and out[1] >= 1
    # spheres only / planes only / nothing / one primitive / identical primitives (equal Morton keys: position breaks ties)
    assert _lbvh_check(emul, sp, pl[:0])[0] == 0
    assert _lbvh_check(emul, sp[:0], pl)[0] == 0
    assert _lbvh_check(emul, sp[:0], pl[:0])[0] == 0
    assert _lbvh_check(emul, sp[:1], pl[:0])[0] == 0
    if len(sp):
        dup = np.repeat(sp[:1], 37)
        rc, out = _lbvh_check(emul, dup, pl[:0])
        assert rc == 0 and out[3] == 37


def test_lbvh_exact_trace_matches_oracle_bit_for_bit(emul, EmuSceneCls, rt):
    """closest hits through the device-built tree == the oracle's exhaustive scan (ids and t bits)"""
    g = golden("test_config_frame.npz")
    es = EmuSceneCls(emul, g["spheres"], g["planes"], g["materials"], builder=1)
    rays = rt.pixel_center_rays(g["camera"])
    ty, ix, t = es.trace(rays, exact=1)
    assert (ty == g["type_brute"]).all() and (ix == g["index_brute"]).all()
    hit = ty >= 0
    assert (bits(t[hit]) == bits(g["t_brute"][hit])).all()
    g = golden("synthetic2k_textured.npz")
    sp, pl, mt, _ = rt.synthetic_scene(2000, seed=1234).arrays()
    es = EmuSceneCls(emul, sp, pl, mt, builder=1)
    ty, ix, t = es.trace(rt.pixel_center_rays(g["trace_camera"]), exact=1)
    assert ((ty != g["type_brute"]) | (ix != g["index_brute"])).sum() == 0
    hit = ty >= 0
    assert (bits(t[hit]) == bits(g["t_brute"][hit])).all()


def test_lbvh_reference_stream_render_is_bit_identical(emul, EmuSceneCls):
    g = golden("test_config_frame.npz")
    es = EmuSceneCls(emul, g["spheres"], g["planes"], g["materials"], builder=1)
    fb, cnt = es.render(g["camera"], rng=1, exact=1, tex_mode=0)
    assert cnt[0] == int(g["segments"][0])
    assert fb.tobytes() == g["fb_sum"].tobytes()


# ---- hoisted primitives (wide_bvh.h): scene-spanning primitives are tested before the traversal, not through the tree

def _big_planes_scene(rt, n_planes):
    """n_planes stacked scene-sized quads (all hoist candidates) + a few small spheres between them"""
    k = np.arange(n_planes, dtype=np.float32)
    base = np.stack([np.full(n_planes, -10, np.float32), np.full(n_planes, -10, np.float32), -1.0 - k], 1)
    u = np.tile(np.array([[20, 0, 0]], np.float32), (n_planes, 1))
    v = np.tile(np.array([[0, 20, 0]], np.float32), (n_planes, 1))
    pl = rt.make_planes(base, u, v, np.zeros(n_planes, np.int32), np.zeros(n_planes, np.int32))
    sp = np.zeros(5, rt.SPHERE)
    sp["center"] = [[0, 0, 0.5], [3, 1, 0.2], [-4, 2, 1.0], [1, -5, 0.3], [6, 6, 0.1]]
    sp["radius"] = [0.5, 0.4, 0.7, 0.3, 0.6]
    mt = np.zeros(1, rt.MATERIAL)
    rng = np.random.default_rng(7)
    rays = np.zeros((4000, 6), np.float32)
    rays[:, :3] = rng.uniform(-8, 8, (4000, 3)) * [1, 1, 0] + [0, 0, 6]
    rays[:, 3:] = rng.normal(size=(4000, 3)) * [0.6, 0.6, 0.2] + [0, 0, -1]
    return sp, pl, mt, rays


@pytest.mark.parametrize("n_planes,n_spheres", [(1, 0), (3, 5), (6, 5), (6, 0)])
@pytest.mark.parametrize("builder", [0, 1])
def test_hoisted_primitives_keep_the_closest_hit(emul, EmuSceneCls, orc, rt, n_planes, n_spheres, builder):
    """all planes are hoist candidates: at most 4 leave the tree (largest first), the tree may even be empty; ids and t stay
    those of the oracle's linear scan"""
    sp, pl, mt, rays = _big_planes_scene(rt, n_planes)
    sp = sp[:n_spheres]
    es = EmuSceneCls(emul, sp, pl, mt, builder=builder)
    ty, ix, t = es.trace(rays, exact=1)
    oy, oi, orec = orc.scene(sp, pl, mt).trace(rays, 0)
    assert (ty == oy).all() and (ix == oi).all()
    hit = ty >= 0
    assert hit.mean() > 0.5 and (bits(t[hit]) == bits(orec[hit, 0])).all()
    out = np.zeros(4, np.int32)
    assert emul.emu_lbvh_check(sp.ctypes.data if len(sp) else None, len(sp), pl.ctypes.data, len(pl), out.ctypes.data) == 0


def _mixed_plane_scene(rt, n=300, seed=3):
    """quads, ELLIPSES (never generated by create_scene, but part of PlaneData: include/plane.h:30-55) and triangles at
    random orientations + spheres; rays from a sphere of origins towards the cloud"""
    rng = np.random.default_rng(seed)
    base = rng.uniform(-4, 4, (n, 3)).astype(np.float32)
    u = rng.normal(size=(n, 3)).astype(np.float32)
    v = np.cross(u, rng.normal(size=(n, 3))).astype(np.float32)
    v *= (np.linalg.norm(u, axis=1) / np.maximum(np.linalg.norm(v, axis=1), 1e-6))[:, None].astype(np.float32)
    pl = rt.make_planes(base, u, v, np.zeros(n, np.int32), (np.arange(n) % 3).astype(np.int32))
    sp = np.zeros(50, rt.SPHERE)
    sp["center"] = rng.uniform(-4, 4, (50, 3))
    sp["radius"] = rng.uniform(0.1, 0.6, 50)
    mt = np.zeros(1, rt.MATERIAL)
    o = rng.normal(size=(6000, 3))
    o = 9 * o / np.linalg.norm(o, axis=1)[:, None]
    rays = np.concatenate([o, rng.uniform(-3, 3, (6000, 3)) - o], 1).astype(np.float32)
    return sp, pl, mt, rays


@pytest.mark.parametrize("builder", [0, 1])
def test_all_three_plane_types_match_the_oracle(emul, EmuSceneCls, orc, rt, builder):
    sp, pl, mt, rays = _mixed_plane_scene(rt)
    es = EmuSceneCls(emul, sp, pl, mt, builder=builder)
    ty, ix, t = es.trace(rays, exact=1)
    oy, oi, orec = orc.scene(sp, pl, mt).trace(rays, 0)
    assert (ty == oy).all() and (ix == oi).all()
    hit = ty >= 0
    assert (bits(t[hit]) == bits(orec[hit, 0])).all()
    for k in range(3):  # every plane type is actually hit
        assert (pl["type"][ix[ty == 1]] == k).sum() > 50


def test_fast_mode_degrades_gracefully_with_camera_distance(emul, EmuSceneCls, rt):
    """fast mode (fp32 leaf tests, unwidened slab test: traverse.cuh RTB_FAST_SLAB) loses precision with |origin| / scene size:
    measured here so that the limit is known -- ids equal to exact mode's except on silhouette pixels at the scene's own camera
    distance, a few per mille 10x further out, a few per cent 40x further out (exact mode keeps every guard)"""
    hs = rt.synthetic_scene(2000, seed=1234)
    sp, pl, mt, _ = hs.arrays()
    es = EmuSceneCls(emul, sp, pl, mt)
    for dist, bound in ((15.0, 1e-3), (150.0, 6e-3), (600.0, 6e-2)):
        cam = rt.build_camera(120, 160, [-dist, 0.3 * dist, 0.3 * dist], [0, 0, 0], 2 * np.degrees(np.arctan(12.0 / dist)), 1, 1)
        rays = rt.pixel_center_rays(cam)
        ty, ix, t = es.trace(rays, exact=1)
        ty2, ix2, t2 = es.trace(rays, exact=0)
        assert (ty >= 0).mean() > 0.3
        mism = ((ty2 != ty) | (ix2 != ix)).mean()
        assert mism < bound, (dist, mism)


@pytest.mark.parametrize("builder", [0, 1])
def test_fast_plane_record_agrees_with_the_reference_arithmetic(emul, EmuSceneCls, rt, builder):
    """fast mode tests planes on the 3 x float4 record (alpha = A.p - a0, type in a0's low mantissa bits; lbvh.cuh): same hits
    as the reference's cross-product form except rays within fp32 noise of a plane's edge, for quads, ellipses and triangles"""
    sp, pl, mt, rays = _mixed_plane_scene(rt)
    es = EmuSceneCls(emul, sp[:0], pl, mt, builder=builder)   # planes only
    ty, ix, t = es.trace(rays, exact=1)
    ty2, ix2, t2 = es.trace(rays, exact=0)
    assert (ty >= 0).mean() > 0.5
    assert ((ty2 != ty) | (ix2 != ix)).mean() < 2e-3
    same = (ty >= 0) & (ty2 == ty) & (ix2 == ix)
    assert np.max(np.abs(t2[same] - t[same]) / t[same]) < 1e-4
    for k in range(3):
        assert (pl["type"][ix2[ty2 == 1]] == k).sum() > 50


def test_sphere_grid_walk_finds_the_tree_walks_hits(emul, EmuSceneCls, orc, rt):
    """the uniform grid over a dense sphere field (scene_pack.h GridAccel, traverse.cuh trav_grid_step; host run of the
    device code): camera rays AND rays that start on surfaces inside the field -- the ones the grid is for -- get the ids
    and t bits of the all-in-the-tree scene in exact mode, the oracle's exhaustive scan agrees on a sample, and the
    production arithmetic differs from exact mode on at most 1e-5 of the rays"""
    hs = rt.synthetic_scene(6000, W=320, H=180, seed=99)
    sp, pl, mt, _ = hs.arrays()
    tree = EmuSceneCls(emul, sp, pl, mt)
    grid = EmuSceneCls(emul, sp, pl, mt, builder=2)
    rays = rt.pixel_center_rays(hs.camera(0))
    ty, ix, t = tree.trace(rays, exact=1)
    hit = ty >= 0
    assert hit.mean() > 0.3
    # secondary rays: from the primary hit points, seeded directions over the whole sphere (half of them leave the field
    # upwards, the others run along it or into the floor)
    rng = np.random.default_rng(4)
    p = rays[hit, :3] + t[hit, None] * rays[hit, 3:]
    d = rng.normal(size=p.shape).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    sec = np.concatenate([p, d], axis=1).astype(np.float32)
    for rr in (rays, sec):
        a = tree.trace(rr, exact=1)
        b = grid.trace(rr, exact=1)
        assert (a[0] == b[0]).all() and (a[1] == b[1]).all() and (bits(a[2]) == bits(b[2])).all()
        f = grid.trace(rr, exact=0)
        assert ((f[0] != b[0]) | (f[1] != b[1])).mean() <= 1e-5
        same = (f[0] == b[0]) & (f[1] == b[1]) & (b[0] >= 0)
        # 1e-5 relative (north_star) on camera rays; a bounce ray's hit a few 1e-3 away is the difference of two numbers of
        # size |oc|, so the bound is relative to the scene's unit there
        assert np.max(np.abs(f[2][same] - b[2][same]) / np.maximum(b[2][same], 1.0 if rr is sec else 0.0)) <= 1e-5
    sel = np.arange(0, len(sec), 37)
    oy, oi, orec = orc.scene(sp, pl, mt).trace(sec[sel], 0)
    by, bi, bt = grid.trace(sec[sel], exact=1)
    mism = (by != oy) | (bi != oi)
    assert mism.mean() <= 2e-3   # the scan's phantom grazing hits outside the boxes (test_gpu_gates_r02.py)
    ok = ~mism & (oy >= 0)
    assert (bits(bt[ok]) == bits(orec[ok, 0])).all()
