"""Device BVH build (csrc/bvh_gpu.cu) on the GPU, through the C-ABI (rt_scene_create_ex with
RT_BVH_GPU_LBVH): the closest hit does not depend on the tree, so every parity gate of the host-built scene must hold
unchanged -- ids and t bit-exact against the oracle / golden fixtures, reference-stream frames bit-identical -- and the
tree the kernels produce must be the tree the same per-element steps produce on the host (tests/emul)."""
import numpy as np
import pytest
from conftest import bits, golden

pytestmark = pytest.mark.gpu


def _gpu_scene(rt, sp, pl, mt, tex=None, tex_mat=-1):
    return rt.Scene(sp, pl, mt, tex, tex_mat, bvh_builder=rt.BVH_GPU_LBVH)


def test_lbvh_primary_ids_bit_exact_golden(rt):
    g = golden("test_config_frame.npz")
    sc = _gpu_scene(rt, g["spheres"], g["planes"], g["materials"])
    assert sc.stats()["bvh_builder"] == rt.BVH_GPU_LBVH
    rays = rt.pixel_center_rays(g["camera"])
    ty, ix, t = sc.trace_primary(rays)
    assert (ty == g["type_brute"]).all() and (ix == g["index_brute"]).all()
    hit = ty >= 0
    assert (bits(t[hit]) == bits(g["t_brute"][hit])).all()
    # whole frame on the reference's RNG stream: same gate as the host-built scene (>= 99.95 % of pixels bit-identical)
    fb, info = sc.render(g["camera"], rng=rt.RNG_REFERENCE, exact=True, tex_mode=rt.TEX_CPU)
    same = (bits(fb) == bits(g["fb_sum"])).all(-1).mean()
    assert info["segments"] == int(g["segments"][0]) and same >= 0.9995, same
    sc.close()
    g = golden("synthetic2k_textured.npz")
    sp, pl, mt, _ = rt.synthetic_scene(2000, seed=1234).arrays()
    sc = _gpu_scene(rt, sp, pl, mt)
    ty, ix, t = sc.trace_primary(rt.pixel_center_rays(g["trace_camera"]))
    assert (ty == g["type_brute"]).all() and (ix == g["index_brute"]).all()
    hit = ty >= 0
    assert (bits(t[hit]) == bits(g["t_brute"][hit])).all()
    sc.close()


@pytest.mark.parametrize("n", [10000, 100000])
def test_lbvh_equals_host_built_scene(rt, orc, n):
    """same ids / t / reference-stream frame as the host-SAH scene, and as the oracle's linear scan on a sample"""
    hs = rt.synthetic_scene(n, W=192, H=108, sqrt_spp=2, depth=8, seed=1234)
    sp, pl, mt, _ = hs.arrays()
    a = rt.Scene(sp, pl, mt)
    b = _gpu_scene(rt, sp, pl, mt)
    cam = rt.build_camera(270, 480, [-15, 0, 4.5], [0, 0, 4.5], 90.0, 1, 1)
    rays = rt.pixel_center_rays(cam)
    ta, tb = a.trace_primary(rays), b.trace_primary(rays)
    assert (ta[0] == tb[0]).all() and (ta[1] == tb[1]).all() and (bits(ta[2]) == bits(tb[2])).all()
    sel = np.arange(0, len(rays), 211)
    ly, li, _ = orc.scene(sp, pl, mt).trace(rays[sel], 0)
    assert (tb[0][sel] == ly).all() and (tb[1][sel] == li).all()
    fa, ia = a.render(hs.camera(0), rng=rt.RNG_REFERENCE, exact=True, tex_mode=rt.TEX_CPU)
    fb, ib = b.render(hs.camera(0), rng=rt.RNG_REFERENCE, exact=True, tex_mode=rt.TEX_CPU)
    assert ia["segments"] == ib["segments"] and fa.tobytes() == fb.tobytes()
    # production stream: fast-math ties may resolve differently between trees; the frames agree to fp32 noise
    pa, ja = a.render(hs.camera(0), seed=3)
    pb, jb = b.render(hs.camera(0), seed=3)
    assert abs(ja["segments"] - jb["segments"]) <= 1e-3 * ja["segments"]
    assert (bits(pa) == bits(pb)).all(-1).mean() > 0.99
    sa, sb = a.stats(), b.stats()
    print("n=%d host SAH: %d nodes depth %d build %.2f ms | device LBVH: %d nodes depth %d stack %d build %.3f ms (+%.3f ms input upload)"
          % (n, sa["wide_nodes"], sa["wide_depth"], sa["build_ms"], sb["wide_nodes"], sb["wide_depth"], sb["max_stack"], sb["build_ms"],
             sb["input_upload_ms"]))
    assert sb["leaves"] == len(sp) + len(pl) and sb["max_stack"] <= 64
    a.close()
    b.close()


@pytest.mark.parametrize("n", [0, 1, 2, 3, 7, 500, 20000])
def test_lbvh_device_tree_is_the_emulated_tree(rt, emul, n):
    """the kernels and the host loops over the same per-element functions (lbvh.cuh) build the same tree"""
    hs = rt.synthetic_scene(n, seed=11)
    sp, pl, mt, _ = hs.arrays()
    out = np.zeros(4, np.int32)
    assert emul.emu_lbvh_check(sp.ctypes.data, len(sp), pl.ctypes.data, len(pl), out.ctypes.data) == 0
    sc = _gpu_scene(rt, sp, pl, mt)
    st = sc.stats()
    assert [st["wide_nodes"], st["wide_depth"], st["max_stack"], st["leaves"]] == list(out), (st, out)
    sc.close()


def test_lbvh_edge_cases(rt):
    mt = np.zeros(1, rt.MATERIAL)
    rays = np.array([[0, 0, 0, 0, 0, -1], [0.2, 0, 0, 0, 0, -1], [5, 5, 0, 0, 0, -1]], np.float32)
    # no primitives at all: every ray misses
    sc = _gpu_scene(rt, np.zeros(0, rt.SPHERE), np.zeros(0, rt.PLANE), mt)
    ty, ix, t = sc.trace_primary(rays)
    assert (ty == -1).all()
    sc.close()
    # one sphere; then 37 identical spheres (equal Morton keys): the reference's linear scan keeps the LAST of equal hits
    sp = np.zeros(1, rt.SPHERE)
    sp["center"] = [0, 0, -5]
    sp["radius"] = 1
    sc = _gpu_scene(rt, sp, np.zeros(0, rt.PLANE), mt)
    ty, ix, t = sc.trace_primary(rays)
    assert list(ty) == [0, 0, -1] and ix[0] == 0 and t[0] == 4.0
    sc.close()
    dup = np.repeat(sp, 37)
    a = rt.Scene(dup, np.zeros(0, rt.PLANE), mt)
    b = _gpu_scene(rt, dup, np.zeros(0, rt.PLANE), mt)
    ra, rb = a.trace_primary(rays), b.trace_primary(rays)
    assert (ra[0] == rb[0]).all() and (ra[1] == rb[1]).all() and (bits(ra[2]) == bits(rb[2])).all()
    assert rb[1][0] == 36
    a.close()
    b.close()


def test_lbvh_build_is_deterministic_and_fast_at_full_size(rt):
    """1M spheres (BASELINE configs[3] primitive count): two builds give the same frame; build time is reported"""
    hs = rt.synthetic_scene(1000000, W=256, H=144, sqrt_spp=1, depth=8, seed=1234)
    sp, pl, mt, _ = hs.arrays()
    a = _gpu_scene(rt, sp, pl, mt)
    b = _gpu_scene(rt, sp, pl, mt)
    fa, ia = a.render(hs.camera(0), seed=5)
    fb, ib = b.render(hs.camera(0), seed=5)
    assert fa.tobytes() == fb.tobytes() and ia["segments"] == ib["segments"]
    st = b.stats()
    print("1M spheres: device build %.3f ms, %d nodes, depth %d, stack %d" % (st["build_ms"], st["wide_nodes"], st["wide_depth"], st["max_stack"]))
    assert st["max_stack"] <= 64 and st["build_ms"] < 200.0
    a.close()
    b.close()


@pytest.mark.parametrize("n_planes,n_spheres", [(1, 0), (3, 5), (6, 5), (6, 0)])
@pytest.mark.parametrize("builder", [0, 1])
def test_hoisted_primitives_keep_the_closest_hit(rt, orc, n_planes, n_spheres, builder):
    """scene-sized quads are hoist candidates: at most 4 leave the tree (it may even end up empty) and are tested before the
    traversal; ids and t stay those of the oracle's linear scan, with the host-built and the device-built tree"""
    from test_emul import _big_planes_scene
    sp, pl, mt, rays = _big_planes_scene(rt, n_planes)
    sp = sp[:n_spheres]
    sc = rt.Scene(sp, pl, mt, bvh_builder=builder)
    assert sc.stats()["num_hoisted"] == min(n_planes, 4)
    ty, ix, t = sc.trace_primary(rays)
    oy, oi, orec = orc.scene(sp, pl, mt).trace(rays, 0)
    assert (ty == oy).all() and (ix == oi).all()
    hit = ty >= 0
    assert hit.mean() > 0.5 and (bits(t[hit]) == bits(orec[hit, 0])).all()
    sc.close()


@pytest.mark.parametrize("builder", [0, 1])
def test_all_three_plane_types_match_the_oracle(rt, orc, builder):
    """quads, ellipses and triangles (include/plane.h:30-55) through both builders' trees: ids and t bits of the oracle's scan"""
    from test_emul import _mixed_plane_scene
    sp, pl, mt, rays = _mixed_plane_scene(rt)
    sc = rt.Scene(sp, pl, mt, bvh_builder=builder)
    ty, ix, t = sc.trace_primary(rays)
    oy, oi, orec = orc.scene(sp, pl, mt).trace(rays, 0)
    assert (ty == oy).all() and (ix == oi).all()
    hit = ty >= 0
    assert (bits(t[hit]) == bits(orec[hit, 0])).all()
    for k in range(3):
        assert (pl["type"][ix[ty == 1]] == k).sum() > 50
    sc.close()


def test_scene_rebuild_swaps_geometry_in_place(rt, orc):
    """rt_scene_rebuild: an animated scene re-submits its primitives each frame; hits follow the new geometry (oracle scan),
    counts may change, materials stay, and a frame rendered after the rebuild equals a freshly created scene's"""
    hs = rt.synthetic_scene(3000, W=96, H=54, sqrt_spp=2, depth=6, seed=21)
    sp, pl, mt, _ = hs.arrays()
    sc = rt.Scene(sp, pl, mt)
    cam = rt.build_camera(135, 240, [-15, 0, 4.5], [0, 0, 4.5], 90.0, 1, 1)
    rays = rt.pixel_center_rays(cam)
    before = sc.trace_primary(rays)
    for step in range(3):
        moved = sp.copy()
        moved["center"][:, 2] += 0.15 * (step + 1) * np.sin(np.arange(len(sp)))[:, None].ravel()
        moved = moved[: len(sp) - 500 * step]
        sc.rebuild(moved, pl)
        st = sc.stats()
        assert st["num_spheres"] == len(moved) and st["bvh_builder"] == rt.BVH_GPU_LBVH and st["max_stack"] <= 64
        ty, ix, t = sc.trace_primary(rays)
        sel = np.arange(0, len(rays), 53)
        oy, oi, orec = orc.scene(moved, pl, mt).trace(rays[sel], 0)
        assert (ty[sel] == oy).all() and (ix[sel] == oi).all()
        hit = oy >= 0
        assert (bits(t[sel][hit]) == bits(orec[hit, 0])).all()
        fresh = rt.Scene(moved, pl, mt, bvh_builder=rt.BVH_GPU_LBVH)
        fa, ia = sc.render(hs.camera(0), rng=rt.RNG_REFERENCE, exact=True, tex_mode=rt.TEX_CPU)
        fb, ib = fresh.render(hs.camera(0), rng=rt.RNG_REFERENCE, exact=True, tex_mode=rt.TEX_CPU)
        assert fa.tobytes() == fb.tobytes() and ia["segments"] == ib["segments"]
        fresh.close()
    assert ((sc.trace_primary(rays)[1] != before[1]).mean()) > 0.01  # the geometry did change
    bad = sp.copy()
    bad["material_idx"][0] = 10 ** 6
    with pytest.raises(rt.RtError):
        sc.rebuild(bad, pl)
    ty2, ix2, _ = sc.trace_primary(rays)  # a refused rebuild leaves the scene as it was
    assert (ty2 == ty).all() and (ix2 == ix).all()
    sc.close()


def test_sphere_grid_scene_matches_the_tree_scene(rt, monkeypatch):
    """RTB_GRID=1 (opt-in, measured slower: DESIGN.md 4.2): the dense sphere field goes into a uniform grid walked by
    trav_grid_step, the rest stays in a small tree.  The closest hit does not depend on the structure: exact-mode ids and t
    bits equal the all-in-the-tree scene's on every pixel-centre ray, the reference-stream frame is byte-identical, and the
    production arithmetic differs on at most 1e-5 of the rays."""
    hs = rt.synthetic_scene(10000, W=480, H=270, sqrt_spp=2, depth=8, seed=1234)
    sp, pl, mt, _ = hs.arrays()
    cam = hs.camera(0)
    tree = rt.Scene(sp, pl, mt)
    monkeypatch.setenv("RTB_GRID", "1")
    grid = rt.Scene(sp, pl, mt)
    monkeypatch.delenv("RTB_GRID")
    st = grid.stats()
    assert tree.stats()["grid_spheres"] == 0 and st["grid_spheres"] >= 9000 and st["grid_cells"] > 1000
    assert st["wide_nodes"] < tree.stats()["wide_nodes"] // 10
    rays = rt.pixel_center_rays(cam)
    a, b = tree.trace_primary(rays), grid.trace_primary(rays)
    assert (a[0] == b[0]).all() and (a[1] == b[1]).all() and (bits(a[2]) == bits(b[2])).all()
    f = grid.trace_primary(rays, exact=False)
    assert ((f[0] != b[0]) | (f[1] != b[1])).mean() <= 1e-5
    fa, ia = tree.render(cam, rng=rt.RNG_REFERENCE, exact=True, tex_mode=rt.TEX_CPU)
    fb, ib = grid.render(cam, rng=rt.RNG_REFERENCE, exact=True, tex_mode=rt.TEX_CPU)
    assert ia["segments"] == ib["segments"] and fa.tobytes() == fb.tobytes()
    pa, _ = tree.render(cam, seed=5)
    pb, _ = grid.render(cam, seed=5)
    assert (bits(pa) == bits(pb)).all(-1).mean() >= 0.999
    tree.close()
    grid.close()
