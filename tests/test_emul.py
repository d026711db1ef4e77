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
