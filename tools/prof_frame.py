#!/usr/bin/env python
"""One frame (or a few) of a named workload through the C-ABI, with overrides -- the short command the ncu captures and
kernel A/B timings are taken on.  Prints kernel ms, Mrays/s and (with --count) the per-ray work counters."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import rtp_b200 as rt


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="spheres100k_1080p_256spp_d32")
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel")
    ap.add_argument("--depth", type=int, default=0)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--kernel", default="mega", choices=["mega", "wavefront", "pool"])
    ap.add_argument("--rng", default="philox", choices=["philox", "reference"])
    ap.add_argument("--exact", action="store_true")
    ap.add_argument("--count", action="store_true")
    ap.add_argument("--top-nodes", type=int, default=-1)
    ap.add_argument("--frame", type=int, default=0)
    ap.add_argument("--builder", default="host", choices=["host", "gpu"], help="BVH builder: host binned SAH / device LBVH")
    ap.add_argument("--chunk", type=int, default=0, help="sample_chunk option (0 auto, -1 whole pixels)")
    a = ap.parse_args()
    builder = rt.BVH_GPU_LBVH if a.builder == "gpu" else rt.BVH_HOST_SAH
    if a.workload == "shipped":  # BASELINE configs[0]: the reference's own config.txt scene (synthetic stand-in for floor.jpg)
        cfg = dict(W=1080, H=720)
        hs = rt.HostScene(rt.shipped_config_text(), texture="")
        sp, pl, mt, _ = hs.arrays()
        scene = rt.Scene(sp, pl, mt, rt.synthetic_texture(), 0, bvh_builder=builder)
    else:
        cfg = rt.CONFIGS[a.workload]
        hs = rt.synthetic_scene(cfg["n"], W=cfg["W"], H=cfg["H"], sqrt_spp=cfg["sqrt_spp"], depth=cfg["depth"], seed=1234, texture="")
        scene = rt.Scene.from_host(hs, bvh_builder=builder)
    cam = hs.camera(a.frame)
    if a.spp:
        cam["samples_per_pixel"] = a.spp
    if a.depth:
        cam["max_depth"] = a.depth
    fb = np.zeros((cfg["H"], cfg["W"], 3), np.float32)
    kw = dict(kernel={"mega": rt.KERNEL_MEGA, "wavefront": rt.KERNEL_WAVEFRONT, "pool": rt.KERNEL_POOL}[a.kernel],
              rng=rt.RNG_PHILOX if a.rng == "philox" else rt.RNG_REFERENCE, exact=a.exact, top_nodes=a.top_nodes, sample_chunk=a.chunk)
    best = None
    for _ in range(a.reps):
        _, info = scene.render(cam, out=fb, **kw)
        if best is None or info["kernel_ms"] < best["kernel_ms"]:
            best = info
    out = dict(workload=a.workload, spp=int(cam["samples_per_pixel"][0]), depth=int(cam["max_depth"][0]), kernel=a.kernel,
               kernel_ms=best["kernel_ms"], segments=best["segments"], mrays_s=best["segments"] / best["kernel_ms"] / 1e3,
               mean=float(fb.mean()), stats=scene.stats())
    if a.count:
        _, ci = scene.render(cam, out=fb, count_work=True, **kw)
        s = max(ci["segments"], 1)
        out["per_ray"] = {k: ci[k] / s for k in ("node_visits", "sphere_tests", "plane_tests", "shaded_hits")}
        out["segments_per_sample"] = ci["segments"] / max(ci["samples"], 1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
