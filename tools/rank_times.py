#!/usr/bin/env python
"""Per-rank kernel time / work of an N-way tile split of the headline frame, rendered rank by rank on ONE GPU (which
rank is the slowest, and why)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import rtp_b200 as rt

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 64
cfg = rt.CONFIGS["spheres100k_1080p_256spp_d32"]
hs = rt.synthetic_scene(cfg["n"], W=cfg["W"], H=cfg["H"], sqrt_spp=cfg["sqrt_spp"], depth=cfg["depth"], seed=1234, texture="")
sc = rt.Scene.from_host(hs)
cam = hs.camera(0)
cam["samples_per_pixel"] = spp
fb = np.zeros((cfg["H"], cfg["W"], 3), np.float32)
for r in range(n):
    sc.render(cam, out=fb, tile_rank=r, tile_nranks=n)
    _, i = sc.render(cam, out=fb, tile_rank=r, tile_nranks=n)
    _, c = sc.render(cam, out=fb, tile_rank=r, tile_nranks=n, count_work=True)
    s = c["segments"]
    print("rank %d: %.2f ms  %.1f Mseg  %d Mrays/s  V %.2f Ts %.2f Tp %.2f  seg/sample %.2f" % (
        r, i["kernel_ms"], s / 1e6, s / i["kernel_ms"] / 1e3, c["node_visits"] / s, c["sphere_tests"] / s, c["plane_tests"] / s,
        s / c["samples"]))
