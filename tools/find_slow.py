#!/usr/bin/env python
"""Locate the most expensive pixels of a frame by timing sub-windows (camera shifted to the window; same rays up to jitter)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import rtp_b200 as rt

cfg = rt.CONFIGS["spheres100k_1080p_256spp_d32"]
hs = rt.synthetic_scene(cfg["n"], W=cfg["W"], H=cfg["H"], sqrt_spp=cfg["sqrt_spp"], depth=cfg["depth"], seed=1234, texture="")
sc = rt.Scene.from_host(hs)
base = hs.camera(0)
base["samples_per_pixel"] = 16


def window(x0, y0, w, h):
    c = base.copy()
    c["pixel00_loc"] = base["pixel00_loc"] + x0 * base["pixel_delta_u"] + y0 * base["pixel_delta_v"]
    c["image_width"], c["image_height"] = w, h
    fb = np.zeros((h, w, 3), np.float32)
    sc.render(c, out=fb)
    _, i = sc.render(c, out=fb, count_work=True)
    return i


col = int(sys.argv[1]) if len(sys.argv) > 1 else 44
res = []
for ty in range(68):
    i = window(col * 16, ty * 16, 16, 16)
    res.append((i["kernel_ms"], ty, i))
i = window(col * 16, 0, 16, 1080)
print("whole column window: %.2f ms, segs %d" % (i["kernel_ms"], i["segments"]))
res.sort(reverse=True)
print("sum of tile times %.1f ms" % sum(r[0] for r in res))
for ms, ty, i in res[:8]:
    print("tile (%d,%d): %.2f ms segs %d V %.1f seg/sample %.2f" % (col, ty, ms, i["segments"], i["node_visits"] / max(i["segments"], 1), i["segments"] / i["samples"]))
ms, ty, _ = res[0]
px = []
for y in range(16):
    for x in range(16):
        i = window(col * 16 + x, ty * 16 + y, 1, 1)
        px.append((i["kernel_ms"], x, y, i))
px.sort(reverse=True)
for ms, x, y, i in px[:5]:
    s = max(i["segments"], 1)
    print("pixel (%d,%d): %.3f ms  segs %d (%.1f/sample)  V %.1f Ts %.1f Tp %.1f per segment" % (
        col * 16 + x, ty * 16 + y, ms, i["segments"], i["segments"] / i["samples"], i["node_visits"] / s, i["sphere_tests"] / s, i["plane_tests"] / s))
