#!/usr/bin/env python
"""Every BASELINE.json config once on one GPU: kernel ms, Mrays/s, per-ray work.  Writes profiles/<name>.json lines."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import rtp_b200 as rt


def run(name, scene, cam, reps):
    W, H = int(cam["image_width"][0]), int(cam["image_height"][0])
    fb = np.zeros((H, W, 3), np.float32)
    best = None
    for _ in range(reps):
        _, i = scene.render(cam, out=fb)
        if best is None or i["kernel_ms"] < best["kernel_ms"]:
            best = i
    st = scene.stats()
    out = dict(config=name, width=W, height=H, spp=int(cam["samples_per_pixel"][0]), max_depth=int(cam["max_depth"][0]),
               spheres=st["num_spheres"], planes=st["num_planes"], wide_nodes=st["wide_nodes"], wide_depth=st["wide_depth"],
               bvh_build_ms=st["build_ms"], device_mb=st["device_bytes"] / 1e6, kernel_ms=best["kernel_ms"], segments=best["segments"],
               mrays_s=best["segments"] / best["kernel_ms"] / 1e3, image_mean=float(fb.mean()), finite=bool(np.isfinite(fb).all()))
    print(json.dumps(out), flush=True)
    return out


def main():
    which = sys.argv[1:] or ["shipped", "spheres10k_1080p_64spp_d16", "spheres100k_1080p_256spp_d32", "spheres1m_4k_64spp_d8",
                             "spheres100k_8k_1024spp_d32"]
    for name in which:
        if name == "shipped":
            hs = rt.HostScene(rt.shipped_config_text(), texture="")
            sp, pl, mt, _ = hs.arrays()
            scene = rt.Scene(sp, pl, mt, rt.synthetic_texture(), 0)
            for frame in (0, 25):
                run("shipped_config_1080x720_2500spp_d50_frame%d" % frame, scene, hs.camera(frame), 2)
        else:
            cfg = rt.CONFIGS[name]
            hs = rt.synthetic_scene(cfg["n"], W=cfg["W"], H=cfg["H"], sqrt_spp=cfg["sqrt_spp"], depth=cfg["depth"], seed=1234, texture="")
            scene = rt.Scene.from_host(hs)
            run(name, scene, hs.camera(0), 1 if "8k" in name else 2)
        scene.close()


if __name__ == "__main__":
    main()
