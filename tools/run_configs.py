#!/usr/bin/env python
"""Every BASELINE.json config once on one GPU: kernel ms, Mrays/s, per-ray work.  Writes profiles/<name>.json lines."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import rtp_b200 as rt


REF = None  # oracle_py.Ref() when --ref is given


def run(name, scene, cam, reps, ref_scene=None):
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
    if REF is not None and ref_scene is not None:
        # north_star: "frame ms vs reference CUDA" -- the unmodified reference render_kernel (oracle/_ref, -arch=sm_100) on the
        # same arrays; its segments are counted by our kernel on the reference's random stream (the reference counts none)
        from oracle import oracle_py as O
        _, iref = scene.render(cam, rng=rt.RNG_REFERENCE)
        O.gpu_render_ref(ref_scene, cam, reps=1)
        _, ms = O.gpu_render_ref(ref_scene, cam, reps=1)
        out["reference_cuda_kernel_ms"] = ms
        out["reference_cuda_mrays_s"] = iref["segments"] / ms / 1e3
        out["speedup_vs_reference_cuda"] = ms / best["kernel_ms"]
    print(json.dumps(out), flush=True)
    return out


def main():
    global REF
    args = [a for a in sys.argv[1:] if a != "--ref"]
    if "--ref" in sys.argv[1:]:
        from oracle import oracle_py as O
        REF = O.Ref() if O.ref_available() else None
    which = args or ["shipped", "spheres10k_1080p_64spp_d16", "spheres100k_1080p_256spp_d32", "spheres1m_4k_64spp_d8",
                             "spheres100k_8k_1024spp_d32"]
    for name in which:
        if name == "shipped":
            hs = rt.HostScene(rt.shipped_config_text(), texture="")
            sp, pl, mt, _ = hs.arrays()
            tex = rt.synthetic_texture()
            scene = rt.Scene(sp, pl, mt, tex, 0)
            rs = REF.scene(sp, pl, mt, tex, 0) if REF is not None else None
            for frame in (0, 25, 50, 75):
                run("shipped_config_1080x720_2500spp_d50_frame%d" % frame, scene, hs.camera(frame), 2, rs)
        else:
            cfg = rt.CONFIGS[name]
            hs = rt.synthetic_scene(cfg["n"], W=cfg["W"], H=cfg["H"], sqrt_spp=cfg["sqrt_spp"], depth=cfg["depth"], seed=1234, texture="")
            scene = rt.Scene.from_host(hs)
            sp, pl, mt, tex = hs.arrays()
            # the 8K / 1024-spp frame would take the reference kernel minutes: its baseline is the headline scene's
            rs = REF.scene(sp, pl, mt) if (REF is not None and "8k" not in name) else None
            run(name, scene, hs.camera(0), 1 if "8k" in name else 2, rs)
        scene.close()


if __name__ == "__main__":
    main()
