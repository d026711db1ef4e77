#!/usr/bin/env python
"""the reference's 100-frame shipped config end to end through the drop-in executable (rt_main),
wall clock incl. file writes -- pipelined (a writer thread encodes / writes frame n while frame n+1 renders) vs --no-pipeline
(the reference's strictly serial loop, src/camera.cu:297-347), frames over 1..N GPUs, every frame's tiles over N GPUs -- next to
the reference CUDA kernel's own time for sample frames (its program also downloads and writes inside its timed region, Q9).
One JSON line per run."""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import rtp_b200 as rt


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=100)
    ap.add_argument("--sqrt-spp", type=int, default=50)
    ap.add_argument("--gpus", type=int, default=rt.device_count())
    ap.add_argument("--ref", action="store_true", help="also time the reference CUDA kernel on frames 0/25/50/75")
    ap.add_argument("--ref-program", type=int, default=0, help="also run the reference program's own gpu_render loop over this many frames")
    ap.add_argument("--only", default="", help="comma-separated run tags to keep (default: all)")
    a = ap.parse_args()
    exe = os.path.join(os.path.dirname(rt.LIB_PATH), "rt_main")
    runs = [("pipelined_1gpu", ["--gpu"]), ("serial_1gpu", ["--gpu", "--no-pipeline"])]
    if a.gpus > 1:
        runs += [("pipelined_frames_over_%dgpus" % a.gpus, ["--gpu", "--gpus", str(a.gpus)]),
                 ("pipelined_tiles_over_%dgpus" % a.gpus, ["--gpu", "--gpus", str(a.gpus), "--split", "tiles"])]
    if a.only:
        runs = [r for r in runs if any(r[0].startswith(k) for k in a.only.split(","))]
    for tag, flags in runs:
        with tempfile.TemporaryDirectory() as td:
            cfg = rt.shipped_config_text(sqrt_spp=a.sqrt_spp, num_frames=a.frames, output=os.path.join(td, "render_%d.png"))
            t0 = time.perf_counter()
            r = subprocess.run([exe] + flags, input=cfg.encode(), capture_output=True, timeout=3000)
            wall = time.perf_counter() - t0
            lines = [l.split("\t") for l in r.stdout.decode().strip().split("\n") if l.count("\t") == 2]
            ms = np.array([float(l[1]) for l in lines]) if lines else np.zeros(0)
            files = len([f for f in os.listdir(td) if f.startswith("render_")])
            print(json.dumps(dict(run=tag, flags=flags, rc=r.returncode, frames=a.frames, files_written=files, wall_s=wall,
                                  ms_per_frame_wall=1e3 * wall / max(a.frames, 1), frame_ms_printed_mean=float(ms.mean()) if len(ms) else None,
                                  frame_ms_printed_max=float(ms.max()) if len(ms) else None, width=1080, height=720, spp=a.sqrt_spp ** 2,
                                  max_depth=50, texture="none (../floor2.jpg of the shipped config is not shipped, SURVEY Q6)",
                                  stdout_lines=len(r.stdout.decode().split("\n")), stderr=r.stderr.decode()[-300:])), flush=True)
    if a.ref_program:
        # the reference PROGRAM's own GPU frame loop (gpu_render, src/camera.cu:290-349) on the same config: its printed per-frame
        # ms (kernel + sync + D2H + per-pixel saver writes, the reference's own event pair) and the wall clock of the whole loop
        from oracle import oracle_py as O
        if O.ref_available():
            with tempfile.TemporaryDirectory() as td:
                cfg = rt.shipped_config_text(sqrt_spp=a.sqrt_spp, num_frames=a.ref_program, output=os.path.join(td, "render_%d.png"))
                rs = O.Ref().scene_from_config(cfg)
                t0 = time.perf_counter()
                lines = O.gpu_render_program_ref(rs)
                wall = time.perf_counter() - t0
                ms = np.array([l[1] for l in lines])
                files = len([f for f in os.listdir(td) if f.startswith("render_")])
                print(json.dumps(dict(run="reference_program_gpu_render", frames=a.ref_program, files_written=files, wall_s=wall,
                                      ms_per_frame_wall=1e3 * wall / max(a.ref_program, 1), frame_ms_printed_mean=float(ms.mean()),
                                      frame_ms_printed_max=float(ms.max()), frame_ms_printed=[round(float(x), 1) for x in ms],
                                      note="unmodified reference gpu_render loop (oracle/_ref, sm_100), its own printed per-frame ms")), flush=True)
    if a.ref:
        from oracle import oracle_py as O
        if O.ref_available():
            hs = rt.HostScene(rt.shipped_config_text(sqrt_spp=a.sqrt_spp), texture="")
            sp, pl, mt, _ = hs.arrays()
            rs = O.Ref().scene(sp, pl, mt)
            O.gpu_render_ref(rs, hs.camera(0), reps=1)
            for f in (0, 25, 50, 75):
                _, ms = O.gpu_render_ref(rs, hs.camera(f), reps=1)
                print(json.dumps(dict(run="reference_cuda_kernel_only", frame=f, kernel_ms=ms,
                                      note="render_kernel alone; the reference program adds D2H + 3 one-byte writes per pixel inside its frame time")), flush=True)


if __name__ == "__main__":
    main()
