#!/usr/bin/env python
"""bench.py -- Mrays/s of the per-pixel render path (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

A step = one frame of the workload (default: BASELINE.json configs[2], the 100k-sphere 1080p / 256 spp / depth 32 scene
the north_star target is quoted on) through the persistent megakernel.  A "ray" is one path segment (one hit_scene call
of the reference).  N>1: launched by torchrun, one rank per GPU; every rank renders the tiles
t % N == rank of the SAME frame (strong scaling) and the packed tiles are gathered to rank 0 over NCCL inside the step.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

DEFAULT_WORKLOAD = "spheres100k_1080p_256spp_d32"
FP32_LANES_PER_SM = 128


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d, "measured"
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


def flops_per_ray(c):
    """56 V + 30 Ts + 36 Tp + 80 H + 20, per-ray counts from a counters run of the same kernel"""
    s = max(c["segments"], 1)
    V, Ts, Tp, H = c["node_visits"] / s, c["sphere_tests"] / s, c["plane_tests"] / s, c["shaded_hits"] / s
    return 56 * V + 30 * Ts + 36 * Tp + 80 * H + 20, 128 * V + 16 * Ts + 64 * Tp + 48 * H, dict(V=V, Ts=Ts, Tp=Tp, H=H)


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md recipe)"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                       "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        rows = [r.strip().split(", ") for r in open(self.f.name) if r.strip()]
        os.unlink(self.f.name)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except ValueError:
                continue
            for n, v in zip(names, r[5:9]):
                if v.strip().lower() == "active":
                    reasons.add(n)
        busy = [s for s in sm if s > 0]
        return {"sm_mhz": float(np.median(busy)) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def workload_scene(rt, name):
    cfg = rt.CONFIGS[name]
    hs = rt.synthetic_scene(cfg["n"], W=cfg["W"], H=cfg["H"], sqrt_spp=cfg["sqrt_spp"], depth=cfg["depth"], seed=1234, texture="")
    return cfg, hs


def cpu_sample(hs, cfg, kind_pref="reference", target_s=12.0):
    """reference CPU path on a bounded sample of the workload: every 8th pixel in x and y, spp scaled to ~target_s."""
    from oracle import oracle_py as O
    sp, pl, mt, _ = hs.arrays()
    threads = os.cpu_count() or 1
    if kind_pref == "reference" and O.ref_available():
        chk, kind = O.Ref(), "reference"
    else:
        chk, kind = O.Oracle(), "port"
    sc = chk.scene(sp, pl, mt)
    cam = hs.camera(0).copy()
    xs = ys = 8
    cam["samples_per_pixel"] = 1
    t0 = time.perf_counter()
    _, segs = sc.render(cam, nthreads=threads, x0=0, xs=xs, y0=0, ys=ys)
    dt = max(time.perf_counter() - t0, 1e-3)
    spp = int(min(max(1, target_s / dt), cfg["sqrt_spp"] ** 2))
    cam["samples_per_pixel"] = spp
    t0 = time.perf_counter()
    _, segs = sc.render(cam, nthreads=threads, x0=0, xs=xs, y0=0, ys=ys)
    dt = time.perf_counter() - t0
    sample = "every 8th pixel in x and y (%dx%d of %dx%d), %d of %d spp, depth %d, %d host threads" % (
        (cfg["W"] + xs - 1) // xs, (cfg["H"] + ys - 1) // ys, cfg["W"], cfg["H"], spp, cfg["sqrt_spp"] ** 2, cfg["depth"], threads)
    return dict(value=segs / dt / 1e6, unit="Mrays/s", cores=threads, kind=kind, sample=sample, seconds=dt, segments=int(segs)), sc, cam


def run_reference_arm(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path (oracle/_ref when it was compiled, else the
    oracle port) on the box's host cores, all threads, bounded sample per step."""
    if rank != 0:
        return
    import rtp_b200 as rt
    cfg, hs = workload_scene(rt, args.workload)
    # size one step to ~6 s so warmup+steps stay within minutes
    base, sc, cam = cpu_sample(hs, cfg, "reference", target_s=6.0)
    threads = base["cores"]
    times, segs_total = [], 0
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        _, segs = sc.render(cam, nthreads=threads, x0=0, xs=8, y0=0, ys=8)
        dt = time.perf_counter() - t0
        if i >= args.warmup:
            times.append(dt); segs_total += segs
    value = segs_total / sum(times) / 1e6
    base.update(value=value)
    line = {"impl": "reference", "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * sum(times) / len(times), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "note": "CPU arm renders a bounded pixel/sample subsample of the workload per step"},
            "cpu_baseline": {k: base[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--kernel", default="mega", choices=["mega", "wavefront"])
    ap.add_argument("--no-baselines", action="store_true", help="skip the cpu_baseline / reference-CUDA legs")
    ap.add_argument("--top-nodes", type=int, default=-1)
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = max(args.warmup, 1)

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    import rtp_b200 as rt
    from rtp_b200_pkg import multi_gpu as mg

    if rt.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device -- this benchmark has no CPU path (" + rt.lib().rt_last_error().decode() + ")")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    cfg, hs = workload_scene(rt, args.workload)
    W, H, spp, depth = cfg["W"], cfg["H"], cfg["sqrt_spp"] ** 2, cfg["depth"]
    scene = rt.Scene.from_host(hs)
    stats = scene.stats()
    cam = hs.camera(0)
    kernel = rt.KERNEL_MEGA if args.kernel == "mega" else rt.KERNEL_WAVEFRONT
    opts = dict(kernel=kernel, seed=0, top_nodes=args.top_nodes)
    frame = mg.TiledFrame(scene, W, H, rank, world, dev)
    stream = torch.cuda.current_stream().cuda_stream
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(timed):
        flush.zero_()  # L2 flush between iterations (outside the event pair)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        frame.render(cam, stream, **opts)
        full = frame.gather(dist, stream)
        e1.record()
        e1.synchronize()
        info = scene.finish()
        return e0.elapsed_time(e1), info, full

    for _ in range(max(args.warmup, 1)):
        step(False)
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    t_wall = time.perf_counter()
    ms_total, kms_total, segs = 0.0, 0.0, 0
    for _ in range(args.steps):
        ms, info, full = step(True)
        ms_total += ms
        kms_total += info["kernel_ms"]
        segs += info["segments"]
    barrier()
    t_wall = time.perf_counter() - t_wall
    clocks = sampler.stop() if sampler else None

    t = torch.tensor([ms_total, kms_total, float(segs)], dtype=torch.float64, device=dev)
    if world > 1:
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone()
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        ms_total, kms_total, segs_all = tmax[0].item(), tmax[1].item(), tsum[2].item()
    else:
        segs_all = float(segs)
    value = segs_all / (ms_total * 1e-3) / 1e6
    if args.kernel == "mega":
        chunked = spp >= 32  # rt_render_opts.sample_chunk auto: render_mega + reduce_chunks
        launches_per_step = 1 + (1 if chunked else 0) + (1 if world > 1 else 0) + (1 if (world > 1 and rank == 0) else 0)  # render, reduce, pack, unpack(rank 0)
    else:
        K = max(1, min(spp, (16 << 20) // max(1, ((W + 15) // 16) * ((H + 15) // 16) * 256 // world)))
        launches_per_step = ((spp + K - 1) // K) * (2 + 2 * depth)

    # ---- end to end through the public C-ABI call with HOST buffers (rt_render: launch + D2H of the sums), every rank
    host_fb = np.zeros((H, W, 3), np.float32)
    e2e_opts = dict(opts, tile_rank=rank, tile_nranks=world)
    scene.render(cam, out=host_fb, **e2e_opts)
    barrier()
    t0 = time.perf_counter()
    e2e_segs = 0
    for _ in range(args.steps):
        _, i2 = scene.render(cam, out=host_fb, **e2e_opts)
        e2e_segs += i2["segments"]
    barrier()
    e2e_t = time.perf_counter() - t0
    te = torch.tensor([e2e_t, float(e2e_segs)], dtype=torch.float64, device=dev)
    if world > 1:
        tm = te.clone(); dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        tsm = te.clone(); dist.all_reduce(tsm, op=dist.ReduceOp.SUM)
        e2e_t, e2e_segs = tm[0].item(), tsm[1].item()
    e2e_value = e2e_segs / e2e_t / 1e6

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (render_mega): FP32 issue, per SURVEY.md section 8d
    peaks, peak_src = load_peaks()
    _, cinfo = scene.render(cam, out=host_fb, count_work=True, **e2e_opts)
    fpr, bpr, per_ray = flops_per_ray(cinfo)
    sm_mhz = float(peaks.get("sm_max_mhz", 1965.0))
    fp32_peak = torch.cuda.get_device_properties(dev).multi_processor_count * FP32_LANES_PER_SM * 2 * sm_mhz * 1e6 / 1e12
    kernel_rays_s = segs_all / world / (kms_total * 1e-3) if kms_total > 0 else 0.0  # per GPU, render kernel only
    achieved = kernel_rays_s * fpr / 1e12
    traffic, traffic_src = None, None
    tp = os.path.join(ROOT, "profiles", "r01_ncu_mega_headline.json")   # dram__bytes_read+write of one render_mega launch (ncu --set full)
    if args.workload == DEFAULT_WORKLOAD and args.kernel == "mega" and os.path.exists(tp):
        traffic, traffic_src = json.load(open(tp))["traffic_bytes_per_launch"], "profiles/r01_ncu_mega_headline.json"
    roofline = {"bound": "fp32_issue", "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s", "frac": achieved / fp32_peak,
                "traffic": traffic, "traffic_source": traffic_src, "algorithmic_bytes_per_launch": bpr * segs_all / world / args.steps, "kernel": "render_mega" if args.kernel == "mega" else "wf_extend",
                "kernel_ms_per_launch": kms_total / args.steps, "flops_per_ray": fpr, "bytes_per_ray": bpr, "per_ray": per_ray,
                "peak_source": "148 SMs x 128 fp32 lanes x 2 x sm_max_mhz of MEASURED_PEAKS.json (%s); HBM %.0f GB/s (%s) is not the bound: "
                               "the working set is L2 resident" % (peak_src, peaks.get("hbm_gbs", 0), peak_src),
                "hbm_algorithmic_gbs": kernel_rays_s * bpr / 1e9}

    line = {"metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "spheres": cfg["n"] + 94, "planes": 105, "width": W, "height": H, "spp": spp,
                       "max_depth": depth, "kernel": args.kernel, "rng": "philox4x32-10", "parallelism": "tiles%d" % world,
                       "l2": "flushed between timed steps (256 MB memset)", "bvh_nodes": stats["wide_nodes"],
                       "bvh_depth": stats["wide_depth"], "bvh_build_ms": stats["build_ms"]},
            "clocks": clocks, "kernel_only_mrays_s": kernel_rays_s / 1e6 * world,
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": 76, "d2h_bytes_per_step": int(W * H * 12 // world) * world,
                    "call": "rt_render (camera struct in, host framebuffer out)"},
            "gpu_launches": launches_per_step * args.steps, "wall_s_timed_region": t_wall, "roofline": roofline}

    if world == 1 and not args.no_baselines:
        try:
            base, _, _ = cpu_sample(hs, cfg, "reference", target_s=12.0)
            line["cpu_baseline"] = {k: base[k] for k in ("value", "unit", "cores", "kind", "sample")}
        except Exception as e:  # the CPU leg must never take the GPU number down
            line["cpu_baseline"] = {"value": None, "unit": "Mrays/s", "cores": 0, "kind": "unavailable", "sample": repr(e)}
        try:
            line["reference_cuda"] = reference_cuda_leg(rt, scene, hs, cam, W, H, spp)
        except Exception as e:
            line["reference_cuda"] = {"unavailable": repr(e)}
        try:
            line["bvh_build"] = bvh_build_leg(rt, hs, cam, stats)
        except Exception as e:
            line["bvh_build"] = {"unavailable": repr(e)}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def bvh_build_leg(rt, hs, cam, host_stats):
    """the device BVH build (rt_scene_create_ex, RT_BVH_GPU_LBVH) next to the host SAH build of the
    same scene -- build time of each, and what the faster-to-build tree costs at render time (one frame, same kernel)."""
    best = None
    for _ in range(3):
        sc = rt.Scene.from_host(hs, bvh_builder=rt.BVH_GPU_LBVH)
        st = sc.stats()
        if best is None or st["build_ms"] < best["build_ms"]:
            best = st
        if _ < 2:
            sc.close()
    sc.render(cam, seed=0)
    _, info = sc.render(cam, seed=0)
    sc.close()
    return {"host_sah_build_ms": host_stats["build_ms"], "gpu_lbvh_build_ms": best["build_ms"], "gpu_lbvh_input_upload_ms": best["input_upload_ms"],
            "host_sah_nodes": host_stats["wide_nodes"], "gpu_lbvh_nodes": best["wide_nodes"], "gpu_lbvh_depth": best["wide_depth"],
            "gpu_lbvh_max_stack": best["max_stack"], "gpu_lbvh_render_mrays_s": info["segments"] / info["kernel_ms"] / 1e3,
            "gpu_lbvh_render_kernel_ms": info["kernel_ms"]}


def reference_cuda_leg(rt, scene, hs, cam, W, H, spp):
    """north_star baseline 1: the reference's own render_kernel (src/camera.cu:17), unmodified, built for sm_100 in
    oracle/_ref, fed the same arrays.  Segments come from our kernel run on the reference's random stream (the reference
    counts primary samples only)."""
    from oracle import oracle_py as O
    if not O.ref_available():
        return {"unavailable": "oracle/_ref not built"}
    ref = O.Ref()
    sp, pl, mt, _ = hs.arrays()
    rs = ref.scene(sp, pl, mt)
    _, info = scene.render(cam, rng=rt.RNG_REFERENCE)
    O.gpu_render_ref(rs, cam, reps=1)  # warm-up
    _, ms = O.gpu_render_ref(rs, cam, reps=2)
    return {"kernel_ms": ms, "mrays_s": info["segments"] / (ms * 1e-3) / 1e6, "segments": info["segments"],
            "note": "unmodified reference render_kernel, -O3 -arch=sm_100, same scene arrays + reference-built BVH"}


if __name__ == "__main__":
    main()
