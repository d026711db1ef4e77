#!/usr/bin/env python
"""bench.py -- Mrays/s of the per-pixel render path (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

A step = one frame of the workload (default: BASELINE.json configs[2], the 100k-sphere 1080p / 256 spp / depth 32 scene
the north_star target is quoted on) through the persistent megakernel, called through the C-ABI (rt_render_tiles).  A "ray"
is one path segment (one hit_scene call of the reference).  N>1: launched by torchrun, one rank per
GPU; every rank renders its interleaved tiles of the SAME frame (strong scaling) and the packed tiles are gathered to
rank 0 with the library's own grouped ncclSend / ncclRecv inside the step.  Prints ONE JSON line on rank 0.
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


_RECORD_FD = None


def emit(line):
    """the one JSON line of the run, on the process's original stdout"""
    sys.stdout.flush()
    data = (json.dumps(line) + "\n").encode()
    if _RECORD_FD is None:
        os.write(1, data)
    else:
        os.write(_RECORD_FD, data)


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
                                       "-lms", "25"], stdout=self.f, stderr=subprocess.DEVNULL)
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
    cfg, hs, tex = rt.workload(name)
    return cfg, hs, tex


def cpu_sample(hs, cfg, tex, kind_pref="reference", target_s=12.0):
    """reference CPU path on a bounded sample of the workload: every 8th pixel in x and y, spp scaled to ~target_s."""
    from oracle import oracle_py as O
    sp, pl, mt, _ = hs.arrays()
    threads = os.cpu_count() or 1
    if kind_pref == "reference" and O.ref_available():
        chk, kind = O.Ref(), "reference"
    else:
        chk, kind = O.Oracle(), "port"
    sc = chk.scene(sp, pl, mt, tex, 0 if tex is not None else -1)
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
    cfg, hs, tex = workload_scene(rt, args.workload)
    # size one step to ~6 s so warmup+steps stay within minutes
    base, sc, cam = cpu_sample(hs, cfg, tex, "reference", target_s=6.0)
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
    emit(line)


def latest_traffic_profile(workload):
    """dram__bytes_read+write of one render_mega launch from the newest committed ncu --set full summary of this workload"""
    import glob
    best = None
    for f in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_mega_*.json"))):
        try:
            d = json.load(open(f))
        except Exception:
            continue
        if d.get("workload", DEFAULT_WORKLOAD) == workload and "traffic_bytes_per_launch" in d:
            best = (d["traffic_bytes_per_launch"], os.path.relpath(f, ROOT))
    return best if best else (None, None)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--kernel", default="mega", choices=["mega", "wavefront"])
    ap.add_argument("--no-baselines", action="store_true", help="skip the cpu_baseline / reference-CUDA / parity legs")
    ap.add_argument("--top-nodes", type=int, default=-1)
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = max(args.warmup, 1)

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    # stdout carries exactly ONE line, the JSON record: native libraries write to fd 1 too (NCCL prints its version banner
    # there when a communicator is created), so fd 1 points at stderr while the benchmark runs and the record goes to the
    # saved descriptor
    global _RECORD_FD
    sys.stdout.flush()
    _RECORD_FD = os.dup(1)
    os.dup2(2, 1)

    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    import rtp_b200 as rt

    if rt.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device -- this benchmark has no CPU path (" + rt.lib().rt_last_error().decode() + ")")
    torch.cuda.set_device(local_rank)
    rt.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    comm = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        comm = rt.Comm.from_torch(dist, dev)   # the library's own NCCL communicator (rt_comm_init_rank), id broadcast by torch

    cfg, hs, tex = workload_scene(rt, args.workload)
    W, H, spp, depth = cfg["W"], cfg["H"], cfg["sqrt_spp"] ** 2, cfg["depth"]
    sp, pl, mt, _ = hs.arrays()
    scene = rt.Scene(sp, pl, mt, tex, 0 if tex is not None else -1)
    stats = scene.stats()
    cam = hs.camera(0)
    kernel = rt.KERNEL_MEGA if args.kernel == "mega" else rt.KERNEL_WAVEFRONT
    opts = dict(kernel=kernel, seed=0, top_nodes=args.top_nodes)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(host_fb=None):
        """one frame through rt_render_tiles: this rank's tiles, NCCL gather to rank 0, unpack (and, with host_fb, the D2H of
        the assembled frame on rank 0).  device_ms = CUDA events on the library's stream around render .. gather + unpack."""
        flush.zero_()              # L2 flush between iterations, outside the timed events
        torch.cuda.synchronize()
        _, info = scene.render_tiles(comm, cam, out=host_fb, want_host=host_fb is not None, **opts)
        return info

    for _ in range(max(args.warmup, 1)):
        step()
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    t_wall = time.perf_counter()
    dev_ms, kms, segs_total, rank_segs = 0.0, 0.0, 0, 0
    for _ in range(args.steps):
        info = step()
        dev_ms += info["device_ms"]
        kms += info["kernel_ms"]
        segs_total += info["segments_total"]       # rank 0: all ranks (the counts travel with the tiles)
        rank_segs += info["segments_rank"]
    barrier()
    t_wall = time.perf_counter() - t_wall
    clocks = sampler.stop() if sampler else None

    t = torch.tensor([dev_ms, kms, float(rank_segs)], dtype=torch.float64, device=dev)
    per_rank = None
    if world > 1:
        allt = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allt, t)
        allt = torch.stack(allt).cpu().numpy()
        ms_total, kms_max = float(allt[:, 0].max()), float(allt[:, 1].max())   # MAX over ranks
        segs_all = float(allt[:, 2].sum())
        k = allt[:, 1] / args.steps
        per_rank = {"kernel_ms": {"min": float(k.min()), "mean": float(k.mean()), "max": float(k.max())},
                    "kernel_ms_by_rank": [float(x) for x in k], "device_ms_by_rank": [float(x) for x in allt[:, 0] / args.steps],
                    "segments_by_rank": [float(x) for x in allt[:, 2] / args.steps]}
    else:
        ms_total, kms_max, segs_all = dev_ms, kms, float(rank_segs)
    value = segs_all / (ms_total * 1e-3) / 1e6
    if args.kernel == "mega":
        chunked = spp >= 32  # rt_render_opts.sample_chunk auto: render_mega + reduce_chunks
        launches_per_step = 1 + (1 if chunked else 0) + (2 if world > 1 else 0)  # render, chunk sum; N>1: segment stash + pack (ranks > 0) / unpack (rank 0)
    else:
        K = max(1, min(spp, (16 << 20) // max(1, ((W + 15) // 16) * ((H + 15) // 16) * 256 // world)))
        launches_per_step = ((spp + K - 1) // K) * (2 + 2 * depth)

    # ---- end to end through the public C-ABI call with HOST buffers: rt_render_tiles with a host framebuffer on rank 0 --
    # launch, this rank's tiles, gather, unpack, ONE download of the assembled frame.  Wall clock, barriers on both sides.
    host_fb = np.zeros((H, W, 3), np.float32) if rank == 0 else None
    step(host_fb)
    barrier()
    t0 = time.perf_counter()
    e2e_segs, e2e_d2h = 0, 0
    for _ in range(args.steps):
        i2 = scene.render_tiles(comm, cam, out=host_fb, want_host=rank == 0, **opts)[1]
        e2e_segs += i2["segments_total"]
        e2e_d2h = i2["d2h_bytes"]
    barrier()
    e2e_t = time.perf_counter() - t0
    if world > 1:
        te = torch.tensor([e2e_t], dtype=torch.float64, device=dev)
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e_t = te.item()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    e2e_value = e2e_segs / e2e_t / 1e6

    # ---- roofline of the dominant kernel (render_mega): FP32 issue, per SURVEY.md section 8d
    peaks, peak_src = load_peaks()
    e2e_opts = dict(opts, tile_rank=rank, tile_nranks=world)
    _, cinfo = scene.render(cam, out=host_fb, count_work=True, **e2e_opts)
    fpr, bpr, per_ray = flops_per_ray(cinfo)
    sm_mhz = float(peaks.get("sm_max_mhz", 1965.0))
    fp32_peak = torch.cuda.get_device_properties(dev).multi_processor_count * FP32_LANES_PER_SM * 2 * sm_mhz * 1e6 / 1e12
    kernel_rays_s = segs_all / world / (kms_max * 1e-3) if kms_max > 0 else 0.0  # per GPU, render kernel only
    achieved = kernel_rays_s * fpr / 1e12
    traffic, traffic_src = latest_traffic_profile(args.workload) if args.kernel == "mega" else (None, None)
    compulsory = stats["device_bytes"] + W * H * 12
    roofline = {"bound": "fp32_issue", "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s", "frac": achieved / fp32_peak,
                "traffic": traffic, "traffic_source": traffic_src, "compulsory_dram_bytes_per_launch": compulsory,
                "algorithmic_bytes_per_launch": bpr * segs_all / world / args.steps, "kernel": "render_mega" if args.kernel == "mega" else "wf_extend",
                "kernel_ms_per_launch": kms_max / args.steps, "flops_per_ray": fpr, "bytes_per_ray": bpr, "per_ray": per_ray,
                "peak_source": "148 SMs x 128 fp32 lanes x 2 x sm_max_mhz of MEASURED_PEAKS.json (%s); HBM %.0f GB/s (%s) is not the bound: "
                               "the working set is L2 resident" % (peak_src, peaks.get("hbm_gbs", 0), peak_src),
                # SURVEY 8d's bytes/ray are what the traversal READS (nodes, primitives, materials) -- served by L1/L2, not by HBM
                "l1_algorithmic_gbs": kernel_rays_s * bpr / 1e9,
                "dram_gbs_measured": (traffic / (kms_max / args.steps * 1e-3) / 1e9) if traffic else None}

    line = {"metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "spheres": stats["num_spheres"], "planes": stats["num_planes"], "width": W, "height": H,
                       "spp": spp, "max_depth": depth, "kernel": args.kernel, "rng": "philox4x32-10", "parallelism": "tiles%d" % world,
                       "floor_texture": "synthetic 2000x1330 RGBA float through the texture unit" if tex is not None else "none",
                       "l2": "flushed between timed steps (256 MB memset)", "bvh_nodes": stats["wide_nodes"],
                       "bvh_depth": stats["wide_depth"], "bvh_build_ms": stats["build_ms"],
                       "call": "rt_render_tiles (C-ABI): render own tiles -> grouped ncclSend/ncclRecv to rank 0 -> unpack"},
            "clocks": clocks, "kernel_only_mrays_s": kernel_rays_s / 1e6 * world,
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": 76, "d2h_bytes_per_step": int(e2e_d2h),
                    "call": "rt_render_tiles (76-byte camera struct in; ONE assembled host framebuffer out on rank 0)"},
            "gpu_launches": launches_per_step * args.steps, "wall_s_timed_region": t_wall, "roofline": roofline}
    if per_rank:
        line["per_rank"] = per_rank

    if world == 1 and not args.no_baselines:
        try:
            base, _, _ = cpu_sample(hs, cfg, tex, "reference", target_s=12.0)
            line["cpu_baseline"] = {k: base[k] for k in ("value", "unit", "cores", "kind", "sample")}
        except Exception as e:  # the CPU leg must never take the GPU number down
            line["cpu_baseline"] = {"value": None, "unit": "Mrays/s", "cores": 0, "kind": "unavailable", "sample": repr(e)}
        try:
            line["parity"] = parity_leg(rt, scene, hs, cam, sp, pl, mt, tex)
        except Exception as e:
            line["parity"] = {"unavailable": repr(e)}
        try:
            _, xi = scene.render(cam, out=host_fb, exact=True, **opts)   # same stream, same chunks, the reference's arithmetic
            line["exact_mode_mrays_s"] = xi["segments"] / xi["kernel_ms"] / 1e3
            line["exact_mode_kernel_ms"] = xi["kernel_ms"]
        except Exception as e:
            line["exact_mode_mrays_s"] = None
            line["exact_mode_error"] = repr(e)
        if W * H * spp <= 4 * 3840 * 2160 * 64:   # the 8K / 1024-spp frame would take the reference kernel minutes
            try:
                line["reference_cuda"] = reference_cuda_leg(rt, scene, hs, cam, sp, pl, mt, tex, kms_max / args.steps)
            except Exception as e:
                line["reference_cuda"] = {"unavailable": repr(e)}
        try:
            line["bvh_build"] = bvh_build_leg(rt, hs, cam, stats)
        except Exception as e:
            line["bvh_build"] = {"unavailable": repr(e)}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def parity_leg(rt, scene, hs, cam, sp, pl, mt, tex):
    """north_star's id / t gate ON THE BENCHMARKED SCENE AND ARITHMETIC: every pixel-centre ray of the frame through
    rt_trace_primary_ex(exact_math=0) -- the traversal + leaf code render_mega is compiled from -- against the checker: the
    oracle's id-recording walk of the reference-built BVH (the oracle only checks here; nothing it computes is timed or
    shipped), and against the library's own exact mode (which tests/ pin bit for bit to the reference's leaf tests)."""
    rays = rt.pixel_center_rays(cam)
    ty, ix, t = scene.trace_primary(rays, exact=False)
    ey, ei, et = scene.trace_primary(rays, exact=True)
    d = (ty != ey) | (ix != ei)
    ok = ~d & (ey >= 0)
    out = {"rays_checked": int(len(rays)), "arithmetic": "production (rt_trace_primary_ex exact_math=0)",
           "id_mismatch_rate_vs_exact_mode": float(d.mean()), "id_mismatches_vs_exact_mode": int(d.sum()),
           "max_rel_t_vs_exact_mode": float((np.abs(t[ok] - et[ok]) / et[ok]).max()) if ok.any() else 0.0}
    try:
        from oracle import oracle_py as O
        osc = O.Oracle().scene(sp, pl, mt)
        by, bi, brec = osc.trace(rays, 1)
        m = (ty != by) | (ix != bi)
        ok = ~m & (by >= 0)
        out.update({"checker": "oracle: walk of the reference-built BVH with the reference leaf tests (include/bvh.h:19-65)",
                    "id_mismatch_rate": float(m.mean()), "id_mismatches": int(m.sum()),
                    "max_rel_t": float((np.abs(t[ok] - brec[ok, 0]) / brec[ok, 0]).max()) if ok.any() else 0.0,
                    "exact_mode_id_mismatches": int(((ey != by) | (ei != bi)).sum())})
        osc.close()
    except Exception as e:
        out["checker"] = "oracle unavailable: " + repr(e)
    return out


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


def reference_cuda_leg(rt, scene, hs, cam, sp, pl, mt, tex, our_kernel_ms):
    """north_star baseline 1: the reference's own render_kernel (src/camera.cu:17), unmodified, built for sm_100 in
    oracle/_ref, fed the same arrays (and the same texture through its own texture object).  Segments come from our kernel
    run on the reference's random stream (the reference counts primary samples only)."""
    from oracle import oracle_py as O
    if not O.ref_available():
        return {"unavailable": "oracle/_ref not built"}
    ref = O.Ref()
    rs = ref.scene(sp, pl, mt, tex, 0 if tex is not None else -1)
    _, info = scene.render(cam, rng=rt.RNG_REFERENCE)
    O.gpu_render_ref(rs, cam, reps=1)  # warm-up
    _, ms = O.gpu_render_ref(rs, cam, reps=2)
    rs.close()
    return {"kernel_ms": ms, "mrays_s": info["segments"] / (ms * 1e-3) / 1e6, "segments": info["segments"],
            "speedup_kernel_ms": ms / our_kernel_ms if our_kernel_ms > 0 else None,
            "note": "unmodified reference render_kernel, -O3 -arch=sm_100, same scene arrays + reference-built BVH"}


if __name__ == "__main__":
    main()
