#!/usr/bin/env python
"""bench.py — samples/s of the path-tracing hot path on N B200s, see the contract in the task statement.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--scene NAME] [--spp S] [--impl reference] [--no-per-scene]

A step = one full-frame render of `spp` camera samples per pixel per GPU in the PRODUCTION arithmetic (RT_PREC_MIXED: fp32 with the
sphere quadratic in fp64).  Headline workload = BASELINE.json configs[1]: `makeRandomScene earthTex 0 1 (randGen 1024)` +
`randomSceneCamera (600,400)` (The Next Week BVH scene), HEAD semantics, depth 50, 1024 spp.  The one JSON line also carries a
`per_scene` block — cfg1, cfg3, cfg4 and the bounded cfg5 samples (90 k and 1 M spheres), each with its own `roofline`,
`cpu_baseline` and `e2e` — because the metric is samples/s PER SCENE.

N > 1 (launched by torchrun, one rank per GPU): the headline `value` is weak scaling over torch.distributed (every rank renders
`spp` samples per pixel, one NCCL reduce of the double framebuffers per step).  Rank 0 ALONE then drives the drop-in path — ONE
rt_render call with n_gpus = N, the library shards and reduces over NVLink itself — for `e2e` (weak) and the `strong` block (the
fixed 1024-spp frame split N ways, with reduce_ms / tail_ms / scene_create_ms), and `sharding_check` compares the N-GPU frame with
the 1-GPU frame.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SCENE_DOC = {
    "random_book1": "cfg1 In One Weekend final scene (makeRandomSceneBookOne, 600x400)",
    "random": "cfg2 The Next Week BVH scene (makeRandomScene: moving spheres, checker, earthmap, glass cuboid; 600x400)",
    "next_week_final": "cfg3 The Next Week final scene (box field, media, instanced sphere cluster; 400x400)",
    "cornell": "cfg4 Rest of Your Life Cornell box, mixture-pdf light sampling (500x500x1000 = app/Main.hs)",
    "synthetic": "cfg5 synthetic sphere grid behind a host-built BVH; BOUNDED sample of the 3840x2160x1024-spp frame (the reference's independent-axis box test visits ~1.8 nodes+leaves per sphere per camera sample, so the stated frame is ~1.5e16 steps)",
}
SCENE_SIZE = {"random_book1": (600, 400), "random": (600, 400), "next_week_final": (400, 400), "cornell": (500, 500), "synthetic": (480, 270)}
SCENE_SPP = {"random_book1": 1024, "random": 1024, "next_week_final": 256, "cornell": 1000, "synthetic": 1}
CPU_SPP = {"random": 100, "random_book1": 100, "cornell": 100, "next_week_final": 16, "synthetic": 1}
FP32_NOMINAL_LANE_OPS = 148 * 128 * 1.965e9  # lane-instructions/s; fallback when profiles/fp32_peak.json is missing


def _json(path):
    try:
        return json.load(open(os.path.join(ROOT, path)))
    except (OSError, ValueError):
        return None


def measured_traffic(scene, W, H, spp, synth_n):
    """DRAM / L2 bytes of ONE renderKernel launch of exactly this workload from the committed ncu captures (profiles/r2_traffic.json,
    falling back to round 1's cfg2 entry); None when this workload was not captured."""
    for f in ("profiles/r2_traffic.json", "profiles/r1_traffic.json"):
        t = (_json(f) or {}).get(scene if scene != "synthetic" else "synthetic_%d" % synth_n)
        if t and (t["width"], t["height"], t["spp"]) == (W, H, spp):
            return t
    return None


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled every 500 ms during the timed regions."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "500"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            time.sleep(0.05)
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


# Shading flops per event (one op = one add / mul / div / compare / sqrt / transcendental, counted off the reference's formulas;
# DESIGN.md 6 has the derivation): camera ray (Lib.hs:1253-1267, 1362-1365), Lambertian scatter incl. onbFromW, cosine direction,
# mixture pdf value and the throughput update (823-836, 367-382, 1329-1333), Metal (837-841), Dielectric (842-859, 887-903), Isotropic
# (861-878), emitted (880-885), htblRandom per light sample (707-724, rect/sphere mix), htblPdfValue per leaf (673-705), checker level
# (498-501), Perlin noise call (441-484; turb's 7 calls per marble lookup are counted by noise_calls), image lookup (504-510).
SHADE_FLOPS = {"paths": 50, "scatter_lambertian": 105, "scatter_metal": 42, "scatter_dielectric": 60, "scatter_isotropic": 36, "emitted": 3,
               "light_samples": 40, "light_pdf_leaves": 20, "checker_evals": 9, "noise_calls": 146, "texels": 12}


def algorithmic_work(c):
    """bytes and FP ops per sample from reference-order traversal counts (+ the shading events)."""
    n = max(c["paths"], 1)
    byts = 32 * c["box_tests"] + 16 * c["sphere_tests"] + 48 * c["moving_tests"] + 32 * c["rect_tests"] + 32 * c["cuboid_tests"] + 16 * c["bounces"] \
        + 4 * c["texels"] + 8 * 16 * c["noise_calls"]
    trav = 24 * c["box_tests"] + 32 * (c["sphere_tests"] + c["moving_tests"]) + 30 * c["sphere_hits"] + 12 * c["rect_tests"] + 72 * c["cuboid_tests"]
    shade = sum(SHADE_FLOPS[k] * c[k] for k in SHADE_FLOPS)
    keys = ("bounces", "box_tests", "sphere_tests", "moving_tests", "rect_tests", "cuboid_tests", "queries", "scatter_lambertian", "scatter_metal",
            "scatter_dielectric", "scatter_isotropic", "emitted", "light_samples", "light_pdf_leaves", "noise_calls", "texels")
    return dict(bytes=byts / n, flops=(trav + shade) / n, flops_traversal=trav / n, flops_shading=shade / n, per_sample={k: c[k] / n for k in keys})


def scene_bytes(d):
    """bytes of the flattened scene description that cross host->device in rt_scene_create"""
    import ctypes as C
    n = d.n_nodes * (48 + 12) + d.n_spheres * 36 + d.n_moving * 76 + d.n_rects * 44 + d.n_cuboids * 52 + d.n_translates * 28 + d.n_rotates * 24
    n += d.n_media * 16 + d.n_materials * 16 + d.n_textures * 44 + d.n_perlin * (768 * 8 + 768 * 4)
    for i in range(d.n_images):
        n += C.cast(d.image_w, C.POINTER(C.c_uint32))[i] * C.cast(d.image_h, C.POINTER(C.c_uint32))[i] * 3
    return int(n)


def cpu_sample(scene, W, H, synth_n, spp, seed, threads):
    """(w, h, spp, seconds) of the oracle<double> in the reference's own RNG plumbing on a bounded sample of the workload"""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O
    import rtb200_loader
    pkg = rtb200_loader.load()
    earth = pkg.load_earth() if scene in ("earth", "random", "next_week_final") else None
    osc = O.OracleScene(scene, earth=earth, synth_n=synth_n)
    w, h = (W, H) if scene != "synthetic" else ((48, 27) if synth_n <= 300 else (24, 14))
    t = time.time()
    osc.render(w, h, spp, depth=50, prec=O.PREC_F64, mode=0, seed=seed, threads=threads)
    return w, h, spp, time.time() - t, osc


def run_reference(args, W, H, rank):
    """--impl reference: the reference's CPU algorithm (oracle<double>, runRender's per-column SplitMix plumbing) on the host cores.
    The Haskell build itself cannot run here (no GHC on either box); see DESIGN.md 9."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O
    import rtb200_loader
    pkg = rtb200_loader.load()
    earth = pkg.load_earth() if args.scene in ("earth", "random", "next_week_final") else None
    osc = O.OracleScene(args.scene, earth=earth, synth_n=args.synth_n)
    cores = os.cpu_count() or 1
    spp = (args.ref_spp or 32) if args.scene != "synthetic" else 1
    w, h = (W, H) if args.scene != "synthetic" else (48, 27)
    times = []
    for i in range(args.warmup + args.steps):
        t = time.time()
        osc.render(w, h, spp, depth=50, prec=O.PREC_F64, mode=0, seed=1000 + i, threads=cores)
        if i >= args.warmup:
            times.append(time.time() - t)
    tot = sum(times)
    v = args.steps * w * h * spp / tot
    sample = "%dx%d, %d spp per step (full frame, reduced spp), depth 50" % (w, h, spp)
    emit({
        "impl": "reference", "metric": "samples/sec", "value": v, "unit": "samples/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * tot / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": SCENE_DOC[args.scene], "scene": args.scene, "width": w, "height": h, "spp_per_step": spp, "max_depth": 50},
        "cpu_baseline": {"value": v, "unit": "samples/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0})


class Ctx:
    pass


def roofline_block(scene, W, H, spp, synth_n, work, kernel_ms, peaks, fp32, fast):
    samples = W * H * spp
    secs = kernel_ms / 1e3
    tr = measured_traffic(scene, W, H, spp, synth_n)
    hbm_peak = peaks["hbm_gbs"] if peaks else 6650.0
    ach_gbs = work["bytes"] * samples / secs / 1e9
    ach_tf = work["flops"] * samples / secs / 1e12
    lane_peak = (fp32["fadd_fmul_lane_inst_per_s"] if fp32 else FP32_NOMINAL_LANE_OPS) / 1e12
    ffma_peak = (fp32["ffma_tflops"] if fp32 else 2 * FP32_NOMINAL_LANE_OPS / 1e12)
    mem = {"algorithmic_bytes_per_sample": work["bytes"], "algorithmic_gbs": ach_gbs, "hbm_peak_gbs": hbm_peak,
           "hbm_peak_source": "MEASURED_PEAKS.json hbm_gbs (burst)" if peaks else "fallback 6.65 TB/s",
           "l2_peak_gbs": fp32["l2_read_gbs"] if fp32 else None,
           "dram_bytes_per_launch": (tr["dram_bytes_read"] + tr["dram_bytes_write"]) if tr else None,
           "lts_bytes_per_launch": tr.get("lts_bytes") if tr else None}
    if tr:
        mem["dram_gbs"] = mem["dram_bytes_per_launch"] / secs / 1e9
        mem["dram_frac_of_hbm_peak"] = mem["dram_gbs"] / hbm_peak
        if tr.get("lts_bytes") and mem["l2_peak_gbs"]:
            mem["l2_gbs"] = tr["lts_bytes"] / (tr.get("kernel_ms", kernel_ms) / 1e3) / 1e9
            mem["l2_frac_of_l2_peak"] = mem["l2_gbs"] / mem["l2_peak_gbs"]
    if work.get("kernel") == "wavefront" and work.get("rounds"):
        # the wavefront frame is ~2 launches per round: DRAM traffic per frame ~ rounds x (one captured full-occupancy round of each kernel);
        # an upper estimate (the drain rounds move less), reported beside the null per-launch `traffic`
        tt = _json("profiles/r2_traffic.json") or {}
        ra, rb = tt.get(scene + "_wfTraverse_round"), tt.get(scene + "_wfShade_round")
        if ra and rb and all((r["width"], r["height"], r["spp"]) == (W, H, spp) for r in (ra, rb)):
            per_round = sum(r["dram_bytes_read"] + r["dram_bytes_write"] for r in (ra, rb))
            mem.update({"wavefront_rounds": work["rounds"], "dram_bytes_per_round_captured": per_round,
                        "round_ms_captured": ra["kernel_ms"] + rb["kernel_ms"], "dram_bytes_per_frame_estimate": per_round * work["rounds"],
                        "dram_gbs_estimate": per_round * work["rounds"] / secs / 1e9,
                        "dram_frac_of_hbm_peak_estimate": per_round * work["rounds"] / secs / 1e9 / hbm_peak,
                        "source": "profiles/r2_ncu_%s_wfTraverseKernel.txt + _wfShadeKernel.txt (one mid-frame round each, --set full)" % scene})
    if scene == "synthetic":
        # the one memory-side config: 1 M spheres = 33.5 MB of nodes + 16 MB of spheres, L2-resident on B200 (126 MB): node / leaf
        # fetch bandwidth against the measured L2 and HBM peaks
        return {"bound": "hbm", "achieved": ach_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": ach_gbs / hbm_peak,
                "traffic": mem["dram_bytes_per_launch"], "kernel": work.get("kernel", "renderKernel"), "kernel_ms": kernel_ms, "peak_source": mem["hbm_peak_source"],
                "l2": {"achieved_algorithmic_gbs": ach_gbs, "peak": mem["l2_peak_gbs"], "frac": ach_gbs / mem["l2_peak_gbs"] if mem["l2_peak_gbs"] else None,
                       "peak_source": "profiles/fp32_peak.json l2_read_gbs (tools/fp32_peak.cu, 48 MB buffer)"},
                "memory": mem, "algorithmic_flops_per_sample": work["flops"], "per_sample": work["per_sample"],
                "note": "scene is L2-resident: algorithmic node/leaf bytes are served by L1/L2, DRAM traffic is the measured `traffic`"}
    return {"bound": "fp32-issue", "achieved": ach_tf, "peak": ffma_peak if fast else lane_peak, "unit": "TFLOP/s",
            "frac": ach_tf / (ffma_peak if fast else lane_peak), "traffic": mem["dram_bytes_per_launch"],
            "peak_source": ("profiles/fp32_peak.json (tools/fp32_peak.cu, measured on this pool's B200): " if fp32 else "nominal 148 SM x 128 lanes x 1.965 GHz: ") +
                           ("FFMA issue rate x 2" if fast else "non-FMA FADD/FMUL issue rate, 1 flop per lane-instruction — the literal build is compiled --fmad=false, "
                            "so no algorithmic op can ride an FMA"),
            "frac_of_ffma_peak": ach_tf / ffma_peak, "ffma_peak_tflops": ffma_peak,
            "kernel": work.get("kernel", "renderKernel"), "kernel_ms": kernel_ms, "algorithmic_flops_per_sample": work["flops"],
            "flops_traversal_per_sample": work["flops_traversal"], "flops_shading_per_sample": work["flops_shading"], "per_sample": work["per_sample"],
            "memory": mem,
            "note": "scene is L1/L2-resident; the binding limit is FP32/ALU issue under divergence (ncu: smsp__issue_active, "
                    "smsp__thread_inst_executed_per_inst_executed in profiles/)"}


def measure(cx, scene, spp, steps, warmup, synth_n=1000, size=None, cpu=True, e2e=True, label=None):
    """One workload: device-timed value (weak scaling over torch.distributed when world > 1), kernel-only time, roofline, e2e through
    the C ABI with host buffers, CPU baseline.  Returns the dict for the JSON line (rank 0) or None."""
    import torch
    import torch.distributed as dist
    pkg, args = cx.pkg, cx.args
    from ray_tracing_b200 import _dist
    W, H = size or SCENE_SIZE[scene]
    world, rank, dev, stream = cx.world, cx.rank, cx.dev, cx.stream
    flags = pkg.FLAG_FAST_MATH if args.fast else 0
    prec = args.precision
    t0 = time.perf_counter()
    hs = pkg.HostScene(scene, synth_n=synth_n)
    desc = hs.desc(W, H)
    host_build_s = time.perf_counter() - t0
    ds = pkg.DeviceScene(desc, cx.local)
    fb = torch.zeros((H, W, 3), dtype=torch.float64, device=dev)
    total_spp = spp * world  # weak scaling: every GPU renders `spp` samples per pixel

    work = None
    if rank == 0:  # algorithmic work per sample (instrumented kernel, small spp; the counters equal the oracle's, tests/test_gpu_strict.py)
        if scene == "synthetic":
            _, st = ds.render(min(W, 96), min(H, 54), 1, counted=True, seed=99, precision=prec, flags=flags)
        else:
            _, st = ds.render(W, H, 2, counted=True, seed=99, precision=prec, flags=flags)
        work = algorithmic_work(st["counters"])
        work["kernel"] = ds.render(W, H, 1, region=(0, 1, 0, min(W, 8)), precision=prec, flags=flags)[1]["kernel"]  # which kernel this scene selects

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    launches = []  # kernels launched per timed step (the wavefront path launches two per round)

    def step(i, ev=None):
        cx.flush.zero_()  # L2 flush between iterations
        fb.zero_()
        s0, n = _dist.shard_samples(total_spp, rank, world)
        if ev:
            ev[0].record()
        st_ = ds.render_device(fb.data_ptr(), W, H, n, stream=stream, depth=50, seed=1 + i, s0=s0, precision=prec, flags=flags, total=total_spp, timed=bool(ev))
        if ev:
            ev[1].record()
            launches.append(st_["launches"])
        _dist.reduce_framebuffer(fb, 0)

    for i in range(warmup):
        step(i)
    sync()
    k_ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        step(100 + i, k_ev[i])
    e1.record()
    sync()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in k_ev]))
    value = steps * W * H * total_spp / (ms / 1e3)
    finite_frac = float(torch.isfinite(fb).all(dim=2).float().mean().item()) if rank == 0 else None

    out = None
    if rank == 0:
        out = {"scene": scene, "workload": SCENE_DOC[scene] if scene != "synthetic" else SCENE_DOC[scene] + "; %d spheres, %dx%d x %d spp" % (synth_n * synth_n, W, H, spp),
               "width": W, "height": H, "spp_per_gpu_per_step": spp, "steps": steps, "warmup": warmup,
               "value": value, "unit": "samples/s", "ms_per_step": ms / steps, "kernel_ms": kernel_ms, "finite_pixel_fraction": finite_frac,
               "roofline": roofline_block(scene, W, H, spp, synth_n, dict(work, rounds=(float(np.mean(launches)) - 1.0) / 2.0 if work.get("kernel") == "wavefront" else None),
                                          kernel_ms, cx.peaks, cx.fp32, args.fast), "gpu_launches": int(sum(launches))}
        if scene == "synthetic":
            out["host_scene_build_s"] = host_build_s

    # ---- e2e: the user-facing call with HOST buffers.  N = 1: rt_scene_create (H2D) + rt_render (D2H of the frame inside), every step.
    #      N > 1: rank 0 alone, ONE rt_render with n_gpus = N (scene replicated by the library on the first call, NVLink reduce, one D2H)
    if e2e:
        cx.cpu_barrier()
        if rank == 0:
            host_out = np.zeros((H, W, 3))
            stats = []

            def e2e_step(i, n_spp):
                d2 = pkg.DeviceScene(desc, cx.local)
                _, st = d2.render(W, H, n_spp, depth=50, seed=500 + i, out=host_out, precision=prec, flags=flags, n_gpus=world)
                d2.destroy()
                return st
            e2e_step(0, total_spp)
            t0 = time.perf_counter()
            for i in range(steps):
                stats.append(e2e_step(1 + i, total_spp))
            e2e_s = time.perf_counter() - t0
            out["e2e"] = {"value": steps * W * H * total_spp / e2e_s, "unit": "samples/s", "h2d_bytes_per_step": scene_bytes(desc.contents) * world,
                          "d2h_bytes_per_step": W * H * 3 * 8, "ms_per_step": 1e3 * e2e_s / steps,
                          "path": "rt_scene_create + rt_render(n_gpus=%d) with host buffers, every step" % world,
                          "kernel_ms": float(np.mean([s["kernel_ms"] for s in stats])), "reduce_ms": float(np.mean([s["reduce_ms"] for s in stats])),
                          "replicate_ms": float(np.mean([s["replicate_ms"] for s in stats])), "tail_ms": float(np.mean([s["tail_ms"] for s in stats])),
                          "reduce_kind": stats[-1]["reduce_kind"], "gpus": stats[-1]["gpus"]}
            if world > 1:
                # strong scaling through the same call: the FIXED frame (spp total) split over the devices; scene kept resident
                d2 = pkg.DeviceScene(desc, cx.local)
                d2.render(W, H, spp, depth=50, seed=700, out=host_out, precision=prec, flags=flags, n_gpus=world)  # replicates
                res = {}
                for n in sorted({1, world}):
                    sts = []
                    d2.render(W, H, spp, depth=50, seed=701, out=host_out, precision=prec, flags=flags, n_gpus=n)
                    t0 = time.perf_counter()
                    for i in range(steps):
                        sts.append(d2.render(W, H, spp, depth=50, seed=710 + i, out=host_out, precision=prec, flags=flags, n_gpus=n)[1])
                    dt = time.perf_counter() - t0
                    res[n] = {"value": steps * W * H * spp / dt, "ms_per_step": 1e3 * dt / steps, "kernel_ms": float(np.mean([s["kernel_ms"] for s in sts])),
                              "reduce_ms": float(np.mean([s["reduce_ms"] for s in sts])), "tail_ms": float(np.mean([s["tail_ms"] for s in sts])),
                              "d2h_ms": float(np.mean([s["d2h_ms"] for s in sts])), "gpu_kernel_ms": sts[-1]["gpu_kernel_ms"]}
                t0 = time.perf_counter()
                d3 = pkg.DeviceScene(desc, cx.local)
                create_ms = 1e3 * (time.perf_counter() - t0)
                d3.destroy()
                out["strong"] = {"scaling": "strong", "frame": "%dx%d x %d spp split over %d GPUs inside one rt_render" % (W, H, spp, world),
                                 "value": res[world]["value"], "unit": "samples/s", "one_gpu_value_same_call": res[1]["value"],
                                 "efficiency": res[world]["value"] / (world * res[1]["value"]), "n": res[world], "one": res[1], "scene_create_ms": create_ms,
                                 "limiter": "per step at N GPUs: kernel_ms (of which tail_ms is the persistent kernel's drain) + reduce_ms + d2h_ms + host overhead "
                                            "(ms_per_step - the three)"}
                # N-GPU frame == 1-GPU frame (small spp, outside every timed region)
                a, _ = d2.render(W, H, 8, depth=50, seed=31, precision=prec, flags=flags, n_gpus=1)
                b, _ = d2.render(W, H, 8, depth=50, seed=31, precision=prec, flags=flags, n_gpus=world)
                okm = np.isfinite(a)
                out["sharding_check"] = {"max_rel_diff": float(np.max(np.abs(a[okm] - b[okm]) / (np.abs(a[okm]) + 1e-3))) if okm.any() else 0.0,
                                         "nan_masks_equal": bool(np.array_equal(np.isnan(a), np.isnan(b))), "spp": 8, "n_gpus": world, "bar": 1e-12}
                d2.destroy()
        cx.cpu_barrier()
        if world > 1:
            # the torch.distributed path's own check: sum of the ranks' shards vs rank 0's single-GPU render of the same small frame
            chk = torch.zeros((H, W, 3), dtype=torch.float64, device=dev)
            _dist.render_sharded(ds, W, H, 8, 50, 31, chk, stream=stream, precision=prec, flags=flags)
            torch.cuda.synchronize()
            if rank == 0:
                a, _ = ds.render(W, H, 8, depth=50, seed=31, precision=prec, flags=flags, n_gpus=1)
                b = chk.cpu().numpy()
                okm = np.isfinite(a)
                out["sharding_check"]["dist_max_rel_diff"] = float(np.max(np.abs(a[okm] - b[okm]) / (np.abs(a[okm]) + 1e-3))) if okm.any() else 0.0
            cx.cpu_barrier()

    if rank == 0 and cpu and world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        ref_spp = args.ref_spp or CPU_SPP[scene]
        w, h, s, dt, _ = cpu_sample(scene, W, H, synth_n, ref_spp, 4242, cores)
        out["cpu_baseline"] = {"value": w * h * s / dt, "unit": "samples/s", "cores": cores, "kind": "port",
                               "sample": "%dx%d at %d spp, depth 50, oracle<double> in runRender's RNG plumbing, %d threads (%.1f s)" % (w, h, s, cores, dt)}
    ds.destroy()
    return out


_OUT = None


def emit(obj):
    """the ONE JSON line, on the process's original stdout"""
    (_OUT or sys.stdout).write(json.dumps(obj) + "\n")
    (_OUT or sys.stdout).flush()


def main():
    # stdout carries ONE JSON line: libraries that write to fd 1 (NCCL prints its version banner there) are pointed at stderr
    global _OUT
    sys.stdout.flush()
    _OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--scene", default="random", choices=list(SCENE_DOC))
    ap.add_argument("--spp", type=int, default=0)
    ap.add_argument("--ref-spp", type=int, default=0, help="spp of the CPU sample (default: 100 for the cpu_baseline leg = perflogs.org:1, 32 per step for --impl reference)")
    ap.add_argument("--synth-n", type=int, default=1000)
    ap.add_argument("--size", default="", help="WxH override (synthetic)")
    ap.add_argument("--fast", action="store_true", help="RT_FLAG_FAST_MATH kernels (not bit-exact; reported separately)")
    ap.add_argument("--precision", type=int, default=1, help="0 fp32 (opt-in, biased on Book-1/2 scenes), 1 = RT_PREC_MIXED the production arithmetic, 2 fp64")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-per-scene", action="store_true", help="only the headline workload (profiling runs)")
    ap.add_argument("--profile", action="store_true", help="under ncu: headline workload only, no e2e loop, no CPU baseline")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    W, H = SCENE_SIZE[args.scene]
    if args.size:
        W, H = (int(x) for x in args.size.split("x"))
    spp = args.spp or SCENE_SPP[args.scene]
    if args.impl == "reference":
        return run_reference(args, W, H, rank)

    import torch
    import torch.distributed as dist
    import rtb200_loader
    cx = Ctx()
    cx.pkg = rtb200_loader.load(); cx.args = args; cx.rank, cx.world, cx.local = rank, world, local
    torch.cuda.set_device(local)
    cx.dev = torch.device("cuda", local)
    gloo = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=cx.dev)
        gloo = dist.new_group(backend="gloo")  # host-side barrier for the rank-0-only sections (an NCCL barrier would spin on the other GPUs)

    def cpu_barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier(group=gloo)
    cx.cpu_barrier = cpu_barrier
    cx.flush = torch.empty(256 << 20, dtype=torch.uint8, device=cx.dev)  # > 126 MB L2
    cx.stream = torch.cuda.current_stream().cuda_stream
    cx.peaks = _json("MEASURED_PEAKS.json")
    cx.fp32 = _json("profiles/fp32_peak.json")

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    if args.profile:
        args.no_per_scene = args.no_cpu_baseline = True
    head = measure(cx, args.scene, spp, args.steps, args.warmup, synth_n=args.synth_n, size=(W, H), e2e=not args.profile)
    # the sampler runs through both timed regions of the headline workload (terminating nvidia-smi stalls the context for up to a few
    # hundred ms, which the wall-clocked end-to-end loop would see) and is stopped only here
    clocks = sampler.stop() if rank == 0 else None

    per_scene = {}
    if not args.no_per_scene and args.scene == "random" and not args.fast:
        small = max(2, min(args.steps, 3))
        if world == 1:
            plan = [("cfg1", "random_book1", 1024, small, 2, 1000, None), ("cfg3", "next_week_final", 256, 2, 1, 1000, None),
                    ("cfg4", "cornell", 1000, small, 2, 1000, None), ("cfg5_90k", "synthetic", 1, 2, 1, 300, (480, 270)),
                    ("cfg5_1m", "synthetic", 1, 1, 1, 1000, (240, 135))]
        else:
            # cfg5 at N > 1: 8 spp of the bounded 240x135 sample.  One path of this scene is ~0.3 s of one warp, so a 1-spp frame (32 400 paths)
            # cannot be split 8 ways (profiles/r2_bench_default_n8_1spp.json: efficiency 0.21, every GPU waits for its slowest path); 8 spp is
            # the smallest frame whose split leaves every device the single-GPU sample's number of paths
            plan = [("cfg3", "next_week_final", 256, 2, 1, 1000, None), ("cfg4", "cornell", 1000, 2, 1, 1000, None),
                    ("cfg5_1m", "synthetic", 8, 1, 1, 1000, (240, 135))]
        for key, sc, s, st, wu, sn, size in plan:
            r = measure(cx, sc, s, st, wu, synth_n=sn, size=size)
            if rank == 0:
                per_scene[key] = r

    if rank == 0:
        e2e = head.pop("e2e", None)
        line = {
            "metric": "samples/sec", "value": head["value"], "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": ["f32", "f32 + f64 sphere quadratic (RT_PREC_MIXED, the production arithmetic)", "f64"][args.precision], "data": "synthetic",
            "config": {"workload": head["workload"], "scene": args.scene, "width": W, "height": H, "spp_per_gpu_per_step": spp, "max_depth": 50,
                       "seed": "randGen 1024", "lights": "HEAD", "parallelism": "spp-sharded x%d + ncclReduce (torch.distributed)" % world,
                       "l2": "flushed between steps (256 MiB write)",
                       "arithmetic": "fast-math" if args.fast else "literal (no FMA contraction, IEEE div/sqrt)", "finite_pixel_fraction": head["finite_pixel_fraction"]},
            "roofline": head["roofline"], "e2e": e2e, "gpu_launches": head["gpu_launches"], "clocks": clocks,
        }
        for k in ("cpu_baseline", "strong", "sharding_check"):
            if k in head:
                line[k] = head[k]
        if per_scene:
            line["per_scene"] = per_scene
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
