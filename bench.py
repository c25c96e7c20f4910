#!/usr/bin/env python
"""bench.py — samples/s of the path-tracing hot path on N B200s (one process per GPU), see the contract in the task statement.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--scene NAME] [--spp S] [--impl reference]

A step = one full-frame render of `spp` camera samples per pixel per GPU (weak scaling: N GPUs render N*spp samples per
pixel, their partial framebuffers are summed onto rank 0 by one NCCL reduce inside the timed region).
Default workload = BASELINE.json configs[1]: `makeRandomScene earthTex 0 1 (randGen 1024)` + `randomSceneCamera (600,400)`
(The Next Week BVH scene: moving spheres, checker + earthmap image textures, glass cuboid), HEAD semantics, depth 50.
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
    "cornell": "cfg4 Rest of Your Life Cornell box, mixture-pdf light sampling (500x500)",
    "synthetic": "cfg5 synthetic 1M spheres behind a host-built BVH (3840x2160)",
}
SCENE_SIZE = {"random_book1": (600, 400), "random": (600, 400), "next_week_final": (400, 400), "cornell": (500, 500), "synthetic": (3840, 2160)}
SCENE_SPP = {"random_book1": 1024, "random": 1024, "next_week_final": 256, "cornell": 1000, "synthetic": 16}
FP32_PEAK_TFLOPS_NOMINAL = 148 * 128 * 2 * 1.965e9 / 1e12  # FFMA = 2 flop; no measured figure in MEASURED_PEAKS.json


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled every 200 ms during the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "200"],
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


def algorithmic_work(counters):
    """bytes and FP ops per sample from reference-order traversal counts."""
    c = counters
    n = max(c["paths"], 1)
    byts = 32 * c["box_tests"] + 16 * c["sphere_tests"] + 48 * c["moving_tests"] + 32 * c["rect_tests"] + 32 * c["cuboid_tests"] + 16 * c["bounces"] \
        + 4 * c["texels"] + 8 * 16 * c["noise_calls"]
    flops = 24 * c["box_tests"] + 32 * (c["sphere_tests"] + c["moving_tests"]) + 30 * c["sphere_hits"] + 12 * c["rect_tests"] + 72 * c["cuboid_tests"]
    return byts / n, flops / n, {k: c[k] / n for k in ("bounces", "box_tests", "sphere_tests", "moving_tests", "rect_tests", "cuboid_tests", "queries")}


def scene_bytes(d):
    """bytes of the flattened scene description that cross host->device in rt_scene_create"""
    n = 0
    n += d.n_nodes * (48 + 12) + d.n_spheres * 36 + d.n_moving * 76 + d.n_rects * 44 + d.n_cuboids * 52 + d.n_translates * 28 + d.n_rotates * 24
    n += d.n_media * 16 + d.n_materials * 16 + d.n_textures * 44 + d.n_perlin * (768 * 8 + 768 * 4)
    import ctypes as C
    for i in range(d.n_images):
        w = C.cast(d.image_w, C.POINTER(C.c_uint32))[i]; h = C.cast(d.image_h, C.POINTER(C.c_uint32))[i]
        n += w * h * 3
    return int(n)
