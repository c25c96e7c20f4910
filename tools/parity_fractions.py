#!/usr/bin/env python
"""Observed same-stream radiance agreement of the CUDA path with the oracle, per scene and arithmetic: the numbers behind the
thresholds of tests/test_gpu_parity.py (VERDICT r1 item 1b asked for them to be recorded, not only asserted).

usage: python tools/parity_fractions.py out.json [--strict] [--spp 8] [--size 60x40]
Writes {scene/variant: {prec: {bit_equal, within_1e-12, within_1e-6, within_1e-3, nan_mask_agree, mean_ratio, max_rel}}}.
--strict renders both sides with RT_FLAG_STRICT_MATH (shared transcendental implementations, beta-order throughput in the oracle).
Test infrastructure: imports oracle/."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle as O  # noqa: E402
import rtb200_loader  # noqa: E402

pkg = rtb200_loader.load()
ap = argparse.ArgumentParser()
ap.add_argument("out")
ap.add_argument("--strict", action="store_true")
ap.add_argument("--spp", type=int, default=8)
ap.add_argument("--size", default="60x40")
ap.add_argument("--scenes", default="")
args = ap.parse_args()
W, H = (int(x) for x in args.size.split("x"))
earth = pkg.load_earth()
scenes = args.scenes.split(",") if args.scenes else pkg.SCENES
variants = [(n, False) for n in scenes] + [(n, True) for n in scenes if n not in ("cornell", "cornell_smoke", "synthetic", "quirks")]
precs = [("f32", pkg.PREC_F32, O.PREC_F32), ("mixed", pkg.PREC_MIXED, O.PREC_MIXED), ("f64", pkg.PREC_F64, O.PREC_F64)]
if hasattr(pkg, "PREC_ROBUST"):
    precs.insert(2, ("robust", pkg.PREC_ROBUST, O.PREC_ROBUST))
flags = getattr(pkg, "FLAG_STRICT_MATH", 0) if args.strict else 0
res = {"config": {"width": W, "height": H, "spp": args.spp, "depth": 50, "seed": 21, "strict_math": bool(args.strict),
                  "rel": "|gpu - oracle| / (|oracle| + 1e-3) per channel of the per-pixel radiance sums; a pixel counts when all three channels pass"}}
for name, lit in variants:
    hs = pkg.HostScene(name, lit=lit, synth_n=30)
    ds = hs.upload(W, H)
    osc = O.OracleScene(name, earth=earth if name in ("earth", "random", "next_week_final") else None, lights_mode=int(lit), synth_n=30)
    key = name + ("+lit" if lit else "")
    res[key] = {}
    for pn, gp, op in precs:
        img, _ = ds.render(W, H, args.spp, depth=50, seed=21, precision=gp, flags=flags)
        kw = dict(strict=True, order="beta") if args.strict else {}
        ref, _ = osc.render(W, H, args.spp, depth=50, prec=op, mode=1, seed=21, **kw)
        ng, nr = np.isnan(img).any(axis=2), np.isnan(ref).any(axis=2)
        ok = ~(ng | nr)
        a, b = img[ok], ref[ok]
        rel = (np.abs(a - b) / (np.abs(b) + 1e-3)).max(axis=1) if ok.any() else np.zeros(0)
        res[key][pn] = {
            "bit_equal": float((a == b).all(axis=1).mean()) if ok.any() else None,
            "within_1e-12": float((rel < 1e-12).mean()) if ok.any() else None,
            "within_1e-6": float((rel < 1e-6).mean()) if ok.any() else None,
            "within_1e-3": float((rel < 1e-3).mean()) if ok.any() else None,
            "nan_mask_agree": float((ng == nr).mean()), "finite_fraction": float(ok.mean()),
            "mean_ratio": float(a.mean() / b.mean()) if ok.any() and b.mean() != 0 else None,
            "max_rel": float(rel.max()) if ok.any() else None,
        }
        print(key, pn, res[key][pn], flush=True)
    ds.destroy()
json.dump(res, open(args.out, "w"), indent=1)
