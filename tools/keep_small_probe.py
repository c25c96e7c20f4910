#!/usr/bin/env python
"""renderKernel on the small-BVH scenes with the library's keep thresholds (8 / 6 for <= 64 nodes) against 12 / 8."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader
pkg = rtb200_loader.load()
for name, lit in (("cornell_smoke", False), ("simple_light", True), ("two_perlin", True), ("earth", True), ("two_spheres", True), ("cornell", False), ("quirks", True)):
    ds = pkg.HostScene(name, lit=lit).upload(400, 400)
    ds.render(400, 400, 4, seed=1, flags=pkg.FLAG_NO_WAVEFRONT)
    out = []
    for kn, kl in (("", ""), ("12", "8"), ("8", "6")):
        for k, v in (("RTB_KEEP_NODE", kn), ("RTB_KEEP_LEAF", kl)):
            if v: os.environ[k] = v
            else: os.environ.pop(k, None)
        ms = min(ds.render(400, 400, 64, seed=2 + i, flags=pkg.FLAG_NO_WAVEFRONT)[1]["kernel_ms"] for i in range(2))
        out.append("%s/%s %.2f ms" % (kn or "lib", kl or "lib", ms))
    print("%-14s %s" % (name, "   ".join(out)), flush=True)
    ds.destroy()
