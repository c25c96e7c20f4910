#!/usr/bin/env python
"""Scheduler thresholds (RTB_RECT_MIN, RTB_RARE_MIN) on scenes they were NOT fitted on: renderKernel time with
the library's choice against every alternative, per scene.  Results never depend on the thresholds; this measures what a wrong choice costs.
usage: python tools/knob_other_scenes.py [spp]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader
pkg = rtb200_loader.load()
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 64
SCENES = [("quirks", True, 300, 200), ("cornell_smoke", False, 300, 300), ("two_perlin", True, 300, 200), ("earth", True, 300, 200), ("simple_light", True, 300, 200),
          ("next_week_final", False, 200, 200), ("cornell", False, 300, 300), ("random", False, 300, 200)]
for name, lit, W, H in SCENES:
    ds = pkg.HostScene(name, lit=lit).upload(W, H)
    ds.render(W, H, 4, seed=1, flags=pkg.FLAG_NO_WAVEFRONT)
    res = {}
    for rect in ("", "2", "3", "4", "6", "8"):
        for rare in ("", "4", "8"):
            if (rect == "") != (rare == ""):
                continue
            for k, v in (("RTB_RECT_MIN", rect), ("RTB_RARE_MIN", rare)):
                if v: os.environ[k] = v
                else: os.environ.pop(k, None)
            ms = min(ds.render(W, H, spp, seed=7 + i, flags=pkg.FLAG_NO_WAVEFRONT)[1]["kernel_ms"] for i in range(2))
            res[(rect or "lib", rare or "lib")] = ms
    base = res[("lib", "lib")]
    best = min(res, key=res.get)
    print("%-16s lit=%d library choice %8.2f ms; best %s %8.2f ms (%+.1f %%); all: %s" % (
        name, lit, base, best, res[best], 100 * (res[best] / base - 1), " ".join("%s/%s=%.1f" % (k[0], k[1], v) for k, v in sorted(res.items()))), flush=True)
    ds.destroy()
