#!/usr/bin/env python
"""Kernel time of the production render kernel under scheduler-knob settings (environment overrides read at every launch).
usage: python tools/knob_sweep.py scene spp 'K1=v K2=v' 'K1=v' ..."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader
pkg = rtb200_loader.load()
name, spp = sys.argv[1], int(sys.argv[2])
W, H = {"random": (600, 400), "random_book1": (600, 400), "cornell": (500, 500), "next_week_final": (400, 400)}.get(name, (400, 400))
hs = pkg.HostScene(name); ds = hs.upload(W, H)
ds.render(W, H, 4, seed=1)
KN = ["RTB_SHADE_MIN", "RTB_KEEP_NODE", "RTB_KEEP_LEAF", "RTB_RECT_MIN", "RTB_RARE_MIN", "RTB_BLOCKS_PER_SM"]
for setting in sys.argv[3:]:
    for k in KN: os.environ.pop(k, None)
    for kv in setting.split():
        if "=" in kv:
            k, v = kv.split("="); os.environ["RTB_" + k] = v
    best = min(ds.render(W, H, spp, seed=2 + i)[1]["kernel_ms"] for i in range(3))
    print("%-18s %-44s %8.2f ms  %7.1f Msamples/s" % (name, setting, best, W * H * spp / best / 1e3), flush=True)
