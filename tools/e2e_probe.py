#!/usr/bin/env python
"""Where the end-to-end call spends its time: rt_scene_create / rt_render (host buffers) / rt_scene_destroy, per step.
usage: python tools/e2e_probe.py scene spp [steps]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import rtb200_loader
pkg = rtb200_loader.load()
name, spp = sys.argv[1], int(sys.argv[2]); steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
W, H = {"random": (600, 400), "random_book1": (600, 400), "cornell": (500, 500), "next_week_final": (400, 400)}.get(name, (400, 400))
hs = pkg.HostScene(name); desc = hs.desc(W, H)
out = np.zeros((H, W, 3))
for i in range(steps + 1):
    t0 = time.perf_counter(); d = pkg.DeviceScene(desc, 0)
    t1 = time.perf_counter(); _, st = d.render(W, H, spp, depth=50, seed=500 + i, out=out)
    t2 = time.perf_counter(); d.destroy()
    t3 = time.perf_counter()
    print("%s step %d: create %.1f ms | render call %.1f ms (kernel %.1f, d2h %.2f) | destroy %.1f ms" %
          (name, i, (t1 - t0) * 1e3, (t2 - t1) * 1e3, st["kernel_ms"], st["d2h_ms"], (t3 - t2) * 1e3), flush=True)
