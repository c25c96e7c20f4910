#!/usr/bin/env python
"""Scheduler statistics of the instrumented render kernel for a scene: python tools/sched_stats.py scene [spp]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader
pkg = rtb200_loader.load()
name = sys.argv[1]; spp = int(sys.argv[2]) if len(sys.argv) > 2 else 8
W, H = {"random": (600, 400), "random_book1": (600, 400), "cornell": (500, 500), "next_week_final": (400, 400)}.get(name, (400, 400))
hs = pkg.HostScene(name); ds = hs.upload(W, H)
_, st = ds.render(W, H, spp, counted=True, seed=3)
s = st["sched"]; it = max(s[15], 1); names = ["node", "sphere", "rect", "rare", "shade"]
print(name, "iterations/sample %.1f" % (it / st["samples"]), "kernel_ms %.1f" % st["kernel_ms"])
for c in range(5):
    print("  %-6s ran %5.1f%% of iterations, lanes served/iter %5.1f, lanes waiting (avg over all iterations) %5.2f" % (names[c], 100 * s[c] / it, s[5 + c] / max(s[c], 1), s[10 + c] / it))
print("  live lanes avg %.1f" % (sum(s[10:15]) / it))
