#!/usr/bin/env python
"""cfg5 probe: scheduler statistics + time of the synthetic sphere-grid scene at a given grid size.
usage: python tools/synth_probe.py synth_n W H spp"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader
pkg = rtb200_loader.load()
n, W, H, spp = (int(x) for x in sys.argv[1:5])
hs = pkg.HostScene("synthetic", synth_n=n); ds = hs.upload(W, H)
ds.render(W, H, 1, seed=1)
_, st = ds.render(W, H, spp, counted=True, seed=3)
s = st["sched"]; it = max(s[15], 1); names = ["node", "sphere", "rect", "rare", "shade"]
c = st.get("counters", {})
ns = st["samples"]
print("synth_n %d %dx%d spp %d: counted kernel %.1f ms, warp iterations/sample %.0f" % (n, W, H, spp, st["kernel_ms"], it / ns))
print("  per sample:", {k: round(v / ns, 1) for k, v in c.items() if v})
for k in range(5):
    print("  %-6s ran %5.1f%% of iterations, lanes served/iter %5.1f, waiting %5.2f" % (names[k], 100 * s[k] / it, s[5 + k] / max(s[k], 1), s[10 + k] / it))
print("  live lanes avg %.1f" % (sum(s[10:15]) / it))
best = min(ds.render(W, H, spp, seed=5 + i)[1]["kernel_ms"] for i in range(2))
print("  production kernel %.1f ms -> %.1f ksamples/s; warp-iteration rate %.2f G/s" % (best, W * H * spp / best, it / best / 1e6))
