#!/usr/bin/env python
"""Kernel time of renderKernel per arithmetic mode on the four reference configs (best of 3, CUDA events inside rt_render).
usage: python tools/prec_probe.py [spp]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader
pkg = rtb200_loader.load()
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 128
SZ = {"random_book1": (600, 400), "random": (600, 400), "next_week_final": (400, 400), "cornell": (500, 500)}
precs = [("f32", pkg.PREC_F32), ("mixed", pkg.PREC_MIXED)] + ([("robust", pkg.PREC_ROBUST)] if hasattr(pkg, "PREC_ROBUST") else [])
out = {}
for name, (W, H) in SZ.items():
    hs = pkg.HostScene(name); ds = hs.upload(W, H)
    n = spp if name != "next_week_final" else max(spp // 4, 8)
    for pn, p in precs:
        ds.render(W, H, 4, seed=1, precision=p)
        best = min(ds.render(W, H, n, seed=2 + i, precision=p)[1]["kernel_ms"] for i in range(3))
        out.setdefault(name, {})[pn] = W * H * n / best / 1e3
        print("%-16s %-7s %8.2f ms %8.1f Msamples/s" % (name, pn, best, W * H * n / best / 1e3), flush=True)
    ds.destroy()
print(json.dumps(out))
