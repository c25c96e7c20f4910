#!/usr/bin/env python
"""renderKernel time per arithmetic (F32 / MIXED / F64) on cfg1-4 at 128 spp: what the fp64 sphere quadratic of the production mode costs."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader
pkg = rtb200_loader.load()
for name, (W, H, n) in {"random_book1": (600, 400, 128), "random": (600, 400, 128), "cornell": (500, 500, 128), "next_week_final": (400, 400, 32)}.items():
    ds = pkg.HostScene(name).upload(W, H)
    for pn, p in (("f32", pkg.PREC_F32), ("mixed", pkg.PREC_MIXED), ("f64", pkg.PREC_F64)):
        nn = n if pn != "f64" else max(n // 8, 4)
        ds.render(W, H, 4, seed=1, precision=p, flags=pkg.FLAG_NO_WAVEFRONT)
        ms = min(ds.render(W, H, nn, seed=2 + i, precision=p, flags=pkg.FLAG_NO_WAVEFRONT)[1]["kernel_ms"] for i in range(2))
        print("%-16s %-5s %8.2f ms %8.1f Msamples/s" % (name, pn, ms, W * H * nn / ms / 1e3), flush=True)
    ds.destroy()
