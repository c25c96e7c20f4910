#!/usr/bin/env python
"""debug build (-DRTB_DEBUG_UNSURE): how often the slab filter hands over to the refined test, per scene"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader
pkg = rtb200_loader.load()
for name, (W, H) in {"random": (600, 400), "random_book1": (600, 400), "cornell": (500, 500), "next_week_final": (400, 400)}.items():
    ds = pkg.HostScene(name).upload(W, H)
    _, st = ds.render(W, H, 8, seed=3, counted=True, flags=pkg.FLAG_NO_WAVEFRONT)
    c = st["counters"]; k = list(c.keys())
    print(name, "box_tests", c["box_tests"], "unsure", c[k[23]], "mode!=0", c[k[22]], "unsure frac %.5f" % (c[k[23]] / max(c["box_tests"], 1)), flush=True)
    ds.destroy()
