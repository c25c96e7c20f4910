#!/usr/bin/env python
"""renderKernel time against the fast-path thresholds of the class vote (RTB_KEEP_NODE / RTB_KEEP_LEAF: run a node / sphere step without a
full vote when at least this many lanes wait for one) after the node step shrank from 75 to 62 instructions."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader
pkg = rtb200_loader.load()
for name, (W, H, n) in {"random": (600, 400, 128), "random_book1": (600, 400, 128), "next_week_final": (400, 400, 32), "cornell": (500, 500, 64)}.items():
    ds = pkg.HostScene(name).upload(W, H)
    ds.render(W, H, 4, seed=1, flags=pkg.FLAG_NO_WAVEFRONT)
    res = {}
    for kn in ("8", "10", "12", "14", "16", "20"):
        for kl in ("6", "8", "10", "12"):
            os.environ["RTB_KEEP_NODE"] = kn; os.environ["RTB_KEEP_LEAF"] = kl
            res[(kn, kl)] = min(ds.render(W, H, n, seed=2 + i, flags=pkg.FLAG_NO_WAVEFRONT)[1]["kernel_ms"] for i in range(2))
    base = res[("12", "8")]
    best = min(res, key=res.get)
    print("%-16s default 12/8 %.2f ms; best %s/%s %.2f ms (%+.1f %%)" % (name, base, best[0], best[1], res[best], 100 * (res[best] / base - 1)))
    print("   " + "  ".join("%s/%s=%.1f" % (k[0], k[1], v) for k, v in sorted(res.items(), key=lambda kv: (int(kv[0][0]), int(kv[0][1])))), flush=True)
    ds.destroy()
