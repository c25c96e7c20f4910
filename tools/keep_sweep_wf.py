#!/usr/bin/env python
"""cfg4 on the wavefront kernels against RTB_KEEP_NODE / RTB_KEEP_LEAF (the traversal kernel shares renderKernel's class vote)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader
pkg = rtb200_loader.load()
W = H = 500
ds = pkg.HostScene("cornell").upload(W, H)
ds.render(W, H, 8, seed=1, flags=pkg.FLAG_WAVEFRONT)
res = {}
for kn in ("4", "6", "8", "12"):
    for kl in ("2", "4", "6", "8"):
        os.environ["RTB_KEEP_NODE"] = kn; os.environ["RTB_KEEP_LEAF"] = kl
        res[(kn, kl)] = min(ds.render(W, H, 256, seed=2 + i, flags=pkg.FLAG_WAVEFRONT)[1]["kernel_ms"] for i in range(2))
        print("wavefront keep node/leaf %s/%s %.2f ms %.1f Msamples/s" % (kn, kl, res[(kn, kl)], W * H * 256 / res[(kn, kl)] / 1e3), flush=True)
ds.destroy()
