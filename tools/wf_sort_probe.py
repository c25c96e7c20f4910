#!/usr/bin/env python
"""cfg4 wavefront kernel time against the shade kernel's material sort (RTB_WF_SORT) and the slots per resident lane, with guided chunks."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader
pkg = rtb200_loader.load()
W = H = 500
ds = pkg.HostScene("cornell").upload(W, H)
ds.render(W, H, 8, seed=1, flags=pkg.FLAG_WAVEFRONT)
for spp in (128, 1000):
    for slots in ("8", "16", "32"):
        for sort in ("0", "1"):
            os.environ["RTB_WF_SORT"] = sort; os.environ["RTB_WF_SLOTS_PER_LANE"] = slots
            rs = [ds.render(W, H, spp, seed=5 + i, flags=pkg.FLAG_WAVEFRONT)[1] for i in range(2)]
            best = min(r["kernel_ms"] for r in rs)
            print("spp %4d slots/lane %-3s sort %s %8.2f ms %7.1f Msamples/s rounds %d" % (spp, slots, sort, best, W * H * spp / best / 1e3, rs[0]["rounds"]), flush=True)
ds.destroy()
