#!/usr/bin/env python
"""Wavefront kernel time on cfg4 against the guided-chunk divisor (RTB_WF_GUIDE: a slot asks for 1 / (guide * P) of the indices left)
at three frame sizes.  usage: python tools/wf_guide_sweep.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader
pkg = rtb200_loader.load()
W = H = 500
ds = pkg.HostScene("cornell").upload(W, H)
ds.render(W, H, 8, seed=1, flags=pkg.FLAG_WAVEFRONT)
for spp in (128, 500, 1000):
    for guide in ("1000000", "1", "2", "4", "8"):   # 1000000: chunks of 1 from the start (no accumulation in the slot)
        os.environ["RTB_WF_GUIDE"] = guide
        rs = [ds.render(W, H, spp, seed=5 + i, flags=pkg.FLAG_WAVEFRONT)[1] for i in range(2)]
        best = min(r["kernel_ms"] for r in rs)
        print("spp %4d guide %-8s %8.2f ms %7.1f Msamples/s rounds %d" % (spp, guide, best, W * H * spp / best / 1e3, rs[0]["rounds"]), flush=True)
ds.destroy()
