#!/usr/bin/env python
"""Wavefront knob sweep on the Cornell box (environment overrides read at every launch).  usage: python tools/wf_sweep.py [spp]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader
pkg = rtb200_loader.load()
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 128
W = H = 500
hs = pkg.HostScene("cornell"); ds = hs.upload(W, H)
ds.render(W, H, 8, seed=1, flags=pkg.FLAG_WAVEFRONT)
base = min(ds.render(W, H, spp, seed=2 + i, flags=pkg.FLAG_NO_WAVEFRONT)[1]["kernel_ms"] for i in range(2))
print("renderKernel %.2f ms %.1f Msamples/s" % (base, W * H * spp / base / 1e3), flush=True)
for slots in (2, 4, 8, 16, 32):
    for item in (4, 8, 16):
        for bps in (8, 6):
            os.environ.update(RTB_WF_SLOTS_PER_LANE=str(slots), RTB_WF_ITEM_MIN=str(item), RTB_WF_BLOCKS_PER_SM=str(bps))
            ds.render(W, H, 8, seed=1, flags=pkg.FLAG_WAVEFRONT)
            rs = [ds.render(W, H, spp, seed=2 + i, flags=pkg.FLAG_WAVEFRONT)[1] for i in range(2)]
            best = min(r["kernel_ms"] for r in rs)
            print("slots/lane %2d item_min %2d blocks/SM %d: %8.2f ms %7.1f Msamples/s rounds %d" % (slots, item, bps, best, W * H * spp / best / 1e3, rs[0]["rounds"]), flush=True)
