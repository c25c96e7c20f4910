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
for slots, item, bps, sort in [(s_, i_, 8, 1) for s_ in (2, 4, 8, 16, 32) for i_ in (4, 8, 16)] + [(4, 8, 6, 1), (8, 8, 6, 1), (4, 8, 8, 0), (8, 8, 8, 0), (16, 8, 8, 0)]:
    if True:
        if True:
            os.environ.update(RTB_WF_SLOTS_PER_LANE=str(slots), RTB_WF_ITEM_MIN=str(item), RTB_WF_BLOCKS_PER_SM=str(bps), RTB_WF_SORT=str(sort))
            ds.render(W, H, 8, seed=1, flags=pkg.FLAG_WAVEFRONT)
            rs = [ds.render(W, H, spp, seed=2 + i, flags=pkg.FLAG_WAVEFRONT)[1] for i in range(2)]
            best = min(r["kernel_ms"] for r in rs)
            print("slots/lane %2d item_min %2d blocks/SM %d sort %d: %8.2f ms %7.1f Msamples/s rounds %d" % (slots, item, bps, sort, best, W * H * spp / best / 1e3, rs[0]["rounds"]), flush=True)
