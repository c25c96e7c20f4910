#!/usr/bin/env python
"""Kernel time of renderKernelCoop on the bounded cfg5 samples (best of 3).  usage: [RTB_CORE_LIB=...] python tools/coop_probe.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader
pkg = rtb200_loader.load()
for sn, (W, H) in ((300, (480, 270)), (1000, (240, 135))):
    ds = pkg.HostScene("synthetic", synth_n=sn).upload(W, H)
    ds.render(W, H, 1, seed=1, flags=pkg.FLAG_COOP, region=(0, 8, 0, W))
    rs = [ds.render(W, H, 1, seed=3 + i, flags=pkg.FLAG_COOP)[1] for i in range(3)]
    best = min(rs, key=lambda r: r["kernel_ms"])
    print("%s synthetic n=%d %dx%d %9.1f ms %8.2f ksamples/s tail %.1f ms" % (os.environ.get("RTB_CORE_LIB", "default"), sn, W, H, best["kernel_ms"], W * H / best["kernel_ms"], best["tail_ms"]), flush=True)
    ds.destroy()
