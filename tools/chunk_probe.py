#!/usr/bin/env python
"""Tail behaviour of small launches: kernel time of the synthetic sphere grid (few, very long paths) and result identity of tiny
launches under RTB_CHUNK_MIN (path indices a warp takes per atomic; results never depend on it).
usage: python tools/chunk_probe.py synth_n W H"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader
pkg = rtb200_loader.load()
n, W, H = (int(x) for x in sys.argv[1:4])
cs = pkg.HostScene("cornell").upload(16, 16)
ref = {}
for cm in (32, 8, 2, 1):
    os.environ["RTB_CHUNK_MIN"] = str(cm)
    for (w, h, spp) in ((1, 1, 7), (3, 2, 1), (16, 16, 5), (16, 16, 64)):
        img = cs.render(w, h, spp, seed=4)[0]
        if cm == 32: ref[(w, h, spp)] = img
        assert np.allclose(img, ref[(w, h, spp)], rtol=1e-12, atol=1e-12), (cm, w, h, spp)
print("tiny launches identical under chunk min 32 / 8 / 2 / 1", flush=True)
hs = pkg.HostScene("synthetic", synth_n=n); ds = hs.upload(W, H)
ds.render(W, H, 1, seed=1)
base = None
for cm in (32, 8, 4, 1):
    os.environ["RTB_CHUNK_MIN"] = str(cm)
    r = [ds.render(W, H, 1, seed=5) for _ in range(2)]
    ms = min(x[1]["kernel_ms"] for x in r)
    if base is None: base = r[0][0]
    assert np.allclose(r[0][0], base, rtol=1e-12, atol=1e-12, equal_nan=True)
    print("synthetic n=%d %dx%d chunk min %2d: %8.1f ms  %7.1f ksamples/s" % (n, W, H, cm, ms, W * H / ms), flush=True)
