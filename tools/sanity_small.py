#!/usr/bin/env python
"""Every scene through rt_render / rt_trace at a tiny size (for compute-sanitizer runs)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader
pkg = rtb200_loader.load()
for name in pkg.SCENES:
    hs = pkg.HostScene(name, synth_n=20); ds = hs.upload(48, 32)
    for prec in (pkg.PREC_F32, pkg.PREC_MIXED, pkg.PREC_F64):
        img, st = ds.render(48, 32, 4, seed=3, precision=prec)
    img, st = ds.render(48, 32, 2, seed=3, counted=True)
    rays = np.random.default_rng(1).normal(size=(256, 7))
    ds.trace(rays, log_stride=8)
    ds.destroy()
    print(name, "ok", float(np.nan_to_num(img).sum()))
