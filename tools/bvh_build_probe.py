#!/usr/bin/env python
"""Scene-construction time of the synthetic sphere grid with makeBVH on the host vs on the device (rt_bvh_build).
usage: python tools/bvh_build_probe.py [synth_n ...]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import rtb200_loader
pkg = rtb200_loader.load()
from ray_tracing_b200._scene import desc_arrays
host = pkg.host()
pkg.core().rt_device_count()
for n in [int(x) for x in sys.argv[1:]] or [300, 1000]:
    t = time.time(); a = pkg.HostScene("synthetic", synth_n=n); th = time.time() - t
    old = host.rth_set_device_bvh(4096)
    pkg.HostScene("synthetic", synth_n=70)  # CUDA context / CUB warm-up outside the timed build (4900 spheres go to the device)
    t = time.time(); b = pkg.HostScene("synthetic", synth_n=n); td = time.time() - t
    host.rth_set_device_bvh(old)
    da, db = desc_arrays(a.desc(64, 36).contents), desc_arrays(b.desc(64, 36).contents)
    same = all(np.array_equal(np.asarray(da[k]), np.asarray(db[k])) for k in da)
    print("synth_n %d (%d spheres): scene build host makeBVH %.2f s | device makeBVH %.2f s (rt_bvh_build %.1f ms incl. copies) | identical arrays: %s"
          % (n, n * n, th, td, host.rth_last_device_bvh_ms(), same), flush=True)
