#!/usr/bin/env python
"""Kernel time of the alternative kernels on the reference configs (best of 3, CUDA events inside rt_render):
renderKernel vs wavefront on cfg1-4, renderKernel vs renderKernelCoop on the bounded cfg5 samples.
usage: python tools/kernel_probe.py [spp] [cfg5: 0/1]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader
pkg = rtb200_loader.load()
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 128
do5 = len(sys.argv) > 2 and sys.argv[2] == "1"
SZ = {"cornell": (500, 500), "random_book1": (600, 400), "random": (600, 400), "next_week_final": (400, 400)}
out = {}
for name, (W, H) in SZ.items():
    hs = pkg.HostScene(name); ds = hs.upload(W, H)
    n = spp if name != "next_week_final" else max(spp // 4, 8)
    for kn, fl in (("renderKernel", pkg.FLAG_NO_WAVEFRONT), ("wavefront", pkg.FLAG_WAVEFRONT)):
        ds.render(W, H, 4, seed=1, flags=fl)
        rs = [ds.render(W, H, n, seed=2 + i, flags=fl)[1] for i in range(3)]
        best = min(r["kernel_ms"] for r in rs)
        out.setdefault(name, {})[kn] = W * H * n / best / 1e3
        print("%-16s %-13s %8.2f ms %8.1f Msamples/s rounds %d tail %.2f ms" % (name, kn, best, W * H * n / best / 1e3, rs[0]["rounds"], rs[0]["tail_ms"]), flush=True)
    ds.destroy()
if do5:
    for sn, (W, H) in ((300, (480, 270)), (1000, (240, 135))):
        hs = pkg.HostScene("synthetic", synth_n=sn); ds = hs.upload(W, H)
        for kn, fl in (("renderKernelCoop", pkg.FLAG_COOP), ("renderKernel", pkg.FLAG_NO_COOP)):
            if sn == 1000 and kn == "renderKernel":
                W2, H2 = 120, 68   # the per-lane kernel needs ~1 s per path here
            else:
                W2, H2 = W, H
            ds.render(W2, H2, 1, seed=1, flags=fl, region=(0, 8, 0, W2))
            r = ds.render(W2, H2, 1, seed=3, flags=fl)[1]
            out.setdefault("synthetic_%d" % sn, {})[kn] = W2 * H2 / r["kernel_ms"]
            print("synthetic n=%d %dx%d %-17s %9.1f ms %8.2f ksamples/s tail %.1f ms" % (sn, W2, H2, kn, r["kernel_ms"], W2 * H2 / r["kernel_ms"], r["tail_ms"]), flush=True)
        ds.destroy()
print(json.dumps(out))
