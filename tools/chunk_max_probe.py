#!/usr/bin/env python
"""renderKernel time against the largest chunk of path indices a warp takes per atomic (RTB_CHUNK_MAX) on cfg3 / cfg2."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader
pkg = rtb200_loader.load()
for name, (W, H, n) in {"next_week_final": (400, 400, 128), "random": (600, 400, 256)}.items():
    ds = pkg.HostScene(name).upload(W, H)
    ds.render(W, H, 4, seed=1, flags=pkg.FLAG_NO_WAVEFRONT)
    for mx in ("1024", "512", "256", "128", "64"):
        os.environ["RTB_CHUNK_MAX"] = mx
        rs = [ds.render(W, H, n, seed=2 + i, flags=pkg.FLAG_NO_WAVEFRONT)[1] for i in range(2)]
        b = min(rs, key=lambda r: r["kernel_ms"])
        print("%-16s chunk max %-5s %8.2f ms %7.1f Msamples/s tail_ms %.1f" % (name, mx, b["kernel_ms"], W * H * n / b["kernel_ms"] / 1e3, b["tail_ms"]), flush=True)
    ds.destroy()
