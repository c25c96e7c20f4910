#!/usr/bin/env python
"""renderKernel time against RTB_CHUNK_DIV (a warp asks for indices-left / (warps * div), at least RTB_CHUNK_TAIL) on cfg3 / cfg2 / cfg4."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader
pkg = rtb200_loader.load()
for name, (W, H, n) in {"next_week_final": (400, 400, 64), "random": (600, 400, 128), "cornell": (500, 500, 64)}.items():
    ds = pkg.HostScene(name).upload(W, H)
    ds.render(W, H, 4, seed=1, flags=pkg.FLAG_NO_WAVEFRONT)
    for div in ("2", "4", "8", "16", "32", "64", "128"):
        os.environ["RTB_CHUNK_DIV"] = div
        rs = [ds.render(W, H, n, seed=2 + i, flags=pkg.FLAG_NO_WAVEFRONT)[1] for i in range(2)]
        b = min(rs, key=lambda r: r["kernel_ms"])
        print("%-16s div %-3s %8.2f ms %7.1f Msamples/s tail_ms %.1f" % (name, div, b["kernel_ms"], W * H * n / b["kernel_ms"] / 1e3, b["tail_ms"]), flush=True)
    ds.destroy()
