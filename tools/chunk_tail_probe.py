#!/usr/bin/env python
"""renderKernel time and tail against the guided-chunk parameters (RTB_CHUNK_TAIL = smallest chunk a warp takes near the end of the index space,
RTB_CHUNK_DIV = a warp asks for indices-left / (warps * div)) on cfg2 / cfg3 / cfg1.  Results never depend on them."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader
pkg = rtb200_loader.load()
for name, (W, H, n) in {"next_week_final": (400, 400, 64), "random": (600, 400, 128), "random_book1": (600, 400, 128)}.items():
    ds = pkg.HostScene(name).upload(W, H)
    ds.render(W, H, 4, seed=1, flags=pkg.FLAG_NO_WAVEFRONT)
    for tail in ("32", "16", "8", "4", "1"):
        for div in ("2", "8"):
            os.environ["RTB_CHUNK_TAIL"] = tail; os.environ["RTB_CHUNK_DIV"] = div
            rs = [ds.render(W, H, n, seed=2 + i, flags=pkg.FLAG_NO_WAVEFRONT)[1] for i in range(2)]
            b = min(rs, key=lambda r: r["kernel_ms"])
            print("%-16s tail %-3s div %s %8.2f ms %7.1f Msamples/s tail_ms %.1f" % (name, tail, div, b["kernel_ms"], W * H * n / b["kernel_ms"] / 1e3, b["tail_ms"]), flush=True)
    ds.destroy()
