#!/usr/bin/env python
"""Adds DRAM bytes / duration of the dominant kernel's captured launch to profiles/r2_traffic.json from the committed ncu SUMMARIES
(profiles/r2_ncu_*.txt, tools/ncu_summary.py output) — for captures whose .ncu-rep stayed on the GPU box (gpurun_out/ is capped at 64 MiB).
usage: python tools/traffic_from_summaries.py key=W,H,spp,summary.txt ..."""
import json, os, re, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
path = os.path.join(ROOT, "profiles", "r2_traffic.json")
out = json.load(open(path))
U = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}
for spec in sys.argv[1:]:
    key, rest = spec.split("=")
    W, H, spp, txt = rest.split(",")
    vals = {}
    for line in open(os.path.join(ROOT, txt)):
        m = re.match(r"(\S+) = ([0-9.eE+-]+) ?(\S*)", line)
        if m:
            vals[m.group(1)] = float(m.group(2)) * U.get(m.group(3), 1)
    out[key] = {"width": int(W), "height": int(H), "spp": int(spp), "report": txt + " (summary of the --set full capture)", "kernels": [],
                "kernel": key, "kernel_ms": vals["gpu__time_duration.sum"], "dram_bytes_read": vals["dram__bytes_read.sum"],
                "dram_bytes_write": vals["dram__bytes_write.sum"], "lts_bytes": None, "l1tex_bytes": None}
    print(key, out[key]["kernel_ms"], out[key]["dram_bytes_read"], out[key]["dram_bytes_write"])
json.dump(out, open(path, "w"), indent=1)
