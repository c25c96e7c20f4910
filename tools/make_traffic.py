#!/usr/bin/env python
"""profiles/r2_traffic.json from the `ncu --set full` captures of the bench workloads: per scene, the DRAM and L2 bytes and the
duration of ONE launch of the dominant kernel (what bench.py reports as roofline.traffic / roofline.memory).
A sixth field `sum` totals every matching launch of the report instead (the wavefront frame of cfg4 is ~1500 launches of two kernels: its
traffic per step is their sum).  Reports may be metrics-only captures (dram / lts / l1tex / duration) at full size where `--set full` is too slow.
usage: python tools/make_traffic.py out.json scene=W,H,spp,report.ncu-rep[,kernel-regex[,sum]] ..."""
import csv, io, json, re, subprocess, sys

out = {}
for spec in sys.argv[2:]:
    key, rest = spec.split("=")
    parts = rest.split(",")
    W, H, spp, rep = int(parts[0]), int(parts[1]), int(parts[2]), parts[3]
    pat = re.compile(parts[4]) if len(parts) > 4 and parts[4] else None
    total = len(parts) > 5 and parts[5] == "sum"
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    if not rows or "Kernel Name" not in rows[0]:
        print("skipping %s: no report at %s" % (key, rep), file=sys.stderr)  # a capture that timed out must not cost the others
        continue
    hdr = rows[0]
    iK = hdr.index("Kernel Name")
    ent = {"width": W, "height": H, "spp": spp, "report": rep.split("/")[-1], "kernels": []}
    for r in rows[2:]:
        if pat and not pat.search(r[iK]):
            continue
        g = lambda n: float(r[hdr.index(n)]) if n in hdr and r[hdr.index(n)] not in ("", "n/a") else None
        units = rows[1]
        def byts(n):
            v = g(n)
            if v is None: return None
            u = units[hdr.index(n)].lower()
            return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "tbyte": 1e12}.get(u, 1)
        def ms(n):
            v = g(n); u = units[hdr.index(n)].lower()
            return v * {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1, "msecond": 1, "s": 1e3, "second": 1e3, "nsecond": 1e-6}.get(u, 1)
        ent["kernels"].append({"kernel": r[iK][:80], "kernel_ms": ms("gpu__time_duration.sum"), "dram_bytes_read": byts("dram__bytes_read.sum"),
                               "dram_bytes_write": byts("dram__bytes_write.sum"), "lts_bytes": (g("lts__t_sectors.sum") or 0) * 32.0 or byts("lts__t_bytes.sum"),  # 32-byte sectors
                               "l1tex_bytes": (g("SM_B.TriageCompute.l1tex__t_sectors.sum") or g("l1tex__t_sectors.sum") or 0) * 32.0 or None})
    if ent["kernels"] and total:
        ks = ent.pop("kernels")
        ent.update({"kernel": "sum over %d launches (%s)" % (len(ks), ", ".join(sorted({k["kernel"].split("<")[0].replace("void ", "") for k in ks}))), "launches": len(ks), "sum": True})
        for f in ("kernel_ms", "dram_bytes_read", "dram_bytes_write", "lts_bytes", "l1tex_bytes"):
            ent[f] = sum((k[f] or 0) for k in ks)
        ent["kernels"] = []
    elif ent["kernels"]:
        k0 = max(ent["kernels"], key=lambda k: k["kernel_ms"])
        ent.update({k: k0[k] for k in ("kernel", "kernel_ms", "dram_bytes_read", "dram_bytes_write", "lts_bytes", "l1tex_bytes")})
    out[key] = ent
json.dump(out, open(sys.argv[1], "w"), indent=1)
print(json.dumps(out, indent=1))
