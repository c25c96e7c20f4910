#!/usr/bin/env python
"""Per-source-line view of an .ncu-rep of renderKernel: instructions, stall samples, no-instruction samples, and the executed
SASS footprint (bytes of distinct executed instructions) per line.  usage: python tools/ncu_by_line.py x.ncu-rep [top]"""
import csv, io, subprocess, sys, os
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
f = None; hdr = None; out = []; cur = None
for r in rows:
    if not r: continue
    if r[0] == "File Path": f = os.path.basename(r[1]); continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No":
        hdr = r; iS = hdr.index("# Samples"); iI = hdr.index("Instructions Executed"); iT = hdr.index("Thread Instructions Executed"); iN = hdr.index("stall_no_inst"); continue
    if r[0]:
        try: int(r[0]); int(r[iS] or 0); int(r[iI] or 0); int(r[iN] or 0)
        except (ValueError, IndexError): cur = None; continue  # a source line whose quotes broke the CSV row
        cur = dict(file=f, line=int(r[0]), src=r[1].strip(), samples=int(r[iS] or 0), inst=int(r[iI] or 0), thr=int(r[iT] or 0), noinst=int(r[iN] or 0), foot=0)
        out.append(cur)
    elif cur is not None and len(r) > iI and r[2].startswith("0x") and r[iI] not in ("-", "") and int(r[iI]) > 0:
        cur["foot"] += 16
ts = sum(o["samples"] for o in out); ti = sum(o["inst"] for o in out); tn = sum(o["noinst"] for o in out); tf = sum(o["foot"] for o in out)
print("total samples %d inst %d no_inst samples %d executed footprint %d B" % (ts, ti, tn, tf))
byfile = {}
for o in out:
    b = byfile.setdefault(o["file"], [0, 0, 0, 0]); b[0] += o["samples"]; b[1] += o["inst"]; b[2] += o["noinst"]; b[3] += o["foot"]
for k, b in byfile.items():
    print("%-28s samples %5.1f%% inst %5.1f%% no_inst %5.1f%% footprint %6d B" % (k, 100 * b[0] / ts, 100 * b[1] / ti, 100 * b[2] / max(tn, 1), b[3]))
for o in sorted(out, key=lambda o: -o["samples"])[:top]:
    print("%-22s %4d smp %4.1f%% inst %4.1f%% noinst %4.1f%% avgthr %4.1f foot %5d  %s" % (o["file"], o["line"], 100 * o["samples"] / ts, 100 * o["inst"] / ti, 100 * o["noinst"] / max(tn, 1),
          o["thr"] / max(o["inst"], 1), o["foot"], o["src"][:70]))
