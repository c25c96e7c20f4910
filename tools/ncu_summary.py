#!/usr/bin/env python
"""Summarise an .ncu-rep of renderKernel: headline metrics + instruction share per SASS block (needs `ncu` on PATH).
usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep [block_instrs]"""
import csv, io, subprocess, sys

rep = sys.argv[1]
blk = int(sys.argv[2]) if len(sys.argv) > 2 else 48
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum", "l1tex__t_bytes.sum",
        "sm__inst_executed_pipe_alu.sum.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.sum.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.sum.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.sum.pct_of_peak_sustained_active",
        "smsp__warps_eligible.avg.per_cycle_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed"]
for h, u, v in zip(hdr, units, vals):
    if h in want or ("issue_stalled" in h and "per_issue_active" in h and float(v or 0) > 0.05):
        print("%s = %s %s" % (h, v, u))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr, data = rows[1], rows[2:]
iA, iS, iI, iT, iN = (hdr.index(x) for x in ("Address", "Source", "Instructions Executed", "Thread Instructions Executed", "# Samples"))
base = int(data[0][iA], 16)
tot = sum(int(r[iI]) for r in data); tt = sum(int(r[iT]) for r in data); ts = sum(int(r[iN]) for r in data)
print("total warp inst %d thread inst %d" % (tot, tt))
for i in range(0, len(data), blk):
    seg = data[i:i + blk]
    I = sum(int(r[iI]) for r in seg); T = sum(int(r[iT]) for r in seg); S = sum(int(r[iN]) for r in seg)
    if I / tot > 0.005:
        print("%05x-%05x inst %5.1f%% thr %5.1f%% avgthr %4.1f samples %5.1f%%" % (int(seg[0][iA], 16) - base, int(seg[-1][iA], 16) - base, 100 * I / tot, 100 * T / tt, T / max(I, 1), 100 * S / ts))
print("-- votes / calls / atomics")
for r in data:
    if any(k in r[iS] for k in ("VOTE", "CALL", "BRX", "ATOM", "RED")) and int(r[iI]) > tot / 5000:
        print("%05x %-44s inst %12s avgthr %.1f" % (int(r[iA], 16) - base, r[iS].strip()[:44], r[iI], int(r[iT]) / max(int(r[iI]), 1)))
