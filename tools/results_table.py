#!/usr/bin/env python
"""Markdown tables of the committed bench lines (profiles/r2_bench_default_n1.json and, when present, the N-GPU lines):
what DESIGN.md section 6 and README.md quote.  usage: python tools/results_table.py  (prints to stdout)"""
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load(name):
    try:
        return json.load(open(os.path.join(ROOT, "profiles", name)))
    except (OSError, ValueError):
        return None


def fmt(x):
    if x is None:
        return "—"
    if x >= 1e6:
        return "%.1f M" % (x / 1e6)
    if x >= 1e3:
        return "%.1f k" % (x / 1e3)
    return "%.0f" % x


def rows(d):
    yield "cfg2 `random` 600×400×1024 (headline)", d
    names = {"cfg1": "cfg1 `random_book1` 600×400×1024", "cfg3": "cfg3 `next_week_final` 400×400×256", "cfg4": "cfg4 `cornell` 500×500×1000",
             "cfg5_90k": "cfg5 90 k spheres, 480×270×1 (bounded sample)", "cfg5_1m": "cfg5 1 M spheres, 240×135×1 (bounded sample)"}
    for k, v in d.get("per_scene", {}).items():
        yield names.get(k, k), v


def main():
    import io, sys
    if "--apply" in sys.argv:
        buf = io.StringIO()
        old, sys.stdout = sys.stdout, buf
        try:
            table()
        finally:
            sys.stdout = old
        for name in ("README.md", "DESIGN.md"):
            path = os.path.join(ROOT, name)
            s = open(path).read()
            a = s.index("<!-- RESULTS:BEGIN"); a = s.index("\n", a) + 1
            b = s.index("<!-- RESULTS:END -->")
            open(path, "w").write(s[:a] + "\nMeasured on one B200 of this pool (`profiles/r2_bench_default_n1.json`, the default `python bench.py`; every figure is recomputable from `profiles/`):\n\n" + buf.getvalue() + "\n" + s[b:])
        return
    table()


def table():
    d = load("r2_bench_default_n1.json")
    if not d:
        print("no bench line")
        return
    print("| workload | kernel | samples/s (device) | ms / step | e2e samples/s (host buffers) | CPU port, all host cores | e2e ÷ CPU | roofline |")
    print("|---|---|---|---|---|---|---|---|")
    for name, v in rows(d):
        r, e, c = v["roofline"], v["e2e"], v["cpu_baseline"]
        if r["bound"] == "hbm":
            roof = "algorithmic node/leaf bytes %.0f GB/s = %.3f of HBM %.0f, %.3f of L2 %.0f; DRAM traffic %s" % (
                r["achieved"], r["frac"], r["peak"], r["l2"]["frac"], r["l2"]["peak"],
                ("%.2f GB / launch" % (r["traffic"] / 1e9)) if r.get("traffic") else "n/a")
        else:
            est = r.get("memory", {}).get("dram_bytes_per_frame_estimate")
            dram = ("%.1f MB / launch" % (r["traffic"] / 1e6)) if r.get("traffic") else (
                "≈ %.0f GB / frame (%.0f rounds × one captured round of both kernels) = %.2f of HBM peak" % (est / 1e9, r["memory"]["wavefront_rounds"], r["memory"]["dram_frac_of_hbm_peak_estimate"]) if est else "n/a")
            roof = "%.2f of %.1f TFLOP/s non-FMA issue = **%.3f** (%.3f of FFMA peak); DRAM %s" % (
                r["achieved"], r["peak"], r["frac"], r.get("frac_of_ffma_peak", 0), dram)
        print("| %s | %s | %s | %.1f | %s | %s (%d threads) | %.0f× | %s |" % (name, r.get("kernel", "?"), fmt(v["value"]), v["ms_per_step"], fmt(e["value"]), fmt(c["value"]), c["cores"],
                                                                  e["value"] / c["value"], roof))
    ck = d.get("clocks", {})
    print("\nClocks during the timed regions: %s MHz of %s, throttle reasons %s; `gpu_launches` %s; L2 %s." % (
        ck.get("sm_mhz"), ck.get("sm_max_mhz"), ck.get("reasons"), d.get("gpu_launches"), d["config"].get("l2")))
    for n in (2, 4, 8):
        m = load("r2_bench_default_n%d.json" % n)
        if not m:
            continue
        print("\n**N = %d** (`torchrun`, one rank per GPU): weak %s samples/s (%.1f ms / step), e2e through ONE `rt_render(n_gpus = %d)` %s samples/s (reduce %s, %.2f ms)." % (
            n, fmt(m["value"]), m["ms_per_step"], n, fmt(m["e2e"]["value"]), m["e2e"].get("reduce_kind"), m["e2e"].get("reduce_ms", 0)))
        sts = [("cfg2 " + m["strong"]["frame"], m["strong"])] if m.get("strong") else []
        sts += [(k + " " + v["strong"]["frame"], v["strong"]) for k, v in m.get("per_scene", {}).items() if v.get("strong")]
        if sts:
            print("\n| strong scaling: the fixed frame split inside one `rt_render` | 1-GPU ms | %d-GPU ms | efficiency | reduce ms | tail ms (N) | per-GPU kernel ms |\n|---|---|---|---|---|---|---|" % n)
            for k, v in sts:
                print("| %s | %.1f | %.1f | **%.3f** | %.2f | %.1f | %s |" % (k, v["one"]["ms_per_step"], v["n"]["ms_per_step"], v["efficiency"], v["n"]["reduce_ms"], v["n"]["tail_ms"],
                                                                       " / ".join("%.0f" % x for x in v["n"]["gpu_kernel_ms"])))
        sc = m.get("sharding_check")
        if sc:
            print("\n`sharding_check` (N-GPU frame vs 1-GPU frame, max relative difference): %s" % json.dumps(sc))


if __name__ == "__main__":
    main()
