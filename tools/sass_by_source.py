#!/usr/bin/env python
"""Static SASS instruction count per source line range of one kernel (nvdisasm -g output on stdin-file).
usage: python tools/sass_by_source.py build/rt_launch_exact.o renderKernelIffLi0ELb0"""
import collections, re, subprocess, sys, tempfile, os, glob
obj, key = sys.argv[1], sys.argv[2]
d = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, capture_output=True)
cub = glob.glob(os.path.join(d, "*.cubin"))[0]
txt = subprocess.run(["nvdisasm", "-g", "-c", cub], capture_output=True, text=True).stdout.splitlines()
on = False; cur = None; cnt = collections.Counter(); inl = collections.Counter()
for ln in txt:
    if ln.startswith(".text.") or "Function :" in ln or re.match(r"^\s*\.section\s+\.text\.", ln):
        on = key in ln
    if not on: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    if re.match(r"^\s+/\*[0-9a-f]{4,6}\*/", ln) and cur:
        cnt[cur] += 1
tot = sum(cnt.values()); print("total", tot)
# bucket per file by 10 lines
b = collections.Counter()
for (f, l), c in cnt.items(): b[(f, l // 10 * 10)] += c
for (f, l), c in sorted(b.items()):
    if c >= 30: print("%-24s %4d-%4d %5d" % (f, l, l + 9, c))
