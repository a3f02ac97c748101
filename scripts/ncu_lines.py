#!/usr/bin/env python3
"""Per-source-line share of samples / executed instructions of one kernel: joins the SASS page of
an .ncu-rep with the line table of the same kernel in the built object (nvdisasm --print-line-info).
Usage: scripts/ncu_lines.py rep object.o kernel_substring [top_n]"""
import csv, io, os, re, subprocess, sys, tempfile, collections
rep, obj, kname = sys.argv[1:4]; topn = int(sys.argv[4]) if len(sys.argv) > 4 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hh = rows[1]; data = rows[2:]
isrc, isamp, iex = hh.index("Source"), hh.index("# Samples"), hh.index("Instructions Executed")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
lines = []
for cb in os.listdir(tmp):
    txt = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(tmp, cb)], capture_output=True, text=True).stdout
    inside = False; cur = ("?", 0)
    for ln in txt.splitlines():
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
        if m: inside = kname in m.group(1); continue
        if not inside: continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m: cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln): lines.append(cur)
print(f"SASS in report {len(data)}, in object {len(lines)}")
n = min(len(data), len(lines))
agg = collections.defaultdict(lambda: [0, 0])
for i in range(n):
    a = agg[lines[i]]; a[0] += int(data[i][isamp] or 0); a[1] += int(data[i][iex] or 0)
ts = sum(a[0] for a in agg.values()) or 1; te = sum(a[1] for a in agg.values()) or 1
srcs = {}
def src(f, l):
    if f not in srcs:
        p = os.path.join(os.path.dirname(os.path.abspath(obj)), "..", f)
        srcs[f] = open(p).read().splitlines() if os.path.exists(p) else []
    return srcs[f][l - 1].strip()[:80] if 0 < l <= len(srcs[f]) else ""
for (f, l), (s, e) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:topn]:
    print(f"{100*s/ts:5.1f}% smp {100*e/te:5.1f}% ins  {f}:{l}  {src(f, l)}")
