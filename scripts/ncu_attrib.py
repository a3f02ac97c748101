#!/usr/bin/env python3
"""Attributes an ncu source page (SASS, csv) to source lines: joins `ncu --page source --csv
--print-source sass` with `nvdisasm -gi -c` of the same cubin (same instruction order) and prints
stall samples and executed instructions per outermost kernel-file line and per innermost line.
    ncu_attrib.py sass.csv kernel.dis [kernel_file_basename]"""
import csv, re, sys, collections

sass_csv, dis, kfile = sys.argv[1], sys.argv[2], (sys.argv[3] if len(sys.argv) > 3 else "stack_psel.cuh")
rows = list(csv.reader(open(sass_csv)))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
H = rows[hdr]
ci = {n: H.index(n) for n in ("Source", "# Samples", "Instructions Executed", "Thread Instructions Executed")}
ins = rows[hdr + 1:]
# disassembly: per instruction, the chain of (file, line)
chain, cur, out, started = [], [], [], False
for line in open(dis):
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', line)
    if m:
        if started: cur = []; started = False
        cur.append((m.group(1).split("/")[-1], int(m.group(2))))
        continue
    if re.match(r"\s*/\*[0-9a-f]{4,}\*/", line):
        started = True
        out.append((list(cur), line.split("*/", 1)[1].strip()))
n = min(len(out), len(ins))
print(f"{len(ins)} profiled instructions, {len(out)} disassembled", file=sys.stderr)
outer, inner, tot_s, tot_i = collections.Counter(), collections.Counter(), 0, 0
outer_i, inner_i = collections.Counter(), collections.Counter()
for k in range(n):
    ch, txt = out[k]
    r = ins[k]
    s, wi = int(r[ci["# Samples"]] or 0), int(r[ci["Instructions Executed"]] or 0)
    tot_s += s; tot_i += wi
    o = next((c for c in reversed(ch) if c[0] == kfile), ch[-1] if ch else ("?", 0))
    i0 = ch[0] if ch else ("?", 0)
    outer[o] += s; outer_i[o] += wi; inner[i0] += s; inner_i[i0] += wi
print(f"total samples {tot_s}, warp instructions {tot_i}")
print("--- by outermost line of", kfile)
for k, v in sorted(outer.items(), key=lambda kv: kv[0][1]):
    if v > tot_s * 0.002 or outer_i[k] > tot_i * 0.002:
        print(f"{k[0]}:{k[1]:4d}  samples {100*v/tot_s:5.1f}%  instr {100*outer_i[k]/tot_i:5.1f}%")
print("--- top innermost lines")
for k, v in inner.most_common(40):
    print(f"{k[0]}:{k[1]:4d}  samples {100*v/tot_s:5.1f}%  instr {100*inner_i[k]/tot_i:5.1f}%")
if len(sys.argv) > 4:   # second grouping: lines of another file anywhere in the chain (e.g. psel.cuh)
    f2 = sys.argv[4]
    g, gi = collections.Counter(), collections.Counter()
    for k in range(n):
        ch, txt = out[k]
        r = ins[k]
        s, wi = int(r[ci["# Samples"]] or 0), int(r[ci["Instructions Executed"]] or 0)
        o = next((c for c in reversed(ch) if c[0] == f2), None)
        if o: g[o] += s; gi[o] += wi
    print("--- by outermost line of", f2)
    for k, v in sorted(g.items(), key=lambda kv: kv[0][1]):
        if v > tot_s * 0.002 or gi[k] > tot_i * 0.002:
            print(f"{k[0]}:{k[1]:4d}  samples {100*v/tot_s:5.1f}%  instr {100*gi[k]/tot_i:5.1f}%")
