#!/usr/bin/env python3
"""Summarise an .ncu-rep (read here, no GPU needed): headline counters of the first kernel in the
report, time per code region (split at BAR.SYNC) and the hottest SASS lines with their top stall
reasons.  Usage: scripts/ncu_summary.py gpurun_out/prof.ncu-rep [top_n]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 15


def page(name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


raw = page("raw")
h, u, v = raw[0], raw[1], raw[2]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "launch__grid_size", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum", "smsp__sass_inst_executed_op_shared_ld.sum",
        "smsp__sass_inst_executed_op_shared_st.sum"]
for i, k in enumerate(h):
    if k in want:
        print(f"{k} = {v[i]} {u[i]}")
for i, k in enumerate(h):
    if k.startswith("smsp__average_warps_issue_stalled") and k.endswith("per_issue_active.ratio") and float(v[i] or 0) > 0.1:
        print(f"{k.replace('smsp__average_warps_issue_stalled_', 'stall ').replace('_per_issue_active.ratio', '')} = {v[i]}")

src = page("source")
hh = src[1]
data = src[2:]
isrc, isamp, iex = hh.index("Source"), hh.index("# Samples"), hh.index("Instructions Executed")
tot = sum(int(r[isamp]) for r in data) or 1
totex = sum(int(r[iex]) for r in data) or 1
print(f"SASS lines {len(data)}, samples {tot}, warp instructions {totex}")
bars = [i for i, r in enumerate(data) if "BAR.SYNC" in r[isrc]]
prev = 0
for b in bars + [len(data)]:
    s = sum(int(r[isamp]) for r in data[prev:b]); e = sum(int(r[iex]) for r in data[prev:b])
    print(f"region SASS[{prev}:{b}]: {100 * s / tot:.1f}% of samples, {100 * e / totex:.1f}% of instructions")
    prev = b
stall = [i for i, k in enumerate(hh) if k.startswith("stall_") and "Not Issued" not in k]
for i in sorted(range(len(data)), key=lambda i: -int(data[i][isamp]))[:topn]:
    r = data[i]
    st = sorted(((int(r[c] or 0), hh[c]) for c in stall), reverse=True)[:2]
    print(f"  [{i}] {100 * int(r[isamp]) / tot:.2f}% x{r[iex]} {r[isrc].strip()[:70]} {st}")
# opcode mix by executed instructions
mix = {}
for r in data:
    t = r[isrc].split()
    if not t:
        continue
    op = t[1] if t[0].startswith("@") and len(t) > 1 else t[0]
    op = op.split(".")[0]
    mix[op] = mix.get(op, 0) + int(r[iex])
print("opcode mix:", ", ".join(f"{k} {100 * n / totex:.1f}%" for k, n in sorted(mix.items(), key=lambda kv: -kv[1])[:14]))
