#!/usr/bin/env python
"""Executed-instruction mix of one `ncu --set full --import-source on` report: opcode histogram (dynamic counts and share of
the stall samples) and the source lines that collected most stall samples.

    python tools/ncu_opmix.py gpurun_out/ncu_x.ncu-rep [n_lines]
"""
import collections
import csv
import re
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 12
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
print(rows[0][1][:110])
hdr, data = rows[1], rows[2:]
ia, ie, ist = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
data = [r for r in data if len(r) > ist and r[ie].isdigit()]
tot = sum(int(r[ie]) for r in data)
ts = max(1, sum(int(r[ist] or 0) for r in data))
ops, stall = collections.Counter(), collections.Counter()
for r in data:
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[ia])
    op = m.group(2).split(".")[0] if m else "?"
    ops[op] += int(r[ie])
    stall[op] += int(r[ist] or 0)
print(f"executed warp instructions: {tot}")
for op, c in ops.most_common(18):
    print(f"  {op:10s} {100 * c / tot:5.1f} % of instructions   {100 * stall[op] / ts:5.1f} % of stall samples")
print("lines with most stall samples:")
for r in sorted(data, key=lambda r: -int(r[ist] or 0))[:top]:
    print(f"  {100 * int(r[ist] or 0) / ts:5.1f} %  {r[ia].strip()[:100]}")
