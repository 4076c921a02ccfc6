"""Top stall sites of an ncu report's source page:  python tools/ncu_stalls.py report.ncu-rep [top]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
lines = out.splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
rows = list(csv.DictReader(io.StringIO("\n".join(lines[start:]))))
stall_cols = [c for c in rows[0].keys() if c.startswith("stall_") and "Not Issued" not in c]
tot = sum(int(r["# Samples"] or 0) for r in rows)
agg = {c: sum(int(r[c] or 0) for r in rows) for c in stall_cols}
print("total samples", tot, {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]})
for r in sorted(rows, key=lambda r: -int(r["# Samples"] or 0))[:top]:
    s = {c[6:]: int(r[c] or 0) for c in stall_cols if int(r[c] or 0) > 0}
    s = dict(sorted(s.items(), key=lambda kv: -kv[1])[:3])
    print(f'{int(r["# Samples"]):7d} {100*int(r["# Samples"])/max(tot,1):5.1f}%  {r["Source"][:70]:70s} {s}')
