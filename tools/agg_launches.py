"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import collections, csv, re, sys
path = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
with open(path) as f:
    lines = [l for l in f if not l.startswith('==')]
agg = collections.defaultdict(lambda: [0, 0.0]); tot = 0.0
for row in csv.DictReader(lines):
    try:
        t = float(row['Metric Value'].replace(',', ''))
    except Exception:
        continue
    u = row['Metric Unit']
    t = t / 1e3 if u == 'ns' else (t * 1e3 if u == 'ms' else t)
    name = re.sub(r'\(.*', '', row['Kernel Name'])[:64]
    agg[name][0] += 1; agg[name][1] += t; tot += t
print(f"total {tot/1e3:.2f} ms over {sum(v[0] for v in agg.values())} launches")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{v[1]:10.1f} us {v[0]:5d}  {100*v[1]/tot:5.1f}%  {k}")
