"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel:
python tools/launch_list.py <launches.csv> [out.md]"""
import collections
import csv
import sys

lines = open(sys.argv[1]).read().splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"ID"'))
rows = list(csv.DictReader(lines[start:]))
agg = collections.OrderedDict()
for r in rows:
    k = r["Kernel Name"]
    a = agg.setdefault(k, [0, 0.0, r["Block Size"], r["Grid Size"]])
    a[0] += 1
    a[1] += float(r["Metric Value"])
tot = sum(a[1] for a in agg.values())
out = [f"launches: {len(rows)}, summed gpu__time_duration: {tot / 1e6:.3f} ms", "",
       "| launches | total ms | share | avg us | grid | block | kernel |", "|---|---|---|---|---|---|---|"]
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    out.append(f"| {a[0]} | {a[1] / 1e6:.3f} | {100 * a[1] / tot:.1f}% | {a[1] / a[0] / 1e3:.1f} | {a[3]} | {a[2]} | `{k[:110]}` |")
text = "\n".join(out)
print(text)
if len(sys.argv) > 2:
    open(sys.argv[2], "a").write(text + "\n")
