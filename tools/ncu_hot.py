"""Per-SASS-instruction stall samples from an ncu source page CSV:
   ncu -i rep --page source --csv > src.csv ; python tools/ncu_hot.py src.csv [top]
Prints the stall-reason totals and the instructions with the most samples."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
ia, isrc, isamp, iex = hdr.index("Address"), hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
data = rows[2:]
tot = sum(int(r[isamp] or 0) for r in data)
print("total samples", tot, "instructions", len(data), "executed warp instr", sum(int(r[iex] or 0) for r in data))
agg = {h: sum(int(r[i] or 0) for r in data) for i, h in stall_cols}
print("stall totals:", ", ".join(f"{h[6:]} {v} ({100 * v / tot:.1f}%)" for h, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v))
byop = {}
for r in data:
    op = r[isrc].split()
    op = (op[1] if op and op[0].startswith("@") else op[0]) if op else "?"
    op = op.split(".")[0] + ("." + op.split(".")[1] if op.split(".")[0] in ("LDS", "STS", "SHFL", "STG", "BAR", "SYNCS") and "." in op else "")
    e = byop.setdefault(op, [0, 0])
    e[0] += int(r[isamp] or 0)
    e[1] += int(r[iex] or 0)
print("by opcode (samples%, executed%):")
ex_tot = sum(v[1] for v in byop.values())
for op, (s, e) in sorted(byop.items(), key=lambda kv: -kv[1][0])[:25]:
    print(f"  {op:12s} {100 * s / tot:5.1f}%  {100 * e / ex_tot:5.1f}%")
print("hottest instructions:")
idx = sorted(range(len(data)), key=lambda i: -int(data[i][isamp] or 0))[:top]
for i in sorted(idx):
    r = data[i]
    st = sorted(((int(r[c] or 0), h[6:]) for c, h in stall_cols), reverse=True)[:3]
    print(f"  #{i:4d} {r[isrc].strip()[:70]:70s} samp {int(r[isamp]):6d} ex {r[iex]:>8s}  " + " ".join(f"{h}:{v}" for v, h in st if v))
