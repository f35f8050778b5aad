"""Static SASS statistics of one kernel in an object/.so: python tools/sass_loop.py <file> <substr of mangled name>
Prints the opcode histogram of the body of the largest backward-branch loop."""
import collections, re, subprocess, sys
f, pat = sys.argv[1], sys.argv[2]
txt = subprocess.run(["cuobjdump", "-sass", f], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", txt)
for fn in funcs[1:]:
    name = fn.split("\n", 1)[0].strip()
    if pat not in name:
        continue
    ins = []
    for m in re.finditer(r"/\*([0-9a-f]{4,})\*/\s+(.*?);", fn):
        ins.append((int(m.group(1), 16), m.group(2).strip()))
    best = None
    for addr, t in ins:
        m = re.search(r"BRA(?:\.\w+)*\s+(?:P\d,\s*)?0x([0-9a-f]+)", t)
        if m and int(m.group(1), 16) < addr:
            span = addr - int(m.group(1), 16)
            if best is None or span > best[2]:
                best = (int(m.group(1), 16), addr, span)
    print(name, "total instr", len(ins), "loop", best and (hex(best[0]), hex(best[1])))
    if not best:
        continue
    body = [t for a, t in ins if best[0] <= a <= best[1]]
    h = collections.Counter()
    for t in body:
        p = t.split()
        op = p[1] if p[0].startswith("@") else p[0]
        q = op.split(".")
        op = q[0] + ("." + q[1] if q[0] in ("IMAD", "LDS", "STS", "LDG", "STG", "SHFL", "LDL", "STL") and len(q) > 1 else "")
        h[op] += 1
    print("loop body instr", len(body))
    print("  ".join(f"{k}:{v}" for k, v in h.most_common(40)))
