"""Regenerate profiles/r02_ptxas.md: `ptxas -v` (registers, spill bytes) of every kernel of every translation unit.
   python tools/ptxas_table.py > profiles/r02_ptxas.md      (CPU only; compiles the four units in parallel, ~3 min)"""
import os
import re
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "stencil_b200", "csrc")
UNITS = ["capi.cu", "resident2d.cu", "warp2d.cu", "tma3db.cu"]
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-Xptxas", "-v"]


def compile_unit(u):
    r = subprocess.run(["nvcc"] + FLAGS + ["-c", u, "-o", f"/tmp/ptxas_{u}.o"], cwd=CSRC, capture_output=True, text=True)
    if r.returncode:
        sys.exit(r.stderr)
    return r.stderr


with ThreadPoolExecutor(len(UNITS)) as ex:
    logs = list(ex.map(compile_unit, UNITS))
rows = {}
for log in logs:
    name = None
    spill = (0, 0)
    for line in log.splitlines():
        m = re.search(r"Compiling entry function '(\S+)'", line)
        if m:
            name = m.group(1)
            continue
        m = re.search(r"(\d+) bytes spill stores, (\d+) bytes spill loads", line)
        if m and name:
            spill = (int(m.group(1)), int(m.group(2)))
        m = re.search(r"Used (\d+) registers", line)
        if m and name:
            rows[name] = (int(m.group(1)),) + spill
            name = None
names = subprocess.run(["c++filt"], input="\n".join(rows), capture_output=True, text=True).stdout.splitlines()
table = []
for mangled, dem in zip(rows, names):
    dem = re.sub(r"^void ", "", dem)
    dem = re.sub(r"\(.*$", "", dem).replace("sb200::", "").replace("(sb200::Shape)", "").replace("(int)", "").replace("(bool)", "")
    dem = dem.replace("<unnamed>::", "").replace("(anonymous namespace)::", "")
    table.append((dem,) + rows[mangled])
table.sort()
print("# r02 — ptxas -v per kernel (registers, spill bytes), `nvcc -gencode arch=compute_100a,code=sm_100a -O3`, end of round 2\n")
print("Regenerate: `python tools/ptxas_table.py > profiles/r02_ptxas.md`.  Template arguments: tma3d `<T, K, SHAPE (0 star, 1 box, "
      "2 separable box), LINKED>`, tma3dr `<T, R, LINKED>`, tma3db `<T, R, PXV, PY, LINKED>`, warp2d `<T, SHAPE, K, LINKED>`,\n"
      "warp2dr `<T, R, K, SHAPE, LINKED>`, line1d `<T, SHAPE, R>`, resident2d `<T, SHAPE, R, CLUSTER>`, vec `<T, NDIM, SHAPE, R>`.\n")
print("| kernel | registers | spill stores (B) | spill loads (B) |\n|---|---|---|---|")
for dem, regs, ss, sl in table:
    print(f"| `{dem}` | {regs} | {ss} | {sl} |")
spilled = [t for t in table if t[2] or t[3]]
print(f"\n{len(table)} kernels, {len(spilled)} with spills" + (": " + ", ".join(f"`{t[0]}` ({t[2]} / {t[3]} B)" for t in spilled) if spilled else "") + ".")
