"""Record the DRAM bytes per launch of the bench's dominant kernel from an ncu raw CSV, together
with the hash of the kernel source it was captured from (bench.py refuses the number once the
source changes):
    python tools/update_traffic.py gpurun_out/ncu_c2k2_raw.csv profiles/r02x_ncu_full_tma3d_f64_k2.md"""
import csv
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
raw, source = sys.argv[1], sys.argv[2]
rows = list(csv.reader(open(raw)))
hdr, units = rows[0], rows[1]
name = hdr.index("Kernel Name")   # ncu prints template arguments as "<double, 2, 0, 0>" or "<double, (int)2, ...>"
r = [x for x in rows[2:] if x[name].replace("(int)", "").replace("(bool)", "").startswith("void sweep_tma3d_kernel<double, 2, 0, 0>")][-1]


def val(name):
    i = hdr.index(name)
    v = float(r[i].replace(",", ""))
    return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[units[i]]


rd, wr = val("dram__bytes_read.sum"), val("dram__bytes_write.sum")
src = "stencil_b200/csrc/tma3d.cuh"
sha = hashlib.sha256(open(os.path.join(ROOT, src), "rb").read()).hexdigest()[:16]
path = os.path.join(ROOT, "profiles", "traffic.json")
d = json.load(open(path))
d["tma3d_f64_star_k2_512"] = {"bytes": rd + wr, "read": rd, "write": wr, "source": source,
                              "kernel_source": src, "kernel_source_sha16": sha}
json.dump(d, open(path, "w"), indent=2)
print(json.dumps(d["tma3d_f64_star_k2_512"], indent=1))
