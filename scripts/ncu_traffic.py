"""DRAM traffic of one captured launch -> profiles/traffic.json (what bench.py scales for `roofline.traffic`).
usage: ncu_traffic.py report.ncu-rep kernel_substring key units_in_launch unit_name [capture_label]
   e.g. ncu_traffic.py gpurun_out/r2e_evaljac.ncu-rep eval_kernel eval_kernel_jac 1536000 frame_view r2e"""
import csv
import json
import os
import subprocess
import sys

rep, pat, key, units, unit = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4]), sys.argv[5]
label = sys.argv[6] if len(sys.argv) > 6 else os.path.basename(rep)
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, u = rows[0], rows[1]


def val(r, name):
    v, unit_ = float(r[h.index(name)].replace(",", "")), u[h.index(name)]
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[unit_]


r = next(r for r in rows[2:] if pat in r[h.index("Kernel Name")])
rd, wr = val(r, "dram__bytes_read.sum"), val(r, "dram__bytes_write.sum")
path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "traffic.json")
tj = json.load(open(path)) if os.path.exists(path) else {}
old = tj.get(key)
tj[key] = {f"dram_bytes_per_{unit}": (rd + wr) / units, "capture": label, "dram_bytes_read": rd, "dram_bytes_write": wr,
           f"{unit}s": units, "grid": r[h.index("Grid Size")]}
if old is not None:
    tj[key]["previous"] = {k: v for k, v in old.items() if k not in ("previous", "recheck_r1j")}
json.dump(tj, open(path, "w"), indent=1)
print(key, tj[key])
