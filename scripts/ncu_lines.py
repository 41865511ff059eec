"""Attribute executed warp-instructions and stall samples of a captured kernel to CUDA source lines.
usage: ncu_lines.py report.ncu-rep lib.so kernel_substring [launch_index]
Joins ncu's SASS page (instruction order) with `nvdisasm -g` line markers of the same kernel."""
import csv
import glob
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict

rep, so, pat = sys.argv[1:4]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks, cur = [], None
for row in csv.reader(out.splitlines()):
    if row and row[0] == "Kernel Name":
        cur = {"name": row[1], "rows": []}
        blocks.append(cur)
    elif cur is not None and row and row[0].startswith("0x"):
        cur["rows"].append(row)
    elif cur is not None and row and row[0] == "Address":
        cur["hdr"] = row
blk = [b for b in blocks if pat in b["name"]][int(sys.argv[4]) if len(sys.argv) > 4 else 0]
hdr = blk["hdr"]
iex, isamp = hdr.index("Instructions Executed"), hdr.index("# Samples")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, capture_output=True)
lines = None
for cubin in glob.glob(os.path.join(tmp, "*.cubin")):
    dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
    # split per function
    parts = re.split(r"\n\s*\.text\.", dis)
    for part in parts:
        head = part.split("\n", 1)[0]
        if pat.replace("<", "I").split("(")[0] in head or all(t in head for t in re.findall(r"\w+", pat)[:1]):
            seq, line = [], ("?", 0)
            for ln in part.splitlines():
                m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
                if m:
                    line = (os.path.basename(m.group(1)), int(m.group(2)))
                    continue
                if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
                    seq.append(line)
            if len(seq) == len(blk["rows"]):
                lines = seq
if lines is None:
    sys.exit("could not match the SASS of %s in %s (%d ncu rows)" % (pat, so, len(blk["rows"])))
agg = defaultdict(lambda: [0, 0])
for row, ln in zip(blk["rows"], lines):
    agg[ln][0] += int(row[iex])
    agg[ln][1] += int(row[isamp])
ti = sum(v[0] for v in agg.values())
ts = sum(v[1] for v in agg.values())
print(f"# {blk['name']}: {ti} warp-instructions, {ts} stall samples")
for ln, v in sorted(agg.items(), key=lambda kv: -kv[1][int(os.environ.get("SORT", "1"))])[:int(os.environ.get("TOP", "45"))]:
    print(f"{ln[0]:>22s}:{ln[1]:<5d} inst {v[0]:>11d} {v[0]/ti:6.1%}   samples {v[1]:>7d} {v[1]/max(ts,1):6.1%}")
