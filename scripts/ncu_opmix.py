"""Executed-instruction mix by SASS opcode for a captured kernel. usage: ncu_opmix.py report.ncu-rep kernel_substring"""
import csv, subprocess, sys
from collections import Counter
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
cur, hdr, c, on = None, None, Counter(), False
for row in csv.reader(out.splitlines()):
    if row and row[0] == "Kernel Name":
        if on: break
        on = sys.argv[2] in row[1]
    elif row and row[0] == "Address":
        hdr = row
    elif on and row and row[0].startswith("0x"):
        sass = row[1].strip()
        if sass.startswith("@"):
            sass = sass.split(None, 1)[1]
        op = sass.split()[0].split(".")[0]
        c[op] += int(row[hdr.index("Instructions Executed")])
tot = sum(c.values())
print("total", tot)
for op, n in c.most_common(30):
    print(f"{op:12s} {n:12d} {n/tot:6.1%}")
