"""Group scripts/ncu_lines.py output of cpe_eval.cu into kernel phases. usage: phase_sum.py lines.txt frames"""
import re, sys
from collections import defaultdict
frames = float(sys.argv[2])
B = [(133, "A load q/cam/meas"), (137, "B sincos"), (141, "C rot"), (226, "D per-frame items"), (251, "E main proj/resid/loss"),
     (274, "F main jacobian"), (10**6, "G tail/reduce")]
INL = {"cpe_internal.cuh": "E main proj/resid/loss"}
ph, sm = defaultdict(int), defaultdict(int)
for ln in open(sys.argv[1]):
    m = re.match(r"\s*(\S+):(\d+)\s+inst\s+(\d+)\s+\S+\s+samples\s+(\d+)", ln)
    if not m:
        continue
    f, l = m.group(1), int(m.group(2))
    if f == "cpe_eval.cu":
        k = next(n for b, n in B if l <= b) if l >= 97 else "helpers(inlined): line %d" % l
    else:
        k = INL.get(f, "inl " + f)
    ph[k] += int(m.group(3)); sm[k] += int(m.group(4))
t, ts = sum(ph.values()), sum(sm.values())
for k in sorted(ph):
    print(f"{k:36s} inst {ph[k]/frames:8.0f}/frame {ph[k]/t:6.1%}  stall samples {sm[k]/ts:6.1%}")
