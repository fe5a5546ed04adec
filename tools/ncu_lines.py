#!/usr/bin/env python3
"""Warp-stall samples of one kernel of an ncu report, aggregated per CUDA source line.
ncu's SASS view carries the samples; the line table comes from nvdisasm -g on the LOCAL object built from the same source
(instruction order is identical).  usage: ncu_lines.py <report.ncu-rep> <object.o> <mangled-name substring> <source.cu> [top]"""
import csv
import re
import subprocess
import sys
import tempfile

import os
rep, obj, key, src = sys.argv[1:5]
obj = os.path.abspath(obj)
top = int(sys.argv[5]) if len(sys.argv) > 5 else 30
tmp = tempfile.mkdtemp()
subprocess.check_call(f"cd {tmp} && cuobjdump -xelf all {obj} > /dev/null", shell=True)
cubin = subprocess.check_output(f"ls {tmp}/*.cubin", shell=True, text=True).split()[0]
sass = subprocess.check_output(["nvdisasm", "-g", "-c", cubin], text=True).split("\n")
fn, cur, seq = None, 0, []
for l in sass:
    m = re.match(r'\s*//## File "(.*?)", line (\d+)', l)
    if m:
        cur = int(m.group(2)); continue
    m = re.match(r"\.text\.(\S+):", l)
    if m:
        fn = m.group(1); continue
    if fn and key in fn and re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", l):
        seq.append(cur)
rows = list(csv.reader(subprocess.check_output(["ncu", "-i", rep, "--page", "source", "--csv"], text=True).split("\n")))
hdr = next(r for r in rows if "# Samples" in r)
data = [r for r in rows[rows.index(hdr) + 1:] if len(r) == len(hdr)]
assert len(data) == len(seq), (len(data), len(seq))
si = hdr.index("# Samples")
cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
agg = {}
for ln, r in zip(seq, data):
    d = agg.setdefault(ln, [0, {}])
    d[0] += int(r[si])
    for c in cols:
        if int(r[c]):
            d[1][hdr[c][6:]] = d[1].get(hdr[c][6:], 0) + int(r[c])
text = open(src).read().split("\n")
tot = sum(v[0] for v in agg.values())
print("samples", tot)
for ln, (n, st) in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
    print(f"{ln:5d} {100 * n / tot:5.1f}%  {text[ln - 1].strip()[:88]:88s} {sorted(st.items(), key=lambda x: -x[1])[:3]}")
