"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump per CUDA source line:
warp instructions executed, share of the kernel, stall samples. Usage: ncu_lines.py dump.csv [file-filter]"""
import csv
import sys
from collections import defaultdict

rows = list(csv.reader(open(sys.argv[1], newline="")))
flt = sys.argv[2] if len(sys.argv) > 2 else ""
inst = defaultdict(int)
samp = defaultdict(int)
src = {}
fname, hdr, line = None, None, None
for r in rows:
    if not r:
        continue
    if r[0] in ("File Name", "File Path"):
        fname = r[1]
        continue
    if r[0] == "Line No":
        hdr = r
        iI, iS = hdr.index("Instructions Executed"), hdr.index("# Samples")
        continue
    if hdr is None or len(r) < len(hdr):
        continue
    if r[0] != "":  # a source-line row carries the sums over its SASS rows (which follow with an empty first column)
        line = (fname or "?", int(r[0]))
        src[line] = r[1]
        try:
            inst[line] += int(r[iI])
            samp[line] += int(r[iS])
        except ValueError:
            pass
tot = sum(inst.values()) or 1
tots = sum(samp.values()) or 1
print(f"total warp instructions {tot}, samples {tots}")
for k in sorted(inst):
    if flt and flt not in k[0]:
        continue
    if inst[k] == 0:
        continue
    print(f"{k[0].split('/')[-1]}:{k[1]:4d} {inst[k]:12d} {100*inst[k]/tot:5.1f}% samp {100*samp[k]/tots:5.1f}%  {src[k].strip()[:110]}")
