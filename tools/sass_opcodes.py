#!/usr/bin/env python
"""Per-kernel SASS opcode histogram of liborbx.so (cuobjdump -sass), for profiles/rNN_sass_opcodes.txt: shows which
Blackwell instructions each kernel actually uses (UTMALDG = TMA box loads, SYNCS = mbarrier, VIMNMX3 = DPX 3-input
min/max, IDP = byte/short dot products, REDUX = warp reductions, ...). Runs without a GPU.
Usage: python tools/sass_opcodes.py [path/to/liborbx.so] > profiles/r02_sass_opcodes.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "orb_slam2_2021_b200", "liborbx.so")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
kern, hist, arch = None, collections.OrderedDict(), set()
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        hist[kern] = collections.Counter()
        continue
    m = re.match(r"\s*arch = (\S+)", line)
    if m:
        arch.add(m.group(1))
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+(?:\.[A-Z0-9_]+)*)", line)
    if m and kern:
        hist[kern][m.group(1)] += 1
NOTE = ("UTMALDG", "SYNCS", "VIMNMX3", "VIMNMX", "IDP", "REDUX", "VABSDIFF4", "PRMT", "SHF", "LDS", "STS", "LDG", "STG", "ATOMS", "ATOMG", "RED",
        "VOTE", "POPC", "SHFL", "MATCH", "BAR", "UBLKCP", "FENCE", "MEMBAR", "NANOSLEEP")
print(f"# {os.path.relpath(so, ROOT)}: cubins for {sorted(arch)}; static instruction counts per kernel (cuobjdump -sass)")
for k, h in hist.items():
    tot = sum(h.values())
    base = collections.Counter()
    for op, n in h.items():
        base[op.split(".")[0]] += n
    print(f"\n## {k}  ({tot} instructions)")
    print("   notable: " + ", ".join(f"{op} {base[op]}" for op in NOTE if base.get(op)))
    full = [f"{op} {n}" for op, n in sorted(h.items(), key=lambda x: -x[1]) if any(op.startswith(t) for t in ("UTMALDG", "SYNCS", "VIMNMX3", "IDP", "REDUX", "VABSDIFF4"))]
    if full:
        print("   forms:   " + ", ".join(full))
    print("   top:     " + ", ".join(f"{op} {n}" for op, n in base.most_common(12)))
