"""Split an `ncu --page source --csv --print-source sass` dump at barriers/exits: warp instructions per kernel phase,
plus the opcode histogram. Usage: ncu_segments.py dump.csv [first_addr_suffix last_addr_suffix] (prints that SASS range)."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1], newline="")))
hdr = rows[1]
iI, iS = hdr.index("Instructions Executed"), hdr.index("# Samples")
body = [r for r in rows[2:] if len(r) >= len(hdr)]
if len(sys.argv) > 3:
    on = False
    for r in body:
        a = r[0][-5:]
        on = on or a == sys.argv[2]
        if on:
            print(a, f"{int(r[iI]):9d} {int(r[iS]):5d}", r[1].strip())
        if a == sys.argv[3]:
            break
    sys.exit(0)
tot = sum(int(r[iI]) for r in body)
seg = samp = 0
ops = {}
print("total warp instructions", tot)
for r in body:
    n, s = int(r[iI]), int(r[iS])
    seg += n
    samp += s
    t = r[1].split()
    op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
    ops[op] = ops.get(op, 0) + n
    if "BAR." in r[1] or "EXIT" in r[1]:
        print(f"  up to {r[0][-5:]} {r[1].strip()[:44]:44s} {seg:11d} {100 * seg / tot:5.1f}%  samples {samp}")
        seg = samp = 0
for k, v in sorted(ops.items(), key=lambda x: -x[1])[:24]:
    print(f"  {k:10s} {v:11d} {100 * v / tot:5.1f}%")
