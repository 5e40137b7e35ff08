#!/usr/bin/env python
"""Copy the outputs of tools/round_measure.sh from gpurun_out/ into profiles/ (round prefix r01): bench lines, launch list,
per-stage traffic, config table, and the `--set full` summary of the last complete step as the new "current build" block of
profiles/r01_ncu_full_summaries.txt. Usage: python tools/refresh_profiles.py "<one-line description of the build>"."""
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
desc = sys.argv[1] if len(sys.argv) > 1 else "current build"
for src, dst in (("bench.json", "r01_bench.json"), ("bench_reference.json", "r01_bench_reference.json"), ("launches.csv", "r01_launches.csv"),
                 ("configs.json", "r01_configs.json")):
    shutil.copy(os.path.join(G, src), os.path.join(P, dst))
with open(os.path.join(P, "r01_traffic.json"), "w") as f:
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "traffic_from_launches.py"), os.path.join(G, "launches.csv"), "64"], stdout=f, check=True, cwd=ROOT)
summ = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), os.path.join(G, "full.ncu-rep"), "--stalls"],
                      capture_output=True, text=True, check=True).stdout.rstrip("\n").split("\n")
ents = [summ[i:i + 2] for i in range(0, len(summ) - 1) if summ[i].startswith("kernel=")]
last = max(i for i, e in enumerate(ents) if "median" in e[0])
first = max(i for i, e in enumerate(ents[:last]) if "resize" in e[0] and (i == 0 or "resize" not in ents[i - 1][0]))
keep = ents[first:last + 1]
path = os.path.join(P, "r01_ncu_full_summaries.txt")
L = open(path).read().split("\n")
hist = "## ---- earlier builds of this round (history) ----"
h = L.index(hist)
old = [l.replace("## current build,", "## build at", 1) for l in L[4:h] if l.strip()]
out = L[:4] + ["## current build, " + desc] + [l for e in keep for l in e] + ["", hist] + old + L[h + 1:]
open(path, "w").write("\n".join(out))
print("kernels in the current block:", len(keep))
