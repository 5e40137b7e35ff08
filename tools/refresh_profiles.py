#!/usr/bin/env python
"""Copy the outputs of tools/round_measure.sh from gpurun_out/ into profiles/ (round prefix R below): bench lines, launch list,
per-stage traffic, config table, and the `--set full` summary of the last complete step as the new "current build" block of
profiles/<R>_ncu_full_summaries.txt. Usage: python tools/refresh_profiles.py "<one-line description of the build>"."""
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
desc = sys.argv[1] if len(sys.argv) > 1 else "current build"
R = "r02"
for src, dst in (("bench.json", R + "_bench.json"), ("bench_reference.json", R + "_bench_reference.json"), ("launches.csv", R + "_launches.csv"),
                 ("configs.json", R + "_configs.json"), ("bench_strong.json", R + "_bench_strong_4096.json"), ("latency.txt", R + "_latency.txt")):
    if os.path.exists(os.path.join(G, src)):
        shutil.copy(os.path.join(G, src), os.path.join(P, dst))
with open(os.path.join(P, R + "_traffic.json"), "w") as f:
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "traffic_from_launches.py"), os.path.join(G, "launches.csv"), "64"], stdout=f, check=True, cwd=ROOT)
summ = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), os.path.join(G, "full.ncu-rep"), "--stalls"],
                      capture_output=True, text=True, check=True).stdout.rstrip("\n").split("\n")
ents = [summ[i:i + 2] for i in range(0, len(summ) - 1) if summ[i].startswith("kernel=")]
last = max(i for i, e in enumerate(ents) if "median" in e[0])
first = max(i for i, e in enumerate(ents[:last]) if "resize" in e[0] and (i == 0 or "resize" not in ents[i - 1][0]))
keep = ents[first:last + 1]
path = os.path.join(P, R + "_ncu_full_summaries.txt")
if not os.path.exists(path):
    open(path, "w").write("# ncu --set full --import-source on --clock-control none, one device-resident step (64 stereo frames = 128 images per launch,\n"
                          "# tuning key run_chunks=1 so that every stage is one launch), summarised by tools/ncu_summary.py --stalls.\n"
                          "# Absolute times are cold-cache and serialised (see profiles/%s_launches.csv for the launch list of the same command and\n"
                          "# profiles/%s_bench.json for the event-timed bench); what matters here is what binds each kernel.\n\n"
                          "## ---- earlier builds of this round (history) ----\n" % (R, R))
L = open(path).read().split("\n")
hist = "## ---- earlier builds of this round (history) ----"
h = L.index(hist)
old = [l.replace("## current build,", "## build at", 1) for l in L[4:h] if l.strip()]
out = L[:4] + ["## current build, " + desc] + [l for e in keep for l in e] + ["", hist] + old + L[h + 1:]
open(path, "w").write("\n".join(out))
print("kernels in the current block:", len(keep))
