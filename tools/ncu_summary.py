#!/usr/bin/env python
"""Summarise an .ncu-rep (raw page) per kernel: duration, DRAM bytes, pipe utilisation, occupancy, top stalls.
Usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep [--stalls]"""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
def col(name):
    return hdr.index(name) if name in hdr else -1
keys = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"), ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_active", "l1%"), ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("sm__inst_issued.avg.pct_of_peak_sustained_active", "issue%"),
        ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu%"), ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma%"),
        ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu%"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smem_wf%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"), ("launch__registers_per_thread", "regs"),
        ("smsp__inst_executed.sum", "winst"), ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "bank_confl")]
units = rows[1]
for r in rows[2:]:
    out = []
    for k, short in keys:
        i = col(k)
        if i < 0: continue
        v = r[i]
        if short == "kernel": v = v.split("(")[0][-28:]
        else:
            try: v = f"{float(v):.4g}"
            except ValueError: pass
            if short in ("dram_rd", "dram_wr", "us"): v += units[i]
        out.append(f"{short}={v}")
    print("  ".join(out))
    if "--stalls" in sys.argv:
        st = [(float(r[i]), h) for i, h in enumerate(hdr) if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio") and r[i]]
        st.sort(reverse=True)
        print("     stalls: " + ", ".join(f"{h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]}={v:.2f}" for v, h in st[:6]))
