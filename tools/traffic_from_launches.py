#!/usr/bin/env python
"""Per-stage DRAM bytes and time share of ONE device-resident step, from an ncu launch list.

  ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
      --csv --log-file gpurun_out/launches.csv python bench.py --frames 64 --steps 1 --warmup 1 --no-cpu-baseline
  python tools/traffic_from_launches.py gpurun_out/launches.csv 64 > profiles/rNN_traffic.json

The longest complete step in the list (launches up to and including an orbx_stereo_median_kernel) is used. bench.py reads the result for roofline.traffic.
"""
import csv
import json
import sys

STAGE = {"orbx_ingest": "pyramid", "orbx_resize": "pyramid", "orbx_fast": "fast", "orbx_octree": "octree", "orbx_count": "octree",
         "orbx_blur": "blur", "orbx_describe": "describe", "orbx_stereo": "stereo"}


def stage_of(kernel):
    kernel = kernel[5:] if kernel.startswith("void ") else kernel  # templated kernels print their return type
    for k, v in STAGE.items():
        if kernel.startswith(k):
            return v
    return None


def main():
    path, frames = sys.argv[1], int(sys.argv[2])
    lines = [l for l in open(path) if l.startswith('"')]
    rows = list(csv.DictReader(lines))
    launches = {}
    for r in rows:
        d = launches.setdefault(int(r["ID"]), {"kernel": r["Kernel Name"]})
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        if r["Metric Name"] == "gpu__time_duration.sum":
            v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1.0)  # -> microseconds
            d["us"] = v
        else:
            v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
            d["bytes"] = d.get("bytes", 0.0) + v
    ids = sorted(launches)
    ours = [i for i in ids if stage_of(launches[i]["kernel"])]
    # split the list into steps (a step ends with its stereo-match launch) and keep the LONGEST one: the end-to-end
    # pipeline of bench.py runs the same kernels on 1/8-size chunks, which must not be mistaken for the full step
    steps, cur = [], []
    for i in ours:
        cur.append(i)
        if "orbx_stereo_median" in launches[i]["kernel"]:
            steps.append(cur)
            cur = []
    dur = [sum(launches[i]["us"] for i in st) for st in steps]
    step = [st for st, d in zip(steps, dur) if d >= 0.7 * max(dur)][-1]  # the last full-size one (warm)
    start, end = step[0], step[-1]
    stages, total_us = {}, sum(launches[i]["us"] for i in step)
    for i in step:
        s = stages.setdefault(stage_of(launches[i]["kernel"]), {"dram_bytes_per_frame": 0.0, "us": 0.0, "launches": 0})
        s["dram_bytes_per_frame"] += launches[i].get("bytes", 0.0) / frames
        s["us"] += launches[i]["us"]
        s["launches"] += 1
    for s in stages.values():
        s["share"] = s["us"] / total_us
        s[f"us_per_{frames}_frames"] = round(s.pop("us"), 3)
    out = {"source": f"{path} (ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum "
                     f"--clock-control none; bench.py --frames {frames}; launches {start}-{end} = one device-resident step)",
           "frames_per_launch": frames, "stages": stages,
           "total_dram_bytes_per_frame": sum(s["dram_bytes_per_frame"] for s in stages.values())}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
