#!/bin/bash
# one bench run, the numbers that matter when comparing two builds: device-resident frames/s, e2e, per-stage ms
python bench.py --no-latency --no-cpu-baseline "$@" 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(round(d['value']), round(d['e2e']['value']), {k: round(v, 4) for k, v in d['roofline']['stage_ms_per_step'].items()})"
