#!/bin/bash
# FAST kernel variants: threads per cell (ORBX_FAST_THREADS = 128 / 64 / 32) x network fed from the widened copies or from
# the raw tile (ORBX_FAST_RAW = 0 / 1): parity subset + device-resident bench
for raw in 0 1; do for nt in 128 64 32; do
  export ORBX_FAST_THREADS=$nt ORBX_FAST_RAW=$raw
  python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "handles or chunk or parameter" 2>&1 | tail -1
  python bench.py --no-cpu-baseline --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('raw=$raw nt=$nt', round(d['value']), round(d['e2e']['value']), d['roofline']['stage_ms_per_step']['fast'])"
done; done
