#!/bin/bash
# FAST kernel: threads per cell sweep (ORBX_FAST_THREADS = 128 / 64 / 32): parity subset + device-resident bench
for nt in 128 64 32; do
  export ORBX_FAST_THREADS=$nt
  python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "handles or chunk" 2>&1 | tail -1
  python bench.py --no-cpu-baseline --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('nt=$nt', round(d['value']), round(d['e2e']['value']), d['roofline']['stage_ms_per_step']['fast'])"
done
