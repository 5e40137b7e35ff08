#!/bin/bash
# End-to-end (host memory to host memory) throughput at N GPUs against the chunk / stream layout of the streaming call.
# Usage (under gpurun --gpus N): tools/e2e_sweep8.sh N
N=${1:-8}
mkdir -p gpurun_out
port=29600
for cfg in "0 0" "2 0" "8 0" "16 0" "4 1" "4 4" "8 4"; do set -- $cfg
  port=$((port+1))
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --no-latency --tune e2e_chunks=$1 --tune e2e_streams=$2 \
      > gpurun_out/e2e_${N}_c$1_s$2.json 2> gpurun_out/e2e_${N}_c$1_s$2.err
  python - <<PY
import json
try:
    j = json.loads(open("gpurun_out/e2e_${N}_c$1_s$2.json").read().strip().splitlines()[-1])
    e = j["e2e"]
    print("N=$N chunks=$1 streams=$2 value", round(j["value"]), "e2e", round(e["value"]), "blocking", round(e["blocking_call"]["value"]),
          "link", round(e["link_ceiling_fps"] or 0), "frac", round(e["frac_of_link"] or 0, 3))
except Exception as ex:
    print("N=$N chunks=$1 streams=$2 failed", ex)
PY
done
