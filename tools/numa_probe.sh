#!/bin/bash
# End-to-end throughput at N GPUs with and without NUMA-local pinned buffers (--tune numa_off=1 / 0), plus the host topology.
# Usage: tools/numa_probe.sh N      (under gpurun --gpus N)
N=${1:-8}
mkdir -p gpurun_out
{
  nvidia-smi topo -m
  lscpu | grep -i -E "numa|socket|model name|^CPU\(s\)"
  for d in $(nvidia-smi --query-gpu=pci.bus_id --format=csv,noheader | tr 'A-Z' 'a-z' | sed 's/^0000//'); do
    echo "$d numa_node=$(cat /sys/bus/pci/devices/$d/numa_node 2>/dev/null)"
  done
  grep -H . /sys/devices/system/node/node*/cpulist 2>/dev/null
  taskset -p $$
} > gpurun_out/topo_${N}gpu.txt 2>&1
for numa in 0 1; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$numa \
      bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline --tune numa_off=$((1-numa)) > gpurun_out/bench_${N}gpu_numa$numa.json 2> gpurun_out/bench_${N}gpu_numa$numa.err
  python - <<PY
import json
try:
    j = json.loads(open("gpurun_out/bench_${N}gpu_numa$numa.json").read().strip().splitlines()[-1])
    print("numa placement=$numa value", round(j["value"]), "e2e", round(j["e2e"]["value"]), "blocking", round(j["e2e"]["blocking_call"]["value"]), j["e2e"].get("host_numa"))
except Exception as e:
    print("numa placement=$numa failed", e)
PY
done
tail -3 gpurun_out/bench_${N}gpu_numa1.err
