#!/bin/bash
# N-GPU bench as the driver launches it (one rank per GPU, torchrun). Usage: tools/multi_gpu_bench.sh N
N=${1:-8}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err
tail -c 2500 gpurun_out/bench_${N}gpu.json
