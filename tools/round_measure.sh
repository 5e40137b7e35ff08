#!/bin/bash
# One gpurun call that regenerates everything under profiles/ for the current build (run on the GPU box, 1 GPU).
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -2 > gpurun_out/pytest_gpu.txt
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err
# config 4 of BASELINE.json as written: 4096 frames per step, strong scaling (here N = 1: all 4096 on one GPU)
python bench.py --scaling strong --total-frames 4096 --steps 5 --warmup 3 --no-cpu-baseline --no-latency > gpurun_out/bench_strong.json 2> gpurun_out/bench_strong.err
python tools/latency_probe.py > gpurun_out/latency.txt 2>&1
# launch list with DRAM bytes (only after the plain run exited 0)
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
    --log-file gpurun_out/launches.csv python bench.py --frames 64 --steps 1 --warmup 1 --no-cpu-baseline --no-latency > gpurun_out/ncu_launches.log 2>&1
# one full capture of every kernel of one device-resident step (single chunk so that every stage is one launch)
ncu --set full --import-source on --clock-control none -k regex:orbx_ -c 40 -o gpurun_out/full \
    python bench.py --frames 64 --steps 1 --warmup 1 --no-cpu-baseline --no-latency --tune run_chunks=1 > gpurun_out/ncu_full.log 2>&1
python tools/measure_configs.py > gpurun_out/configs.json 2> gpurun_out/configs.err
tail -1 gpurun_out/pytest_gpu.txt
cat gpurun_out/bench.json | cut -c1-400
