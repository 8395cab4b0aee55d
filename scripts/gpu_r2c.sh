#!/bin/bash
mkdir -p gpurun_out
BENCH_ARGS="--sharded none --sweep none" bash scripts/gpu_launches.sh r2c
python scripts/launch_summary.py gpurun_out/launches_r2c.csv 5
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --sharded none --sweep none"
ncu --set full --import-source on --clock-control none -k "regex:^tile_" -s 9 -c 3 -f -o gpurun_out/tile_r2c $CMD > gpurun_out/ncu_tile_r2c.log 2>&1
tail -3 gpurun_out/ncu_tile_r2c.log
ls -la gpurun_out/*.ncu-rep
