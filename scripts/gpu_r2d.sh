#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py -m gpu -q -x > gpurun_out/r2d_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/r2d_pytest.log
bash scripts/gpu_variants.sh
BENCH_ARGS="--sharded none --sweep none" bash scripts/gpu_launches.sh r2d
python scripts/launch_summary.py gpurun_out/launches_r2d.csv 5
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --sharded none --sweep none"
ncu --set full --import-source on --clock-control none -k "regex:^tile_(insert|mark|family)" -s 9 -c 3 -f -o gpurun_out/tile_r2d $CMD > gpurun_out/ncu_tile_r2d.log 2>&1
tail -2 gpurun_out/ncu_tile_r2d.log
