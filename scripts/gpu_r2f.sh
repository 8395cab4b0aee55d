#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py tests/test_gpu_fullsize.py -m gpu -q -x > gpurun_out/r2f_pytest.log 2>&1
echo "pytest rc=$?"; tail -12 gpurun_out/r2f_pytest.log
bash scripts/gpu_variants.sh
BENCH_ARGS="--sharded none --sweep none" bash scripts/gpu_launches.sh r2f
python scripts/launch_summary.py gpurun_out/launches_r2f.csv 5
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --sharded none --sweep none"
ncu --set full --import-source on --clock-control none -k "regex:^tile_(local|family|place)" -s 12 -c 4 -f -o gpurun_out/tile_r2f $CMD > gpurun_out/ncu_tile_r2f.log 2>&1
tail -2 gpurun_out/ncu_tile_r2f.log
for w in c3 c4; do
  python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --sharded none --sweep none --workload $w --pairs 20000000 > gpurun_out/r2f_bench_$w.json 2> gpurun_out/r2f_bench_$w.err
  python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2f_bench_$w.json') if l.startswith('{')][0])
print('$w', d['value']/1e9, d['ms_per_step'], d['roofline']['phase_ms_per_step'])
PY
done
