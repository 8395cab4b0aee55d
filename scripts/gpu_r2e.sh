#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py tests/test_gpu_fullsize.py -m gpu -q -x > gpurun_out/r2e_pytest.log 2>&1
echo "pytest rc=$?"; tail -12 gpurun_out/r2e_pytest.log
bash scripts/gpu_variants.sh
BENCH_ARGS="--sharded none --sweep none" bash scripts/gpu_launches.sh r2e
python scripts/launch_summary.py gpurun_out/launches_r2e.csv 5
for w in c3 c4; do
  python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --sharded none --sweep none --workload $w --pairs 20000000 > gpurun_out/r2e_bench_$w.json 2> gpurun_out/r2e_bench_$w.err
  python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2e_bench_$w.json') if l.startswith('{')][0])
print('$w', d['value']/1e9, d['ms_per_step'], d['roofline']['phase_ms_per_step'])
PY
done
