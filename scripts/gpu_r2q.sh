#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py tests/test_gpu_stream.py tests/test_gpu_fullsize.py -m gpu -q -x > gpurun_out/r2q_pytest.log 2>&1
echo "pytest rc=$?"; tail -8 gpurun_out/r2q_pytest.log
for g in pair tile; do
  CCGPU_GROUP=$g python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu-baseline --sharded none --sweep none > gpurun_out/r2q_bench_$g.json 2> gpurun_out/r2q_bench_$g.err
  echo "$g rc=$?"; python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2q_bench_$g.json') if l.startswith('{')][0])
print('$g', d['value']/1e9, d['ms_per_step'], d['roofline']['phase_ms_per_step'], d['counts'])
PY
done
for w in c3 c4; do
  python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --sharded none --sweep none --workload $w --pairs 20000000 > gpurun_out/r2q_bench_$w.json 2> gpurun_out/r2q_bench_$w.err
  python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2q_bench_$w.json') if l.startswith('{')][0])
print('$w', d['value']/1e9, d['ms_per_step'], d['roofline']['phase_ms_per_step'])
PY
done
