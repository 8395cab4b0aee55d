#!/bin/bash
# packed result rows: tests, then the bench line (e2e legs)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_sharded.py -m gpu -q -k "planes or compact or short_pair or qualities or stay_as" > gpurun_out/r2w_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r2w_pytest.log
python bench.py --gpus 1 --steps 10 --warmup 3 --sharded none --sweep none --no-cpu-baseline > gpurun_out/r2w_bench.json 2> gpurun_out/r2w_bench.err; echo "bench rc=$?"; tail -n 2 gpurun_out/r2w_bench.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2w_bench.json') if l.startswith('{')][0])
print(d['ms_per_step'], json.dumps(d['e2e'])[:1800])
PY
