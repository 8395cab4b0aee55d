#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py -m gpu -q -x > gpurun_out/r2g_pytest.log 2>&1
echo "pytest rc=$?"; tail -6 gpurun_out/r2g_pytest.log
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --sharded none --sweep none > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err
echo "bench rc=$?"; tail -3 gpurun_out/r2g_bench.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2g_bench.json') if l.startswith('{')][0])
print(d['value']/1e9, d['ms_per_step'], d['roofline']['phase_ms_per_step'])
e=d['e2e']; print({k:v for k,v in e.items()})
PY
