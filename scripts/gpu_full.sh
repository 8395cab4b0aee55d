#!/bin/bash
# the round-end sequence on one GPU: whole -m gpu suite, smoke, the default bench line, launch list
TAG=${1:-r2}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/${TAG}_pytest.log
python __graft_entry__.py smoke 2>&1 | tail -1
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench rc=$?"; tail -2 gpurun_out/${TAG}_bench.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/${TAG}_bench.json') if l.startswith('{')][0])
print('value %.3f G reads/s, %.3f ms/step, e2e %.1f M reads/s, roofline %.3f' % (d['value']/1e9, d['ms_per_step'], d['e2e']['value']/1e6, d['roofline']['frac']))
print(d['roofline']['phase_ms_per_step'])
print('cpu', d['cpu_baseline'])
for k,v in (d.get('sharded') or {}).items(): print(k, v.get('ms_per_step'), v.get('value'), v.get('failed'))
print('sweep fails', [c for c in d['sweep_c5']['cases'] if 'failed' in c])
PY
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_ref.json 2>/dev/null; tail -c 600 gpurun_out/${TAG}_ref.json
BENCH_ARGS="--sharded none --sweep none" bash scripts/gpu_launches.sh $TAG > /dev/null
python scripts/launch_summary.py gpurun_out/launches_$TAG.csv 5
