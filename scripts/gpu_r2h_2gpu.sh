#!/bin/bash
# two GPUs: the two-GPU parity tests, then the bench line as the driver launches it
mkdir -p gpurun_out
nvidia-smi -L
python -m pytest tests/test_gpu_sharded.py -m gpu -q -x -k "two_gpus or stats_allreduce" > gpurun_out/r2h_pytest.log 2>&1
echo "pytest rc=$?"; tail -5 gpurun_out/r2h_pytest.log
NCCL_DEBUG=INFO python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2h_bench2.json 2> gpurun_out/r2h_bench2.err
echo "bench rc=$?"; tail -4 gpurun_out/r2h_bench2.err | cut -c1-300
grep -c "ncclCommInitRank\|Init COMPLETE" gpurun_out/r2h_bench2.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2h_bench2.json') if l.startswith('{')][0])
print(d['n_gpus'], d['value']/1e9, d['ms_per_step'], d['e2e']['value']/1e6, d['collective'])
for k,v in (d.get('sharded') or {}).items():
    print(k, {x:v[x] for x in v if x not in ('workload','parity_rank0')})
    print('  parity', v.get('parity_rank0'))
PY
