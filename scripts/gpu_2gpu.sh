#!/bin/bash
# two GPUs: the multi-GPU tests (sharded datasets through NCCL, chunked walk over two contexts), then the bench line
TAG=${1:-r2}
mkdir -p gpurun_out
python -m pytest tests/test_gpu_sharded.py tests/test_gpu_stream.py -m gpu -q > gpurun_out/${TAG}_pytest2.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest2.log
NCCL_DEBUG=WARN python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/${TAG}_bench2.json 2> gpurun_out/${TAG}_bench2.err
echo "bench rc=$?"
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/${TAG}_bench2.json') if l.startswith('{')][0])
print(d['n_gpus'], 'value %.2f G' % (d['value']/1e9), d['ms_per_step'], 'e2e %.1f M' % (d['e2e']['value']/1e6), d['collective']['comm'])
for k,v in (d.get('sharded') or {}).items(): print(k, v.get('ms_per_step'), v.get('value'), v.get('failed'), v.get('skipped'))
PY
