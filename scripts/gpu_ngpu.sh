#!/bin/bash
# N GPUs: the bench line as the driver launches it (torchrun, one rank per GPU): bash scripts/gpu_ngpu.sh <tag> <N> [N ...]
TAG=$1; shift
mkdir -p gpurun_out
for n in "$@"; do
NCCL_DEBUG=WARN python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/${TAG}_bench$n.json 2> gpurun_out/${TAG}_bench$n.err
echo "bench $n rc=$?"; tail -3 gpurun_out/${TAG}_bench$n.err | cut -c1-300
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/${TAG}_bench$n.json') if l.startswith('{')][0])
print(d['n_gpus'], 'value %.2f G' % (d['value']/1e9), d['ms_per_step'], 'e2e %.1f M' % (d['e2e']['value']/1e6), d['e2e']['ms_per_step'], d['collective']['comm'])
for k,v in (d.get('sharded') or {}).items():
    print(k, {x:v[x] for x in v if x not in ('workload','parity_rank0','reads_per_rank','bands_per_rank')})
    print('  parity', json.dumps(v.get('parity_rank0'))[:400])
PY
done
