#!/bin/bash
# round 2, first GPU call: the whole -m gpu suite, then the default bench line (with the sharded legs)
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r2a_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
tail -5 gpurun_out/r2a_pytest.log
NCCL_DEBUG=WARN python bench.py --steps 20 --warmup 5 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err
echo "bench rc=$?"
tail -c 3000 gpurun_out/r2a_bench.json
tail -5 gpurun_out/r2a_bench.err
