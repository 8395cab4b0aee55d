#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/r2i_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/r2i_pytest.log
bash scripts/gpu_variants.sh
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --sharded none --sweep none"
ncu --set full --import-source on --clock-control none -k "regex:^vote_small_kernel" -s 6 -c 1 -f -o gpurun_out/vote_r2i $CMD > gpurun_out/ncu_vote_r2i.log 2>&1
tail -2 gpurun_out/ncu_vote_r2i.log
