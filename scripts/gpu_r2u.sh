#!/bin/bash
# final-build validation on one GPU: the re-vote test under both builds, then the whole suite + bench
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -q -k "voting_a_batch_again or stagewise" > gpurun_out/r2u_revote.log 2>&1; echo "revote rc=$?"; tail -3 gpurun_out/r2u_revote.log
CCGPU_LIB=$PWD/variants/checks.so python -m pytest tests/test_gpu_parity.py -m gpu -q -k "voting_a_batch_again or stagewise or fuzz" > gpurun_out/r2u_revote_checks.log 2>&1; echo "revote checks rc=$?"; tail -3 gpurun_out/r2u_revote_checks.log
bash scripts/gpu_full.sh r2u
