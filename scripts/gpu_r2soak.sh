#!/bin/bash
mkdir -p gpurun_out
python scripts/gpu_soak.py 5000 300 > gpurun_out/r2z_soak.log 2>&1; echo "soak rc=$?"; tail -n 2 gpurun_out/r2z_soak.log
rm -f variants/checks.so; bash scripts/build_variant.sh checks -DCCGPU_CHECKS > /dev/null 2>&1
CCGPU_LIB=$PWD/variants/checks.so python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py tests/test_gpu_stream.py tests/test_stage_commands.py -m gpu -q > gpurun_out/r2z_checks_pytest.log 2>&1; echo "checks pytest rc=$?"; tail -n 2 gpurun_out/r2z_checks_pytest.log
CCGPU_LIB=$PWD/variants/checks.so python scripts/gpu_soak.py 6000 60 > gpurun_out/r2z_checks_soak.log 2>&1; echo "checks soak rc=$?"; tail -n 1 gpurun_out/r2z_checks_soak.log
