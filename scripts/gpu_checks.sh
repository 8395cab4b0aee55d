#!/bin/bash
# the parity suite and the small soak cases through a library built with -DCCGPU_CHECKS (device-side bounds asserts)
mkdir -p gpurun_out
[ -f variants/checks.so ] || bash scripts/build_variant.sh checks -DCCGPU_CHECKS
CCGPU_LIB=$PWD/variants/checks.so python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py tests/test_gpu_stream.py -m gpu -q > gpurun_out/checks_pytest.log 2>&1
echo "checks pytest rc=$?"; tail -3 gpurun_out/checks_pytest.log
CCGPU_LIB=$PWD/variants/checks.so python scripts/sanitize_cases.py > gpurun_out/checks_cases.log 2>&1; echo "checks cases rc=$?"; tail -2 gpurun_out/checks_cases.log
CCGPU_LIB=$PWD/variants/checks.so python scripts/gpu_soak.py 2000 60 > gpurun_out/checks_soak.log 2>&1; echo "checks soak rc=$?"; tail -1 gpurun_out/checks_soak.log
