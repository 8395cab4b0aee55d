#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/r2m_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2m_pytest.log
bash scripts/gpu_variants.sh
bash scripts/gpu_checks.sh
bash scripts/gpu_ncu_step.sh r2m
python scripts/cli_timing.py 150000 > gpurun_out/r2m_cli_timing.log 2>&1; echo "cli rc=$?"; tail -6 gpurun_out/r2m_cli_timing.log
