#!/bin/bash
# round-2 evidence: full ncu capture of one step, soak in all three grouping modes, CLI timing
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/r2l_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2l_pytest.log
bash scripts/gpu_variants.sh
bash scripts/gpu_ncu_step.sh r2l
python scripts/gpu_soak.py 100 300 > gpurun_out/r2l_soak_tile.log 2>&1; echo "soak tile rc=$?"; tail -1 gpurun_out/r2l_soak_tile.log
CCGPU_GROUP=hash python scripts/gpu_soak.py 500 120 > gpurun_out/r2l_soak_hash.log 2>&1; echo "soak hash rc=$?"; tail -1 gpurun_out/r2l_soak_hash.log
CCGPU_GROUP=sort python scripts/gpu_soak.py 700 120 > gpurun_out/r2l_soak_sort.log 2>&1; echo "soak sort rc=$?"; tail -1 gpurun_out/r2l_soak_sort.log
python scripts/sanitize_cases.py > gpurun_out/r2l_cases.log 2>&1; echo "cases rc=$?"; tail -2 gpurun_out/r2l_cases.log
python scripts/cli_timing.py 150000 > gpurun_out/r2l_cli_timing.log 2>&1; echo "cli rc=$?"; cat gpurun_out/r2l_cli_timing.log
