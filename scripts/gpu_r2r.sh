#!/bin/bash
# pair grouping: whole suite, soak, the bounds-checked build, launch list + ncu step capture, bench
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r2r_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2r_pytest.log
python scripts/gpu_soak.py 3000 300 > gpurun_out/r2r_soak_pair.log 2>&1; echo "soak pair rc=$?"; tail -1 gpurun_out/r2r_soak_pair.log
python scripts/sanitize_cases.py > gpurun_out/r2r_cases.log 2>&1; echo "cases rc=$?"; tail -1 gpurun_out/r2r_cases.log
bash scripts/gpu_checks.sh
BENCH_ARGS="--sharded none --sweep none" bash scripts/gpu_launches.sh r2r > /dev/null; python scripts/launch_summary.py gpurun_out/launches_r2r.csv 5
bash scripts/gpu_ncu_step.sh r2r
