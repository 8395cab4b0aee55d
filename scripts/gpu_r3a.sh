#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/r3a_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 2 gpurun_out/r3a_pytest.log
python scripts/gpu_phases.py c3 200000000 8
python scripts/gpu_phases.py c3 200000000
