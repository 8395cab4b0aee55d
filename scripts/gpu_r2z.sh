#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_stage_commands.py tests/test_bam_cli.py -m gpu -q > gpurun_out/r2z_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 5 gpurun_out/r2z_pytest.log
