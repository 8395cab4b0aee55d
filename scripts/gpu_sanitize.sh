#!/bin/bash
# compute-sanitizer over the small cases of scripts/sanitize_cases.py; logs -> gpurun_out/sanitize_*.log
mkdir -p gpurun_out
python -c "from oracle import oracle; oracle.build()"
for tool in memcheck racecheck synccheck initcheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python scripts/sanitize_cases.py > gpurun_out/sanitize_$tool.log 2>&1
  echo "$tool rc=$? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' gpurun_out/sanitize_$tool.log | tail -1)"
done
for g in hash sort; do
  CCGPU_GROUP=$g timeout 900 compute-sanitizer --tool memcheck --print-limit 20 python scripts/sanitize_cases.py > gpurun_out/sanitize_memcheck_$g.log 2>&1
  echo "memcheck $g rc=$? : $(grep -E 'ERROR SUMMARY' gpurun_out/sanitize_memcheck_$g.log | tail -1)"
done
