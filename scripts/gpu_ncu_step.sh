#!/bin/bash
# Full ncu capture of every kernel of ONE step of the bench command (after warm-up): bash scripts/gpu_ncu_step.sh <tag> [skip] [count]
TAG=${1:-r1}; SKIP=${2:-84}; CNT=${3:-21}
OUT=gpurun_out; mkdir -p $OUT
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$CMD > $OUT/plain_step_$TAG.log 2>&1 &&
ncu --set full --clock-control none -k "regex:^(group_|dup_|scan|domain_check|classify|vote|rescue|duplex)" -s $SKIP -c $CNT -f -o $OUT/step_$TAG $CMD > $OUT/ncu_step_$TAG.log 2>&1
tail -2 $OUT/ncu_step_$TAG.log
