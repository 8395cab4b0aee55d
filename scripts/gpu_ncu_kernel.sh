#!/bin/bash
# Full ncu capture of one kernel of the bench command.  Usage: bash scripts/gpu_ncu_kernel.sh <kernel-regex> <tag> [skip] [count]
K=${1:-vote_kernel}; TAG=${2:-r1}; SKIP=${3:-3}; CNT=${4:-2}
OUT=gpurun_out; mkdir -p $OUT
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$CMD > $OUT/plain_ncu_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:$K" -s $SKIP -c $CNT -f -o $OUT/prof_$TAG $CMD > $OUT/ncu_full_$TAG.log 2>&1
tail -3 $OUT/ncu_full_$TAG.log
ls -la $OUT/prof_$TAG.ncu-rep
