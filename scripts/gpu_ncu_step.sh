#!/bin/bash
# Full ncu capture of every kernel of ONE step of the bench command (after warm-up): bash scripts/gpu_ncu_step.sh <tag> [skip] [count]
# (19 launches per step with the tile grouping: 3 warm-up steps are skipped by default)
TAG=${1:-r2}; SKIP=${2:-57}; CNT=${3:-19}
OUT=gpurun_out; mkdir -p $OUT
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --sharded none --sweep none"
$CMD > $OUT/plain_step_$TAG.log 2>&1 &&
ncu --set full --import-source on --clock-control none -k "regex:^(group_|tile_|pair_|edge_|table_prepare|word_scan|dup_|scan|domain_check|classify|dcssc_row|vote|rescue|duplex)" -s $SKIP -c $CNT -f -o $OUT/step_$TAG $CMD > $OUT/ncu_step_$TAG.log 2>&1
tail -2 $OUT/ncu_step_$TAG.log
