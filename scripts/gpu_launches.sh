#!/bin/bash
# ncu launch list (device time of every libccgpu kernel) of the short bench command
TAG=${1:-r1}; OUT=gpurun_out; mkdir -p $OUT
KREGEX='regex:^(group_|tile_|pair_|edge_|table_prepare|word_scan|dcssc_row|dup_|expand_|coord_|hash|radix_|os_|scan|boundary|fix_mixed_runs|mark_first|creation_rank|family_build|domain_check|join_build|join_probe|classify|vote|rescue|duplex|agree|read_fam)'
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline ${BENCH_ARGS:-}"
$CMD > $OUT/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "$KREGEX" -c 900 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_$TAG.log 2>&1
tail -2 $OUT/ncu_$TAG.log
