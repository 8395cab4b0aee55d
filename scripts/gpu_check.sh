#!/bin/bash
# GPU-box routine: parity tests, bench, reference arm, ncu launch list of the bench command.
# Usage (under gpurun): bash scripts/gpu_check.sh [tag]
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
KREGEX='regex:^(group_|dup_|expand_|coord_|hash|os_|scan|boundary|fix_mixed_runs|mark_first|creation_rank|family_build|domain_check|join_build|join_probe|classify|vote|rescue|duplex|agree|read_fam)'
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee $OUT/pytest_$TAG.log
python bench.py 2> $OUT/bench_$TAG.err | tee $OUT/bench_$TAG.json
python bench.py --impl reference --steps 2 --warmup 1 2> $OUT/bench_ref_$TAG.err | tee $OUT/bench_ref_$TAG.json
BENCH_SMALL="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$BENCH_SMALL > $OUT/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "$KREGEX" -c 600 --csv \
    --log-file $OUT/launches_$TAG.csv $BENCH_SMALL > $OUT/ncu_$TAG.log 2>&1
tail -3 $OUT/ncu_$TAG.log
tail -c 600 $OUT/bench_$TAG.err
