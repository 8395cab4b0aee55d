#!/bin/bash
# gpurun with retries while the pod answers "transient" (nothing charged): scripts/gpurun_retry.sh [gpurun args] -- cmd
for i in 1 2 3 4 5 6 7 8 9 10; do
  out=$(/usr/local/graft/bin/gpurun "$@" 2>&1)
  echo "$out"
  echo "$out" | grep -q "status=transient\|status=busy\|no box\|retry in a few" || exit 0
  sleep 120
done
