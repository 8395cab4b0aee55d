#!/bin/bash
# round-end sequence on one GPU + the ncu capture of one step
TAG=${1:-r2y}
bash scripts/gpu_full.sh $TAG
bash scripts/gpu_ncu_step.sh $TAG 51 17
