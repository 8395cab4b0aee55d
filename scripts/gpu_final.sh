#!/bin/bash
# round-end sequence on one GPU + the ncu capture of one step (18 launches per step on the pair path)
TAG=${1:-r3b}
bash scripts/gpu_full.sh $TAG
bash scripts/gpu_ncu_step.sh $TAG 54 18
