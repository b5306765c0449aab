#!/bin/bash
# ncu launch list of a short bench run (run under gpurun). usage: scripts/launches.sh <tag>
set -u
TAG=${1:-r1}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-other"
$CMD > gpurun_out/plain_${TAG}.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv \
    --log-file gpurun_out/launches_${TAG}.csv $CMD > gpurun_out/ncu_launches_${TAG}.log 2>&1
tail -2 gpurun_out/ncu_launches_${TAG}.log
