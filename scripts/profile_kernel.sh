#!/bin/bash
# ncu --set full capture of one kernel of the bench run (run under gpurun).
# usage: scripts/profile_kernel.sh <kernel-regex> <tag> [skip]
set -u
KRE=$1; TAG=$2; SKIP=${3:-1}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-other"
$CMD > gpurun_out/plain_${TAG}.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:${KRE} -s ${SKIP} -c 1 \
    -o gpurun_out/prof_${TAG} -f $CMD > gpurun_out/ncu_${TAG}.log 2>&1
tail -2 gpurun_out/ncu_${TAG}.log
