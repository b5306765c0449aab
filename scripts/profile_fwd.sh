#!/bin/bash
# ncu --set full capture of the headline forward kernel (run under gpurun). usage: scripts/profile_fwd.sh <tag>
set -u
TAG=${1:-r1}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-extras"
$CMD > gpurun_out/plain_fwd_${TAG}.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:irbfn_fwd_kernel -s 3 -c 1 \
    -o gpurun_out/prof_fwd_${TAG} -f $CMD > gpurun_out/ncu_fwd_${TAG}.log 2>&1
tail -2 gpurun_out/ncu_fwd_${TAG}.log
