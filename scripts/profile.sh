#!/bin/bash
# ncu evidence for one round (run under gpurun; writes gpurun_out/):
#   launches_<tag>.csv   every launch of a short bench run with its device time
#   prof_<tag>.ncu-rep   --set full capture of the forward, backward-P and backward-X kernels
# usage: scripts/profile.sh <tag>
set -u
TAG=${1:-r1}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-other"
$CMD > gpurun_out/plain_${TAG}.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
    --log-file gpurun_out/launches_${TAG}.csv $CMD > gpurun_out/ncu_launches_${TAG}.log 2>&1
$CMD > gpurun_out/plain2_${TAG}.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:irbfn_fwd_kernel -s 3 -c 1 \
    -o gpurun_out/prof_fwd_${TAG} -f $CMD > gpurun_out/ncu_fwd_${TAG}.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:irbfn_bwd_p_kernel -s 3 -c 1 \
    -o gpurun_out/prof_bwdp_${TAG} -f $CMD > gpurun_out/ncu_bwdp_${TAG}.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:irbfn_bwd_x_kernel -s 1 -c 1 \
    -o gpurun_out/prof_bwdx_${TAG} -f $CMD > gpurun_out/ncu_bwdx_${TAG}.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:irbfn_jac_kernel -s 1 -c 1 \
    -o gpurun_out/prof_jac_${TAG} -f $CMD > gpurun_out/ncu_jac_${TAG}.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:nearest_row_kernel -s 1 -c 1 \
    -o gpurun_out/prof_nearest_${TAG} -f $CMD > gpurun_out/ncu_nearest_${TAG}.log 2>&1
ls -la gpurun_out/
