#!/bin/bash
# ncu --set full on single modes of the pipe micro-benchmark (build/peak <mode>); run under gpurun.
set -u
for m in "$@"; do
  ./build/peak $m > gpurun_out/peak_plain_$m.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:kernel -s 1 -c 1 \
      -o gpurun_out/peak_mode$m -f ./build/peak $m > gpurun_out/ncu_peak_$m.log 2>&1
done
ls gpurun_out
