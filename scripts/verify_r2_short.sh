#!/bin/bash
# Short round-2 verification on one B200 (run under gpurun): GPU tests, smoke, both bench arms, one full ncu capture
# of the small-batch latency kernel (B = 400 on the C2 shape).  usage: scripts/verify_r2_short.sh <tag>
set -u
TAG=${1:-r2}
O=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > $O/pytest_gpu_${TAG}.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke_${TAG}.log 2>&1; echo smoke=$? >> $O/smoke_${TAG}.log
python bench.py > $O/bench_${TAG}_n1.json 2> $O/bench_${TAG}_n1.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_${TAG}_reference.json 2> $O/bench_${TAG}_ref.err
LAT_B=400 timeout 120 ncu --set full --clock-control none --import-source on -k regex:irbfn_fwd_small_kernel -s 2 -c 1 \
    -o $O/prof_fwdsmall_${TAG} -f python scripts/dev_latency.py --profile > $O/ncu_fwdsmall_${TAG}.log 2>&1
tail -n 3 $O/pytest_gpu_${TAG}.log $O/smoke_${TAG}.log
cut -c1-300 $O/bench_${TAG}_n1.json; tail -c 1200 $O/bench_${TAG}_n1.json; cut -c1-300 $O/bench_${TAG}_reference.json
tail -n 2 $O/ncu_fwdsmall_${TAG}.log
