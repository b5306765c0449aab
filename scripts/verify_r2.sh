#!/bin/bash
# Round-2 verification on one B200 (run under gpurun): GPU tests, smoke, both bench arms, launch list, full ncu
# captures of the headline forward and of the gate-sparse backward kernels.  usage: scripts/verify_r2.sh <tag>
set -u
TAG=${1:-r2}
O=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > $O/pytest_gpu_${TAG}.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke_${TAG}.log 2>&1; echo smoke=$? >> $O/smoke_${TAG}.log
python bench.py > $O/bench_${TAG}_n1.json 2> $O/bench_${TAG}_n1.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_${TAG}_reference.json 2> $O/bench_${TAG}_ref.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $O/launches_${TAG}.csv $CMD > $O/ncu_launches_${TAG}.log 2>&1
bash scripts/profile_fwd.sh ${TAG}
# gate-sparse parameter-gradient kernel (C3 shape) and gate-sparse input-gradient kernel (C4 shape)
ncu --set full --clock-control none --import-source on -k regex:irbfn_bwd_p_kernel -s 170 -c 1 -o $O/prof_bwdps_${TAG} -f \
    python scripts/dev_bwd_sparse.py > $O/ncu_bwdps_${TAG}.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:irbfn_bwd_x_kernel -s 50 -c 1 -o $O/prof_bwdxs_${TAG} -f \
    python scripts/dev_bwdx_sparse.py > $O/ncu_bwdxs_${TAG}.log 2>&1
tail -n 3 $O/pytest_gpu_${TAG}.log $O/smoke_${TAG}.log
cut -c1-400 $O/bench_${TAG}_n1.json; cut -c1-300 $O/bench_${TAG}_reference.json
tail -n 2 $O/ncu_bwdps_${TAG}.log $O/ncu_bwdxs_${TAG}.log
