#!/bin/bash
# Round-end verification on one B200 (run under gpurun): GPU tests, smoke, both bench arms, compute-sanitizer.
set -u
O=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo smoke=$? >> $O/smoke.log
python bench.py > $O/bench_n1.json 2> $O/bench_n1.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
timeout 420 compute-sanitizer --tool memcheck --error-exitcode 9 python scripts/dev_sanitize.py > $O/memcheck.log 2>&1; echo rc=$? >> $O/memcheck.log
timeout 420 compute-sanitizer --tool racecheck --error-exitcode 9 python scripts/dev_sanitize.py > $O/racecheck.log 2>&1; echo rc=$? >> $O/racecheck.log
tail -3 $O/pytest_gpu.log $O/smoke.log $O/memcheck.log $O/racecheck.log
cut -c1-500 $O/bench_n1.json; cut -c1-500 $O/bench_ref.json
