#!/bin/bash
# Round-end verification on one B200 (run under gpurun): GPU tests, smoke, both bench arms.
# (compute-sanitizer is refused on this pool's boxes, so scripts/dev_sanitize.py runs plain: every kernel on
# small ragged shapes, errors surface through the C ABI's return codes.)
set -u
O=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo smoke=$? >> $O/smoke.log
python bench.py > $O/bench_n1.json 2> $O/bench_n1.err
python scripts/dev_sanitize.py > $O/all_kernels.log 2>&1; echo rc=$? >> $O/all_kernels.log
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
tail -n 3 $O/pytest_gpu.log $O/smoke.log
cut -c1-500 $O/bench_n1.json; cut -c1-500 $O/bench_ref.json
