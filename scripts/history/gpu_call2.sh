#!/bin/bash
# round 2, call 2: first run of the restructured step (conv12 fwd, dgrad/wgrad without dY, conv2 reduce on the
# branch, 3xTF32 fc1) + the new tests
O=gpurun_out/r2c2; mkdir -p $O
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.txt 2>&1; echo "smoke rc=$?" >> $O/status.txt
timeout 1500 python -m pytest tests -m gpu -q -rA -p no:cacheprovider > $O/pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> $O/status.txt
timeout 300 python bench.py --steps 200 --warmup 10 > $O/bench_fp32fc.json 2> $O/bench_fp32fc.err; echo "bench rc=$?" >> $O/status.txt
B200_TC=15 timeout 300 python bench.py --steps 200 --warmup 10 > $O/bench_tf32fc.json 2> $O/bench_tf32fc.err; echo "bench tf32 rc=$?" >> $O/status.txt
B200_TRACE=6 timeout 120 python bench.py --steps 5 --warmup 3 --no-graph > $O/trace_eager.json 2> $O/trace_eager.err
B200_FORCE_DDP=1 timeout 300 python bench.py --steps 200 --warmup 10 --comm twoshot > $O/bench_forceddp.json 2> $O/bench_forceddp.err; echo "bench forceddp rc=$?" >> $O/status.txt
cat $O/status.txt; tail -3 $O/smoke.txt; grep -E "passed|failed" $O/pytest_gpu.txt | tail -3; grep -E "^FAILED|^ERROR" $O/pytest_gpu.txt | head -40
