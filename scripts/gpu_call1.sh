#!/bin/bash
# round 2, call 1: baseline of the round-1 engine on this box + first execution of the v4 drafts
O=gpurun_out/r2c1; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
timeout 300 python bench.py --steps 100 --warmup 10 > $O/bench_base.json 2> $O/bench_base.err
echo "bench rc=$?" >> $O/status.txt
B200_DRAFT=1 timeout 240 python -m pytest tests/test_gpu_kernels.py -q -k "dgrad_draft_v4" > $O/draft_dgrad.txt 2>&1
echo "draft dgrad rc=$?" >> $O/status.txt
B200_DRAFT=1 timeout 240 python -m pytest tests/test_gpu_kernels.py -q -k "conv12_fwd_draft" > $O/draft_fwd.txt 2>&1
echo "draft fwd rc=$?" >> $O/status.txt
B200_TRACE=6 timeout 120 python bench.py --steps 5 --warmup 3 --no-graph > $O/trace_eager.json 2> $O/trace_eager.err
echo "trace rc=$?" >> $O/status.txt
timeout 900 python -m pytest tests -m gpu -q > $O/pytest_gpu.txt 2>&1
echo "pytest rc=$?" >> $O/status.txt
tail -5 $O/pytest_gpu.txt; cat $O/status.txt; tail -3 $O/draft_dgrad.txt; tail -3 $O/draft_fwd.txt
