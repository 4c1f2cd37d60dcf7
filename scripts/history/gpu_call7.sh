#!/bin/bash
O=gpurun_out/${1:-r2c7}; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_engine.py -m gpu -q -x -p no:cacheprovider > $O/pytest.txt 2>&1; echo "pytest rc=$?" >> $O/status.txt
for tc in 31 15; do
B200_TC=$tc timeout 300 python bench.py --steps 200 --warmup 10 > $O/bench_$tc.json 2> $O/bench_$tc.err; echo "bench $tc rc=$?" >> $O/status.txt
done
B200_FWD_EW=8 B200_TC=15 timeout 300 python bench.py --steps 200 --warmup 10 > $O/bench_15_ew8.json 2> $O/bench_15_ew8.err
B200_EXTRA_NVCC_FLAGS=-DB200_STAMPS timeout 300 python scripts/kernel_timeline.py fwd 200 > $O/timeline_fwd.txt 2>&1
cat $O/status.txt; tail -3 $O/pytest.txt
python - <<PY
import json
for tc in ('31','15','15_ew8'):
    try:
        d=json.load(open('$O/bench_%s.json'%tc)); print(tc, round(d['ms_per_step']*1000,1),'us flushed', round(d['steady_state']['ms_per_step']*1000,1),'steady', d['launches_per_step'])
    except Exception as e: print(tc,'ERR',e)
PY
grep -E "^CTA   0|^CTA 104|mean" $O/timeline_fwd.txt
