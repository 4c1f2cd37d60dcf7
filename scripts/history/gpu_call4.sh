#!/bin/bash
O=gpurun_out/${1:-r2c4}; mkdir -p $O
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.txt 2>&1; echo "smoke rc=$?" >> $O/status.txt
timeout 1500 python -m pytest tests -m gpu -q -rA -p no:cacheprovider > $O/pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> $O/status.txt
for tc in 31 15; do
B200_TC=$tc timeout 300 python bench.py --steps 200 --warmup 10 > $O/bench_$tc.json 2> $O/bench_$tc.err; echo "bench $tc rc=$?" >> $O/status.txt
done
bash scripts/gpu_ncu_launches.sh ${1:-r2c4} > /dev/null 2>&1
cat $O/status.txt; tail -2 $O/smoke.txt; grep -E "passed|failed" $O/pytest_gpu.txt | tail -3; grep -E "^FAILED|^ERROR" $O/pytest_gpu.txt | head -40
python - <<PY
import json
for tc in (31,15):
    try:
        d=json.load(open('$O/bench_%d.json'%tc)); print(tc, round(d['ms_per_step']*1000,1),'us flushed', round(d['steady_state']['ms_per_step']*1000,1),'steady', d['launches_per_step'])
    except Exception as e: print(tc,'ERR',e)
PY
