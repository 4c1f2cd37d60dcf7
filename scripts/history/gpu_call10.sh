#!/bin/bash
O=gpurun_out/${1:-r2c10}; mkdir -p $O
B200_TC=15 timeout 300 python bench.py --steps 200 --warmup 10 > $O/prio_15.json 2> $O/prio_15.err
B200_TC=15 B200_FLAT_PRIORITY=1 timeout 300 python bench.py --steps 200 --warmup 10 > $O/flat_15.json 2> $O/flat_15.err
B200_TC=15 B200_NO_BRANCH=1 timeout 300 python bench.py --steps 200 --warmup 10 > $O/serial_15.json 2> $O/serial_15.err
B200_TC=31 timeout 300 python bench.py --steps 200 --warmup 10 > $O/prio_31.json 2> $O/prio_31.err
B200_TC=15 B200_TRACE=30 timeout 300 python bench.py --steps 40 --warmup 10 > $O/trace1.json 2> $O/trace1.err
bash scripts/gpu_ncu_launches.sh ${1:-r2c10} > /dev/null 2>&1
python scripts/show_multi.py $O; grep "b200 trace" $O/trace1.err; python scripts/parse_launches.py $O/launches_15.csv
