#!/bin/bash
# per-kernel durations of the step (serialised by ncu, warm caches): SHARES, not absolutes
O=gpurun_out/${1:-ncu}; mkdir -p $O
for tc in 31 15; do
  B200_TC=$tc timeout 300 python bench.py --steps 6 --warmup 3 > $O/plain_$tc.log 2>&1 &&
  B200_TC=$tc timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -s 60 -c 120 --csv \
      --log-file $O/launches_$tc.csv python bench.py --steps 6 --warmup 3 > $O/ncu_$tc.log 2>&1
  echo "tc=$tc rc=$?" >> $O/status.txt
done
cat $O/status.txt
