#!/bin/bash
O=gpurun_out/${1:-b4096}; mkdir -p $O
for tc in 31 15; do
B200_TC=$tc timeout 300 python bench.py --steps 30 --warmup 5 --batch 4096 --epochs 0 > $O/b4096_$tc.json 2> $O/b4096_$tc.err
done
timeout 600 python bench.py --impl reference --steps 30 --warmup 5 --batch 4096 --epochs 1 > $O/ref_b4096.json 2> $O/ref_b4096.err
timeout 600 python bench.py --impl reference --steps 100 --warmup 10 > $O/ref_1.json 2> $O/ref_1.err; echo "ref rc=$?" >> $O/status.txt
B200_TC=31 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -s 40 -c 60 --csv \
      --log-file $O/launches_b4096.csv python bench.py --steps 3 --warmup 3 --batch 4096 --epochs 0 > $O/ncu.log 2>&1
python scripts/show_multi.py $O; python scripts/parse_launches.py $O/launches_b4096.csv
python - <<PY
import json
for n in ('ref_1','ref_b4096'):
    try:
        j=json.loads([l for l in open('$O/%s.json'%n).read().splitlines() if l.startswith('{')][-1])
        print(n, round(j['ms_per_step']*1000,1),'us; e2e', round(j['e2e']['ms_per_step']*1000,1), '; job', j.get('epoch20_job'))
    except Exception as e: print(n,'ERR',e)
PY
