#!/bin/bash
O=gpurun_out/${1:-bench1}; mkdir -p $O
timeout 600 python bench.py --impl reference --steps 100 --warmup 10 > $O/ref_1.json 2> $O/ref_1.err; echo "ref rc=$?" >> $O/status.txt
timeout 600 python bench.py --steps 300 --warmup 20 > $O/ours_1.json 2> $O/ours_1.err; echo "ours rc=$?" >> $O/status.txt
B200_TC=15 timeout 600 python bench.py --steps 300 --warmup 20 > $O/ours_1_tf32.json 2> $O/ours_1_tf32.err; echo "ours tf32 rc=$?" >> $O/status.txt
timeout 600 python bench.py --steps 30 --warmup 5 --batch 4096 --epochs 2 > $O/ours_1_b4096.json 2> $O/ours_1_b4096.err; echo "ours b4096 rc=$?" >> $O/status.txt
cat $O/status.txt; tail -3 $O/ours_1.err
python - <<PY
import json
for n in ('ref_1','ours_1','ours_1_tf32','ours_1_b4096'):
    try:
        j=json.loads([l for l in open('$O/%s.json'%n).read().splitlines() if l.startswith('{')][-1])
        print(n, round(j['ms_per_step']*1000,1),'us; e2e', round(j['e2e']['ms_per_step']*1000,1), '; job', j.get('epoch20_job',{}).get('device_s'), j.get('epoch20_job',{}).get('wall_s'), j.get('epoch20_job',{}).get('last_test_line'), j.get('checks'))
    except Exception as e: print(n,'ERR',e)
PY
