#!/bin/bash
O=gpurun_out/${1:-r2c15}; mkdir -p $O
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 1200 python -m pytest tests -m gpu -q -rA -p no:cacheprovider > $O/pytest.txt 2>&1; echo "pytest rc=$?" >> $O/status.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.txt 2>&1; echo "smoke rc=$?" >> $O/status.txt
timeout 300 python bench.py --steps 300 --warmup 20 > $O/b1.json 2> $O/b1.err
for comm in auto twoshot; do
  timeout 300 $RUN --master-port 29611 bench.py --gpus 2 --steps 300 --warmup 20 --comm $comm > $O/b2_${comm}.json 2> $O/b2_${comm}.err
done
grep -E "passed|failed" $O/pytest.txt | tail -1; grep -E "^FAILED|HUNG_PEER|DDP-vs-single" $O/pytest.txt | cut -c1-330; tail -2 $O/smoke.txt; python scripts/show_multi.py $O
