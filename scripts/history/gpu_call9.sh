#!/bin/bash
O=gpurun_out/${1:-r2c9}; mkdir -p $O
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests -m gpu -q -x -p no:cacheprovider > $O/pytest.txt 2>&1; echo "pytest rc=$?" >> $O/status.txt
for tc in 31 15; do
B200_TC=$tc timeout 300 python bench.py --steps 200 --warmup 10 > $O/bench1_$tc.json 2> $O/bench1_$tc.err; echo "bench $tc rc=$?" >> $O/status.txt
done
B200_TRACE=30 timeout 300 python bench.py --steps 40 --warmup 10 > $O/trace1.json 2> $O/trace1.err
for comm in auto twoshot; do
  timeout 300 $RUN --master-port 29611 bench.py --gpus 2 --steps 200 --warmup 10 --comm $comm > $O/ours_2_$comm.json 2> $O/ours_2_$comm.err; echo "ours $comm rc=$?" >> $O/status.txt
done
B200_TRACE=30 timeout 300 $RUN --master-port 29612 bench.py --gpus 2 --steps 40 --warmup 10 --comm twoshot > $O/trace2.json 2> $O/trace2.err
tail -2 $O/pytest.txt; python scripts/show_multi.py $O; grep "b200 trace" $O/trace1.err
