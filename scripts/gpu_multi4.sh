#!/bin/bash
O=gpurun_out/${1:-m4}; N=4; mkdir -p $O
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $RUN --master-port 29611 bench.py --gpus $N --steps 300 --warmup 20 > $O/ours_${N}_auto.json 2> $O/ours_${N}_auto.err; echo "ours rc=$?" >> $O/status.txt
timeout 600 $RUN --master-port 29613 bench.py --gpus $N --steps 100 --warmup 10 --impl reference > $O/ref_$N.json 2> $O/ref_$N.err; echo "ref rc=$?" >> $O/status.txt
timeout 400 python -m pytest tests/test_gpu_comm.py -m gpu -q -rA -k "multi_gpu and 4 and (auto or twoshot)" -p no:cacheprovider > $O/pytest_comm.txt 2>&1; echo "pytest rc=$?" >> $O/status.txt
grep -E "passed|failed" $O/pytest_comm.txt | tail -1; python scripts/show_multi.py $O
