#!/bin/bash
O=gpurun_out/${1:-r2c12}; mkdir -p $O
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_gpu_engine.py tests/test_gpu_comm.py -m gpu -q -x -p no:cacheprovider > $O/pytest.txt 2>&1; echo "pytest rc=$?" >> $O/status.txt
for ord in 1 0; do
B200_CONV2_ORDER=$ord timeout 300 python bench.py --steps 300 --warmup 10 > $O/b1_ord$ord.json 2> $O/b1_ord$ord.err
B200_CONV2_ORDER=$ord B200_TC=15 timeout 300 python bench.py --steps 300 --warmup 10 > $O/b1_tf32_ord$ord.json 2> $O/b1_tf32_ord$ord.err
for comm in auto twoshot; do
  B200_CONV2_ORDER=$ord timeout 300 $RUN --master-port 29611 bench.py --gpus 2 --steps 300 --warmup 10 --comm $comm > $O/b2_${comm}_ord$ord.json 2> $O/b2_${comm}_ord$ord.err
done
done
tail -2 $O/pytest.txt; python scripts/show_multi.py $O
