#!/bin/bash
O=gpurun_out/${1:-r2c13}; mkdir -p $O
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_gpu_comm.py -m gpu -q -x -p no:cacheprovider > $O/pytest.txt 2>&1; echo "pytest rc=$?" >> $O/status.txt
for comm in auto twoshot oneshot; do
  B200_CONV2_ORDER=0 timeout 300 $RUN --master-port 29611 bench.py --gpus 2 --steps 300 --warmup 10 --comm $comm > $O/b2_${comm}_ord0.json 2> $O/b2_${comm}_ord0.err
done
timeout 300 $RUN --master-port 29611 bench.py --gpus 2 --steps 300 --warmup 10 --comm twoshot > $O/b2_twoshot_ord1.json 2> $O/b2_twoshot_ord1.err
tail -2 $O/pytest.txt; python scripts/show_multi.py $O
