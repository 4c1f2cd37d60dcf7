#!/bin/bash
O=gpurun_out/${1:-regsab}; mkdir -p $O; touch $O/status.txt
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
for regs in 104 128; do
  B200_DGRAD_REGS=$regs timeout 300 python bench.py --steps 300 --warmup 20 --epochs 0 > $O/b1_regs$regs.json 2> $O/b1_regs$regs.err
  B200_DGRAD_REGS=$regs timeout 300 $RUN --master-port 29611 bench.py --gpus 2 --steps 300 --warmup 20 --epochs 0 > $O/b2_regs$regs.json 2> $O/b2_regs$regs.err
done
python scripts/show_multi.py $O
