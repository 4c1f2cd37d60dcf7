#!/bin/bash
# 8-GPU validation + measurements:  scripts/gpu_multi8.sh <outdir>
O=gpurun_out/${1:-m8}; N=8; mkdir -p $O
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi topo -m > $O/topo.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_comm.py -m gpu -q -rA -k "multi_gpu and 8 and not oneshot" -p no:cacheprovider > $O/pytest_comm.txt 2>&1; echo "pytest rc=$?" >> $O/status.txt
for comm in auto twoshot; do
  timeout 300 $RUN --master-port 29611 bench.py --gpus $N --steps 300 --warmup 20 --comm $comm > $O/ours_${N}_$comm.json 2> $O/ours_${N}_$comm.err; echo "ours $comm rc=$?" >> $O/status.txt
done
timeout 600 $RUN --master-port 29613 bench.py --gpus $N --steps 100 --warmup 10 --impl reference > $O/ref_$N.json 2> $O/ref_$N.err; echo "ref rc=$?" >> $O/status.txt
timeout 300 $RUN --master-port 29614 bench.py --gpus $N --steps 30 --warmup 5 --batch 4096 --epochs 20 > $O/ours_${N}_b4096.json 2> $O/ours_${N}_b4096.err; echo "ours b4096 rc=$?" >> $O/status.txt
timeout 600 $RUN --master-port 29615 bench.py --gpus $N --steps 30 --warmup 5 --batch 4096 --epochs 20 --impl reference > $O/ref_${N}_b4096.json 2> $O/ref_${N}_b4096.err; echo "ref b4096 rc=$?" >> $O/status.txt
B200_TRACE=30 timeout 300 $RUN --master-port 29612 bench.py --gpus $N --steps 40 --warmup 10 --epochs 0 > $O/trace_$N.json 2> $O/trace_$N.err
timeout 300 $RUN --master-port 29616 benchmarks/allreduce_bench.py --max-mb 64 --iters 10 --out $O/allreduce_w8.json > $O/allreduce.log 2>&1; echo "allreduce rc=$?" >> $O/status.txt
cat $O/status.txt; grep -E "passed|failed" $O/pytest_comm.txt | tail -2; grep -E "^FAILED|DDP-vs-single" $O/pytest_comm.txt | cut -c1-150 | head
python scripts/show_multi.py $O
