#!/bin/bash
# multi-GPU validation + measurements on N GPUs of one box:  scripts/gpu_multi.sh <outdir> <N>
O=gpurun_out/${1:-multi}; N=${2:-2}; mkdir -p $O
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi topo -m > $O/topo.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_comm.py -m gpu -q -rA -k "multi_gpu" -p no:cacheprovider > $O/pytest_comm.txt 2>&1; echo "pytest rc=$?" >> $O/status.txt
for comm in ${COMMS:-auto twoshot oneshot}; do
  timeout 300 $RUN --master-port 29611 bench.py --gpus $N --steps 200 --warmup 10 --comm $comm > $O/ours_${N}_$comm.json 2> $O/ours_${N}_$comm.err; echo "ours $comm rc=$?" >> $O/status.txt
done
B200_TRACE=8 timeout 300 $RUN --master-port 29612 bench.py --gpus $N --steps 6 --warmup 4 --no-graph > $O/trace_$N.json 2> $O/trace_$N.err
timeout 600 $RUN --master-port 29613 bench.py --gpus $N --steps 100 --warmup 10 --impl reference > $O/ref_$N.json 2> $O/ref_$N.err; echo "ref rc=$?" >> $O/status.txt
cat $O/status.txt; grep -E "passed|failed" $O/pytest_comm.txt | tail -2; grep -E "^FAILED|DDP-vs-single" $O/pytest_comm.txt | head -20
python - <<PY
import json,glob
for f in sorted(glob.glob('$O/*_${N}*.json')):
    try:
        d=json.load(open(f)); print(f.split('/')[-1], round(d['ms_per_step']*1000,1),'us flushed', round(d['steady_state']['ms_per_step']*1000,1),'steady', d['config'].get('comm'), d['clocks']['reasons'])
    except Exception as e: print(f,'ERR',e)
PY
grep "b200 trace" $O/trace_$N.err | head -20
