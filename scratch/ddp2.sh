#!/bin/bash
b() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus 2 --steps 300 --warmup 20 --comm $1 2>&1 | tail -1; }
echo "comm tests: $(timeout 500 python -m pytest tests/test_gpu_comm.py -m gpu -q 2>&1 | tail -1)"
b auto > gpurun_out/ours_2_auto_v2.json; cut -c90-330 gpurun_out/ours_2_auto_v2.json; grep -o '"comm": "[^"]*"' gpurun_out/ours_2_auto_v2.json
echo "twoshot: $(b twoshot | grep -o 'ms_per_step": [0-9.]*' | tr '\n' ' ')"
echo "auto again: $(b auto | grep -o 'ms_per_step": [0-9.]*' | tr '\n' ' ')"
