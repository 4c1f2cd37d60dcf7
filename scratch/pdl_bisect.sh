#!/bin/bash
# which kernels break bit-identity under PDL?
run() { B200_PDL=1 B200_PDL_MASK=$1 timeout 100 python -m pytest tests/test_gpu_engine.py -m gpu -q -k "graph_replay_equals_eager" 2>&1 | tail -1; }
echo "all: $(B200_PDL=1 timeout 100 python -m pytest tests/test_gpu_engine.py -m gpu -q -k graph_replay_equals_eager 2>&1 | tail -1)"
for i in 0 1 2 3 4 5 6 7 8 9 10 11 12; do echo "bit $i: $(run $((1<<i)))"; done
