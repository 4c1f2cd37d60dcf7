#!/bin/bash
# ncu --set full of selected kernels of the step (after the same command ran clean without ncu)
O=gpurun_out/${1:-ncufull}; K=${2:-conv12}; mkdir -p $O
timeout 300 python bench.py --steps 4 --warmup 3 --no-graph > $O/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$K" -s ${3:-6} -c ${4:-2} -o $O/prof -f \
    python bench.py --steps 4 --warmup 3 --no-graph > $O/ncu.log 2>&1
echo "rc=$?"; tail -3 $O/ncu.log
