#!/bin/bash
# single-GPU evidence for profiles/: bench of both arms, per-launch durations, ncu --set full of the step's kernels and
# of the communication kernels (data-parallel path with one rank, and four emulated ranks in one cooperative launch)
O=gpurun_out/${1:-prof1}; mkdir -p $O
timeout 600 python bench.py --impl reference --steps 100 --warmup 10 > $O/ref_1.json 2> $O/ref_1.err; echo "ref rc=$?" >> $O/status.txt
timeout 600 python bench.py --steps 300 --warmup 20 > $O/ours_1.json 2> $O/ours_1.err; echo "ours rc=$?" >> $O/status.txt
B200_TC=15 timeout 600 python bench.py --steps 300 --warmup 20 > $O/ours_1_tf32fc.json 2> $O/ours_1_tf32fc.err; echo "ours tf32 rc=$?" >> $O/status.txt
timeout 600 python bench.py --steps 30 --warmup 5 --batch 4096 > $O/ours_1_b4096.json 2> $O/ours_1_b4096.err; echo "ours b4096 rc=$?" >> $O/status.txt
timeout 600 python bench.py --impl reference --steps 30 --warmup 5 --batch 4096 > $O/ref_1_b4096.json 2> $O/ref_1_b4096.err; echo "ref b4096 rc=$?" >> $O/status.txt
CMD="python bench.py --steps 4 --warmup 3 --epochs 0"
timeout 300 $CMD > $O/plain_graph.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -s 60 -c 120 --csv --log-file $O/step_launches_warm.csv $CMD > $O/ncu_launches.log 2>&1
echo "launches rc=$?" >> $O/status.txt
CMD2="python bench.py --steps 4 --warmup 3 --epochs 0 --no-graph"
timeout 300 $CMD2 > $O/plain_eager.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -s 39 -c 13 -f -o $O/step_full $CMD2 > $O/ncu_step.log 2>&1
echo "ncu step rc=$?" >> $O/status.txt
CMD3="python bench.py --steps 4 --warmup 3 --epochs 0 --no-graph --comm twoshot"
B200_FORCE_DDP=1 timeout 300 $CMD3 > $O/plain_ddp.log 2>&1 &&
B200_FORCE_DDP=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"fused_allreduce|conv2_reduce|conv_finalize" -s 9 -c 3 -f -o $O/comm_full $CMD3 > $O/ncu_comm.log 2>&1
echo "ncu comm rc=$?" >> $O/status.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1 &&
timeout 900 ncu --set full --clock-control none -k regex:"fused_allreduce" -c 3 -f -o $O/comm_emulated python -c "import __graft_entry__ as g; g.smoke()" > $O/ncu_emul.log 2>&1
echo "ncu emulated rc=$?" >> $O/status.txt
cat $O/status.txt; ls -la $O | head -40
