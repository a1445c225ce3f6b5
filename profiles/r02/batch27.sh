#!/bin/bash
# r02 batch 27: why voices are deferred (ZL_B200_WALK_PROF) + ncu of plan_kernel and walk_kernel.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; : > $O/b27_summary.txt
S="--steps 4 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity"
ZL_B200_WALK_PROF=1 timeout 300 python bench.py $S > $O/b27_prof.json 2> $O/b27_prof.err; grep "walk prof" $O/b27_prof.err >> $O/b27_summary.txt
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"walk_kernel" -s 40 -c 1 -o $O/b27_walk_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity > $O/b27_ncu_walk.log 2>&1; echo "ncu walk rc=$?" >> $O/b27_summary.txt
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"plan_kernel" -s 40 -c 1 -o $O/b27_plan_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity > $O/b27_ncu_plan.log 2>&1; echo "ncu plan rc=$?" >> $O/b27_summary.txt
cat $O/b27_summary.txt
