#!/bin/bash
# r02 batch 41: ncu of cont_kernel and planq_kernel.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; : > $O/b41_summary.txt
for k in cont_kernel planq_kernel; do
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"$k" -s 40 -c 1 -o $O/b41_$k -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity > $O/b41_ncu_$k.log 2>&1; echo "ncu $k rc=$?" >> $O/b41_summary.txt
done
cat $O/b41_summary.txt
