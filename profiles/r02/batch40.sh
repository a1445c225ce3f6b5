#!/bin/bash
# r02 batch 40: run continuation - how many voices are carried (ZL_B200_WALK_PROF) and the CTA residency trace.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; : > $O/b40_summary.txt
S="--steps 4 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity"
ZL_B200_WALK_PROF=1 timeout 300 python bench.py $S > $O/b40_prof.json 2> $O/b40_prof.err; grep "prof" $O/b40_prof.err >> $O/b40_summary.txt
ZL_B200_LIB=$PWD/build/v_trace.so ZL_B200_CTA_TRACE=$O/b40_trace_live.txt timeout 300 python bench.py $S --no-kernel-events > $O/b40_live.json 2> $O/b40_live.err
python profiles/r02/cta_trace.py $O/b40_trace_live.txt >> $O/b40_summary.txt 2>&1
cat $O/b40_summary.txt
