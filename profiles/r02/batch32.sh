#!/bin/bash
# r02 batch 32: per-CTA residency trace (diagnostic build -DZL_CTA_TRACE=1) of one gather launch, live and alone.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; : > $O/b32_summary.txt
S="--steps 4 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity --no-kernel-events"
ZL_B200_LIB=$PWD/build/v_trace.so ZL_B200_CTA_TRACE=$O/b32_trace_live.txt timeout 300 python bench.py $S > $O/b32_live.json 2> $O/b32_live.err; echo "live rc=$?" >> $O/b32_summary.txt
ZL_B200_LIB=$PWD/build/v_trace.so ZL_B200_NO_OVERLAP=1 ZL_B200_CTA_TRACE=$O/b32_trace_alone.txt timeout 300 python bench.py $S > $O/b32_alone.json 2> $O/b32_alone.err; echo "alone rc=$?" >> $O/b32_summary.txt
for v in live alone; do echo "== $v" >> $O/b32_summary.txt; python profiles/r02/cta_trace.py $O/b32_trace_$v.txt >> $O/b32_summary.txt 2>&1; done
cat $O/b32_summary.txt
