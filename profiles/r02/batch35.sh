#!/bin/bash
# r02 batch 35: CTA residency trace with capped planner grids.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; : > $O/b35_summary.txt
T="--steps 6 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity"
for c in "2 2" "4 2" "1 1"; do
  set -- $c
  echo "== plan $1/SM walk $2/SM" >> $O/b35_summary.txt
  ZL_B200_PLAN_CTAS=$1 ZL_B200_WALK_CTAS=$2 ZL_B200_LIB=$PWD/build/v_trace.so ZL_B200_CTA_TRACE=$O/b35_trace_$1_$2.txt timeout 300 python bench.py $T > $O/b35_$1_$2.json 2> $O/b35_$1_$2.err
  python - "$O/b35_$1_$2.json" <<'PY' >> $O/b35_summary.txt
import json,sys
j=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=j["roofline"]
print(f'ms/step {j["ms_per_step"]:.3f} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f}')
PY
  python profiles/r02/cta_trace.py $O/b35_trace_$1_$2.txt >> $O/b35_summary.txt 2>&1
done
cat $O/b35_summary.txt
