#!/bin/bash
# r02 batch 17: thin-and-slow planner beside the gather - grid caps per SM for plan_kernel x walk_kernel.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
B="python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity"
: > $O/b17_summary.txt
for p in 0 3 4 6 8; do for w in 0 2 4 8; do
  ZL_B200_PLAN_CTAS=$p ZL_B200_WALK_CTAS=$w timeout 300 $B > $O/b17_p${p}_w${w}.json 2> $O/b17_p${p}_w${w}.err
  python - "plan $p/SM walk $w/SM" $O/b17_p${p}_w${w}.json <<'PY' >> $O/b17_summary.txt
import json,sys
try:
    j=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=j["roofline"]
    print(f'{sys.argv[1]:24s} ms/step {j["ms_per_step"]:.3f} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} frac {r["frac"]:.3f} e2e {j["e2e"]["value"]:.4g}')
except Exception as ex:
    print(sys.argv[1], "FAILED", ex)
PY
done; done
cat $O/b17_summary.txt
