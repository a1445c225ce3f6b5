#!/bin/bash
# r02 batch 52: room reserved for the planner instead of taken by it: gather built for 24 / 22 CTAs per SM (registers for 2-3
# plan_kernel CTAs stay free), planner at low / same priority as persistent grids that fit that room.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; : > $O/b52_summary.txt
S="--steps 12 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity"
run() { tag=$1; shift; env "$@" timeout 600 python bench.py $S > $O/b52_$tag.json 2> $O/b52_$tag.err
  python - $tag <<'PY' >> $O/b52_summary.txt
import json,sys
try:
    j=json.loads(open(f"gpurun_out/b52_{sys.argv[1]}.json").read().strip().splitlines()[-1]); r=j["roofline"]
    print(f'{sys.argv[1]}: ms/step {j["ms_per_step"]:.3f} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f}')
except Exception as ex: print(sys.argv[1], "failed", ex)
PY
}
L24=ZL_B200_LIB=$PWD/build/v_m24.so; L22=ZL_B200_LIB=$PWD/build/v_m22.so
run base ZL_X=1
run m24_low_p2w2 $L24 ZL_B200_PLAN_PRIORITY=low ZL_B200_PLAN_CTAS=2 ZL_B200_WALK_CTAS=2
run m24_same_p2w2 $L24 ZL_B200_PLAN_PRIORITY=same ZL_B200_PLAN_CTAS=2 ZL_B200_WALK_CTAS=2
run m24_high_p2w2 $L24 ZL_B200_PLAN_CTAS=2 ZL_B200_WALK_CTAS=2
run m22_low_p3w2 $L22 ZL_B200_PLAN_PRIORITY=low ZL_B200_PLAN_CTAS=3 ZL_B200_WALK_CTAS=2
run m22_high_p3w2 $L22 ZL_B200_PLAN_CTAS=3 ZL_B200_WALK_CTAS=2
run m22_low_p3w4 $L22 ZL_B200_PLAN_PRIORITY=low ZL_B200_PLAN_CTAS=3 ZL_B200_WALK_CTAS=4
cat $O/b52_summary.txt
