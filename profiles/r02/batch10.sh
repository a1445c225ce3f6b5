#!/bin/bash
# r02 batch 10: planner confined to K SMs (instruction-cache hypothesis) - sweep K and the per-SM CTA counts; parity suite at the best setting.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
B="python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity"
sum() { python - "$1" "$2" <<'PY'
import json,sys
try:
    j=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=j["roofline"]
    print(f'{sys.argv[1]:34s} ms/step {j["ms_per_step"]:.3f} vf/s {j["value"]:.4g} e2e {j["e2e"]["value"]:.4g} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} frac {r["frac"]:.3f} whole {r["whole_step_frac"]:.3f}')
except Exception as ex:
    print(sys.argv[1], "FAILED", ex)
PY
}
: > $O/b10_summary.txt
run() { tag=$1; shift; env "$@" timeout 300 $B > $O/b10_$tag.json 2> $O/b10_$tag.err; sum "$tag" $O/b10_$tag.json >> $O/b10_summary.txt; }
run k0 ZL_B200_PLANNER_SMS=0
run k12 ZL_B200_PLANNER_SMS=12
run k16 ZL_B200_PLANNER_SMS=16
run k24 ZL_B200_PLANNER_SMS=24
run k32 ZL_B200_PLANNER_SMS=32
run k48 ZL_B200_PLANNER_SMS=48
run k74 ZL_B200_PLANNER_SMS=74
run k148 ZL_B200_PLANNER_SMS=148
run k24_p6w6 ZL_B200_PLANNER_SMS=24 ZL_B200_PLANNER_PLAN_CTAS=6 ZL_B200_PLANNER_WALK_CTAS=6
run k24_p4w13 ZL_B200_PLANNER_SMS=24 ZL_B200_PLANNER_PLAN_CTAS=4 ZL_B200_PLANNER_WALK_CTAS=13
run k24_p12w4 ZL_B200_PLANNER_SMS=24 ZL_B200_PLANNER_PLAN_CTAS=12 ZL_B200_PLANNER_WALK_CTAS=4
ZL_B200_PLANNER_SMS=24 timeout 1800 python -m pytest tests -m gpu -q -x > $O/b10_pytest_k24.log 2>&1; echo "pytest k24 rc=$?" >> $O/b10_summary.txt; tail -3 $O/b10_pytest_k24.log >> $O/b10_summary.txt
cat $O/b10_summary.txt
