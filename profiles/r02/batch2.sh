#!/bin/bash
# r02 batch 2: full GPU suite after the CREDUX fix, walk_kernel experiments, ncu --set full of walk / plan / gather.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
B="python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-latency --no-thumb"
sum() { python - "$1" "$2" <<'PY'
import json,sys
try:
    j=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=j["roofline"]
    print(f'{sys.argv[1]:28s} ms/step {j["ms_per_step"]:.3f} vf/s {j["value"]:.4g} e2e {j["e2e"]["value"]:.4g} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} frac {r["frac"]:.3f} whole {r["whole_step_frac"]:.3f}')
except Exception as ex:
    print(sys.argv[1], "FAILED", ex)
PY
}
timeout 1800 python -m pytest tests -m gpu -q > $O/b2_pytest.log 2>&1; echo "pytest rc=$?" > $O/b2_summary.txt
tail -5 $O/b2_pytest.log >> $O/b2_summary.txt
timeout 300 $B > $O/b2_default.json 2> $O/b2_default.err; sum default $O/b2_default.json >> $O/b2_summary.txt
for c in 1 2 4 8; do
  ZL_B200_WALK_CTAS=$c timeout 300 $B > $O/b2_walkcap$c.json 2> $O/b2_walkcap$c.err; sum "walk cap $c/SM" $O/b2_walkcap$c.json >> $O/b2_summary.txt
done
ZL_B200_PLAN_CTAS=2 timeout 300 $B > $O/b2_plancap2.json 2> $O/b2_plancap2.err; sum "plan cap 2/SM" $O/b2_plancap2.json >> $O/b2_summary.txt
rm -f $O/b2_timeline.txt; ZL_B200_TIMELINE=$O/b2_timeline.txt timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-kernel-events > $O/b2_tl.json 2> $O/b2_tl.err
# ncu --set full: one launch each of walk, plan, gather (skip the warm-up launches)
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"walk_kernel" -s 40 -c 2 -o $O/b2_walk_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-latency --no-thumb > $O/b2_ncu_walk.log 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"gather_kernel" -s 40 -c 1 -o $O/b2_gather_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-latency --no-thumb > $O/b2_ncu_gather.log 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"plan_kernel" -s 40 -c 1 -o $O/b2_plan_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-latency --no-thumb > $O/b2_ncu_plan.log 2>&1
cat $O/b2_summary.txt
