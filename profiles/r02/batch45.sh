#!/bin/bash
# r02 batch 45: home-made memcheck (compute-sanitizer is closed on this pool): the -DZL_BOUNDS_CHECK=1 build under the sanitizer
# case list, the whole GPU suite and the bench workload; every address formed from planner data is tested against its object.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; : > $O/b45_summary.txt
export ZL_B200_LIB=$PWD/build/v_bounds.so
rm -f $O/b45_cases.txt $O/b45_suite.txt $O/b45_bench.txt
ZL_B200_CTA_TRACE=$O/b45_cases.txt timeout 600 python profiles/r02/sanitize_cases.py > $O/b45_cases.out 2>&1; echo "cases rc=$?" >> $O/b45_summary.txt
tail -1 $O/b45_cases.txt >> $O/b45_summary.txt
ZL_B200_CTA_TRACE=$O/b45_suite.txt timeout 2400 python -m pytest tests -m gpu -q > $O/b45_pytest.log 2>&1; echo "suite rc=$?" >> $O/b45_summary.txt
tail -2 $O/b45_pytest.log >> $O/b45_summary.txt
python - <<'PY' >> $O/b45_summary.txt
import re
tot=0; viol=0; n=0
for l in open("gpurun_out/b45_suite.txt"):
    m=re.match(r"bounds checks (\d+); violations: (.*)", l)
    if m:
        n+=1; tot=max(tot,int(m.group(1))); viol=max(viol,sum(int(x) for x in re.findall(r" (\d+)", m.group(2))))
print(f"suite: {n} contexts destroyed, largest cumulative check count {tot}, largest violation total {viol}")
PY
ZL_B200_CTA_TRACE=$O/b45_bench.txt timeout 900 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-latency --no-thumb > $O/b45_bench.json 2> $O/b45_bench.err; echo "bench rc=$?" >> $O/b45_summary.txt
tail -1 $O/b45_bench.txt >> $O/b45_summary.txt
cat $O/b45_summary.txt
