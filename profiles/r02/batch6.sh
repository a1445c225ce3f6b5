#!/bin/bash
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
timeout 1800 python -m pytest tests -m gpu -q -x > $O/b6_pytest.log 2>&1; echo "pytest rc=$?" > $O/b6_summary.txt
tail -3 $O/b6_pytest.log >> $O/b6_summary.txt
timeout 300 $B > $O/b6_default.json 2> $O/b6_default.err; sum default $O/b6_default.json >> $O/b6_summary.txt
timeout 300 $B > $O/b6_default2.json 2> $O/b6_default2.err; sum "default (again)" $O/b6_default2.json >> $O/b6_summary.txt
ZL_B200_NO_OVERLAP=1 timeout 300 $B > $O/b6_noovl.json 2> $O/b6_noovl.err; sum "one stream" $O/b6_noovl.json >> $O/b6_summary.txt
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:"plan_kernel|walk_kernel|gather_kernel|master_mix" -s 64 -c 64 --csv --log-file $O/b6_launches.csv python bench.py --steps 1 --warmup 2 --no-cpu-baseline --no-latency --no-thumb > $O/b6_ncu.log 2>&1
cat $O/b6_summary.txt
