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
timeout 1800 python -m pytest tests -m gpu -q -x > $O/b5_pytest.log 2>&1; echo "pytest rc=$?" > $O/b5_summary.txt
tail -3 $O/b5_pytest.log >> $O/b5_summary.txt
timeout 300 $B > $O/b5_default.json 2> $O/b5_default.err; sum default $O/b5_default.json >> $O/b5_summary.txt
for v in walk_m16 walk_m24 walk_m32; do
  ZL_B200_LIB=$PWD/build/variants/$v.so timeout 300 $B > $O/b5_$v.json 2> $O/b5_$v.err; sum $v $O/b5_$v.json >> $O/b5_summary.txt
done
rm -f $O/b5_timeline.txt; ZL_B200_TIMELINE=$O/b5_timeline.txt timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-kernel-events > $O/b5_tl.json 2> $O/b5_tl.err
cat $O/b5_summary.txt; grep "walk queue" $O/b5_timeline.txt | head -16
