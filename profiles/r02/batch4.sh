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
: > $O/b4_summary.txt
timeout 300 $B > $O/b4_default.json 2> $O/b4_default.err; sum default $O/b4_default.json >> $O/b4_summary.txt
ZL_B200_DEBUG_SKIP=4 timeout 300 $B > $O/b4_skip4.json 2> $O/b4_skip4.err; sum "walk launched, queues empty" $O/b4_skip4.json >> $O/b4_summary.txt
ZL_B200_DEBUG_SKIP=2 timeout 300 $B > $O/b4_skip2.json 2> $O/b4_skip2.err; sum "walk not launched" $O/b4_skip2.json >> $O/b4_summary.txt
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"walk_kernel" -s 40 -c 3 -o $O/b4_walk_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-latency --no-thumb > $O/b4_ncu_walk.log 2>&1
cat $O/b4_summary.txt
