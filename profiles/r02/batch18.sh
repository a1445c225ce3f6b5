#!/bin/bash
# r02 batch 18: what does a co-resident kernel cost the gather?  walk_kernel replaced (after the warm-up) by stand-ins of the
# same shape: nothing / a 4-instruction spin loop / a 38 KB straight-line body.  Results of these runs are wrong by design.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
B="python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity"
: > $O/b18_summary.txt
run() { tag=$1; shift; env "$@" ZL_BENCH_NO_CHECKS=1 timeout 300 $B > $O/b18_$tag.json 2> $O/b18_$tag.err
  python - "$tag" $O/b18_$tag.json <<'PY' >> $O/b18_summary.txt
import json,sys
try:
    j=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=j["roofline"]
    print(f'{sys.argv[1]:34s} ms/step {j["ms_per_step"]:.3f} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f}')
except Exception as ex:
    print(sys.argv[1], "FAILED", ex)
PY
}
run default ZL_B200_DEBUG_SKIP=0
run no_walk ZL_B200_DEBUG_SKIP=6
run walk_is_small_spin ZL_B200_DEBUG_SKIP=14
run walk_is_big_body ZL_B200_DEBUG_SKIP=22
run no_plan_no_walk ZL_B200_DEBUG_SKIP=7
run plan_is_small_spin ZL_B200_DEBUG_SKIP=15
run plan_is_big_body ZL_B200_DEBUG_SKIP=23
cat $O/b18_summary.txt; tail -2 $O/b18_walk_is_big_body.err
