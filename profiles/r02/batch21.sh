#!/bin/bash
# r02 batch 21: stand-in with walk-like memory traffic; the workload without quirk Q3 (no release-window walkers).
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
B="python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity"
: > $O/b21_summary.txt
run() { tag=$1; shift; env "$@" timeout 300 $B $EXTRA > $O/b21_$tag.json 2> $O/b21_$tag.err
  python - "$tag" $O/b21_$tag.json <<'PY' >> $O/b21_summary.txt
import json,sys
try:
    j=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=j["roofline"]
    print(f'{sys.argv[1]:34s} ms/step {j["ms_per_step"]:.3f} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} vf/step {j["voice_frames_per_step"]:.4g}')
except Exception as ex:
    print(sys.argv[1], "FAILED", ex)
PY
}
EXTRA=""
run default X=0
run no_walk ZL_B200_DEBUG_SKIP=6
run walk_is_small_spin ZL_B200_DEBUG_SKIP=14
run walk_is_mem_spin ZL_B200_DEBUG_SKIP=38
EXTRA="--quirks 3"
run quirks3_no_Q3 X=0
EXTRA="--quirks 0"
run quirks0 X=0
cat $O/b21_summary.txt
