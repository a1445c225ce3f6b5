#!/bin/bash
# r02 batch 43: plan_kernel CTA size (ZL_B200_PLAN_THREADS) on the final build.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; : > $O/b43_summary.txt
S="--steps 12 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity"
for v in 32 64 128; do
  ZL_B200_PLAN_THREADS=$v timeout 600 python bench.py $S > $O/b43_$v.json 2> $O/b43_$v.err
  python - $v <<'PY' >> $O/b43_summary.txt
import json,sys
j=json.loads(open(f"gpurun_out/b43_{sys.argv[1]}.json").read().strip().splitlines()[-1]); r=j["roofline"]
print(f'plan threads {sys.argv[1]}: ms/step {j["ms_per_step"]:.3f} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f}')
PY
done
cat $O/b43_summary.txt
