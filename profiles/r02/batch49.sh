#!/bin/bash
# r02 batch 49: persistent plan_kernel grids that fill the SMs (ZL_B200_PLAN_CTAS = 10 / 12 / 16 CTAs per SM, grid-stride over the bus groups).
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; : > $O/b49_summary.txt
S="--steps 12 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity"
for c in 0 10 12 16; do
  ZL_B200_PLAN_CTAS=$c timeout 600 python bench.py $S > $O/b49_$c.json 2> $O/b49_$c.err
  python - $c <<'PY' >> $O/b49_summary.txt
import json,sys
j=json.loads(open(f"gpurun_out/b49_{sys.argv[1]}.json").read().strip().splitlines()[-1]); r=j["roofline"]
print(f'plan CTAs/SM {sys.argv[1]}: ms/step {j["ms_per_step"]:.3f} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f}')
PY
done
cat $O/b49_summary.txt
