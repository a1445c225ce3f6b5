#!/bin/bash
# r02 batch 38: ncu of plan_kernel after the tile-loop change; bench line with the NVML clock sampler.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; : > $O/b38_summary.txt
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"plan_kernel" -s 40 -c 1 -o $O/b38_plan_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity > $O/b38_ncu_plan.log 2>&1; echo "ncu plan rc=$?" >> $O/b38_summary.txt
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity > $O/b38_bench.json 2> $O/b38_bench.err; echo "bench rc=$?" >> $O/b38_summary.txt
python - <<'PY' >> $O/b38_summary.txt
import json
j=json.loads(open("gpurun_out/b38_bench.json").read().strip().splitlines()[-1]); print("clocks", j["clocks"], "ms/step", j["ms_per_step"])
PY
cat $O/b38_summary.txt
