#!/bin/bash
# r02 batch 25: ncu --set full of the new gather (window-relative chained positions, packed M/S, elect-issued TMA) + slot/occupancy variants.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"gather_kernel" -s 40 -c 1 -o $O/b25_gather_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity > $O/b25_ncu_gather.log 2>&1; echo "ncu full rc=$?" > $O/b25_summary.txt
S="--steps 12 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity"
for v in s3m22 m24 m21; do
  ZL_B200_LIB=$PWD/build/v_$v.so timeout 600 python bench.py $S > $O/b25_$v.json 2> $O/b25_$v.err; echo "$v rc=$?" >> $O/b25_summary.txt
done
python - <<'PY' >> $O/b25_summary.txt
import json
for n in ("s3m22", "m24", "m21"):
    try:
        j=json.loads(open(f"gpurun_out/b25_{n}.json").read().strip().splitlines()[-1]); r=j["roofline"]
        print(f'{n}: ms/step {j["ms_per_step"]:.3f} vf/s {j["value"]:.4g} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} frac {r["frac"]:.3f} whole {r["whole_step_frac"]:.3f}')
    except Exception as ex: print(n, "parse failed", ex)
PY
cat $O/b25_summary.txt
