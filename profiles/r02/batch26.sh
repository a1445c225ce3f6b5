#!/bin/bash
# r02 batch 26: L2 prefetch of a tile's later windows at tile start (ZL_L2PF=1 bulk, =2 per line) vs the in-tree build.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; : > $O/b26_summary.txt
S="--steps 12 --warmup 3 --no-cpu-baseline --no-latency --no-thumb"
for v in base pf1 pf2; do
  L=$PWD/build/v_$v.so; [ $v = base ] && L=$PWD/libzl_b200/libzl_b200.so
  ZL_B200_LIB=$L timeout 600 python bench.py $S > $O/b26_$v.json 2> $O/b26_$v.err; echo "$v rc=$?" >> $O/b26_summary.txt
done
python - <<'PY' >> $O/b26_summary.txt
import json
for n in ("base", "pf1", "pf2"):
    try:
        j=json.loads(open(f"gpurun_out/b26_{n}.json").read().strip().splitlines()[-1]); r=j["roofline"]; p=j.get("parity_at_size") or {}
        print(f'{n}: ms/step {j["ms_per_step"]:.3f} vf/s {j["value"]:.4g} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} frac {r["frac"]:.3f} whole {r["whole_step_frac"]:.3f} parity {p.get("masters_bit_equal")} {p.get("peaks_bit_equal")} {p.get("voice_table_bit_equal")}')
    except Exception as ex: print(n, "parse failed", ex)
PY
cat $O/b26_summary.txt
