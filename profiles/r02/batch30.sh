#!/bin/bash
# r02 batch 30: every render kernel asks for the same (largest) shared-memory carve-out vs the driver's own per-kernel choice.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; : > $O/b30_summary.txt
S="--steps 12 --warmup 3 --no-cpu-baseline --no-latency --no-thumb"
for v in same driver c50; do
  E=""; [ $v = driver ] && E="ZL_B200_CARVEOUT=0"; [ $v = c50 ] && E="ZL_B200_CARVEOUT=50"
  env $E timeout 600 python bench.py $S > $O/b30_$v.json 2> $O/b30_$v.err; echo "$v rc=$?" >> $O/b30_summary.txt
done
python - <<'PY' >> $O/b30_summary.txt
import json
for n in ("same", "driver", "c50"):
    try:
        j=json.loads(open(f"gpurun_out/b30_{n}.json").read().strip().splitlines()[-1]); r=j["roofline"]; p=j.get("parity_at_size") or {}
        print(f'{n}: ms/step {j["ms_per_step"]:.3f} vf/s {j["value"]:.4g} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} frac {r["frac"]:.3f} parity {p.get("masters_bit_equal")} {p.get("peaks_bit_equal")} {p.get("voice_table_bit_equal")}')
    except Exception as ex: print(n, "parse failed", ex)
PY
cat $O/b30_summary.txt
