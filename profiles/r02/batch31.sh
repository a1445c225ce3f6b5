#!/bin/bash
# r02 batch 31: window queue running across tile boundaries (ZL_PRESTAGE): GPU suite + bench; A/B against ZL_PRESTAGE=0.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -q > $O/b31_pytest.log 2>&1; echo "pytest rc=$?" > $O/b31_summary.txt
tail -6 $O/b31_pytest.log >> $O/b31_summary.txt
S="--steps 12 --warmup 3 --no-cpu-baseline --no-latency --no-thumb"
for v in pre nopre; do
  L=$PWD/build/v_$v.so; [ $v = pre ] && L=$PWD/libzl_b200/libzl_b200.so
  ZL_B200_LIB=$L timeout 600 python bench.py $S > $O/b31_$v.json 2> $O/b31_$v.err; echo "$v rc=$?" >> $O/b31_summary.txt
done
python - <<'PY' >> $O/b31_summary.txt
import json
for n in ("pre", "nopre"):
    try:
        j=json.loads(open(f"gpurun_out/b31_{n}.json").read().strip().splitlines()[-1]); r=j["roofline"]; p=j.get("parity_at_size") or {}
        print(f'{n}: ms/step {j["ms_per_step"]:.3f} vf/s {j["value"]:.4g} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} frac {r["frac"]:.3f} parity {p.get("masters_bit_equal")} {p.get("peaks_bit_equal")} {p.get("voice_table_bit_equal")}')
    except Exception as ex: print(n, "parse failed", ex)
PY
cat $O/b31_summary.txt
