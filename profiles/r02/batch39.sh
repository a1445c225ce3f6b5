#!/bin/bash
# r02 batch 39: run continuation (cont_kernel + planq_kernel): GPU suite with it on and off, bench with it on and off.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -q -x > $O/b39_pytest.log 2>&1; echo "pytest (cont on) rc=$?" > $O/b39_summary.txt
tail -15 $O/b39_pytest.log >> $O/b39_summary.txt
ZL_B200_CONT=0 timeout 2400 python -m pytest tests -m gpu -q > $O/b39_pytest_off.log 2>&1; echo "pytest (cont off) rc=$?" >> $O/b39_summary.txt
tail -3 $O/b39_pytest_off.log >> $O/b39_summary.txt
S="--steps 12 --warmup 3 --no-cpu-baseline --no-latency --no-thumb"
for v in on off; do
  E="ZL_B200_CONT=1"; [ $v = off ] && E="ZL_B200_CONT=0"
  env $E timeout 600 python bench.py $S > $O/b39_$v.json 2> $O/b39_$v.err; echo "$v rc=$?" >> $O/b39_summary.txt
done
python - <<'PY' >> $O/b39_summary.txt
import json
for n in ("on", "off"):
    try:
        j=json.loads(open(f"gpurun_out/b39_{n}.json").read().strip().splitlines()[-1]); r=j["roofline"]; p=j.get("parity_at_size") or {}
        print(f'{n}: ms/step {j["ms_per_step"]:.3f} vf/s {j["value"]:.4g} e2e {j["e2e"]["value"]:.4g} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} frac {r["frac"]:.3f} parity {p.get("masters_bit_equal")} {p.get("peaks_bit_equal")} {p.get("voice_table_bit_equal")} vf/step {j["voice_frames_per_step"]} launches {j["gpu_launches"]}')
    except Exception as ex: print(n, "parse failed", ex)
PY
cat $O/b39_summary.txt
