#!/bin/bash
# r02 batch 28: JACK port mix folded into the gather (last bus of a session to finish a tile sums it): GPU suite, bench, A/B against the separate kernel.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -q > $O/b28_pytest.log 2>&1; echo "pytest rc=$?" > $O/b28_summary.txt
tail -6 $O/b28_pytest.log >> $O/b28_summary.txt
timeout 900 python bench.py > $O/b28_bench.json 2> $O/b28_bench.err; echo "bench rc=$?" >> $O/b28_summary.txt
S="--steps 12 --warmup 3 --no-cpu-baseline --no-latency --no-thumb"
ZL_B200_FUSED_MIX=0 timeout 600 python bench.py $S > $O/b28_sep.json 2> $O/b28_sep.err; echo "separate mix rc=$?" >> $O/b28_summary.txt
python - <<'PY' >> $O/b28_summary.txt
import json
for n in ("bench", "sep"):
    try:
        j=json.loads(open(f"gpurun_out/b28_{n}.json").read().strip().splitlines()[-1]); r=j["roofline"]; p=j.get("parity_at_size") or {}
        print(f'{n}: ms/step {j["ms_per_step"]:.3f} vf/s {j["value"]:.4g} e2e {j["e2e"]["value"]:.4g} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} frac {r["frac"]:.3f} whole {r["whole_step_frac"]:.3f} parity {p.get("masters_bit_equal")} {p.get("peaks_bit_equal")} {p.get("voice_table_bit_equal")} launches {j.get("gpu_launches")}')
        print("   latency", j.get("latency"))
    except Exception as ex: print(n, "parse failed", ex)
PY
cat $O/b28_summary.txt
