#!/bin/bash
# r02 batch 33/34: planner - lean fast-path tile loop (running sums instead of per-tile conversions) [33: also voice records requested before the command phase, dropped]: GPU suite, bench, CTA trace.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -q > $O/b34_pytest.log 2>&1; echo "pytest rc=$?" > $O/b34_summary.txt
tail -6 $O/b34_pytest.log >> $O/b34_summary.txt
S="--steps 12 --warmup 3 --no-cpu-baseline --no-latency --no-thumb"
timeout 600 python bench.py $S > $O/b34_bench.json 2> $O/b34_bench.err; echo "bench rc=$?" >> $O/b34_summary.txt
python - <<'PY' >> $O/b34_summary.txt
import json
for n in ("bench",):
    try:
        j=json.loads(open(f"gpurun_out/b34_{n}.json").read().strip().splitlines()[-1]); r=j["roofline"]; p=j.get("parity_at_size") or {}
        print(f'{n}: ms/step {j["ms_per_step"]:.3f} vf/s {j["value"]:.4g} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} frac {r["frac"]:.3f} parity {p.get("masters_bit_equal")} {p.get("peaks_bit_equal")} {p.get("voice_table_bit_equal")}')
    except Exception as ex: print(n, "parse failed", ex)
PY
T="--steps 4 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity --no-kernel-events"
ZL_B200_LIB=$PWD/build/v_trace.so ZL_B200_CTA_TRACE=$O/b34_trace_live.txt timeout 300 python bench.py $T > $O/b34_live.json 2> $O/b34_live.err
python profiles/r02/cta_trace.py $O/b34_trace_live.txt >> $O/b34_summary.txt 2>&1
cat $O/b34_summary.txt
