#!/bin/bash
# r02 batch 42: cont_kernel with shared-memory staged, coalesced traffic: GPU suite, bench on/off, ncu of cont_kernel, CTA trace.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -q -x > $O/b42_pytest.log 2>&1; echo "pytest (cont on) rc=$?" > $O/b42_summary.txt
tail -12 $O/b42_pytest.log >> $O/b42_summary.txt
S="--steps 12 --warmup 3 --no-cpu-baseline --no-latency --no-thumb"
for v in on off; do
  E="ZL_B200_CONT=1"; [ $v = off ] && E="ZL_B200_CONT=0"
  env $E timeout 600 python bench.py $S > $O/b42_$v.json 2> $O/b42_$v.err; echo "$v rc=$?" >> $O/b42_summary.txt
done
python - <<'PY' >> $O/b42_summary.txt
import json
for n in ("on", "off"):
    try:
        j=json.loads(open(f"gpurun_out/b42_{n}.json").read().strip().splitlines()[-1]); r=j["roofline"]; p=j.get("parity_at_size") or {}
        print(f'{n}: ms/step {j["ms_per_step"]:.3f} vf/s {j["value"]:.4g} e2e {j["e2e"]["value"]:.4g} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} frac {r["frac"]:.3f} parity {p.get("masters_bit_equal")} {p.get("peaks_bit_equal")} {p.get("voice_table_bit_equal")}')
    except Exception as ex: print(n, "parse failed", ex)
PY
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"cont_kernel" -s 40 -c 1 -o $O/b42_cont_kernel -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity > $O/b42_ncu_cont.log 2>&1; echo "ncu cont rc=$?" >> $O/b42_summary.txt
T="--steps 4 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity --no-kernel-events"
ZL_B200_LIB=$PWD/build/v_trace.so ZL_B200_CTA_TRACE=$O/b42_trace_live.txt timeout 300 python bench.py $T > $O/b42_live.json 2> $O/b42_live.err
python profiles/r02/cta_trace.py $O/b42_trace_live.txt >> $O/b42_summary.txt 2>&1
cat $O/b42_summary.txt
