#!/bin/bash
# r02 batch 24: gather with window-relative chained positions / envelopes, packed M/S pan, elect-issued TMA:
# full GPU suite + bench; ZL_KT=256 variant of the same source (bench only).
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -q > $O/b24_pytest.log 2>&1; echo "pytest rc=$?" > $O/b24_summary.txt
tail -6 $O/b24_pytest.log >> $O/b24_summary.txt
timeout 900 python bench.py > $O/b24_bench.json 2> $O/b24_bench.err; echo "bench rc=$?" >> $O/b24_summary.txt
S="--steps 12 --warmup 3 --no-cpu-baseline --no-latency --no-thumb"
for v in kt256; do
  ZL_B200_LIB=$PWD/build/v_$v.so timeout 600 python bench.py $S > $O/b24_$v.json 2> $O/b24_$v.err; echo "$v rc=$?" >> $O/b24_summary.txt
done
python - <<'PY' >> $O/b24_summary.txt
import json
for n in ("bench", "kt256"):
    try:
        j=json.loads(open(f"gpurun_out/b24_{n}.json").read().strip().splitlines()[-1]); r=j["roofline"]
        print(f'{n}: ms/step {j["ms_per_step"]:.3f} vf/s {j["value"]:.4g} e2e {j["e2e"]["value"]:.4g} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} frac {r["frac"]:.3f} whole {r["whole_step_frac"]:.3f}')
        print("  parity", j.get("parity_at_size"))
    except Exception as ex: print(n, "parse failed", ex)
PY
cat $O/b24_summary.txt
